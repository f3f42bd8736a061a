timeout 300 python -m pytest tests/test_scenario.py -m gpu -x -q 2>&1 | tail -40 > gpurun_out/pytest_scn.log
