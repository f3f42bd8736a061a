#!/bin/bash
# Round-end style GPU pass: parity tests, smoke, both bench arms, per-config timings,
# then (each only after the plain run exited 0) the ncu launch list and one --set full capture.
# compute-sanitizer is closed on this pool (it answers "closed ... stays closed"); the
# bounds-checked build (scripts/checked_build.sh) stands in for memcheck.
# usage: scripts/gpu_round.sh TAG     outputs: gpurun_out/*_TAG.*
T=${1:-r2}
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$T.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu_$T.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke_$T.log 2>&1; echo "smoke rc=$?" >> $O/smoke_$T.log
timeout 600 python bench.py > $O/bench_$T.json 2> $O/bench_$T.err || exit 1
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref_$T.json 2> $O/bench_ref_$T.err
timeout 600 python scripts/bench_configs.py 1 2 3 4 R > $O/configs_$T.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file $O/launches_$T.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-legs --sustained-seconds 0 > $O/ncu_l_$T.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_world_step -s 4 -c 1 \
  -o $O/prof_$T -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-legs --sustained-seconds 0 > $O/ncu_full_$T.log 2>&1
