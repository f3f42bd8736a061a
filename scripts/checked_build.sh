#!/bin/bash
# Stand-in for compute-sanitizer memcheck (closed on this pool): builds libumrs with
# -DUMRS_CHECKED (every shared-memory index and pointer offset of the fan kernels asserted
# against the allocated length, device asserts print file:line and fail the launch) and runs
# the whole GPU test suite against it.   usage (GPU box): scripts/checked_build.sh [LOG]
LOG=${1:-gpurun_out/checked_pytest.log}
make -C ut_multirobot_sim_b200/csrc OUT=../../build/libumrs_checked.so EXTRA=-DUMRS_CHECKED > /dev/null || exit 1
UMRS_LIB=build/libumrs_checked.so timeout 1800 python -m pytest tests -m gpu -x -q > $LOG 2>&1
echo "checked-build pytest rc=$?" >> $LOG
