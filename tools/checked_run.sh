#!/bin/bash
# Stand-in for compute-sanitizer (closed on this pool, profiles/r02a_sanitizer_memcheck.log): the GPU tests and the
# compact workload of tools/sanitize_case.py against the CHECKED library (`make checked`: every gather index is
# range-checked on the device, the first offender's source line is returned as an error by the next C-ABI call).
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
export ICP_B200_LIB=$PWD/iterative-closest-point_b200/lib/variants/libicp_b200_checked.so
[ -f "$ICP_B200_LIB" ] || { echo "build it first: make -C iterative-closest-point_b200 checked"; exit 2; }
timeout 900 python tools/sanitize_case.py > gpurun_out/checked_sanitize_case.log 2>&1; echo "rc=$?" >> gpurun_out/checked_sanitize_case.log
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tree.py tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/checked_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/checked_pytest.log
tail -n 3 gpurun_out/checked_sanitize_case.log gpurun_out/checked_pytest.log
