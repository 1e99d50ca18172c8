#!/bin/bash
# Round 2, 1-GPU call: the no-allocation entry point's test alone in a fresh process, then the final verification of the
# committed tree (tools/r02_call9.sh: whole GPU suite - which now ends with that test -, smoke, the driver's bench
# commands, C++ harnesses, checked build, launch list of bench.py).
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_no_alloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_no_alloc_alone.log
tail -n 15 gpurun_out/pytest_no_alloc_alone.log
bash tools/r02_call9.sh
