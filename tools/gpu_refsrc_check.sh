#!/bin/bash
# Runs the reference-source binaries (bin/ref_icp, bin/ref_icp_gpu, bin/ref_bench: make refsrc) on the GPU box without
# Python: ref_icp / ref_icp_gpu must write the same output.txt as bin/icp, ref_bench must print its 14 captures.
# Usage (from the repo root): gpurun --timeout 120 -- bash tools/gpu_refsrc_check.sh
set -u
R=${GRAFT_REPO_ROOT:-$(pwd)}
B=$R/iterative-closest-point_b200/bin
O=$R/gpurun_out
mkdir -p "$O" /tmp/refsrc_chk/data_students
cd /tmp/refsrc_chk
for n in cow_ref cow_tr1; do xz -dc "$R/tests/golden/clouds/$n.txt.xz" > data_students/$n.txt; done
{
for exe in icp ref_icp ref_icp_gpu; do
    mkdir -p $exe
    (cd $exe && timeout 60 "$B/$exe" ../data_students/cow_ref.txt ../data_students/cow_tr1.txt 10 > stdout.txt 2> stderr.txt; echo "$exe rc=$? iterations=$(grep -c 'iteration number' stderr.txt)")
done
cmp icp/output.txt ref_icp/output.txt && cmp icp/output.txt ref_icp_gpu/output.txt && cmp icp/stderr.txt ref_icp/stderr.txt && echo "OUTPUTS IDENTICAL"
ICP_BENCH_MIN_TIME=${ICP_BENCH_MIN_TIME:-0.1} timeout 100 "$B/ref_bench" 2> bench_stderr.txt
echo "ref_bench rc=$? stderr_lines=$(wc -l < bench_stderr.txt)"
grep -v "^\[ICP\]\|^\[load\]" bench_stderr.txt | head -5
} > "$O/refsrc_check.txt" 2>&1
cat "$O/refsrc_check.txt"
