#!/bin/bash
# compute-sanitizer passes (racecheck / memcheck in the test strategy) over the kernels
# with hand-rolled synchronisation: the cp.async.bulk + mbarrier ring SpMM (B200_SPMM_PATH=tma),
# the peer-table gather SpMM (all ranks' blocks in one process), the tcgen05 GEMM (mbarrier
# pipelines, TMEM), the static-scratch attention projection, plus the default SpMM / softmax / GAT.
# Usage (GPU box): bash profiles/run_sanitizer.sh   -> gpurun_out/r2_sanitizer_{memcheck,racecheck}.txt
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
SEL='tests/test_gpu_golden.py::test_spmm_widths_vs_oracle tests/test_gpu_golden.py::test_hub_rows_are_segmented tests/test_gpu_partition.py::test_direct_gather_one_process tests/test_zz_no_alloc.py tests/test_gpu_nn.py::test_edge_softmax_hub_row tests/test_gpu_gat.py'
for tool in memcheck racecheck; do
  out=gpurun_out/r2_sanitizer_$tool.txt
  : > $out
  echo "=== $tool: pytest $SEL  (LDG and TMA-ring SpMM paths are both parametrised inside)" >> $out
  timeout 2400 /usr/local/cuda/bin/compute-sanitizer --tool $tool --print-limit 5 \
    python -m pytest $SEL -q -p no:cacheprovider > gpurun_out/r2_sanitizer_${tool}_pytest.raw 2>&1
  echo "exit code $?" >> $out
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|Error|hazard" gpurun_out/r2_sanitizer_${tool}_pytest.raw | tail -n 8 >> $out
  tail -n 3 gpurun_out/r2_sanitizer_${tool}_pytest.raw >> $out
  echo "=== $tool: tcgen05 3xTF32 GEMM (profiles/gemm_check.py --quick)" >> $out
  timeout 900 /usr/local/cuda/bin/compute-sanitizer --tool $tool --print-limit 5 \
    python profiles/gemm_check.py --quick > gpurun_out/r2_sanitizer_${tool}_gemm.raw 2>&1
  echo "exit code $?" >> $out
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|worst_err|Error|hazard" gpurun_out/r2_sanitizer_${tool}_gemm.raw | tail -n 6 >> $out
  tail -n 3 gpurun_out/r2_sanitizer_${tool}_gemm.raw >> $out
done
cat gpurun_out/r2_sanitizer_memcheck.txt gpurun_out/r2_sanitizer_racecheck.txt
