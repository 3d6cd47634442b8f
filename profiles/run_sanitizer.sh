#!/bin/bash
# compute-sanitizer passes (racecheck / memcheck in the test strategy) over the kernels
# with hand-rolled synchronisation or peer addressing: the cp.async.bulk + mbarrier ring SpMM
# (B200_SPMM_PATH=tma), the partitioned aggregation in all four halo modes (push scatter, pull,
# two-slot TAB gather, addend pass; all ranks' buffers in one process), the tcgen05 GEMMs
# (mbarrier pipelines, TMEM; forward and MN-major weight gradient), the static-scratch attention
# projection, the lazily rebuilt in-CSR, plus the default SpMM / softmax / GAT.
# Usage (GPU box): bash profiles/run_sanitizer.sh -> gpurun_out/r2_sanitizer_{memcheck,racecheck}.txt
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
MEM_SEL='tests/test_gpu_golden.py::test_spmm_widths_vs_oracle tests/test_gpu_golden.py::test_hub_rows_are_segmented tests/test_gpu_partition.py::test_direct_gather_one_process tests/test_zz_no_alloc.py tests/test_gpu_index.py::test_index_memory_is_two_csrs tests/test_gpu_gat.py'
MEM_K='not 256 and not 4-range'
RACE_SEL='tests/test_gpu_golden.py::test_spmm_widths_vs_oracle tests/test_gpu_nn.py::test_edge_softmax_hub_row tests/test_gpu_gat.py tests/test_zz_no_alloc.py'
RACE_K='not 602'
for tool in memcheck racecheck; do
  out=gpurun_out/r2_sanitizer_$tool.txt
  : > $out
  if [ $tool = memcheck ]; then SEL=$MEM_SEL; K=$MEM_K; else SEL=$RACE_SEL; K=$RACE_K; fi
  echo "=== $tool: pytest $SEL -k '$K'" >> $out
  timeout 500 /usr/local/cuda/bin/compute-sanitizer --tool $tool --print-limit 5 \
    python -m pytest $SEL -k "$K" -q -p no:cacheprovider > gpurun_out/r2_sanitizer_${tool}_pytest.raw 2>&1
  echo "exit code $?" >> $out
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|Error|hazard" gpurun_out/r2_sanitizer_${tool}_pytest.raw | tail -n 8 >> $out
  echo "=== $tool: tcgen05 3xTF32 GEMMs (profiles/gemm_check.py --quick --wgrad)" >> $out
  timeout 300 /usr/local/cuda/bin/compute-sanitizer --tool $tool --print-limit 5 \
    python profiles/gemm_check.py --quick --wgrad > gpurun_out/r2_sanitizer_${tool}_gemm.raw 2>&1
  echo "exit code $?" >> $out
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|worst_err|Error|hazard" gpurun_out/r2_sanitizer_${tool}_gemm.raw | tail -n 6 >> $out
done
cat gpurun_out/r2_sanitizer_memcheck.txt gpurun_out/r2_sanitizer_racecheck.txt
