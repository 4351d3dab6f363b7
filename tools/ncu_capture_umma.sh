#!/bin/bash
# Re-capture only the two tensor-GEMM workloads (after a change to umma_gemm.cu); see tools/ncu_capture.sh.
set -u
OUT=gpurun_out/ncu_r02
mkdir -p $OUT
for w in hgemm linear_s8; do
  ncu --set full --clock-control none --import-source on -k regex:umma_gemm_kernel --launch-skip 4 -c 1 -f -o $OUT/r02_$w \
      python bench.py --workload $w --steps 2 --warmup 1 --quick --no-cpu --no-also > $OUT/ncu_$w.log 2>&1
  echo "$w rc=$?"
done
