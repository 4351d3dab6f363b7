#!/bin/bash
# Run ON THE GPU BOX (gpurun): one `ncu --set full` capture of the dominant kernel of every workload, plus the
# launch list of a plain bench run.  Reports land in gpurun_out/ncu_r02/; summarise them here afterwards with
#   python tools/ncu_summary.py gpurun_out/ncu_r02
# Numbers printed by a program running under ncu are never bench values.
set -u
OUT=gpurun_out/ncu_r02
mkdir -p $OUT
python bench.py --steps 2 --warmup 1 --quick --no-cpu --no-also > $OUT/plain.json 2> $OUT/plain.err || exit 1   # must exit 0 without ncu first
for w in dot sum scan sgemm hgemm linear_s8; do
  case $w in
    dot|sum) k=reduce_kernel;; scan) k=scan_kernel;; sgemm) k=sgemm_kernel;; *) k=umma_gemm_kernel;;
  esac
  ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip 4 -c 1 -f -o $OUT/r02_$w \
      python bench.py --workload $w --steps 2 --warmup 1 --quick --no-cpu --no-also > $OUT/ncu_$w.log 2>&1
  echo "$w rc=$?"
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/r02_launches_dot.csv \
    python bench.py --steps 2 --warmup 1 --quick --no-cpu --no-also > $OUT/ncu_launches.log 2>&1
echo "launch list rc=$?"
