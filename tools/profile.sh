#!/bin/bash
# Run on the GPU box under gpurun: plain run first, then the ncu launch list and one --set full
# capture per hot kernel (B200_PROFILING.md).  Outputs land in gpurun_out/.
set -o pipefail
CMD="python bench.py --steps 1 --warmup 1 --decode 4 --no-e2e --no-cpu-baseline --no-extra"
mkdir -p gpurun_out
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv \
    --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
for K in ${KERNELS:-swa_decode_split_kernel swa_prefill_kernel gla_prefill_kernel gla_decode_kernel}; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 2 -c 1 \
      -o gpurun_out/prof_$K -f $CMD > gpurun_out/ncu_$K.log 2>&1
  echo "$K rc=$?"
done
ls -la gpurun_out/
