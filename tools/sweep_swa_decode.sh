#!/bin/bash
# GPU box: SWA decode split-size sweep (HYMETA_SWA_DEC_TPS = tiles per CTA)
for t in 0 2 4 6 8 11 16 22; do
  echo "== TPS=$t"; HYMETA_SWA_DEC_TPS=$t python tools/bench_kernels.py --which swa_decode 2>&1 | grep swa_decode
done
