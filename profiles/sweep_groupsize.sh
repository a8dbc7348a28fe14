#!/bin/bash
# lanes per series of k_ma_only (FOREMAST_CUDA_MA_GROUP) against the history length, config 5 slice
for nh in 60 120 240 480; do for gsz in 2 4 8 16; do
  echo -n "n_hist=$nh G=$gsz  "
  FOREMAST_CUDA_MA_GROUP=$gsz timeout 120 python bench.py --workload c5 --n-hist $nh --compact --no-cpu --no-e2e --no-secondary --steps 10 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print('%.4f ms  frac %.3f' % (r['kernel_ms'], r['frac']))"
done; done
