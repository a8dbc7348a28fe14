# k_canary_count: trips between two CTA barriers
for sy in $SYNCS; do
  echo -n "count_sync=$sy  "
  FOREMAST_CUDA_COUNT_SYNC=$sy timeout 200 python bench.py --no-cpu --no-e2e --no-secondary --steps 5 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%.3f ms  frac %.4f' % (d['ms_per_step'], d['roofline']['frac']))"
done
