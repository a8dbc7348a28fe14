# k_canary_count: multiset slots per warp (more slots = fewer overflow rounds, fewer resident warps)
for sl in $SLOTS; do for st in $STAGES; do
  if [ $sl = default ]; then unset FOREMAST_CUDA_SLOTS; else export FOREMAST_CUDA_SLOTS=$sl; fi
  if [ $st = default ]; then unset FOREMAST_CUDA_STAGES; else export FOREMAST_CUDA_STAGES=$st; fi
  echo -n "slots=$sl stages=$st  "
  timeout 200 python bench.py --no-cpu --no-e2e --no-secondary --steps 5 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%.3f ms  frac %.4f' % (d['ms_per_step'], d['roofline']['frac']))"
done; done
