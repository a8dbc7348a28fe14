# k_ma_only variants (FOREMAST_CUDA_LIB) on the config-5 slice without roll-up
for v in $VARIANTS; do for nh in $NHS; do
  if [ $v = default ]; then unset FOREMAST_CUDA_LIB; else export FOREMAST_CUDA_LIB=$PWD/foremast_1_b200/lib/var_$v.so; fi
  echo -n "$v n_hist=$nh  "; timeout 120 python bench.py --workload c5 --n-hist $nh --compact --no-cpu --no-e2e --no-secondary --steps 10 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print('%.4f ms  frac %.3f' % (r['kernel_ms'], r['frac']))"
done; done
