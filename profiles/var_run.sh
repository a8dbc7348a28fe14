# tuning runs: kernel variants built as foremast_1_b200/lib/var_<name>.so (FOREMAST_CUDA_LIB selects one)
for v in $VARIANTS; do for sy in $SYNCS; do
  if [ $v = default ]; then unset FOREMAST_CUDA_LIB; else export FOREMAST_CUDA_LIB=$PWD/foremast_1_b200/lib/var_$v.so; fi
  echo -n "$v sync=$sy  "; FOREMAST_CUDA_GROUPS_SYNC=$sy timeout 120 python profiles/bench_groups.py | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['kernel_ms'])"
done; done
