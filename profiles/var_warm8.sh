# warm end-to-end on N GPUs (torchrun): pipeline depth (library variants) x chunk size
N=${N:-8}
for v in $VARIANTS; do for mb in $CHUNKS; do
  if [ $v = default ]; then unset FOREMAST_CUDA_LIB; else export FOREMAST_CUDA_LIB=$PWD/foremast_1_b200/lib/var_$v.so; fi
  if [ $mb = default ]; then unset FOREMAST_CUDA_CHUNK_MB; else export FOREMAST_CUDA_CHUNK_MB=$mb; fi
  echo -n "$v chunk=$mb  "
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 3 --warmup 3 --no-secondary --no-cpu 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); w=d['e2e']['warm']; print('value %.1f M  cold %.2f M  warm %.1f M  %.2f ms/step  kernels %.2f ms' % (d['value']/1e6, d['e2e']['value']/1e6, w['value']/1e6, w['ms_per_step'], w['kernel_ms_per_step']))"
done; done
