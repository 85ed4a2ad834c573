# strong-scaling lines of the default workload on one 8-GPU box: bash tools/scale_run.sh <outdir>
D=gpurun_out/$1; mkdir -p $D; cd /root/repo
for n in 8 4 2; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) bench.py --gpus $n --steps 20 --warmup 5 --no-cpu --no-sampling > $D/bench_${n}gpu.log 2>&1
  grep '^{' $D/bench_${n}gpu.log | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], d['ms_per_step'], d['value'], 'e2e', d['e2e']['value'], d['roofline']['per_kernel_ms'])"
done
