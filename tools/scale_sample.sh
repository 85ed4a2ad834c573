# sampling curve of BASELINE config 5 (8192 sketches x 300 steps sharded over N GPUs) on one 8-GPU box: bash tools/scale_sample.sh <outdir>
D=gpurun_out/$1; mkdir -p $D; cd /root/repo
for n in 8 4 2; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + n)) bench.py --workload sample_8192 --gpus $n --steps 10 --warmup 3 --no-cpu > $D/sample_${n}gpu.log 2>&1
  grep '^{' $D/sample_${n}gpu.log | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], d['ms_per_step'], d['value'], 'e2e', d['e2e']['value'])"
done
