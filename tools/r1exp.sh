# experiments on the gru / rnn tensor-core kernels: prefetch settings
D=gpurun_out/$1; mkdir -p $D; cd /root/repo
for pf in 2 -1 1 2L1; do
  for m in gru rnn; do
    SKRNN_R1_PREFETCH=$pf SKRNN_TC_DEBUG=1 timeout 120 python tools/run_one.py 256 2 100 300 bf16 $m > $D/dbg_${m}_pf$pf.log 2>&1
    SKRNN_R1_PREFETCH=$pf timeout 300 python bench.py --workload train_$m --no-cpu --no-sampling --steps 20 --warmup 5 > $D/bench_${m}_pf$pf.log 2>&1
  done
done
