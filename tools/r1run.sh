D=gpurun_out/$1; mkdir -p $D; cd /root/repo
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_config_shapes.py -m gpu -x -q -k "gru or rnn" > $D/tests.log 2>&1; echo "tests rc $?" >> $D/tests.log
for m in gru rnn; do
  SKRNN_TC_DEBUG=1 timeout 120 python tools/run_one.py 256 2 100 300 bf16 $m > $D/dbg_$m.log 2>&1
  SKRNN_R1_CLUSTER=0 SKRNN_TC_DEBUG=1 timeout 120 python tools/run_one.py 256 2 100 300 bf16 $m > $D/dbg_${m}_cl1.log 2>&1
  timeout 300 python bench.py --workload train_$m --no-cpu --no-sampling --steps 20 --warmup 5 > $D/bench_$m.log 2>&1
  SKRNN_R1_CLUSTER=0 timeout 300 python bench.py --workload train_$m --no-cpu --no-sampling --steps 20 --warmup 5 > $D/bench_${m}_cl1.log 2>&1
done
tail -5 $D/tests.log
