# experiments on the gru / rnn tensor-core kernels: how the operand tiles are published
D=gpurun_out/$1; mkdir -p $D; cd /root/repo
for pm in 0 1; do
  SKRNN_R1_PUBLISH=$pm timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_config_shapes.py -m gpu -x -q -k "gru or rnn" > $D/tests_pm$pm.log 2>&1
  for m in gru rnn; do
    SKRNN_R1_PUBLISH=$pm SKRNN_TC_DEBUG=1 timeout 120 python tools/run_one.py 256 2 100 300 bf16 $m > $D/dbg_${m}_pm$pm.log 2>&1
    SKRNN_R1_PUBLISH=$pm timeout 300 python bench.py --workload train_$m --no-cpu --no-sampling --steps 20 --warmup 5 > $D/bench_${m}_pm$pm.log 2>&1
  done
done
