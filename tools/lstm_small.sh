# LSTM recurrent kernels at small batch tiles: parity + timing with and without the rank-to-rank exchanges
D=gpurun_out/$1; mkdir -p $D; cd /root/repo
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_config_shapes.py -m gpu -x -q -k "lstm or bf16 or config" > $D/tests.log 2>&1; tail -4 $D/tests.log
for pm in "32 16" "32 0" "0 0" "32 32"; do
  set -- $pm
  SKRNN_TC_PUSH=$1 SKRNN_TC_PUSH_BWD=$2 timeout 300 python bench.py --workload train_kanji --no-cpu --no-sampling --steps 20 --warmup 5 > $D/bench_kanji_push$1_$2.log 2>&1
  SKRNN_TC_PUSH=$1 SKRNN_TC_PUSH_BWD=$2 SKRNN_TC_DEBUG=1 timeout 120 python tools/run_one.py 256 2 100 300 bf16 lstm > $D/dbg_kanji_push$1_$2.log 2>&1
done
for f in $D/bench_kanji_push*.log; do echo -n "$f "; grep '^{' $f | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['per_kernel_ms'])"; done
