#!/bin/bash
# Round-2 evidence run on one B200 (under gpurun): tests, smoke, bench lines, train.py time/batch, ncu launch list and
# full captures.  Everything lands in gpurun_out/<first argument, default r02p>/.
O=gpurun_out/${1:-r02p}; mkdir -p $O
python -m pytest tests -m gpu -q > $O/tests.log 2>&1; tail -3 $O/tests.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -2 $O/smoke.log
python bench.py --steps 20 --warmup 5 > $O/bench_train_h512_l3_b800_bf16.json 2> $O/bench_default.err; tail -c 300 $O/bench_default.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference_arm.json 2>> $O/bench_default.err
for w in train_kanji train_gru train_rnn sample_8192 sample_10; do
  python bench.py --workload $w --steps 20 --warmup 5 > $O/bench_$w.json 2> $O/bench_$w.err; tail -c 200 $O/bench_$w.err
done
python bench.py --workload train_kanji --precision f32 --steps 10 --warmup 3 --no-cpu > $O/bench_train_kanji_f32.json 2>/dev/null
python tools/train_time.py 60 > $O/train_time.txt 2>&1; cat $O/train_time.txt
# launch list (cold-cache, serialised: compare shares)
python bench.py --steps 2 --warmup 3 --no-cpu --no-sampling > $O/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $O/launches_train_h512_l3_b800_bf16.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-sampling > $O/ncu_launches.log 2>&1
# full captures of the kernels of one config-4 step (second iteration), of a GRU step, and of the throughput sampler
python tools/run_one.py 512 3 800 300 > $O/plain_c4.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"lstm_bwd_pp|lstm_fwd_pp|head_tc|clip_adam" -s 8 -c 8 -o $O/c4 \
    python tools/run_one.py 512 3 800 300 > $O/ncu_c4.log 2>&1
python tools/run_one.py 256 2 100 300 bf16 gru > $O/plain_gru.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"rnn1_" -s 4 -c 4 -o $O/gru \
    python tools/run_one.py 256 2 100 300 bf16 gru > $O/ncu_gru.log 2>&1
python tools/run_one.py 256 2 100 300 bf16 rnn > $O/plain_rnn.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"rnn1_" -s 4 -c 4 -o $O/rnn \
    python tools/run_one.py 256 2 100 300 bf16 rnn > $O/ncu_rnn.log 2>&1
for n in 1 10 28 64 512; do python tools/sample_time.py $n 300; done > $O/sample_latency.txt 2>&1; cat $O/sample_latency.txt
python tools/sample_time.py 10 300 gru >> $O/sample_latency.txt 2>&1; python tools/sample_time.py 10 300 rnn >> $O/sample_latency.txt 2>&1
python tools/sample_time.py 4 60 > $O/plain_samplecl.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"sample_cluster" -s 2 -c 1 -o $O/samplecl \
    python tools/sample_time.py 4 60 > $O/ncu_samplecl.log 2>&1
python tools/run_sample.py 8192 40 > $O/plain_sample.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"sample_tp_kernel" -s 1 -c 1 -o $O/sampler \
    python tools/run_sample.py 8192 40 > $O/ncu_sampler.log 2>&1
python tools/sample_time.py 8192 40 > $O/plain_stream.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"sample_kernel" -s 2 -c 1 -o $O/stream \
    python tools/sample_time.py 8192 40 > $O/ncu_stream.log 2>&1
for n in 8192 4096 1024; do python tools/tp_time.py $n 300; done > $O/sample_throughput.txt 2>&1
SKRNN_SAMPLE_TP=0 python tools/tp_time.py 8192 300 >> $O/sample_throughput.txt 2>&1
SKRNN_TP_DEBUG=1 python tools/tp_time.py 8192 300 >> $O/sample_throughput.txt 2>&1
for m in lstm gru rnn; do python tools/sample_time.py 8192 300 $m | tail -1; done >> $O/sample_throughput.txt 2>&1; cat $O/sample_throughput.txt
# gpurun brings back at most 64 MiB: keep the raw-metrics pages, drop the big reports
for r in c4 gru rnn sampler samplecl stream; do
  [ -f $O/$r.ncu-rep ] && ncu -i $O/$r.ncu-rep --page raw --csv > $O/${r}_raw.csv 2>/dev/null && rm -f $O/$r.ncu-rep
done
ls -la $O | tail -40
