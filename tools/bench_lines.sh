# the bench lines of every workload on one GPU: bash tools/bench_lines.sh <outdir>
O=gpurun_out/$1; mkdir -p $O; cd /root/repo
python bench.py --steps 20 --warmup 5 > $O/bench_train_h512_l3_b800_bf16.json 2> $O/bench_default.err
for w in train_kanji train_gru train_rnn sample_8192 sample_10; do
  python bench.py --workload $w --steps 20 --warmup 5 > $O/bench_$w.json 2> $O/bench_$w.err
done
python tools/train_time.py 60 > $O/train_time.txt 2>&1
for f in $O/bench_*.json; do echo -n "$f: "; grep '^{' $f | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d.get('ms_per_step'), d.get('value'), 'e2e', (d.get('e2e') or {}).get('value'), (d.get('roofline') or {}).get('per_kernel_ms'))"; done; cat $O/train_time.txt
