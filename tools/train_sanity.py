# a few hundred training steps through train.py on a synthetic corpus, per cell: the loss must fall and stay finite
#   python tools/train_sanity.py [batches]
import os, re, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sketch_rnn_b200 import utils
n = sys.argv[1] if len(sys.argv) > 1 else "300"
d = tempfile.mkdtemp()
utils.make_synthetic_dataset(os.path.join(d, "data", "kanji.cpkl"), num_sketches=11000, seed=0)
for model in ("lstm", "gru", "rnn"):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "train.py"), "--model", model, "--max_batches", n, "--save_every", "100000"],
                       cwd=d, env=dict(os.environ, PYTHONPATH=ROOT), capture_output=True, text=True)
    losses = [float(m.group(1)) for m in re.finditer(r"cost = ([-0-9.einfa]+)", r.stdout)]
    if len(losses) < 20:
        print(model, "no losses parsed:", r.stdout[-400:], r.stderr[-400:])
        continue
    first, last = sum(losses[:10]) / 10, sum(losses[-10:]) / 10
    ok = all(l == l and abs(l) < 1e6 for l in losses) and last < first
    tail = [l for l in r.stdout.splitlines() if l.startswith("mean time/batch")]
    print("%s: %d steps, loss %.4f -> %.4f, %s; %s" % (model, len(losses), first, last, "ok" if ok else "NOT DECREASING / NOT FINITE", tail[-1] if tail else ""))
