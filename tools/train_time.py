# train.py's own timed region (next_batch + one step, reference train.py:83-87) on a synthetic kanji-sized corpus:
#   python tools/train_time.py [batches]      -> prints the mean time/batch with and without the prefetching loader
import os, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sketch_rnn_b200 import utils
n = sys.argv[1] if len(sys.argv) > 1 else "60"
d = tempfile.mkdtemp()
utils.make_synthetic_dataset(os.path.join(d, "data", "kanji.cpkl"), num_sketches=11000, seed=0)
for extra in ([], ["--no_prefetch"]):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "train.py"), "--max_batches", n, "--save_every", "100000"] + extra,
                       cwd=d, env=dict(os.environ, PYTHONPATH=ROOT), capture_output=True, text=True)
    tail = [l for l in r.stdout.splitlines() if l.startswith("mean time/batch")]
    print("train.py (lstm H256 L2 B100 T300, bf16)%s: %s" % (" --no_prefetch" if extra else "", tail[-1] if tail else r.stderr[-500:]))
