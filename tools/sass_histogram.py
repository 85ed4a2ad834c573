# opcode histogram of the built objects (what proves the Blackwell-native instructions): python tools/sass_histogram.py
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
bd = os.path.join(ROOT, "sketch_rnn_b200", "build")
want = re.compile(r"^(UTC\w*MMA|UTCBAR|LDTM|STTM|UTMALDG|UTMASTG|UBLKCP|STAS|UTMAPF|UTMACMDFLUSH|SYNCS|UCGABAR_\w+|HMMA|LDGSTS|MUFU|FENCE|MEMBAR|REDG?|ATOMG?|CCTL)")
COLS = ["UTC*MMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "STAS", "SYNCS", "UCGABAR", "LDGSTS", "FFMA2", "HMMA"]
print("| object | " + " | ".join(COLS) + " |")
print("|---|" + "---:|" * len(COLS))
for f in sorted(os.listdir(bd)):
    if not f.endswith(".o"):
        continue
    out = subprocess.run(["cuobjdump", "-sass", os.path.join(bd, f)], capture_output=True, text=True).stdout
    c = collections.Counter()
    for line in out.splitlines():
        m = re.search(r"/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d\s+)?([A-Z][A-Z0-9_]*)", line)
        if m:
            op = m.group(1)
            if op.startswith("UTC") and op.endswith("MMA"): c["UTC*MMA"] += 1
            elif op.startswith("UCGABAR"): c["UCGABAR"] += 1
            else: c[op] += 1
    print("| %s | %s |" % (f, " | ".join(str(c.get(k, 0)) for k in COLS)))
print()
print("UTC*MMA = tcgen05.mma; LDTM / STTM = tcgen05.ld / st; UTMALDG / UTMASTG = TMA tensor load / store; UBLKCP = bulk copy")
print("(shared memory -> peer CTA's shared memory in rnn1_tc); STAS = st.async (store + transaction count on the receiver's")
print("mbarrier, sampler_cl); SYNCS = mbarrier operations; UCGABAR = barrier.cluster; LDGSTS = cp.async (the streaming sampler's")
print("weight ring); FFMA2 = two fp32 fused multiply-adds per instruction (same sampler); HMMA = legacy mma.sync (none).")
