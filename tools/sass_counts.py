"""Per-kernel counts of the SASS mnemonics that prove the Blackwell paths in libnerfemitter.so (B200_PROFILING.md):
UTCHMMA / UTCQMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st (TMEM), UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk,
UTMALDG = cp.async.bulk.tensor, REDG = red.global, SYNCS = mbarrier.  Usage: python tools/sass_counts.py > profiles/rN_sass_counts.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "nerf_emitter_b200", "libnerfemitter.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
KEYS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UBLKCP", "UTMALDG", "REDG", "SYNCS", "HFMA2", "FFMA", "LDG", "total"]
counts = collections.OrderedDict()
cur = None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", name)
        counts[cur] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and cur is not None:
        op = m.group(1)
        counts[cur]["total"] += 1
        for k in KEYS:
            if op.startswith(k):
                counts[cur][k] += 1
print("SASS mnemonic counts per kernel of %s (sm_100a)" % os.path.relpath(so, ROOT))
print("%-70s" % "kernel" + "".join("%9s" % k for k in KEYS))
for name, c in counts.items():
    print("%-70s" % name[:70] + "".join("%9d" % c[k] for k in KEYS))
