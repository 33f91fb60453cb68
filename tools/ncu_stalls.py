"""Top stall sites of a kernel in an .ncu-rep (run here, no GPU): python tools/ncu_stalls.py file.ncu-rep [top=40] [view=sass|cuda]
sass: per SASS instruction; cuda: per CUDA source line (needs --import-source on at capture time)."""
import csv, io, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
view = sys.argv[3] if len(sys.argv) > 3 else "sass"
cmd = ["ncu", "-i", rep, "--page", "source", "--csv"] + (["--print-source", "cuda"] if view == "cuda" else [])
raw = subprocess.run(cmd, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
start = [i for i, r in enumerate(rows) if r and r[0] in ("Address", "#", "Line")]
hdr = rows[start[0]]
body = rows[start[0] + 1:]
col = {h: i for i, h in enumerate(hdr)}
samp = col.get("# Samples", col.get("Warp Stall Sampling (All Samples)"))
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[samp] or 0) for r in body if len(r) > samp and (r[samp] or "0").isdigit())
print("total samples", tot)
agg = {s: sum(int(r[col[s]] or 0) for r in body if len(r) > col[s] and (r[col[s]] or "0").isdigit()) for s in stalls}
print("by reason:", {k: "%.1f%%" % (100.0 * v / max(tot, 1)) for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
best = sorted([r for r in body if len(r) > samp and (r[samp] or "0").isdigit()], key=lambda r: -int(r[samp]))[:top]
for r in best:
    why = sorted(((int(r[col[s]] or 0), s) for s in stalls), reverse=True)[:2]
    print("%6d %5.1f%%  %-70s %s" % (int(r[samp]), 100.0 * int(r[samp]) / max(tot, 1), r[col["Source"]].strip()[:70],
                                   ", ".join("%s %d" % (s[6:], n) for n, s in why if n)))
