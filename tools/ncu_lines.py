"""Warp-stall samples of one kernel per CUDA source line (outermost line of the kernel's own file, so inlined helpers
are charged to their call site).  Run here, no GPU:  python tools/ncu_lines.py file.ncu-rep lib.so kernel_substring [top]
The .so must be the build that was profiled (SASS offsets are matched one to one)."""
import csv, io, os, re, subprocess, sys, tempfile, collections
rep, so, kname = sys.argv[1], os.path.abspath(sys.argv[2]), sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, capture_output=True)
line_of = {}          # demangled function name -> {offset: ((file, line), sass)}
for f in sorted(os.listdir(tmp)):
    txt = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
    if kname not in txt:
        continue
    cur_fun, outer, pending = None, None, False
    for ln in txt.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", ln)
        if m:
            cur_fun = m.group(1)
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)( inlined at.*)?$', ln)
        if m:
            # a group of marker lines precedes an instruction: innermost first, then each inlined-at site, outermost last
            if os.environ.get("INNER"):
                if not pending:
                    outer = (os.path.basename(m.group(1)), int(m.group(2)))
            elif not m.group(3):
                outer = (os.path.basename(m.group(1)), int(m.group(2)))
            pending = True
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            pending = False
        if m and cur_fun and kname in cur_fun:
            line_of.setdefault(cur_fun, {})[int(m.group(1), 16)] = (outer, m.group(2).strip())
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
# several kernels may be in the report: take the blocks whose "Kernel Name" matches
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "rows": []}
        blocks.append(cur)
    elif cur is not None and r and r[0] == "Address":
        cur["hdr"] = r
    elif cur is not None and cur["hdr"] and r and r[0].startswith("0x"):
        cur["rows"].append(r)
def norm(x):
    x = re.sub(r"\((int|bool|unsigned int)\)", "", x)          # ncu prints template arguments as <(int)2>, c++filt as <2>
    return re.sub(r"\(.*", "", x).replace("nemit::", "").replace("void ", "").replace(" ", "")
demangled = {norm(subprocess.run(["c++filt", f], capture_output=True, text=True).stdout.strip()): f for f in line_of}
for b in blocks:
    key = norm(b["name"])
    if kname not in key or key not in demangled:
        continue
    lo = line_of[demangled[key]]
    hdr = b["hdr"]
    col = {h: i for i, h in enumerate(hdr)}
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    base = int(b["rows"][0][0], 16)
    per = collections.defaultdict(lambda: collections.Counter())
    tot = 0
    for r in b["rows"]:
        off = int(r[0], 16) - base
        n = int(r[col["# Samples"]] or 0)
        tot += n
        where = lo.get(off, ((None, 0), ""))[0] or ("?", 0)
        per[where]["n"] += n
        per[where]["inst"] += int(r[col["Instructions Executed"]] or 0)
        for s in stalls:
            per[where][s] += int(r[col[s]] or 0)
    tot_inst = sum(c["inst"] for c in per.values())
    print("==", b["name"], "total samples", tot, "warp instructions", tot_inst)
    src = {}
    for (fn, ln), c in sorted(per.items(), key=lambda kv: -kv[1]["n"])[:top]:
        if fn not in src:
            p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "nerf_emitter_b200", "csrc", fn or "")
            src[fn] = open(p).read().splitlines() if os.path.isfile(p) else []
        text = src[fn][ln - 1].strip()[:90] if 0 < ln <= len(src[fn]) else ""
        why = sorted(((c[s], s[6:]) for s in stalls), reverse=True)[:3]
        print("%6d %5.1f%% inst %5.1f%%  %s:%d  %-92s %s" % (c["n"], 100.0 * c["n"] / max(tot, 1), 100.0 * c["inst"] / max(tot_inst, 1), fn, ln, text,
                                               ", ".join("%s %d" % (s, n) for n, s in why if n)))
