"""Static SASS instruction count per CUDA source line of one kernel (outermost line of the kernel's own file, inlined helpers
charged to their call site).  Run here, no GPU:  python tools/sass_lines.py lib.so kernel_substring [file.cu] [top]"""
import collections, os, re, subprocess, sys, tempfile
so, kname = os.path.abspath(sys.argv[1]), sys.argv[2]
own = sys.argv[3] if len(sys.argv) > 3 else None
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, capture_output=True)
for f in sorted(os.listdir(tmp)):
    txt = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
    cur, outer, cnt = None, None, {}
    group = []
    for ln in txt.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", ln)
        if m:
            cur = m.group(1); continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)( inlined at.*)?$', ln)
        if m:
            group.append((os.path.basename(m.group(1)), int(m.group(2)))); continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m and cur and kname in cur:
            if group:
                cand = [g for g in group if own is None or g[0] == own]
                outer = group[0] if os.environ.get("INNER") else (cand[-1] if cand else group[-1])
                group = []
            cnt.setdefault(cur, collections.Counter())[outer] += 1
        elif m:
            group = []
    for fun, c in cnt.items():
        name = subprocess.run(["c++filt", fun], capture_output=True, text=True).stdout.strip()[:100]
        print("==", name, sum(c.values()), "instructions")
        src = {}
        for (fn, l), n in c.most_common(top):
            p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "nerf_emitter_b200", "csrc", fn or "")
            if fn not in src:
                src[fn] = open(p).read().splitlines() if os.path.isfile(p) else []
            text = src[fn][l - 1].strip()[:100] if 0 < l <= len(src[fn]) else ""
            print("%6d  %s:%d  %s" % (n, fn, l, text))
