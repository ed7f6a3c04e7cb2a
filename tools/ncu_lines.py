"""Join ncu's per-instruction SASS metrics (--page source --csv) with nvdisasm line info, aggregate per source line.
usage: python tools/ncu_lines.py <report.ncu-rep> <mangled-kernel-name-substring> [top_n]
Needs the same librt_b200.so that was profiled."""
import csv, io, os, re, subprocess, sys, tempfile, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, kern = sys.argv[1], sys.argv[2]
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 45
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "opengl_raytracer_b200/csrc/librt_b200.so")], cwd=tmp, check=True, capture_output=True)
dis = subprocess.run(["nvdisasm", "--print-line-info-inline", os.path.join(tmp, "rt_kernels.sm_100a.cubin")], capture_output=True, text=True).stdout.splitlines()
# locate the function, collect (file, line) per instruction in order
start = next(i for i, l in enumerate(dis) if l.startswith(".text.") and kern in l)
lines, cur = [], ("?", 0)
for l in dis[start + 1:]:
    if l.startswith("\t.section") or l.startswith(".text."):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        # innermost frame of an inlined chain is listed first; keep the first seen until the next instruction
        if cur[2] if len(cur) > 2 else False:
            continue
        cur = (os.path.basename(m.group(1)), int(m.group(2)), True)
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        lines.append((cur[0], cur[1]))
        cur = (cur[0], cur[1])
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[h]
ix, it, isamp = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
body = [r for r in rows[h + 1:] if len(r) > it]
if len(body) != len(lines):
    print(f"warning: {len(body)} profiled instructions vs {len(lines)} disassembled", file=sys.stderr)
agg = collections.defaultdict(lambda: [0, 0, 0, 0])
tot = tot_t = tot_s = 0
for (f, ln), r in zip(lines, body):
    n, t, s = int(r[ix]), int(r[it]), int(r[isamp])
    a = agg[(f, ln)]
    a[0] += n; a[1] += t; a[2] += s; a[3] += 1
    tot += n; tot_t += t; tot_s += s
print(f"total warp-inst {tot:,}  thread-inst {tot_t:,}  lanes/inst {tot_t / tot:.2f}  samples {tot_s}")
src_cache = {}
def src(f, ln):
    for d in ("opengl_raytracer_b200/csrc",):
        p = os.path.join(ROOT, d, f)
        if os.path.isfile(p):
            if p not in src_cache: src_cache[p] = open(p).read().splitlines()
            return src_cache[p][ln - 1].strip()[:90] if 0 < ln <= len(src_cache[p]) else ""
    return ""
for (f, ln), (n, t, s, k) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:topn]:
    print(f"{100 * n / tot:5.1f}% inst {100 * s / max(tot_s, 1):5.1f}% smp  lanes {t / max(n, 1):5.1f}  sass {k:3d}  {f}:{ln:<4d} {src(f, ln)}")
