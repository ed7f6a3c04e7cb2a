"""Join ncu's per-instruction SASS metrics (--page source --csv) with nvdisasm line info, aggregate per source line.
usage: python tools/ncu_lines.py <report.ncu-rep> <mangled-kernel-name-substring> [top_n] [--outer] [--cubin rt_kernels] [--filter kernel-regex]
       [--sections name:first-last,name:first-last,...]
Default: the innermost inlined frame of each instruction (which arithmetic helper).  --outer: the outermost frame (the
line of the kernel's own file that the instruction was inlined into: which part of the kernel); with --sections the
outer lines are summed over named line ranges.  Needs the same librt_b200.so that was profiled."""
import csv, io, os, re, subprocess, sys, tempfile, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
skip = {sys.argv[i + 1] for i, a in enumerate(sys.argv[:-1]) if a in ("--cubin", "--sections", "--filter")}
args = [a for a in sys.argv[1:] if not a.startswith("--") and a not in skip]
rep, kern = args[0], args[1]
topn = int(args[2]) if len(args) > 2 else 45
outer = "--outer" in sys.argv
cubin_name = sys.argv[sys.argv.index("--cubin") + 1] if "--cubin" in sys.argv else "rt_kernels"
sections = []
if "--sections" in sys.argv:
    for item in sys.argv[sys.argv.index("--sections") + 1].split(","):
        name, rng = item.split(":")
        a, b = rng.split("-")
        sections.append((name, int(a), int(b)))
args = [a for a in args if a not in (cubin_name,)]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "opengl_raytracer_b200/csrc/librt_b200.so")], cwd=tmp, check=True, capture_output=True)
dis = subprocess.run(["nvdisasm", "--print-line-info-inline", os.path.join(tmp, cubin_name + ".sm_100a.cubin")], capture_output=True, text=True).stdout.splitlines()
# locate the function, collect (file, line) per instruction in order
start = next(i for i, l in enumerate(dis) if l.startswith(".text.") and kern in l)
lines = []
group, last = [], ("?", 0)
fresh = False            # a new group of //## lines started since the last instruction
for l in dis[start + 1:]:
    if l.startswith("\t.section") or l.startswith(".text."):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        if not fresh:
            group = []
            fresh = True
        group.append((os.path.basename(m.group(1)), int(m.group(2))))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        if fresh and group:
            last = group[-1] if outer else group[0]
        fresh = False
        lines.append(last)
flt = ["-k", "regex:" + sys.argv[sys.argv.index("--filter") + 1]] if "--filter" in sys.argv else []   # reports with several kernels
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + flt, capture_output=True, text=True).stdout
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
    key = (f, ln)
    if sections:
        key = next(((name, 0) for name, a, b in sections if a <= ln <= b), ("other:" + f, 0))
    a = agg[key]
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
    where = f"{f}:{ln:<4d} {src(f, ln)}" if not sections else f
    print(f"{100 * n / tot:5.1f}% inst {100 * s / max(tot_s, 1):5.1f}% smp  lanes {t / max(n, 1):5.1f}  sass {k:3d}  {where}")
