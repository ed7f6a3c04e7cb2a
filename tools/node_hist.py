"""Analysis tool (CPU): where in the tree do node expansions happen?  Uses an instrumented build of the oracle
(-DRT_ORACLE_HIST).  Prints the share of internal-node expansions that fall on the first K nodes in breadth-first
order — the quantity that sizes an on-chip (shared-memory) copy of the top of the tree.
usage: python tools/node_hist.py <scene> <w> <h> <spp>"""
import ctypes, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle
from opengl_raytracer_b200 import configs
from opengl_raytracer_b200.scene import Scene
lib_path = "/tmp/librt_oracle_hist.so"
subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-fno-fast-math", "-pthread", "-DRT_ORACLE_HIST",
                "-shared", "-o", lib_path, os.path.join(ROOT, "oracle", "rt_oracle.cpp")], check=True)
oracle.PORT_LIB = lib_path
name, w, h, spp = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
cfg = configs.scene_configs()[name].scaled(w, h, spp)
tris, boxes, idx = Scene(configs.build_objects(cfg.objects)).pack()
hist = np.zeros(len(boxes), dtype=np.uint64)
lib = ctypes.CDLL(lib_path)
lib.oracle_set_node_hist.argtypes = [ctypes.c_void_p]
lib.oracle_set_node_hist(hist.ctypes.data)
_, _, cnt = oracle.render_pass(tris, boxes, idx, cfg.params(), want_counters=True)
# breadth-first rank of internal nodes, as rt_scene_upload numbers them
ca, cb = boxes["childA"], boxes["childB"]
order, q = [], [0]
while q:
    nq = []
    for i in q:
        if ca[i] == -1: continue
        order.append(i)
        for c in (ca[i], cb[i]):
            if c >= 0: nq.append(c)
    q = nq
order = np.array(order)
byrank = hist[order].astype(np.float64)
tot = byrank.sum()
print(f"{name}: {len(order)} internal nodes, {int(tot)} expansions, counters internal={cnt['internal']} rays={cnt['rays']}")
cum = np.cumsum(byrank) / tot
for k in (64, 128, 256, 512, 1024, 1536, 2048, 3072, 4096, 8192, 16384):
    if k <= len(order):
        print(f"  first {k:6d} nodes ({k * 64 / 1024:7.1f} KB as 64-B records, {k*48/1024:7.1f} KB as 48-B): {100 * cum[k - 1]:5.1f} % of expansions")
srt = np.sort(byrank)[::-1]; cs = np.cumsum(srt) / tot
for k in (512, 1024, 2048, 4096):
    if k <= len(order):
        print(f"  hottest {k:6d} nodes: {100 * cs[k - 1]:5.1f} %")
