import sys, time, torch
sys.path.insert(0, "/root/repo")
import numpy as np
from opengl_raytracer_b200 import configs, rtlib, distributed
from opengl_raytracer_b200.scene import Scene
dev = torch.device("cuda:0")
cfg = configs.scene_configs()["cfg4_open"]
tris, boxes, indices = Scene(configs.build_objects(cfg.objects)).pack()
ctx = rtlib.Context(0); ctx.use_torch_stream()
host = {}
for key, arr in (("tris", tris), ("boxes", boxes), ("indices", indices)):
    t_ = torch.empty(arr.nbytes, dtype=torch.uint8).pin_memory()
    t_.numpy()[:] = arr.view(np.uint8).reshape(-1)
    host[key] = t_
def upload():
    ctx._check(ctx.lib.rt_scene_upload(ctx.handle, host["tris"].data_ptr(), len(tris), host["boxes"].data_ptr(), len(boxes), host["indices"].data_ptr(), indices.size))
upload()
params = cfg.params(); ctx.set_params(params); ctx.target_create(cfg.width, cfg.height)
out_host = torch.empty((cfg.height, cfg.width, 4), dtype=torch.float32).pin_memory()
def t(fn, n=10):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    h0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); h1 = time.perf_counter()
    return round((h1 - h0) * 1e3 / n, 3)
print("upload wall ms", t(upload))
print("read_accum wall ms", t(lambda: ctx._check(ctx.lib.rt_read_accum_f32(ctx.handle, out_host.data_ptr()))))
dst = torch.empty((cfg.height, cfg.width, 4), dtype=torch.float32, device=dev)
print("torch d2h 132MB pinned", t(lambda: out_host.copy_(dst, non_blocking=True)))
print("pass", t(lambda: ctx.render_pass(0, 0, 0, flags=rtlib.PASS_SKIP_MASKED), 5))
