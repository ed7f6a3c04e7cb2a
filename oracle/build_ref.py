"""Compile the REFERENCE's own shaders/fragment.glsl as C++ -> oracle/_ref/libfragment_ref.so.

TEST INFRASTRUCTURE ONLY.  The reference's hot path is a GLSL fragment shader; no graphics API exists on
either machine, but the shader source is plain C-family code: with the type shim in oracle/glsl_shim.h
and the mechanical rewrites below it compiles unchanged as C++17.  The shader
is read from where it lies under /root/reference, rewritten into a scratch directory OUTSIDE the
repo's tracked tree (a temp dir), and only the resulting shared object lands in oracle/_ref/ (git-ignored;
it travels to the GPU box with gpurun).  No reference source is copied into the repository.

Rewrites (1-8 mechanical, 9 hand-sequenced because C++ leaves operand evaluation order open while the
pinned semantics draw left to right):
 1 drop ``#version``                       2 ``in``/``out`` varyings -> thread_local globals
 3 ``uniform T x;`` -> ``static T x;``     4 shader-scope mutable globals -> thread_local
 5 the three SSBO blocks -> const pointers 6 ``inout uint`` -> ``uint&``
 7 ``void main()`` -> ``void shader_main()``
 8 every unsuffixed real literal gets an ``f`` suffix (GLSL literals are fp32)
 9 ``vec3(R(), R(), R())`` (fragment.glsl:221) -> brace-init (left-to-right by the C++ standard) and the
   jitter line (fragment.glsl:398) -> two named draws, camRight first.
"""
from __future__ import annotations

import os
import re
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
OUT_DIR = os.path.join(HERE, "_ref")
OUT_LIB = os.path.join(OUT_DIR, "libfragment_ref.so")
DEFAULT_SHADER = "/root/reference/shaders/fragment.glsl"

_REAL = re.compile(r"(?<![\w.])((?:\d+\.\d*|\.\d+)(?:[eE][-+]?\d+)?|\d+[eE][-+]?\d+)(?![\w.])")


def translate(glsl: str) -> str:
    src = glsl
    src = re.sub(r"^#version.*$", "", src, flags=re.M)                                        # 1
    src = re.sub(r"^in\s+vec2\s+uv\s*;", "static thread_local vec2 uv;", src, flags=re.M)       # 2
    src = re.sub(r"^out\s+vec4\s+color\s*;", "static thread_local vec4 color;", src, flags=re.M)
    src = re.sub(r"^uniform\s+", "static ", src, flags=re.M)                                   # 3
    src = re.sub(r"^(uint\s+seed\s*;|vec3\s+skyColor\s*;|float\s+seed2\s*=)", r"static thread_local \1",
                 src, flags=re.M)                                                              # 4
    src, n_blocks = re.subn(r"layout\s*\(\s*std430[^)]*\)\s*buffer\s+\w+\s*\{\s*(\w+)\s+(\w+)\s*\[\s*\]\s*;\s*\}\s*;",
                            r"static const \1* \2;", src)                                      # 5
    assert n_blocks == 3, f"expected 3 SSBO blocks, rewrote {n_blocks}"
    src = re.sub(r"\binout\s+uint\s+(\w+)", r"uint& \1", src)                                   # 6
    src, n_main = re.subn(r"\bvoid\s+main\s*\(\s*\)", "void shader_main()", src)                # 7
    assert n_main == 1
    src = _REAL.sub(lambda m: m.group(1) + "f", src)                                           # 8
    src, n_a = re.subn(r"vec3\(\s*RandomValue\(seed\)\s*,\s*RandomValue\(seed\)\s*,\s*RandomValue\(seed\)\s*\)",
                       "vec3{RandomValue(seed), RandomValue(seed), RandomValue(seed)}", src)    # 9
    src, n_b = re.subn(
        r"dir\s*\+=\s*\(\s*camRight\s*\*\s*RandomValue\(seed\)\s*\+\s*camUp\s*\*\s*RandomValue\(seed\)\s*\)\s*\*\s*jitterAmount\s*;",
        "{ float jitterRight = RandomValue(seed); float jitterUp = RandomValue(seed); "
        "dir += (camRight * jitterRight + camUp * jitterUp) * jitterAmount; }", src)
    assert n_a == 1 and n_b == 1, "hand-sequenced RNG lines not found"
    return src


DRIVER = r'''
#include "glsl_shim.h"
static thread_local vec4 gl_FragCoord;
#include "fragment_translated.inc"
#include <atomic>
#include <cstdlib>
#include <thread>
#include <vector>
#include "rt_b200.h"

static_assert(sizeof(Triangle) == 144 && sizeof(BoundingBox) == 48, "std430 strides");

static int ref_threads() {
    const char* env = std::getenv("ORACLE_THREADS");
    int n = env ? std::atoi(env) : int(std::thread::hardware_concurrency());
    return n < 1 ? 1 : n;
}

extern "C" int ref_num_threads(void) { return ref_threads(); }

// One draw of the full-screen quad (main.py:376-394): every fragment of the W x H viewport.
extern "C" int ref_render_pass(const void* tris_, uint32_t, const void* boxes_, uint32_t n_boxes,
                               const uint32_t* indices_, uint32_t, const rt_params* p, int width, int height,
                               const float* prev_rgba, float* next_rgba) {
    tris = static_cast<const Triangle*>(tris_);
    boundingBoxes = static_cast<const BoundingBox*>(boxes_);
    triangleIndices = indices_;
    tileX = p->tileX; tileY = p->tileY; tileSizeX = p->tileSizeX; tileSizeY = p->tileSizeY;
    numTilesX = p->numTilesX; numTilesY = p->numTilesY; total_frames = p->total_frames;
    lambertian = p->lambertian != 0; skyBrightness = p->skyBrightness; fov = p->fov;
    xStep = p->xStep; yStep = p->yStep; dirStartX = p->dirStartX; dirStartY = p->dirStartY;
    camPos = vec3(p->camPos[0], p->camPos[1], p->camPos[2]);
    camRight = vec3(p->camRight[0], p->camRight[1], p->camRight[2]);
    camUp = vec3(p->camUp[0], p->camUp[1], p->camUp[2]);
    camForward = vec3(p->camForward[0], p->camForward[1], p->camForward[2]);
    nBounces = p->nBounces; rays_per_pixel = p->rays_per_pixel; jitterAmount = p->jitterAmount;
    frameNumber = p->frameNumber; trisCount = p->trisCount; boundingBoxCount = int(n_boxes);
    prevFrame.texels = prev_rgba; prevFrame.width = width; prevFrame.height = height;

    std::atomic<int> next_row{0};
    auto worker = [&]() {
        for (;;) {
            int py = next_row.fetch_add(1);
            if (py >= height) break;
            for (int px = 0; px < width; px++) {
                gl_FragCoord = vec4(float(px) + 0.5f, float(py) + 0.5f, 0.0f, 1.0f);
                uv = vec2((float(px) + 0.5f) / float(width), (float(py) + 0.5f) / float(height));
                shader_main();
                float* o = next_rgba + 4 * (size_t(py) * size_t(width) + size_t(px));
                o[0] = color.x; o[1] = color.y; o[2] = color.z; o[3] = color.w;
            }
        }
    };
    std::vector<std::thread> pool;
    int n = ref_threads();
    for (int t = 1; t < n; t++) pool.emplace_back(worker);
    worker();
    for (auto& th : pool) th.join();
    return 0;
}
'''


def build(shader_path: str = DEFAULT_SHADER, quiet: bool = False) -> str | None:
    """Build oracle/_ref/libfragment_ref.so; returns its path, or None when the reference is absent."""
    if not os.path.isfile(shader_path):
        return OUT_LIB if os.path.isfile(OUT_LIB) else None
    os.makedirs(OUT_DIR, exist_ok=True)
    if os.path.isfile(OUT_LIB) and all(os.path.getmtime(OUT_LIB) > os.path.getmtime(p) for p in
                                       (shader_path, __file__, os.path.join(HERE, "glsl_shim.h"))):
        return OUT_LIB
    with open(shader_path) as f:
        translated = translate(f.read())
    with tempfile.TemporaryDirectory(prefix="fragref_") as scratch:
        with open(os.path.join(scratch, "fragment_translated.inc"), "w") as f:
            f.write(translated)
        with open(os.path.join(scratch, "ref_driver.cpp"), "w") as f:
            f.write(DRIVER)
        cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-fno-fast-math", "-pthread", "-shared",
               "-I", HERE, "-I", os.path.join(HERE, "..", "include"), "-I", scratch,
               "-o", OUT_LIB, os.path.join(scratch, "ref_driver.cpp")]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        if proc.returncode != 0:
            raise RuntimeError("reference shader did not compile as C++:\n" + proc.stderr[-4000:])
    if not quiet:
        print(f"built {OUT_LIB}")
    return OUT_LIB


if __name__ == "__main__":
    print(build(*(sys.argv[1:2])))
