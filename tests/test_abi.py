"""The C-ABI library loads and exports every symbol include/rt_b200.h declares; without a GPU the render
path fails loudly instead of falling back to anything."""
import ctypes
import os
import re

import pytest

from conftest import ROOT
from opengl_raytracer_b200 import rtlib


def _declared():
    text = open(os.path.join(ROOT, "include", "rt_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    if not os.path.isfile(rtlib.LIB_PATH):
        pytest.skip("librt_b200.so not built (python -c 'import __graft_entry__ as g; g.build()')")
    names = _declared()
    assert len(names) >= 30
    lib = ctypes.CDLL(rtlib.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    bound = rtlib.load_library()
    assert bound.rt_version().startswith(b"librt_b200")


def test_params_struct_matches_header():
    # field order of struct rt_params == ctypes mirror (same names as the shader's uniforms)
    text = open(os.path.join(ROOT, "include", "rt_b200.h")).read()
    body = re.search(r"typedef struct rt_params \{(.*?)\} rt_params;", text, flags=re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        decl = re.sub(r"^(int32_t|float)\s+", "", decl)
        names += [re.sub(r"\[\d+\]", "", n.strip()) for n in decl.split(",")]
    assert names == [f[0] for f in rtlib.RtParams._fields_]
    assert ctypes.sizeof(rtlib.RtParams) == 4 * (6 + 2 + 6 + 12 + 2 + 1 + 1 + 2)


def test_no_cpu_fallback_without_gpu():
    if not os.path.isfile(rtlib.LIB_PATH):
        pytest.skip("librt_b200.so not built")
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError):
        rtlib.Context(0)
    with pytest.raises(RuntimeError):
        rtlib.load_library(os.path.join(ROOT, "does_not_exist.so"))


def test_every_entry_point_rejects_a_null_context():
    """Argument validation happens before anything touches CUDA, so it can be checked without a GPU: every function that
    takes the context refuses NULL (error code, null pointer or no-op) instead of dereferencing it."""
    if not os.path.isfile(rtlib.LIB_PATH):
        pytest.skip("librt_b200.so not built")
    text = open(os.path.join(ROOT, "include", "rt_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    protos = re.findall(r"^\s*(int|void\s*\*|const char\s*\*|void)\s+(rt_[a-z0-9_]+)\s*\(\s*rt_ctx\s*\*", text, flags=re.M)
    assert len(protos) >= 40, len(protos)
    lib = ctypes.CDLL(rtlib.LIB_PATH)
    null = ctypes.c_void_p(None)
    for ret, name in protos:
        fn = getattr(lib, name)
        kind = ret.replace(" ", "")
        fn.restype = ctypes.c_int if kind == "int" else None if kind == "void" else ctypes.c_void_p
        out = fn(null, null, null, null, null, null, null)        # the other arguments are never looked at
        if name in ("rt_last_pass_launches", "rt_peer_is_open"):
            assert out == 0, name                                   # queries: "none"
        elif kind == "int":
            assert out != 0, f"{name}(NULL) returned RT_OK"
        elif name == "rt_last_error":
            assert ctypes.string_at(out), "rt_last_error(NULL) names the problem"      # a static message, not a crash
        elif kind != "void":
            assert not out, f"{name}(NULL) returned a pointer"
