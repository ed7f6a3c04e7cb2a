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
