"""Run the REFERENCE's own host code (object.py / scene.py / *.pyx under /root/reference) on a scene
description and dump the three SSBO payloads it would upload.  Test infrastructure only.

The reference imports PyOpenGL, which is not installed and cannot work here (no graphics stack), so a
stub ``OpenGL.GL`` module that records ``glBufferData`` payloads is put on sys.path first.
Nothing from the reference is copied: its modules are imported from where they lie, with the cwd set
to the reference root because ``Mesh`` opens ``models/<dir>`` relative to the cwd (object.py:9).
pyximport builds go to $HOME/.pyxbld; the caller points HOME at a scratch directory inside the repo.

usage: ref_host_dump.py <scene.json> <out.npz> [reference_root]
"""
import json
import os
import sys
import tempfile


def main():
    spec_path, out_path = os.path.abspath(sys.argv[1]), os.path.abspath(sys.argv[2])
    ref_root = sys.argv[3] if len(sys.argv) > 3 else "/root/reference"
    stub_root = tempfile.mkdtemp(prefix="glstub_")
    os.makedirs(os.path.join(stub_root, "OpenGL"))
    open(os.path.join(stub_root, "OpenGL", "__init__.py"), "w").close()
    with open(os.path.join(stub_root, "OpenGL", "GL.py"), "w") as f:
        f.write(
            "import numpy as _np\n"
            "GL_SHADER_STORAGE_BUFFER = 0x90D2\nGL_DYNAMIC_DRAW = 0x88E8\n"
            "uploads = []\n_next = [0]\n"
            "def glGenBuffers(n):\n    _next[0] += 1\n    return _next[0]\n"
            "def glBindBuffer(target, buf):\n    pass\n"
            "def glBindBufferBase(target, index, buf):\n    uploads[-1]['binding'] = index\n"
            "def glBufferData(target, nbytes, data, usage):\n"
            "    uploads.append({'nbytes': int(nbytes), 'data': _np.array(data, copy=True)})\n"
            "def glDeleteBuffers(n, bufs):\n    pass\n"
        )
    sys.path.insert(0, stub_root)
    sys.path.insert(0, ref_root)
    os.chdir(ref_root)

    import numpy as np
    import OpenGL.GL as gl
    import object as ref_object          # noqa: the reference's module
    import scene as ref_scene            # noqa: the reference's module

    with open(spec_path) as f:
        specs = json.load(f)
    objects = []
    for spec in specs:
        spec = dict(spec)
        kind = spec.pop("kind")
        if kind == "Mesh":
            objects.append(ref_object.Mesh(spec.pop("pos"), spec.pop("eulers"), spec.pop("dirPath"), **spec))
        else:
            objects.append(ref_object.Rect(spec.pop("size"), spec.pop("pos"), spec.pop("eulers"), **spec))
    sc = ref_scene.Scene(objects)
    sc.send()
    by_binding = {u["binding"]: u for u in gl.uploads}
    np.savez(out_path,
             tris=np.frombuffer(by_binding[0]["data"].tobytes(), dtype=np.uint8),
             boxes=np.frombuffer(by_binding[1]["data"].tobytes(), dtype=np.uint8),
             indices=np.frombuffer(by_binding[2]["data"].tobytes(), dtype=np.uint8),
             total_triangles=sc.total_triangles, total_boxes=sc.total_boxes)


if __name__ == "__main__":
    main()
