"""Wavefront OBJ reader — host surface, replaces the reference's ``loadObject.pyx``.

Behaviour contract (what the reference's reader does, cited per rule; /root/reference paths):

* a line is split on single spaces only, empty tokens dropped (``loadObject.pyx:23-29``);
  tabs are NOT separators, so ``"v\\t1 2 3"`` is ignored like the reference ignores it;
* ``v`` keeps the LAST three tokens (``loadObject.pyx:113-118``) so ``v x y z w`` drops x;
* ``vt`` keeps the first two tokens, ``vn`` the first three (``:120-131``);
* ``f`` is fan-triangulated around its first corner (``:53-67``);
* a corner is ``p/t/n``, ``p//n``, ``p/t/``, ``p/t`` or ``p`` (``:69-111``); a missing normal is
  (0,0,1), a missing texcoord (0,0); v is flipped (``1 - v``) after lookup (``:109``);
* indices are ``int(tok) - 1`` looked up with Python list indexing against the lists as they are
  when the face is read, so a relative index ``-1`` reaches the second-to-last entry, exactly as
  in the reference (``:82-104``);
* everything else (``o``, ``g``, ``s``, ``usemtl``, ``mtllib``, comments) is skipped.

The result is a flat list of 8-float vertices ``[px,py,pz,nx,ny,nz,u,v]`` (Python floats), three per
triangle, in file order.  This is host preprocessing: it runs once per mesh and is not on the hot path.
"""
from __future__ import annotations

_DEFAULT_NORMAL = (0.0, 0.0, 1.0)
_DEFAULT_UV = (0.0, 0.0)


def _corner(token, positions, normals, texcoords):
    parts = token.split("/")
    p = positions[int(parts[0]) - 1]
    uv = _DEFAULT_UV
    n = _DEFAULT_NORMAL
    if len(parts) == 3:
        if parts[1] != "":
            uv = texcoords[int(parts[1]) - 1]
        if parts[2] != "":
            n = normals[int(parts[2]) - 1]
    elif len(parts) == 2:
        uv = texcoords[int(parts[1]) - 1]
    return [p[0], p[1], p[2], n[0], n[1], n[2], uv[0], 1 - uv[1]]


def loadObj(filePath, verbose=False):
    """Parse ``filePath`` and return the triangle-soup vertex list (see module docstring)."""
    positions, texcoords, normals, out = [], [], [], []
    with open(filePath, "r") as handle:
        for raw in handle:
            tokens = [t for t in raw.strip().split(" ") if t != ""]
            if not tokens:
                continue
            tag = tokens[0]
            if tag == "v":
                positions.append((float(tokens[-3]), float(tokens[-2]), float(tokens[-1])))
            elif tag == "vt":
                texcoords.append((float(tokens[1]), float(tokens[2])))
            elif tag == "vn":
                normals.append((float(tokens[1]), float(tokens[2]), float(tokens[3])))
            elif tag == "f":
                corners = tokens[1:]
                for k in range(len(corners) - 2):
                    out.append(_corner(corners[0], positions, normals, texcoords))
                    out.append(_corner(corners[1 + k], positions, normals, texcoords))
                    out.append(_corner(corners[2 + k], positions, normals, texcoords))
    if verbose:
        print(f"{filePath}: {len(out) // 3:,} triangles")
    return out
