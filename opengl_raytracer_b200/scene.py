"""Scene packing — host surface, replaces the reference's ``scene.py``.

``Scene(objects)`` concatenates the objects' world-space vertex streams, builds the 144-byte
``tris[]`` records and the per-triangle material, computes float32 centroids and runs the BVH
builder (``/root/reference/scene.py:9-143``).  ``Scene.send()`` flattens the node list into 48-byte
``boundingBoxes[]`` records plus the ``triangleIndices[]`` array (``scene.py:145-221``) and hands the
three buffers to the device library — the three ``glBufferData`` calls of ``scene.py:223-236`` become
one ``rt_scene_upload`` (``include/rt_b200.h``).  Like the reference, ``send()`` drops the host copies
afterwards (``scene.py:247-267``); ``pack()`` returns them without uploading, for tests and the oracle.

Layouts (std430, ``fragment.glsl:52-66,83-104``):

``tris[]`` 144 B: v0.pos@0 v0.normal@16 v1.pos@32 v1.normal@48 v2.pos@64 v2.normal@80 color@96
emission_color@112 surface(emission, roughness)@128, every vec3 padded to 16 B.
``boundingBoxes[]`` 48 B: numTriangles u32@0 triangleOffset u32@4 childA i32@8 childB i32@12
posMin@16 posMax@32.  Leaf <=> childA == -1.
"""
from __future__ import annotations

import time

import numpy as np

from . import boundingBoxes

vertexStruct = np.dtype([
    ("pos", np.float32, 3), ("_pad0", np.float32),
    ("normal", np.float32, 3), ("_pad1", np.float32),
], align=False)

triangleStruct = np.dtype([
    ("v0", vertexStruct), ("v1", vertexStruct), ("v2", vertexStruct),
    ("color", np.float32, 3), ("_pad_color", np.float32),
    ("emission_color", np.float32, 3), ("_pad_emc", np.float32),
    ("surface", np.float32, 2), ("_pad_final", np.float32, 2),
], align=False)

boundingBoxStruct = np.dtype([
    ("numTriangles", np.uint32), ("triangleOffset", np.uint32),
    ("childA", np.int32), ("childB", np.int32),
    ("posMin", np.float32, 3), ("_pad2", np.float32),
    ("posMax", np.float32, 3), ("_pad3", np.float32),
])

assert triangleStruct.itemsize == 144 and boundingBoxStruct.itemsize == 48


class Scene:
    vertexStruct = vertexStruct
    triangleStruct = triangleStruct
    boundingBoxStruct = boundingBoxStruct

    def __init__(self, objects: list, verbose=False):
        self.objects = objects
        self.verbose = verbose
        streams = {"pos": [], "normals": [], "uvs": []}
        vertex_counts = []
        self.colors, self.emission_colors, self.surfaces = [], [], []
        for obj in objects:
            p = np.asarray(obj.pos, dtype=np.float32)
            streams["pos"].append(p)
            streams["normals"].append(np.asarray(obj.normals, dtype=np.float32))
            streams["uvs"].append(np.asarray(obj.uvs, dtype=np.float32))
            vertex_counts.append(p.shape[0])
            self.colors.append(np.asarray(obj.color, dtype=np.float32))
            self.emission_colors.append(np.asarray(obj.emission_color, dtype=np.float32))
            self.surfaces.append([obj.emission, obj.roughness])

        def stacked(name, width):
            return np.vstack(streams[name]) if streams[name] else np.zeros((0, width), dtype=np.float32)

        self.pos, self.normals, self.uvs = stacked("pos", 3), stacked("normals", 3), stacked("uvs", 2)

        n_tris = self.pos.shape[0] // 3          # leftover vertices are ignored (scene.py:82-84)
        self.tris = np.zeros(n_tris, dtype=triangleStruct)
        self.total_triangles = n_tris
        if n_tris == 0:
            self.poses = np.zeros((0, 3), dtype=np.float32)
            self.indices, self.totalBoxes, self.leaves = [], [], []
            return

        centroid = None
        for k, name in enumerate(("v0", "v1", "v2")):
            corner = self.pos[k::3][:n_tris]
            self.tris[name]["pos"] = corner
            self.tris[name]["normal"] = self.normals[k::3][:n_tris]
            third = corner / 3                    # float32; summed v0/3 + v1/3 + v2/3 (scene.py:93-97)
            centroid = third if centroid is None else centroid + third
        self.poses = centroid

        # material of the object whose vertex range holds the triangle's first vertex (scene.py:116-133)
        starts = np.concatenate(([0], np.cumsum(vertex_counts)))
        owner = np.searchsorted(starts, np.arange(n_tris) * 3, side="right") - 1
        owner = np.clip(owner, 0, len(vertex_counts) - 1)
        self.tris["color"] = np.vstack(self.colors).astype(np.float32)[owner]
        self.tris["emission_color"] = np.vstack(self.emission_colors).astype(np.float32)[owner]
        self.tris["surface"] = np.vstack(self.surfaces).astype(np.float32)[owner]

        started = time.time()
        self.indices, self.totalBoxes, self.leaves = self.getBoundingBoxesCompiled()
        self.bvh_seconds = time.time() - started
        if verbose:
            print(f"BVH: {len(self.totalBoxes):,} nodes in {self.bvh_seconds:.2f} s")

    def getBoundingBoxesCompiled(self):
        builder = boundingBoxes.BoundingBoxes()
        return builder.getBoundingBoxes(self.total_triangles, self.tris, self.poses, verbose=self.verbose)

    # ------------------------------------------------------------------ packing
    def pack(self):
        """Return ``(tris, boxes, indices)`` exactly as the reference's ``send()`` uploads them."""
        n = len(self.totalBoxes)
        boxes = np.zeros(n, dtype=boundingBoxStruct)
        child_a = np.full(n, -1, dtype=np.int32)
        child_b = np.full(n, -1, dtype=np.int32)
        for i, node in enumerate(self.totalBoxes):
            # a node without childB is a leaf to the shader even if it has a childA (scene.py:168-179)
            if "childB" in node:
                child_b[i] = node["childB"]
                if "childA" in node:
                    child_a[i] = node["childA"]
        if n:
            boxes["numTriangles"] = np.array([b["numTriangles"] for b in self.totalBoxes], dtype=np.uint32)
            boxes["triangleOffset"] = np.array([b["triangleOffset"] for b in self.totalBoxes], dtype=np.uint32)
            boxes["posMin"] = np.array([b["posMin"] for b in self.totalBoxes], dtype=np.float32)
            boxes["posMax"] = np.array([b["posMax"] for b in self.totalBoxes], dtype=np.float32)
        boxes["childA"] = child_a
        boxes["childB"] = child_b
        return np.ascontiguousarray(self.tris), boxes, np.array(self.indices, dtype=np.uint32)

    def send(self, context=None):
        """Upload the three scene buffers to the device (reference ``scene.py:145-267``)."""
        from . import rtlib

        ctx = context if context is not None else rtlib.current_context()
        tris, boxes, indices = self.pack()
        self.total_boxes = len(boxes)
        ctx.scene_upload(tris, boxes, indices)
        self._context = ctx
        if self.verbose and self.leaves:
            sizes = [len(leaf[0]) for leaf in self.leaves]
            print(f"triangles {len(tris):,}  nodes {len(boxes):,}  last-level leaf size "
                  f"min {min(sizes)} mean {np.mean(sizes):.1f} max {max(sizes)}")
        for name in ("tris", "objects", "totalBoxes", "indices", "leaves", "pos", "normals", "colors",
                     "uvs", "emission_colors", "surfaces", "poses"):
            self.__dict__.pop(name, None)

    def clearMemory(self):
        ctx = self.__dict__.pop("_context", None)
        if ctx is not None:
            ctx.scene_free()
