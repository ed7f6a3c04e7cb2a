"""BVH builder — host surface, replaces the reference's ``boundingBoxes.pyx``.

The tree is the hot path's *input contract*: node order decides the traversal order and therefore
which triangle wins an exact-``t`` tie, so the build below reproduces the reference's tree exactly
(checked byte for byte against ``/root/reference/boundingBoxes.pyx`` when it is mounted):

* depth: ``ceil(log2 N) - 1`` levels, every node of a level is split (``boundingBoxes.pyx:27,45-94``);
* split: triangle centroids along the axis of largest centroid extent, ``> mean`` goes to childB and
  the rest to childA (``:162-175`` — float32 ``mean``/``ptp`` through the same numpy calls);
* numbering: children are appended in breadth-first order; a child's stored index is
  ``parent + visited + skipped`` with the reference's running counters (``:69-92``), which equals its
  position in the list whenever the ``<= mean`` half is non-empty;
* a node missing either child is a leaf: it owns ``numTriangles`` entries of ``indices`` starting at
  ``triangleOffset`` (``:121-130``).  A node whose only child is childA keeps that (unreachable) chain
  in the list, and the chain's last node re-adds the same triangles to ``indices`` — kept as is;
* bounds: min/max over the three vertex streams of the node's triangles (``:134-160``).

One-off host preprocessing (about 2 s for 38 k triangles), not the hot path.
"""
from __future__ import annotations

import math

import numpy as np


class BoundingBoxes:
    def __init__(self):
        pass

    def getBoundingBoxes(self, total_triangles, tris, poses, verbose=False):
        self.total_triangles = total_triangles
        self.tris = tris
        self.poses = poses
        self._corner_streams = tuple(tris[v]["pos"] for v in ("v0", "v1", "v2"))

        everything = np.array(list(range(total_triangles)))
        nodes = [{"box": everything}]
        frontier = [[everything, 0]]
        levels = math.ceil(math.log(total_triangles, 2)) - 1

        visited = 0      # the reference's ``idx``
        skipped = 0      # the reference's ``minus`` (counts empty halves, negative)
        for level in range(levels):
            if verbose:
                print(f"level {level + 1}/{levels}: {len(nodes):,} nodes")
            parents, frontier = frontier, []
            for members, parent in parents:
                visited += 1
                slot_low = parent + visited + skipped
                slot_high = slot_low + 1
                low, high = self.sliceBoundingBox(members)
                for key, half, slot in (("childA", low, slot_low), ("childB", high, slot_high)):
                    if len(half) >= 1:
                        nodes[parent][key] = slot
                        frontier.append([half, slot])
                        nodes.append({"box": half})
                    else:
                        skipped -= 1

        indices = []
        cursor = 0
        for node in nodes:
            members = node["box"]
            node["posMin"], node["posMax"] = self.getBoundingBoxCorners(members)
            if "childA" in node and "childB" in node:
                node["numTriangles"] = np.uint32(0)
                node["triangleOffset"] = np.uint32(0)
            else:
                node["numTriangles"] = np.uint32(len(members))
                node["triangleOffset"] = np.uint32(cursor)
                cursor += len(members)
                indices.extend(members.tolist())
        return indices, nodes, frontier

    def getBoundingBoxCorners(self, box):
        if len(box) == 0:
            raise ValueError("box must contain at least one triangle")
        rows = np.asarray(box, dtype=np.intp)
        lo = hi = None
        for stream in self._corner_streams:
            corners = stream[rows].astype(np.float32, copy=False)
            smin, smax = corners.min(axis=0), corners.max(axis=0)
            lo = smin if lo is None else np.minimum(lo, smin)
            hi = smax if hi is None else np.maximum(hi, smax)
        return lo, hi

    def sliceBoundingBox(self, indices):
        """Return (``<= mean`` half, ``> mean`` half) of ``indices`` along the widest centroid axis."""
        if indices.size == 0:
            return np.empty(0, dtype=np.int32), np.empty(0, dtype=np.int32)
        centroids = self.poses[indices]
        axis = np.argmax(np.ptp(centroids, axis=0))
        above = centroids[:, axis] > centroids.mean(axis=0)[axis]
        return indices[~above], indices[above]
