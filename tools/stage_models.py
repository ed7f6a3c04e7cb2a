"""Stage the reference's OBJ assets into ./models (git-ignored, travels to the GPU box with gpurun).

The assets are data inputs named by BASELINE.json's configs (models/sphere, models/ground,
models/stanford_minidragon, models/dragon, models/knight); they are not source code and are kept out
of the git history.  ``Mesh(dirPath=...)`` resolves ``models/<dirPath>/*.obj`` like the reference
(``/root/reference/object.py:9,17-23``).  Four OBJ files are absent from the reference checkout
(``.MISSING_LARGE_BLOBS``) and cannot be staged; SURVEY §8d lists the substitutions.
"""
import os
import shutil
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def stage(src_root="/root/reference/models", dst_root=None, quiet=False):
    dst_root = dst_root or os.path.join(REPO, "models")
    if not os.path.isdir(src_root):
        return 0
    copied = 0
    for entry in sorted(os.listdir(src_root)):
        src_dir = os.path.join(src_root, entry)
        if not os.path.isdir(src_dir):
            continue
        for name in sorted(os.listdir(src_dir)):
            if not name.endswith(".obj"):
                continue
            dst_dir = os.path.join(dst_root, entry)
            dst = os.path.join(dst_dir, name)
            src = os.path.join(src_dir, name)
            if os.path.exists(dst) and os.path.getsize(dst) == os.path.getsize(src):
                continue
            os.makedirs(dst_dir, exist_ok=True)
            shutil.copyfile(src, dst)
            copied += 1
            if not quiet:
                print(f"staged {dst}")
    return copied


if __name__ == "__main__":
    stage(*sys.argv[1:2])
