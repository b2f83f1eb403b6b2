"""Recipe for oracle/_ref/: a verbatim, git-ignored copy of the reference files the checker and the timed CPU
baseline execute — TEST / BASELINE INFRASTRUCTURE, NOT PRODUCT.

    python oracle/make_ref.py            # needs the reference tree (TDSSF_REFERENCE or /root/reference)

The reference (CharlesChang012/3D-Semantic-Segmentation-Fusion @ e5103fd) is a Python script tree, not a package:
there is nothing to compile.  Its hot-path modules are pure torch and import on a machine without open3d / spconv /
wandb, so the *unmodified* files are what `bench.py --impl reference`, the `cpu_baseline` leg and the
"reference loops run unchanged" tests execute.  /root/reference does not exist on the GPU box; oracle/_ref/ is listed in
.gitignore (the reference's sources never enter this repository's history) and not in .gpurunignore, so the copy made
here by `__graft_entry__.build()` travels to the box like the built libtdssf.so does.

Nothing under 3d-semantic-segmentation-fusion_b200/ reads oracle/_ref/ (tests/test_abi.py greps for it).
MANIFEST.json records the sha256 of every copied file; oracle/ref_loader.py refuses a tree whose hashes differ.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")

# file -> the SURVEY §8 row it serves
FILES = {
    "utils/fusion_model.py": "a1-a6, a12: multi_camera_projector, scale_pixel_coords, FeatureFusionModel",
    "utils/losses.py": "a8: CELSLoss",
    "utils/evaluation.py": "a11: confusion matrix and metrics",
    "LovaszSoftmax/pytorch/lovasz_losses.py": "a9, a10: lovasz_softmax_flat, lovasz_grad",
    "utils/train.py": "b: the training loop that must run unchanged over the drop-ins (utils/train.py:13)",
    "utils/test.py": "b: test_model / test_sample (utils/test.py:13,93)",
    "LovaszSoftmax/LICENSE": "licence of the vendored Lovasz code",
    "LICENSE": "licence of the reference",
}


def sha256(path: str) -> str:
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def reference_root() -> str | None:
    root = os.environ.get("TDSSF_REFERENCE", "/root/reference")
    return root if os.path.isfile(os.path.join(root, "utils", "fusion_model.py")) else None


def make(dest: str = DEST, verbose: bool = True) -> bool:
    """Copies the files; returns False (and leaves `dest` untouched) when the reference tree is not mounted."""
    root = reference_root()
    if root is None:
        if verbose:
            print("oracle/make_ref.py: reference tree not found (TDSSF_REFERENCE or /root/reference); oracle/_ref/ left as is")
        return False
    manifest = {"source": root, "files": {}}
    for rel in FILES:
        src = os.path.join(root, rel)
        if not os.path.isfile(src):
            continue
        dst = os.path.join(dest, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        manifest["files"][rel] = sha256(dst)
    with open(os.path.join(dest, "MANIFEST.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)
    if verbose:
        print(f"oracle/_ref: {len(manifest['files'])} files copied from {root}")
    return True


if __name__ == "__main__":
    sys.exit(0 if make() else 1)
