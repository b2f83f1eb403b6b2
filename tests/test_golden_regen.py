"""The golden recipe is reproducible: tests/golden/make_golden.py, run here against the mounted reference tree, writes
files bit-identical to the committed ones, and every module it executed came from the reference tree (not from the
drop-in package of the same name).  Skipped where the reference is not mounted (GPU box)."""
import glob
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import make_ref, ref_loader  # noqa: E402

needs_ref = pytest.mark.skipif(make_ref.reference_root() is None, reason="reference tree not mounted")


@needs_ref
def test_loader_resolves_to_the_reference_tree():
    root = make_ref.reference_root()
    m = ref_loader.load_hot_path(root)
    for mod in (m.fusion_model, m.losses, m.evaluation, m.lovasz_losses):
        assert os.path.abspath(mod.__file__).startswith(os.path.abspath(root) + os.sep)
    # the reference's own statement of the module tree (utils/fusion_model.py:144-150)
    model = m.fusion_model.FeatureFusionModel(None, None, 64, 384, 16, device="cpu")
    assert [k for k in model.state_dict()] == [f"mlp.{i}.{p}" for i in (0, 2, 4) for p in ("weight", "bias")]
    assert m.losses.lovasz_softmax_flat is m.lovasz_losses.lovasz_softmax_flat


@needs_ref
def test_goldens_regenerate_bit_identically(tmp_path, golden_dir):
    out = str(tmp_path / "golden")
    r = subprocess.run([sys.executable, os.path.join(golden_dir, "make_golden.py"), "--out", out],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout + r.stderr
    committed = sorted(glob.glob(os.path.join(golden_dir, "*.npz")))
    assert len(committed) == 13
    assert sorted(os.path.basename(f) for f in glob.glob(os.path.join(out, "*.npz"))) == [os.path.basename(f) for f in committed]
    for f in committed:
        a, b = np.load(f), np.load(os.path.join(out, os.path.basename(f)))
        assert sorted(a.files) == sorted(b.files), f
        for k in a.files:
            assert a[k].dtype == b[k].dtype and a[k].shape == b[k].shape and a[k].tobytes() == b[k].tobytes(), (f, k)


@needs_ref
def test_make_ref_copies_verbatim(tmp_path):
    dest = str(tmp_path / "_ref")
    assert make_ref.make(dest, verbose=False)
    root = make_ref.reference_root()
    for rel in ("utils/fusion_model.py", "utils/losses.py", "utils/train.py", "utils/test.py",
                "LovaszSoftmax/pytorch/lovasz_losses.py"):
        assert open(os.path.join(dest, rel), "rb").read() == open(os.path.join(root, rel), "rb").read()
