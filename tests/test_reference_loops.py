"""Row b on the GPU: the reference's UNMODIFIED loops — utils/train.py:train_model, utils/test.py:test_model and
test_sample, loaded by file path from the reference tree (build container) or from its verbatim copy oracle/_ref/
(GPU box) — run over the drop-in modules (FeatureFusionModel, CELSLoss, evaluation) with a synthetic `dataloaders`
dict and stub encoders in front of them.  wandb and utils.plot are recording stand-ins."""
import copy
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import harness  # noqa: E402
from oracle import ref_loader  # noqa: E402
from tdssf import synth  # noqa: E402


@pytest.fixture(scope="module")
def loops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    if ref_loader.ref_root() is None:
        pytest.skip("neither the reference tree nor oracle/_ref/ is present (run __graft_entry__.build() where the reference is mounted)")
    import utils.evaluation as dropin_eval
    assert "3d-semantic-segmentation-fusion_b200" in dropin_eval.__file__
    return ref_loader.load_loops(dropin_eval)


def _host(x):
    return np.asarray(x.detach().cpu() if torch.is_tensor(x) else x, dtype=np.float64)


def _setup(n_batches=4, B=2, V=3000, seed=70):
    from utils.fusion_model import FeatureFusionModel
    from utils.losses import CELSLoss
    sweeps, batches = harness.make_batches(n_batches, B, V, seed=seed)
    enc, pcd = harness.TokenLookupEncoder(sweeps), harness.IdentityVoxels().cuda()
    model = FeatureFusionModel(enc, pcd, 64, 384, 16, device="cuda").cuda()
    model.load_state_dict({**synth.make_mlp_weights(seed), "pcd_encoder.scale": torch.ones(64)}, strict=True)
    crit = CELSLoss(weight=torch.tensor(synth.CLASS_WEIGHTS).cuda(), ignore_index=0, lamda_lovasz=1.0)
    return batches, enc, pcd, model, crit


def test_reference_train_model_runs_unchanged(loops, tmp_path, capsys):
    """utils/train.py:13 train_model: forward (:83) -> criterion (:86) -> zero_grad / backward / step (:90-92) ->
    .item() (:94) -> validation through utils.evaluation.evaluate (:127) -> best-weights checkpoint (:131-136)."""
    batches, enc, pcd, model, crit = _setup()
    loaders = {"train": batches[:3], "val": batches[3:]}
    opt = torch.optim.AdamW(model.parameters(), lr=2e-3)
    w0 = copy.deepcopy(model.state_dict())
    tr_hist, val_hist = loops.train.train_model(loaders, enc, pcd, model, opt, crit, "cuda", save_dir=str(tmp_path),
                                               num_epochs=4, fusion_model_name="3DSSF", config={})
    out = capsys.readouterr().out
    assert "EVALUATION METRICS" in out and "Best val Acc" in out
    assert len(tr_hist) == 4 and len(val_hist) == 4
    train_loss = [e["train/loss"] for e in loops.wandb.logged if "train/loss" in e]
    assert len(train_loss) == 4 * 3
    # running mean of the epoch at its last step: the loss goes down under the drop-in forward / loss / backward
    assert train_loss[-1] < train_loss[2] - 0.05, train_loss
    assert any(not torch.equal(w0[k].cpu(), v.cpu()) for k, v in model.state_dict().items() if k.startswith("mlp."))
    assert not torch.equal(w0["pcd_encoder.scale"].cpu(), model.state_dict()["pcd_encoder.scale"].cpu()), \
        "the point-feature gradient must reach the pcd encoder (SURVEY F8)"
    ckpt = torch.load(os.path.join(str(tmp_path), "3DSSF.pth"), map_location="cpu")
    assert sorted(ckpt) == sorted(model.state_dict())          # the reference's checkpoint keys (a12)
    val = [e for e in loops.wandb.logged if "val/mean_IoU" in e]
    assert len(val) == 4 and all(0.0 <= e["val/acc"] <= 1.0 for e in val)


def test_reference_test_model_and_sample_match_dropin_loops(loops, tmp_path, capsys):
    """utils/test.py:13 test_model and :93 test_sample, unmodified, against the device-resident drop-in loops."""
    from utils import test as dropin_loops
    batches, enc, pcd, model, crit = _setup(n_batches=3, seed=80)
    loaders = {"test": batches}
    want = loops.test.test_model(loaders, enc, pcd, model, crit, "cuda")
    got = dropin_loops.test_model(loaders, enc, pcd, model, crit, "cuda")
    assert set(got) == set(want)
    for k in want:   # the reference returns overall_acc as a CUDA tensor (total_corrects.double() / total_points)
        assert np.allclose(_host(got[k]), _host(want[k]), rtol=1e-6, atol=1e-9), k

    config = {"test_params": {"checkpoint_path": str(tmp_path)}}
    s_want = loops.test.test_sample(config, loaders, enc, pcd, model, crit, "cuda")
    assert len(loops.plot_calls) == 1 and loops.plot_calls[0]["save_dir"] == str(tmp_path)   # utils/test.py:162-171
    s_got = dropin_loops.test_sample(config, loaders, enc, pcd, model, crit, "cuda")
    assert set(s_got) == set(s_want)
    assert torch.equal(s_got["predictions"], s_want["predictions"]) and torch.equal(s_got["labels"], s_want["labels"])
    assert np.array_equal(s_got["points"], s_want["points"])
    for k in ("loss", "overall_acc", "mean_iou", "mean_per_class_acc", "precision", "recall", "f1"):
        assert np.allclose(_host(s_got[k]), _host(s_want[k]), rtol=1e-6, atol=1e-9), k
    assert "EVALUATION METRICS" in capsys.readouterr().out
