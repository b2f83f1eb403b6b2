"""GPU parity tests (run on the B200 box with `-m gpu`): every CUDA kernel is called through the C ABI
(ctypes -> libtdssf.so) and compared with the CPU oracle on the same seeded inputs and with the golden
vectors produced by the reference itself (tests/golden/).  Nothing here reads /root/reference.

Tolerances (north_star): bit-exact for pixel coordinates, masks, patch indices, camera records, confusion
matrices and sort-dependent integer results; 1e-5 relative for fp32 features / logits / loss."""
import hashlib
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from oracle import np_oracle as O  # noqa: E402
from tdssf import synth  # noqa: E402

FUSION_CASES = ["fusion_dinov3", "fusion_dinov2", "fusion_hires", "fusion_fp16tok"]


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from tdssf import ops as _ops, _lib
    _lib.check(_lib.lib.tdssf_check_device(), "device")
    return _ops


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"), allow_pickle=False)


def _regen(g):
    dt = getattr(torch, str(g["tok_dtype"]))
    sw = synth.make_sweeps(int(g["B"]), int(g["V"]), str(g["encoder"]), seed=int(g["seed"]), tok_dtype=dt)
    h = hashlib.sha256()
    for a in (sw.points, sw.point_feat, sw.tokens.float(), torch.stack(sw.K), torch.stack(sw.Rt), sw.labels):
        h.update(np.ascontiguousarray(a.numpy()).tobytes())
    assert h.hexdigest() == str(g["input_checksum"]), "synthetic generator drifted from the golden inputs"
    return sw


def _dev_inputs(sw):
    d = "cuda"
    return dict(points=sw.points.to(d), K=torch.stack(sw.K).to(d), Rt=torch.stack(sw.Rt).to(d),
                hw=sw.image_sizes.float().to(d), tokens=sw.tokens.to(d), feat=sw.point_feat.to(d))


def _rec_fields(rec):
    """(B,V,4) int32 records -> patch (B,V,6), mask (B,V), count (B,V)."""
    r = rec.cpu().numpy().view(np.uint16).reshape(rec.shape[0], rec.shape[1], 8)
    return r[..., :6].astype(np.int64), (r[..., 6] & 0xff).astype(np.int64), (r[..., 6] >> 8).astype(np.int64)


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


# ---------------------------------------------------------------------------------------------- a1-a3
@pytest.mark.parametrize("name", FUSION_CASES)
def test_project_bit_exact(ops, golden_dir, name):
    g = _load(golden_dir, name)
    sw = _regen(g)
    d = _dev_inputs(sw)
    out = ops.project(d["points"][:, :, :3], d["K"], d["Rt"], d["hw"], sw.encoder, want_flat=True, want_rec=True)
    torch.cuda.synchronize()
    valid = out["valid"].cpu().numpy().astype(bool)
    assert np.array_equal(valid, g["valid"])
    assert np.array_equal(_bits(out["pix"].cpu().numpy()), _bits(g["pixel_coords"])), "pixel coords not bit-exact"
    assert np.array_equal(_bits(out["depth"].cpu().numpy()), _bits(g["depth"]))
    assert np.array_equal(out["flat_idx"].cpu().numpy(), g["flat_idx"])
    patch, mask, count = _rec_fields(out["rec"])
    assert np.array_equal(patch, np.transpose(g["flat_idx"], (0, 2, 1)))
    bits = (g["valid"].astype(np.int64) << np.arange(6)[None, :, None]).sum(1)
    assert np.array_equal(mask, bits) and np.array_equal(count, g["valid"].sum(1))
    # the oracle restatement agrees too (it is what full-size runs are checked against)
    pix_o, depth_o, valid_o = O.project(sw.points[:, :, :3].numpy(), torch.stack(sw.K).numpy(),
                                        torch.stack(sw.Rt).numpy(), sw.image_sizes.numpy())
    assert np.array_equal(valid, valid_o) and np.array_equal(_bits(out["pix"].cpu().numpy()), _bits(pix_o))


def test_project_kat_and_strided_view(ops, golden_dir):
    g = _load(golden_dir, "kat_projection")
    pts = torch.from_numpy(g["points"]).cuda()
    out = ops.project(pts, torch.from_numpy(g["K"]).cuda(), torch.from_numpy(g["Rt"]).cuda(),
                      torch.from_numpy(g["image_sizes"]).float().cuda())
    assert np.array_equal(out["valid"].cpu().numpy().astype(bool), g["valid"])
    assert np.array_equal(out["pix"].cpu().numpy(), g["pixel_coords"])
    assert np.array_equal(out["depth"].cpu().numpy(), g["depth"])


def test_projector_dropin_signature_and_errors(ops):
    from utils.fusion_model import multi_camera_projector, scale_pixel_coords
    sw = synth.make_sweeps(2, 1000, "dinov3", seed=7)
    raw = sw.points.cuda()
    # lists of CPU tensors (train/test loops) and numpy arrays are both accepted (utils/fusion_model.py:36-46)
    pix, depth, valid = multi_camera_projector(raw[:, :, :3], sw.K, [r.numpy() for r in sw.Rt], sw.image_sizes.cuda())
    assert pix.shape == (2, 6, 1000, 2) and depth.shape == (2, 6, 1000) and valid.dtype == torch.bool
    pix_o, depth_o, valid_o = O.project(sw.points[:, :, :3].numpy(), torch.stack(sw.K).numpy(),
                                        torch.stack(sw.Rt).numpy(), sw.image_sizes.numpy())
    assert np.array_equal(valid.cpu().numpy(), valid_o) and np.array_equal(_bits(pix.cpu().numpy()), _bits(pix_o))
    scaled = scale_pixel_coords(pix, sw.image_sizes.cuda(), (640, 1138))
    ref = O.scale_pixel_coords(pix_o, sw.image_sizes.numpy().astype(np.float32), (640, 1138))
    assert np.array_equal(_bits(scaled.cpu().numpy()), _bits(ref))
    sq = scale_pixel_coords(pix, torch.tensor([900, 1600]), 224)  # 1-D sizes broadcast, int target
    assert np.array_equal(_bits(sq.cpu().numpy()), _bits(O.scale_pixel_coords(pix_o, np.array([900, 1600]), 224)))
    with pytest.raises(ValueError):
        multi_camera_projector(raw[:, :, :3], sw.K, [r[:, :3] for r in sw.Rt], sw.image_sizes)
    with pytest.raises(ValueError):
        multi_camera_projector(raw[:, :, :3], sw.K, sw.Rt, sw.image_sizes[:1])
    with pytest.raises(ValueError):
        scale_pixel_coords(pix, torch.zeros(2, 3), 224)
    with pytest.raises(RuntimeError):
        multi_camera_projector(sw.points[:, :, :3], sw.K, sw.Rt, sw.image_sizes)  # CPU points: no fallback


# ---------------------------------------------------------------------------------------------- a4-a5
@pytest.mark.parametrize("name", FUSION_CASES)
def test_gather_mean(ops, golden_dir, name):
    g = _load(golden_dir, name)
    sw = _regen(g)
    d = _dev_inputs(sw)
    rec = ops.project(d["points"], d["K"], d["Rt"], d["hw"], sw.encoder, want_pix=False, want_depth=False,
                      want_valid=False, want_rec=True)["rec"]
    feat = ops.gather_mean(d["tokens"], rec).cpu().numpy()
    stride = int(g["stride"])
    if sw.tokens.dtype == torch.float32:
        assert np.array_equal(feat[:, ::stride], g["img_feat_sub"])  # camera-order sum + one divide: exact
    else:
        np.testing.assert_allclose(feat[:, ::stride], g["img_feat_sub"], rtol=2e-3, atol=2e-3)
    # points seen by no camera -> zeros; a point seen by two cameras -> mean of the two tokens
    hits = g["valid"].sum(1)
    assert np.all(feat[hits == 0] == 0)
    b, v = np.argwhere(hits == 2)[0]
    cams = np.nonzero(g["valid"][b, :, v])[0]
    tok = sw.tokens.float().numpy()
    want = (tok[b, cams[0], g["flat_idx"][b, cams[0], v]] + tok[b, cams[1], g["flat_idx"][b, cams[1], v]]) / 2
    np.testing.assert_allclose(feat[b, v], want, rtol=2e-3 if sw.tokens.dtype != torch.float32 else 0, atol=1e-3 if sw.tokens.dtype != torch.float32 else 0)


def test_gather_bf16_tokens(ops):
    sw = synth.make_sweeps(1, 3000, "dinov2", seed=9, tok_dtype=torch.bfloat16)
    d = _dev_inputs(sw)
    out = ops.project(d["points"], d["K"], d["Rt"], d["hw"], sw.encoder, want_flat=True, want_rec=True)
    feat = ops.gather_mean(d["tokens"], out["rec"]).cpu().numpy()
    ref = O.gather_mean(sw.tokens.float().numpy(), out["flat_idx"].cpu().numpy(), out["valid"].cpu().numpy().astype(bool))
    np.testing.assert_allclose(feat, ref, rtol=1e-2, atol=1e-2)  # double rounding to bf16 inside the mean


# ---------------------------------------------------------------------------------------------- a6
@pytest.mark.parametrize("name", FUSION_CASES)
def test_forward_fp32_dropin(ops, golden_dir, name, monkeypatch):
    """FeatureFusionModel.forward through the drop-in module, fp32 kernels, vs the reference's voxel_scores."""
    from utils.fusion_model import FeatureFusionModel
    monkeypatch.setenv("TDSSF_MLP", "fp32")
    g = _load(golden_dir, name)
    sw = _regen(g)

    class Pcd(torch.nn.Module):
        def devoxelize(self, s):
            return s, None

    model = FeatureFusionModel(sw.encoder, Pcd(), 64, 384, 16, device="cuda").cuda()
    model.load_state_dict(synth.make_mlp_weights(int(g["seed"])), strict=True)
    with torch.no_grad():
        # the reference loops hand CPU patch tokens and CPU calibration lists
        out = model(sw.tokens, sw.point_feat.cuda(), sw.points.cuda(), None, sw.image_sizes.cuda(), sw.K, sw.Rt)
    ref = g["voxel_scores"]
    err = np.abs(out.cpu().numpy() - ref).max() / np.abs(ref).max()
    assert err < (1e-5 if sw.tokens.dtype == torch.float32 else 3e-3), err


def test_mlp_backward_fp32(ops, golden_dir):
    g = _load(golden_dir, "mlp_backward")
    w = {k: v.cuda().requires_grad_(True) for k, v in synth.make_mlp_weights(int(g["seed"])).items()}
    x_pt = torch.from_numpy(g["x_pt"]).cuda().requires_grad_(True)
    x_img = torch.from_numpy(g["x_img"]).cuda()
    out = ops.mlp_f32(x_pt, x_img, *[w[k] for k in ("mlp.0.weight", "mlp.0.bias", "mlp.2.weight", "mlp.2.bias",
                                                    "mlp.4.weight", "mlp.4.bias")])
    assert np.abs(out.detach().cpu().numpy() - g["logits"]).max() / np.abs(g["logits"]).max() < 1e-5
    (out * torch.from_numpy(g["G"]).cuda()).sum().backward()
    for k, p in w.items():
        ref = g["grad_" + k.replace("mlp.", "")]
        err = np.abs(p.grad.cpu().numpy() - ref).max() / np.abs(ref).max()
        assert err < 1e-4, (k, err)
    err = np.abs(x_pt.grad.cpu().numpy() - g["dx_pt"]).max() / np.abs(g["dx_pt"]).max()
    assert err < 1e-4, err


# ---------------------------------------------------------------------------------------------- a7
def test_devoxelize(ops, golden_dir):
    g = _load(golden_dir, "devox")
    pts, vox = torch.from_numpy(g["pts"]).cuda(), torch.from_numpy(g["vox"]).cuda()
    idx = ops.devox_nn(pts, vox).cpu().numpy()
    exact = np.stack([O.devox_nn(g["pts"][b][:, :3], g["vox"][b][:, :3]) for b in range(pts.shape[0])])
    for b in range(idx.shape[0]):
        best, second = O.devox_gap(g["pts"][b][:, :3], g["vox"][b][:, :3])
        ambiguous = (second - best) < 1e-4
        assert not ((idx[b] != exact[b]) & ~ambiguous).any()
        assert not ((idx[b] != g["idx"][b]) & ~((second - best) < 1e-3)).any()  # vs the reference's cdist (H7)
    scores = torch.from_numpy(g["scores"]).cuda().requires_grad_(True)
    out = ops.devox_gather(scores, torch.from_numpy(exact).cuda())
    assert np.array_equal(out.detach().cpu().numpy(), np.take_along_axis(g["scores"], exact[..., None], axis=1))
    gsum = torch.randn_like(out)
    (out * gsum).sum().backward()
    ref = np.zeros_like(g["scores"])
    for b in range(exact.shape[0]):
        np.add.at(ref[b], exact[b], gsum[b].cpu().numpy())
    np.testing.assert_allclose(scores.grad.cpu().numpy(), ref, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("case", ["voxelised", "sparse", "padded", "degenerate"])
@pytest.mark.parametrize("cell", [0.1, 0.5, 4.0])
def test_devox_grid_equals_brute_force(ops, case, cell):
    """The grid-accelerated nearest-voxel search returns the brute-force indices bit for bit, for any cell size:
    voxel centres of a real voxelisation, a sparse cloud (most points fall back to the full scan), zero-padded voxel
    rows, and duplicates / non-finite coordinates / a single voxel."""
    gen = torch.Generator().manual_seed(11)
    B, P = 2, 6000
    pts = synth.make_points(B, P, gen)[..., :3].contiguous()
    if case == "voxelised":
        vs = 0.1
        vox = torch.stack([torch.unique(torch.floor(pts[b] / vs), dim=0)[:4000] * vs + vs / 2 for b in range(B)])
    elif case == "sparse":
        vox = synth.make_points(B, 700, gen)[..., :3].contiguous()
    elif case == "padded":
        vox = torch.zeros(B, 5000, 3)
        vox[:, :3000] = pts[:, :3000] + 0.03
        pts[0, :50] *= 0.01           # points near the origin, where the padding rows sit
    else:
        vox = pts[:, :1].clone() if cell == 0.1 else pts[:, :2000].repeat(1, 2, 1)   # one voxel / every voxel twice
        pts[0, 5] = float("nan")
        pts[1, 7, 0] = float("inf")
        vox = vox.clone()
        if vox.shape[1] > 10:
            vox[0, 3, 1] = float("nan")
    pts_d, vox_d = pts.cuda(), vox.cuda()
    brute = ops.devox_nn(pts_d, vox_d, cell=None)
    grid = ops.devox_nn(pts_d, vox_d, cell=cell)
    torch.cuda.synchronize()
    assert torch.equal(grid, brute)
    # and the brute force is the exact nearest neighbour wherever it is unambiguous
    b = 0
    finite = torch.isfinite(pts[b]).all(-1).numpy()
    exact = O.devox_nn(pts[b].numpy()[finite], np.nan_to_num(vox[b].numpy(), nan=1e30))
    best, second = O.devox_gap(pts[b].numpy()[finite], np.nan_to_num(vox[b].numpy(), nan=1e30))
    clear = (second - best) > 1e-4
    assert np.array_equal(brute[b].cpu().numpy()[finite][clear], exact[clear])


def test_lidar_encoder_contract(ops):
    """forward() pads/caches like utils/lidar.py:27-81 and devoxelize() reads that cache (zero padding searched)."""
    from utils.lidar import LiDARFeatureEncoder

    class Backbone(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.lin = torch.nn.Linear(4, 64)

        def forward(self, d):
            return self.lin(d["feat"])

    def voxelizer(cloud, size):  # deterministic stand-in for Open3D: one voxel per occupied cell, first point wins
        grid = torch.floor(cloud[:, :3] / size).int()
        _, inv = torch.unique(grid, dim=0, return_inverse=True)
        first = torch.full((int(inv.max()) + 1,), cloud.shape[0], device=cloud.device, dtype=torch.long)
        first = first.scatter_reduce(0, inv, torch.arange(cloud.shape[0], device=cloud.device), reduce="amin")
        return cloud[first], grid[first]

    gen = torch.Generator().manual_seed(3)
    pts = synth.make_points(2, 1500, gen).cuda()
    pts[1, 700:] = pts[1, :800].clone()  # second sweep has fewer occupied voxels -> padding
    enc = LiDARFeatureEncoder({"dataset_params": {"lidar": {"voxel_size": 2.0}}}, backbone=Backbone(), voxelizer=voxelizer).cuda()
    feats, raw, coords, mask = enc(pts)
    assert feats.shape[:2] == raw.shape[:2] == coords.shape[:2] == mask.shape and raw.shape[2] == 4
    assert mask[0].sum() != mask[1].sum() and float(raw[~mask].abs().sum()) == 0.0
    scores = torch.randn(2, raw.shape[1], 16, device="cuda")
    out, idx = enc.devoxelize(scores)
    assert out.shape == (2, 1500, 16) and idx.dtype == torch.int64
    ref, ref_idx = O.devoxelize(scores.cpu().numpy(), pts.cpu().numpy(), raw.cpu().numpy())
    assert (idx.cpu().numpy() == ref_idx).mean() > 0.999
    same = idx.cpu().numpy() == ref_idx
    assert np.array_equal(out.cpu().numpy()[same], ref[same])


# ---------------------------------------------------------------------------------------------- a8-a10
@pytest.mark.parametrize("name", ["cels_equal", "cels_longer_mask", "cels_shorter_mask", "cels_nomask"])
def test_cels_loss_dropin(ops, golden_dir, name, capsys):
    from utils.losses import CELSLoss
    g = _load(golden_dir, name)
    crit = CELSLoss(weight=torch.from_numpy(g["class_weights"]).cuda(), ignore_index=0, lamda_lovasz=float(g["lam"]))
    logits = torch.from_numpy(g["logits"]).cuda().requires_grad_(True)
    mask = torch.from_numpy(g["mask"]).cuda() if bool(g["has_mask"]) else None
    total, ce, lov, preds, gt = crit(logits, torch.from_numpy(g["labels"]).cuda(), mask)
    total.backward()
    for got, key in ((total, "total"), (ce, "ce"), (lov, "lovasz")):
        assert abs(got.item() - float(g[key])) <= 1e-5 * abs(float(g[key])), (key, got.item(), float(g[key]))
    assert np.array_equal(gt.cpu().numpy(), g["gt"])
    assert np.array_equal(preds.cpu().numpy(), g["preds"])
    ref = g["dlogits"]
    err = np.abs(logits.grad.cpu().numpy() - ref).max() / np.abs(ref).max()
    assert err <= 3e-5, err
    if "mask" in name and name != "cels_nomask":
        assert ("Warning" in capsys.readouterr().out) == (name != "cels_equal")


def test_cels_vs_oracle_and_sort_properties(ops):
    """Larger seeded case against the numpy oracle + properties of the Lovasz sort: the loss is invariant to
    row order, heavy ties (few distinct logits) are handled, absent classes are skipped."""
    gen = torch.Generator().manual_seed(31)
    n = 60_000
    logits = torch.randn(1, n, 16, generator=gen) * 1.5
    labels = synth.make_labels(1, n, gen)
    labels[labels == 2] = 11  # class 2 absent -> skipped by classes='present'
    mask = torch.rand(1, n, generator=gen) > 0.05
    w = torch.tensor(synth.CLASS_WEIGHTS)
    total, ce, lov, preds, gt = ops.cels(logits.cuda(), labels.cuda(), mask.cuda(), w.cuda(), 1.0, 0)
    t_o, ce_o, lov_o, preds_o, gt_o = O.cels_loss(logits.numpy(), labels.numpy(), mask.numpy(), w.numpy(), 0, 1.0)
    assert abs(ce.item() - ce_o) < 1e-5 * abs(ce_o) and abs(lov.item() - lov_o) < 1e-5 * abs(lov_o)
    assert np.array_equal(gt.cpu().numpy(), gt_o)
    assert (preds.cpu().numpy() != preds_o).mean() < 1e-4  # expf differs from numpy's exp by <= 1 ulp
    perm = torch.randperm(n, generator=gen)
    total_p = ops.cels(logits[:, perm].cuda(), labels[:, perm].cuda(), mask[:, perm].cuda(), w.cuda(), 1.0, 0)[0]
    assert abs(total_p.item() - total.item()) < 1e-5 * abs(total.item())
    # ties: logits quantised to 3 values -> thousands of equal keys per class; stable order = ascending index
    q = torch.randint(0, 3, (1, 5000, 16), generator=gen).float()
    lab = synth.make_labels(1, 5000, gen)
    x = q.cuda().requires_grad_(True)
    out = ops.cels(x, lab.cuda(), None, w.cuda(), 1.0, 0)
    out[0].backward()
    t_o, _, lov_o, _, _, grad_o = O.cels_loss(q.numpy(), lab.numpy(), None, w.numpy(), 0, 1.0, return_grad=True)
    assert abs(out[2].item() - lov_o) < 1e-5 * abs(lov_o)
    assert np.abs(x.grad.cpu().numpy() - grad_o).max() <= 3e-5 * np.abs(grad_o).max()


def test_cels_scan_paths_agree(ops, monkeypatch):
    """The radix sort derives its scatter offsets inside the scatter kernel up to 256 tiles per class and with a
    separate scan launch beyond; both paths must give the same bits (loss, class ids, gradient)."""
    gen = torch.Generator().manual_seed(33)
    n = 30_000
    logits = (torch.randn(2, n, 16, generator=gen) * 2).cuda()
    labels = synth.make_labels(2, n, gen).cuda()
    w = torch.tensor(synth.CLASS_WEIGHTS).cuda()

    def run():
        x = logits.clone().requires_grad_(True)
        out = ops.cels(x, labels, None, w, 1.0, 0)
        out[0].backward()
        return [t.detach().clone() for t in out] + [x.grad.clone()]

    a = run()
    monkeypatch.setenv("TDSSF_CELS_SCAN", "separate")
    b = run()
    for x, y in zip(a, b):
        assert torch.equal(x, y)


@pytest.mark.parametrize("scale", [1.0, 6.0])
def test_cels_key_selection_is_exact(ops, monkeypatch, scale):
    """Only the foreground keys and the background keys not below the smallest foreground error are sorted (everything
    behind the last foreground element has Jaccard gradient exactly 0); sorting all 16 N keys like the reference must give
    the same bits - loss, class ids and gradient - for flat (scale 1) and confident (scale 6) predictions, with ties."""
    gen = torch.Generator().manual_seed(35)
    n = 40_000
    logits = torch.randn(2, n, 16, generator=gen) * scale
    labels = synth.make_labels(2, n, gen)
    logits[0, :n // 2] += 2.0 * scale * torch.nn.functional.one_hot(labels[0, :n // 2].clamp(min=1) - 1, 16)  # partly trained
    logits[1, ::7] = torch.round(logits[1, ::7])                                                            # ties
    logits, labels, w = logits.cuda(), labels.cuda(), torch.tensor(synth.CLASS_WEIGHTS).cuda()

    def run():
        x = logits.clone().requires_grad_(True)
        out = ops.cels(x, labels, None, w, 1.0, 0)
        out[0].backward()
        return [t.detach().clone() for t in out] + [x.grad.clone()]

    a = run()
    monkeypatch.setenv("TDSSF_CELS_SELECT", "all")
    b = run()
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    t_o, _, lov_o, _, _, grad_o = O.cels_loss(logits.cpu().numpy(), labels.cpu().numpy(), None, w.cpu().numpy(), 0, 1.0, return_grad=True)
    assert abs(a[2].item() - lov_o) < 1e-5 * abs(lov_o)
    # fp32 softmax backward against the float64 oracle (the bits of the two device paths are compared above)
    assert np.abs(a[5].cpu().numpy() - grad_o).max() <= 1e-3 * np.abs(grad_o).max()


def test_cels_empty_selection(ops):
    from utils.losses import CELSLoss
    crit = CELSLoss(weight=torch.ones(16).cuda())
    logits = torch.randn(1, 10, 16).cuda()
    total, ce, lov, preds, gt = crit(logits, torch.zeros(1, 10, dtype=torch.long).cuda(), None)  # all ignored
    assert preds.numel() == 0 and gt.numel() == 0 and tuple(lov.shape) == (0, 16) and torch.isnan(ce)


# ---------------------------------------------------------------------------------------------- a11
def test_confusion_golden_and_accumulation(ops, golden_dir):
    from utils import evaluation as E
    g = _load(golden_dir, "confusion")
    preds, labels = torch.from_numpy(g["preds"]), torch.from_numpy(g["labels"])
    cm = E.compute_confusion_matrix(preds, labels, 16)  # CPU tensors in, like the reference's loops
    assert cm.device.type == "cpu" and cm.dtype == torch.int64 and np.array_equal(cm.numpy(), g["cm"])
    iou, miou = E.compute_iou(cm)
    assert np.array_equal(iou.numpy(), g["iou"]) and abs(miou.item() - float(g["miou"])) < 1e-7
    pr, rc, f1 = E.precision_recall_f1(cm)
    assert abs(pr.item() - float(g["precision"])) < 1e-7 and abs(f1.item() - float(g["f1"])) < 1e-6
    assert abs(E.overall_accuracy(cm).item() - float(g["oa"])) < 1e-7
    # device-resident accumulation over shards == one shot (integer sum, order free)
    acc = torch.zeros(16, 16, dtype=torch.int64, device="cuda")
    for part in range(4):
        E.compute_confusion_matrix(preds[part::4].cuda(), labels[part::4].cuda(), 16, out=acc)
    assert np.array_equal(acc.cpu().numpy(), g["cm"])
    with pytest.raises(RuntimeError):
        E.compute_confusion_matrix(torch.tensor([0, 3]), torch.tensor([1, 2]), 16)


# ---------------------------------------------------------------------------------------------- full size
def test_full_size_config2_properties(ops):
    """BASELINE config 2 shapes (8 sweeps x 34 720 points, 40x71x384 tokens): the oracle is too slow for the
    whole batch, so check one sweep against it and size-independent properties on the rest."""
    sw = synth.make_sweeps(8, 34720, "dinov3", seed=2002)
    d = _dev_inputs(sw)
    out = ops.project(d["points"], d["K"], d["Rt"], d["hw"], sw.encoder, want_flat=True, want_rec=True)
    pix_o, _, valid_o = O.project(sw.points[3:4, :, :3].numpy(), torch.stack(sw.K)[3:4].numpy(),
                                  torch.stack(sw.Rt)[3:4].numpy(), sw.image_sizes[3:4].numpy())
    assert np.array_equal(out["valid"][3:4].cpu().numpy().astype(bool), valid_o)
    assert np.array_equal(_bits(out["pix"][3:4].cpu().numpy()), _bits(pix_o))
    patch, mask, count = _rec_fields(out["rec"])
    assert patch.max() < 2840 and np.array_equal(count, np.array([bin(m).count("1") for m in range(64)])[mask])
    feat = ops.gather_mean(d["tokens"], out["rec"])
    # linearity: gather(2*tokens) == 2*gather(tokens) exactly (powers of two commute with fp32 rounding)
    feat2 = ops.gather_mean(d["tokens"] * 2, out["rec"])
    assert torch.equal(feat2, feat * 2)
    assert float(feat[torch.from_numpy(count == 0).cuda()].abs().max()) == 0.0


# ---------------------------------------------------------------------------------------------- f3
def test_evaluation_loop_dropin(ops, capsys):
    """utils.test.test_model (device-resident confusion matrix and counters) returns exactly what the reference's loop
    structure returns (per-batch predictions appended on the host, utils/test.py:64-87, then utils.evaluation.evaluate)."""
    from utils import test as eval_loop
    from utils.evaluation import evaluate
    from utils.fusion_model import FeatureFusionModel
    from utils.losses import CELSLoss
    import harness
    n_batches, B, V = 3, 2, 3000
    sweeps, batches = harness.make_batches(n_batches, B, V, seed=50)
    loaders = {"test": batches}
    enc, pcd = harness.TokenLookupEncoder(sweeps, device="cuda"), harness.IdentityVoxels().cuda()
    model = FeatureFusionModel(enc, pcd, device="cuda").cuda()
    model.load_state_dict({**synth.make_mlp_weights(50), "pcd_encoder.scale": torch.ones(64)}, strict=True)
    crit = CELSLoss(weight=torch.tensor(synth.CLASS_WEIGHTS).cuda())
    got = eval_loop.test_model(loaders, enc, pcd, model, crit, "cuda")
    assert "EVALUATION METRICS" in capsys.readouterr().out

    # the reference's loop structure, with the same modules
    model.eval()
    tot_loss, correct, points, preds_all, gt_all = 0.0, 0, 0, [], []
    with torch.no_grad():
        for images, sizes, pts, labels, mask, K, Rt in loaders["test"]:
            tokens = torch.stack([enc(images[:, v])["patch_features"] for v in range(6)], dim=1)
            feat, raw, _, _ = pcd(pts.cuda())
            out = model(tokens, feat, raw, None, sizes.cuda(), K, Rt)
            total, _, _, preds, gt = crit(out, labels.cuda(), mask=mask.cuda())
            tot_loss += total.item()
            correct += int((preds == gt).sum())
            points += gt.size(0)
            preds_all.append(preds.cpu())
            gt_all.append(gt.cpu())
    want = evaluate(torch.cat(preds_all), torch.cat(gt_all), 16, tot_loss, float(correct), points, n_batches)
    assert set(got) == set(want)
    assert got["overall_acc"] == want["overall_acc"]           # integer counters: exact
    for k in want:   # fp32 means of the same integer matrix, reduced on the device here and on the host there
        assert np.allclose(np.asarray(got[k], dtype=np.float64), np.asarray(want[k], dtype=np.float64), rtol=1e-6, atol=1e-9), k


def test_cels_counted_graph_replay_equals_synchronous_call(ops):
    """tdssf_cels_count + tdssf_cels_fwd_bwd_counted (no host synchronisation, grids sized for n_rows, CUDA-graph replay
    when the arguments repeat) gives bit-identical scalars, predictions, labels and gradients to tdssf_cels_fwd_bwd on the
    same buffers, for fresh data in every call (first call plain, second captured, later ones replayed), including a call
    in which no row survives."""
    import ctypes as C
    from tdssf import _lib
    from tdssf.ops import _ptr, _stream
    lib = _lib.lib
    n_rows = 9000
    gen = torch.Generator().manual_seed(77)
    x = torch.empty(n_rows, 16, device="cuda")
    lab = torch.empty(n_rows, dtype=torch.int64, device="cuda")
    mask = torch.empty(n_rows, dtype=torch.uint8, device="cuda")
    cw = torch.tensor(synth.CLASS_WEIGHTS, device="cuda")
    ws_bytes = lib.tdssf_cels_workspace_bytes(n_rows)
    bufs = lambda: dict(work=torch.empty(ws_bytes, dtype=torch.uint8, device="cuda"), scal=torch.empty(4, device="cuda"),
                        preds=torch.zeros(n_rows, dtype=torch.int64, device="cuda"), gt=torch.zeros(n_rows, dtype=torch.int64, device="cuda"),
                        dlog=torch.empty(n_rows, 16, device="cuda"))
    a, b = bufs(), bufs()
    n_dev = torch.zeros(1, dtype=torch.int32, device="cuda")
    for call in range(5):
        x.copy_(torch.randn(n_rows, 16, generator=gen) * 2)
        lab.copy_(synth.make_labels(1, n_rows, gen).view(-1) if call != 3 else torch.zeros(n_rows, dtype=torch.int64))
        mask.copy_((torch.rand(n_rows, generator=gen) > 0.1).to(torch.uint8))
        nv = C.c_int64(0)
        _lib.check(lib.tdssf_cels_fwd_bwd(_ptr(x), _ptr(lab), _ptr(mask), n_rows, _ptr(cw), 1.0, 0, _ptr(a["scal"]), C.byref(nv),
                                          _ptr(a["preds"]), _ptr(a["gt"]), _ptr(a["dlog"]), _ptr(a["work"]), ws_bytes, _stream(x)))
        _lib.check(lib.tdssf_cels_count(_ptr(lab), _ptr(mask), n_rows, 0, _ptr(n_dev), _ptr(b["work"]), ws_bytes, _stream(x)))
        _lib.check(lib.tdssf_cels_fwd_bwd_counted(_ptr(x), _ptr(lab), _ptr(mask), n_rows, _ptr(cw), 1.0, 0, _ptr(b["scal"]),
                                                  _ptr(b["preds"]), _ptr(b["gt"]), _ptr(b["dlog"]), _ptr(b["work"]), ws_bytes, _stream(x)))
        n = int(nv.value)
        assert int(n_dev.item()) == n and (n == 0) == (call == 3)
        assert torch.equal(a["scal"].view(torch.int32), b["scal"].view(torch.int32)), (call, a["scal"], b["scal"])
        assert torch.equal(a["preds"][:n], b["preds"][:n]) and torch.equal(a["gt"][:n], b["gt"][:n])
        assert torch.equal(a["dlog"].view(torch.int32), b["dlog"].view(torch.int32)), call
