"""Pins the numpy oracle (oracle/np_oracle.py) against outputs of the reference itself
(tests/golden/*.npz, produced by tests/golden/make_golden.py from /root/reference).  CPU only."""
import hashlib
import os

import numpy as np
import pytest
import torch

from oracle import np_oracle as O
from tdssf import synth

FUSION_CASES = ["fusion_dinov3", "fusion_dinov2", "fusion_hires", "fusion_fp16tok"]


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"), allow_pickle=False)


def _checksum(sw):
    h = hashlib.sha256()
    for a in (sw.points, sw.point_feat, sw.tokens.float(), torch.stack(sw.K), torch.stack(sw.Rt), sw.labels):
        h.update(np.ascontiguousarray(a.numpy()).tobytes())
    return h.hexdigest()


def regen(g):
    dt = getattr(torch, str(g["tok_dtype"]))
    sw = synth.make_sweeps(int(g["B"]), int(g["V"]), str(g["encoder"]), seed=int(g["seed"]), tok_dtype=dt)
    assert _checksum(sw) == str(g["input_checksum"]), "synthetic generator drifted from the golden inputs"
    return sw


@pytest.mark.parametrize("name", FUSION_CASES)
def test_projection_bit_exact(golden_dir, name):
    g = _load(golden_dir, name)
    sw = regen(g)
    pix, depth, valid = O.project(sw.points[:, :, :3].numpy(), torch.stack(sw.K).numpy(), torch.stack(sw.Rt).numpy(),
                                  sw.image_sizes.numpy())
    assert np.array_equal(valid, g["valid"])
    assert np.array_equal(pix.view(np.uint32), g["pixel_coords"].view(np.uint32)), "pixel coords not bit-exact"
    assert np.array_equal(depth.view(np.uint32), g["depth"].view(np.uint32))


@pytest.mark.parametrize("name", FUSION_CASES)
def test_patch_index_bit_exact(golden_dir, name):
    g = _load(golden_dir, name)
    sw = regen(g)
    enc = sw.encoder
    scaled = O.scale_pixel_coords(g["pixel_coords"], sw.image_sizes.numpy().astype(np.float32),
                                  (enc.resize_height, enc.resize_width))
    flat = O.patch_index(scaled, enc.patch_size, enc.resize_height, enc.resize_width)
    assert np.array_equal(flat, g["flat_idx"].astype(np.int64))


@pytest.mark.parametrize("name", FUSION_CASES)
def test_gather_mean_and_logits(golden_dir, name):
    g = _load(golden_dir, name)
    sw = regen(g)
    enc = sw.encoder
    w = {k: v.numpy() for k, v in synth.make_mlp_weights(int(g["seed"])).items()}
    tok = sw.tokens.numpy() if sw.tokens.dtype != torch.bfloat16 else sw.tokens.float().numpy()
    logits, img, flat, valid = O.fusion_forward(tok, sw.point_feat.numpy(), sw.points.numpy(), sw.image_sizes.numpy(),
                                                torch.stack(sw.K).numpy(), torch.stack(sw.Rt).numpy(), w,
                                                enc.patch_size, enc.resize_height, enc.resize_width)
    stride = int(g["stride"])
    ref_img = g["img_feat_sub"]
    got = img[:, ::stride].astype(np.float32)
    if sw.tokens.dtype == torch.float32:
        # camera-order fp32 sum then one division: exact up to the sign of zero
        assert np.array_equal(got, ref_img)
    else:
        np.testing.assert_allclose(got, ref_img, rtol=2e-3, atol=2e-3)
    tol = 1e-5 if sw.tokens.dtype == torch.float32 else 3e-3
    ref = g["voxel_scores"]
    err = np.abs(logits - ref).max() / np.abs(ref).max()
    assert err < tol, f"logits rel err {err}"


def test_kat_projection(golden_dir):
    g = _load(golden_dir, "kat_projection")
    pix, depth, valid = O.project(g["points"], g["K"], g["Rt"], g["image_sizes"])
    assert np.array_equal(valid, g["valid"])
    assert np.array_equal(pix, g["pixel_coords"])
    assert np.array_equal(depth, g["depth"])
    # hand-checked facts: u == W and v == H are outside, u == 0 is inside, z <= 0 / NaN are invalid
    v0 = valid[0, 0]
    assert v0[0] and v0[1] and not v0[2] and not v0[3] and not v0[4]
    assert not v0[8] and not v0[9] and not v0[12]


def test_projector_shape_errors():
    xyz = np.zeros((2, 5, 3), np.float32)
    K = np.zeros((2, 6, 3, 3), np.float32)
    Rt = np.zeros((2, 6, 4, 4), np.float32)
    with pytest.raises(ValueError):
        O.project(xyz, K, Rt[:, :, :3], np.array([[900, 1600]] * 2))
    with pytest.raises(ValueError):
        O.project(xyz, K, Rt, np.array([[900, 1600]]))
    with pytest.raises(ValueError):
        O.scale_pixel_coords(np.zeros((2, 6, 5, 2), np.float32), np.zeros((2, 3)), 224)


def test_mlp_backward(golden_dir):
    g = _load(golden_dir, "mlp_backward")
    w = {k: v.numpy() for k, v in synth.make_mlp_weights(int(g["seed"])).items()}
    logits = O.mlp_forward(g["x_pt"], g["x_img"], w)
    assert np.abs(logits - g["logits"]).max() / np.abs(g["logits"]).max() < 1e-5
    grads, dx = O.mlp_backward(g["x_pt"], g["x_img"], w, g["G"])
    for k, v in grads.items():
        ref = g["grad_" + k.replace("mlp.", "")]
        assert np.abs(v - ref).max() / np.abs(ref).max() < 1e-4, k
    assert np.abs(dx - g["dx_pt"]).max() / np.abs(g["dx_pt"]).max() < 1e-4


@pytest.mark.parametrize("name", ["cels_equal", "cels_longer_mask", "cels_shorter_mask", "cels_nomask"])
def test_cels_loss(golden_dir, name):
    g = _load(golden_dir, name)
    mask = g["mask"] if bool(g["has_mask"]) else None
    total, ce, lov, preds, gt, grad = O.cels_loss(g["logits"], g["labels"], mask, g["class_weights"], 0,
                                                  float(g["lam"]), return_grad=True)
    assert abs(ce - g["ce"]) <= 1e-5 * abs(g["ce"])
    assert abs(lov - g["lovasz"]) <= 1e-5 * abs(g["lovasz"])
    assert abs(total - g["total"]) <= 1e-5 * abs(g["total"])
    assert np.array_equal(gt, g["gt"])
    assert np.array_equal(preds, g["preds"])
    ref = g["dlogits"]
    grad = grad[:, :ref.shape[1]] if grad.shape[1] >= ref.shape[1] else np.pad(
        grad, ((0, 0), (0, ref.shape[1] - grad.shape[1]), (0, 0)))
    # the reference's fp32 Jaccard first differences (jac[k]-jac[k-1]) cancel ~3 digits, and a 1-ulp
    # difference in exp() can swap neighbours in the sort: 3e-5 of the largest gradient covers both
    assert np.abs(grad - ref).max() <= 3e-5 * np.abs(ref).max()


def test_lovasz_small(golden_dir):
    g = _load(golden_dir, "lovasz_small")
    assert np.array_equal(O.lovasz_grad(np.array([1, 0, 1, 0], np.float32)), g["grad_1010"])
    assert np.array_equal(O.lovasz_grad(np.array([1], np.float32)), g["grad_single"])
    # rational known answers: G=2 -> jaccard = [1/2, 2/3, 1, 1] -> differences [1/2, 1/6, 1/3, 0]
    np.testing.assert_allclose(g["grad_1010"], [0.5, 1 / 6, 1 / 3, 0.0], rtol=1e-6, atol=1e-7)
    loss = O.lovasz_softmax_flat(g["probas"], g["labels"])
    assert abs(loss - g["loss"]) < 1e-6
    empty = O.lovasz_softmax_flat(np.zeros((0, 3), np.float32), np.zeros(0, np.int64))
    assert tuple(empty.shape) == tuple(g["empty_shape"])


def test_confusion(golden_dir):
    g = _load(golden_dir, "confusion")
    cm = O.confusion_matrix(g["preds"], g["labels"], 16)
    assert np.array_equal(cm, g["cm"])
    iou, miou = O.compute_iou(cm)
    acc, macc = O.per_class_accuracy(cm)
    pr, rc, f1 = O.precision_recall_f1(cm)
    assert np.array_equal(iou, g["iou"]) and np.array_equal(acc, g["acc"])
    for a, b in [(miou, g["miou"]), (macc, g["macc"]), (O.overall_accuracy(cm), g["oa"]), (pr, g["precision"]),
                 (rc, g["recall"]), (f1, g["f1"])]:
        assert abs(float(a) - float(b)) < 1e-6
    # a permutation is the identity confusion matrix
    perm = np.arange(1, 17)
    assert np.array_equal(O.confusion_matrix(perm, perm, 16), np.eye(16, dtype=np.int64))


def test_devox(golden_dir):
    g = _load(golden_dir, "devox")
    out, idx = O.devoxelize(g["scores"], g["pts"], g["vox"])
    ref_idx = g["idx"]
    differ = idx != ref_idx
    # H7: the reference's fp32 cdist expansion may flip near-ties; everywhere else indices are equal
    for b in range(idx.shape[0]):
        best, second = O.devox_gap(g["pts"][b][:, :3], g["vox"][b][:, :3])
        ambiguous = (second - best) < 1e-3
        assert not (differ[b] & ~ambiguous).any()
        d_ref = np.linalg.norm(g["pts"][b][:, :3].astype(np.float64) - g["vox"][b][ref_idx[b], :3], axis=1)
        assert np.all(d_ref - best < 1e-3)
    same = ~differ
    assert np.array_equal(out[same], g["out"][same])
    assert differ.mean() < 0.01


@pytest.mark.parametrize("name", ["fusion_dinov3", "fusion_dinov2"])
def test_torch_port_matches_reference(golden_dir, name):
    """The timed CPU baseline (oracle/torch_port.py) reproduces the reference's outputs."""
    from oracle import torch_port as T
    g = _load(golden_dir, name)
    sw = regen(g)
    enc = sw.encoder
    w = synth.make_mlp_weights(int(g["seed"]))
    with torch.no_grad():
        out, img, flat, valid, _ = T.fusion_forward(sw.tokens, sw.point_feat, sw.points, sw.image_sizes, sw.K, sw.Rt, w,
                                                    enc.patch_size, enc.resize_height, enc.resize_width,
                                                    return_parts=True)
    assert np.array_equal(valid.numpy(), g["valid"])
    assert np.array_equal(flat.numpy(), g["flat_idx"])
    assert np.array_equal(img[:, ::int(g["stride"])].numpy(), g["img_feat_sub"])
    assert np.abs(out.numpy() - g["voxel_scores"]).max() <= 1e-5 * np.abs(g["voxel_scores"]).max()
