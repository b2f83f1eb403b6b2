"""Generates the golden vectors under tests/golden/ by RUNNING THE REFERENCE ITSELF.

    python tests/golden/make_golden.py            # needs /root/reference (build container only)

The reference (CharlesChang012/3D-Semantic-Segmentation-Fusion @ e5103fd) is loaded unmodified, BY FILE PATH,
from /root/reference through oracle/ref_loader.py: utils/fusion_model.py (multi_camera_projector,
scale_pixel_coords, FeatureFusionModel), utils/losses.py (CELSLoss), LovaszSoftmax/pytorch/lovasz_losses.py,
utils/evaluation.py.  (A plain `import utils.fusion_model` would resolve to this repository's drop-in package of
the same name; the loader asserts every module's __file__ lies under the reference root.)
tests/test_golden_regen.py re-runs this script into a temporary directory and requires bit-identical files.
utils/lidar.py cannot be imported (open3d / spconv absent), so the devoxelize golden re-executes its
torch statements (utils/lidar.py:147-155) verbatim on CPU.

Large inputs (patch tokens) are not stored: they are regenerated from the seed by
`tdssf.synth.make_sweeps`, and a checksum of every regenerated input is stored so that drift of the
generator is detected instead of silently compared.  Environment of the committed files:
torch 2.11.0+cu128 CPU, 8 threads.
"""
import contextlib
import hashlib
import io
import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "3d-semantic-segmentation-fusion_b200")]
from tdssf import synth  # noqa: E402  (seeded input generators only: no kernel is involved)
from oracle import ref_loader  # noqa: E402

OUT = HERE   # --out redirects (tests/test_golden_regen.py)
REFM = None  # the reference's modules, set by main()


def checksum(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        a = a.detach().cpu().numpy() if torch.is_tensor(a) else np.asarray(a)
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def sweeps_checksum(sw):
    return checksum(sw.points, sw.point_feat, sw.tokens.float(), torch.stack(sw.K), torch.stack(sw.Rt),
                    sw.labels)


class StubPcd(torch.nn.Module):
    """devoxelize = identity: the golden captures voxel_scores (utils/fusion_model.py:204)."""

    def devoxelize(self, s):
        return s, None


def run_fusion_case(name, B, V, encoder, seed, tok_dtype=torch.float32, stride=16):
    FeatureFusionModel, multi_camera_projector, scale_pixel_coords = (
        REFM.fusion_model.FeatureFusionModel, REFM.fusion_model.multi_camera_projector, REFM.fusion_model.scale_pixel_coords)
    sw = synth.make_sweeps(B, V, encoder, seed=seed, tok_dtype=tok_dtype)
    enc = sw.encoder
    w = synth.make_mlp_weights(seed)
    model = FeatureFusionModel(enc, StubPcd(), 64, 384, 16, device="cpu")
    model.load_state_dict(w, strict=True)
    with torch.no_grad():
        pix, depth, valid = multi_camera_projector(sw.points[:, :, :3], sw.K, sw.Rt, sw.image_sizes)
        # lines 162-188 of utils/fusion_model.py, executed with the reference's own helper
        origin = torch.as_tensor(sw.image_sizes, dtype=pix.dtype)
        scaled = scale_pixel_coords(pix, origin, (enc.resize_height, enc.resize_width))
        gh, gw = enc.resize_height // enc.patch_size, enc.resize_width // enc.patch_size
        pxy = (scaled / float(enc.patch_size)).long()
        pxy[..., 0] = pxy[..., 0].clamp(0, gw - 1)
        pxy[..., 1] = pxy[..., 1].clamp(0, gh - 1)
        flat = (pxy[..., 1] * gw + pxy[..., 0]).clamp(0, gh * gw - 1)
        # image feature exactly as lines 184-201 compute it (hooked through the first Linear's input)
        captured = {}
        hook = model.mlp[0].register_forward_hook(lambda m, i, o: captured.__setitem__("x", i[0].detach()))
        with contextlib.redirect_stdout(io.StringIO()):
            scores = model(sw.tokens, sw.point_feat, sw.points, None, sw.image_sizes, sw.K, sw.Rt)
        hook.remove()
    fused_in = captured["x"]
    img_feat = fused_in[..., 64:]
    np.savez_compressed(
        os.path.join(OUT, name + ".npz"),
        B=B, V=V, seed=seed, encoder=encoder, tok_dtype=str(tok_dtype).replace("torch.", ""),
        input_checksum=sweeps_checksum(sw), stride=stride,
        pixel_coords=pix.numpy(), depth=depth.numpy(), valid=valid.numpy(),
        flat_idx=flat.numpy().astype(np.int32),
        img_feat_sub=img_feat[:, ::stride].float().numpy(),
        img_feat_dtype=str(img_feat.dtype).replace("torch.", ""),
        concat_dtype=str(fused_in.dtype).replace("torch.", ""),
        voxel_scores=scores.numpy(),
    )
    hits = valid.sum(1)
    print(f"{name}: valid pairs {int(valid.sum())}, pts with >=1 cam {(hits > 0).float().mean():.3f}, "
          f"mean hits {hits.float().mean():.3f}, 2+ cams {(hits > 1).float().mean():.4f}")


def run_kat_projection():
    """Hand-checkable projection cases: identity extrinsic, pinhole K, points on the image borders."""
    multi_camera_projector = REFM.fusion_model.multi_camera_projector
    f, cx, cy, H, W = 1000.0, 800.0, 450.0, 900, 1600
    K = torch.tensor([[f, 0, cx], [0, f, cy], [0, 0, 1.0]]).repeat(2, 1, 1)
    Rt = torch.eye(4).repeat(2, 1, 1)
    Rt[1, 0, 3] = 0.5  # second camera shifted in x
    z = 10.0
    pts = []
    for u, v in [(0.0, 0.0), (W - 1.0, H - 1.0), (float(W), 10.0), (10.0, float(H)), (-0.5, 5.0), (800.0, 450.0),
                 (1599.5, 899.5), (0.25, 0.75)]:
        pts.append([(u - cx) * z / f, (v - cy) * z / f, z])
    pts += [[0.0, 0.0, 0.0], [0.0, 0.0, -5.0], [1.0, 1.0, 1e-13], [1e6, 0.0, 1.0], [float("nan"), 0.0, 1.0],
            [0.0, 0.0, float("inf")]]
    pts = torch.tensor(pts, dtype=torch.float32)[None]
    with torch.no_grad():
        pix, depth, valid = multi_camera_projector(pts, [K], [Rt], torch.tensor([[H, W]]))
    np.savez_compressed(os.path.join(OUT, "kat_projection.npz"), points=pts.numpy(), K=K.numpy()[None],
                        Rt=Rt.numpy()[None], image_sizes=np.array([[H, W]]), pixel_coords=pix.numpy(),
                        depth=depth.numpy(), valid=valid.numpy())
    print("kat_projection: valid =", valid[0].int().tolist())


def run_mlp_backward(seed=5, B=2, V=192):
    FeatureFusionModel = REFM.fusion_model.FeatureFusionModel
    gen = torch.Generator().manual_seed(seed)
    x_pt = torch.randn(B, V, 64, generator=gen, requires_grad=True)
    x_img = torch.randn(B, V, 384, generator=gen)
    G = torch.randn(B, V, 16, generator=gen)
    model = FeatureFusionModel(None, None, 64, 384, 16, device="cpu")
    w = synth.make_mlp_weights(seed)
    model.load_state_dict(w, strict=True)
    out = model.mlp(torch.cat([x_pt, x_img], dim=-1))
    (out * G).sum().backward()
    np.savez_compressed(
        os.path.join(OUT, "mlp_backward.npz"), seed=seed, x_pt=x_pt.detach().numpy(), x_img=x_img.numpy(),
        G=G.numpy(), logits=out.detach().numpy(), dx_pt=x_pt.grad.numpy(),
        **{"grad_" + k: p.grad.numpy() for k, p in model.mlp.named_parameters()})
    print("mlp_backward: |dW1| =", float(model.mlp[0].weight.grad.norm()))


def run_cels(name, B, P, seed, mask_mode="equal", lam=1.0, scale=2.0):
    CELSLoss = REFM.losses.CELSLoss
    gen = torch.Generator().manual_seed(seed)
    logits = (torch.randn(B, P, 16, generator=gen) * scale).requires_grad_(True)
    labels = synth.make_labels(B, P, gen)
    if mask_mode == "none":
        mask = None
    else:
        Pm = {"equal": P, "longer": P + 37, "shorter": P - 29}[mask_mode]
        mask = torch.rand(B, Pm, generator=gen) > 0.1
    crit = CELSLoss(weight=torch.tensor(synth.CLASS_WEIGHTS), ignore_index=0, lamda_lovasz=lam)
    with contextlib.redirect_stdout(io.StringIO()):
        total, ce, lov, preds, gt = crit(logits, labels, mask)
    total.backward()
    np.savez_compressed(
        os.path.join(OUT, name + ".npz"), logits=logits.detach().numpy(), labels=labels.numpy(),
        mask=(mask.numpy() if mask is not None else np.zeros(0, dtype=bool)), has_mask=mask is not None,
        lam=lam, total=total.item(), ce=ce.item(), lovasz=lov.item(), preds=preds.numpy(), gt=gt.numpy(),
        dlogits=logits.grad.numpy(), class_weights=np.array(synth.CLASS_WEIGHTS, dtype=np.float32))
    print(f"{name}: total {total.item():.6f} ce {ce.item():.6f} lovasz {lov.item():.6f} N {preds.numel()}")


def run_lovasz_small():
    lovasz_grad, lovasz_softmax_flat = REFM.lovasz_losses.lovasz_grad, REFM.lovasz_losses.lovasz_softmax_flat
    cases = {}
    gt_sorted = torch.tensor([1.0, 0.0, 1.0, 0.0])
    cases["grad_1010"] = lovasz_grad(gt_sorted).numpy()
    cases["grad_single"] = lovasz_grad(torch.tensor([1.0])).numpy()
    probas = torch.tensor([[0.7, 0.2, 0.1], [0.1, 0.8, 0.1], [0.3, 0.3, 0.4], [0.25, 0.5, 0.25]])
    labels = torch.tensor([0, 1, 1, 0])  # class 2 absent -> skipped
    cases["probas"] = probas.numpy()
    cases["labels"] = labels.numpy()
    cases["loss"] = np.float32(lovasz_softmax_flat(probas, labels).item())
    empty = lovasz_softmax_flat(torch.zeros(0, 3), torch.zeros(0, dtype=torch.long))
    cases["empty_shape"] = np.array(empty.shape)
    np.savez_compressed(os.path.join(OUT, "lovasz_small.npz"), **cases)
    print("lovasz_small: loss", cases["loss"], "grad", cases["grad_1010"])


def run_confusion(seed=11, n=50_000):
    E = REFM.evaluation
    compute_confusion_matrix, compute_iou, per_class_accuracy, overall_accuracy, precision_recall_f1 = (
        E.compute_confusion_matrix, E.compute_iou, E.per_class_accuracy, E.overall_accuracy, E.precision_recall_f1)
    gen = torch.Generator().manual_seed(seed)
    labels = synth.make_labels(1, n, gen).view(-1)
    labels = labels[labels != 0]
    preds = torch.where(torch.rand(labels.numel(), generator=gen) < 0.7, labels,
                        torch.randint(1, 17, (labels.numel(),), generator=gen))
    cm = compute_confusion_matrix(preds, labels, 16)
    iou, miou = compute_iou(cm)
    acc, macc = per_class_accuracy(cm)
    pr, rc, f1 = precision_recall_f1(cm)
    np.savez_compressed(os.path.join(OUT, "confusion.npz"), preds=preds.numpy(), labels=labels.numpy(),
                        cm=cm.numpy(), iou=iou.numpy(), miou=miou.item(), acc=acc.numpy(), macc=macc.item(),
                        oa=overall_accuracy(cm).item(), precision=pr.item(), recall=rc.item(), f1=f1.item())
    print("confusion: mIoU", miou.item())


def run_devox(seed=13, B=2, P=3000, V=2200, pad=100):
    """utils/lidar.py:147-155 re-executed: cdist over all V_max rows (zero padding included), argmin, index."""
    gen = torch.Generator().manual_seed(seed)
    pts = synth.make_points(B, P, gen)
    pts[:, :6, :3] = torch.randn(B, 6, 3, generator=gen) * 0.2  # a few points nearest to the zero padding (F7)
    vox = torch.zeros(B, V, 4)
    for b in range(B):
        sel = torch.randperm(P, generator=gen)[: V - pad]
        vox[b, : V - pad] = pts[b, sel] + torch.randn(V - pad, 4, generator=gen) * 0.05
    scores = torch.randn(B, V, 16, generator=gen)
    idx = torch.zeros(B, P, dtype=torch.long)
    out = torch.zeros(B, P, 16)
    for b in range(B):
        d = torch.cdist(pts[b][:, :3], vox[b][:, :3])
        idx[b] = torch.argmin(d, dim=1)
        out[b] = scores[b, :V][idx[b]]
    np.savez_compressed(os.path.join(OUT, "devox.npz"), pts=pts.numpy(), vox=vox.numpy(), scores=scores.numpy(),
                        idx=idx.numpy(), out=out.numpy())
    print("devox: points mapped to a padded (zero) voxel:", int((idx >= V - pad).sum()))


def main(out=None):
    global OUT, REFM
    root = ref_loader.ref_root(allow_copy=False)  # the mounted tree only: goldens are never made from a copy
    if root is None:
        raise SystemExit("reference tree not found: goldens can only be regenerated where the reference is mounted")
    REFM = ref_loader.load_hot_path(root)
    for mod in (REFM.fusion_model, REFM.losses, REFM.evaluation, REFM.lovasz_losses):
        assert os.path.abspath(mod.__file__).startswith(os.path.abspath(root) + os.sep), mod.__file__
    OUT = out or HERE
    os.makedirs(OUT, exist_ok=True)
    warnings.filterwarnings("ignore", category=SyntaxWarning)
    torch.manual_seed(0)
    run_kat_projection()
    run_fusion_case("fusion_dinov3", B=2, V=4096, encoder="dinov3", seed=1)
    run_fusion_case("fusion_dinov2", B=1, V=4096, encoder="dinov2", seed=2)
    run_fusion_case("fusion_hires", B=1, V=6000, encoder="dinov3_hires", seed=3)
    run_fusion_case("fusion_fp16tok", B=1, V=2048, encoder="dinov3", seed=4, tok_dtype=torch.float16)
    run_mlp_backward()
    run_cels("cels_equal", B=2, P=3000, seed=21)
    run_cels("cels_longer_mask", B=2, P=1500, seed=22, mask_mode="longer", lam=1.5)
    run_cels("cels_shorter_mask", B=1, P=2000, seed=23, mask_mode="shorter")
    run_cels("cels_nomask", B=1, P=1000, seed=24, mask_mode="none")
    run_lovasz_small()
    run_confusion()
    run_devox()


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None, help="write the .npz files here instead of tests/golden/")
    main(ap.parse_args().out)
