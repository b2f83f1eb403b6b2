"""CPU oracle of the 3DSSF cross-modal fusion hot path — TEST INFRASTRUCTURE, NOT PRODUCT.

A numpy restatement of the reference's algorithm for every row of SURVEY.md §8(a), with every
function citing the reference file:line it follows (paths relative to the reference tree,
CharlesChang012/3D-Semantic-Segmentation-Fusion @ e5103fd).  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may import this
package; the product (`3d-semantic-segmentation-fusion_b200/`) never does and fails loudly when
its CUDA library is missing.

Pinning.  The reference has no golden vectors or tests for this path, so the oracle
is pinned against outputs of the reference itself: `tests/golden/make_golden.py` imports the
reference's own modules from /root/reference in the build container, runs them on seeded inputs
and commits inputs + outputs as `tests/golden/*.npz`; `tests/test_oracle_golden.py` checks every
function here against those files (bit-exact for projection, masks, patch indices, confusion
matrices; documented tolerances for floating-point reductions).

fp32 evaluation order is spelled out explicitly so the result does not depend on
the BLAS of the machine that runs the oracle.
"""
from __future__ import annotations

import numpy as np

F32 = np.float32


def _fma(a, b, c):
    """fp32 fused multiply-add: the fp64 product of two fp32 numbers is exact, so one fp64 add and one
    rounding to fp32 reproduces fma up to double-rounding ties (probability ~2^-29 per op)."""
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(F32)


# ------------------------------------------------------------------------------------------------
# a1  multi_camera_projector                                     utils/fusion_model.py:14-86
# ------------------------------------------------------------------------------------------------
def project(xyz, K, Rt, image_sizes):
    """xyz (B,V,3) f32; K (B,NC,3,3); Rt (B,NC,4,4); image_sizes (B,2) (H,W).
    Returns pixel_coords (B,NC,V,2) f32, depth (B,NC,V) f32, valid (B,NC,V) bool.

    Shape checks follow utils/fusion_model.py:45,53.  Per camera (:63-84):
      cam  = Rt @ [x y z 1]   — torch's batched 4x4 mat-vec on CPU: ((r0*x + r1*y) + r2*z) + r3*1,
                                 each product and sum rounded to fp32 (no FMA)            (:68)
      pix  = K @ cam_xyz       — MKL sgemm: fma(k2, z, fma(k1, y, k0*x))                   (:72)
      u,v  = pix_xy / clamp(pix_z, 1e-12)                                                  (:73-75)
      valid = z>0 & 0<=u<W & 0<=v<H ; invalid -> coords -1, depth -1                       (:77-84)
    """
    xyz = np.asarray(xyz, dtype=F32)
    K = np.asarray(K, dtype=F32)
    Rt = np.asarray(Rt, dtype=F32)
    B, V, _ = xyz.shape
    NC = K.shape[1]
    if K.shape[1:] != (NC, 3, 3):
        raise ValueError(f"Expected shape (*, {(NC, 3, 3)}) but got {K.shape}.")
    if Rt.shape[1:] != (NC, 4, 4):
        raise ValueError(f"Expected shape (*, {(NC, 4, 4)}) but got {Rt.shape}.")
    hw = np.asarray(image_sizes, dtype=F32)
    if hw.shape != (B, 2):
        raise ValueError("image_sizes must be of shape (B, 2) where each entry is (H, W).")
    img_h = hw[:, 0].reshape(B, 1)
    img_w = hw[:, 1].reshape(B, 1)
    x, y, z = xyz[..., 0], xyz[..., 1], xyz[..., 2]
    one = F32(1.0)

    pix_out = np.full((B, NC, V, 2), -1.0, dtype=F32)
    depth = np.zeros((B, NC, V), dtype=F32)
    valid_out = np.zeros((B, NC, V), dtype=bool)
    for c in range(NC):
        cam = []
        for i in range(3):
            r = Rt[:, c, i, :].reshape(B, 4, 1)
            acc = r[:, 0] * x + r[:, 1] * y
            acc = acc + r[:, 2] * z
            acc = acc + r[:, 3] * one
            cam.append(acc.astype(F32))
        pix = []
        for i in range(3):
            k = K[:, c, i, :].reshape(B, 3, 1)
            acc = (np.broadcast_to(k[:, 0], cam[0].shape) * cam[0]).astype(F32)
            acc = _fma(np.broadcast_to(k[:, 1], cam[1].shape), cam[1], acc)
            acc = _fma(np.broadcast_to(k[:, 2], cam[2].shape), cam[2], acc)
            pix.append(acc)
        with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
            denom = np.where(np.isnan(pix[2]), pix[2], np.maximum(pix[2], F32(1e-12))).astype(F32)
            u = (pix[0] / denom).astype(F32)
            v = (pix[1] / denom).astype(F32)
            ok = (cam[2] > 0) & (u >= 0) & (u < img_w) & (v >= 0) & (v < img_h)
        valid_out[:, c] = ok
        pix_out[:, c, :, 0] = np.where(ok, u, F32(-1.0))
        pix_out[:, c, :, 1] = np.where(ok, v, F32(-1.0))
        depth[:, c] = np.where(ok, cam[2], F32(-1.0))
    return pix_out, depth, valid_out


# ------------------------------------------------------------------------------------------------
# a2  scale_pixel_coords                                        utils/fusion_model.py:89-127
# ------------------------------------------------------------------------------------------------
def scale_pixel_coords(pixel_coords, origin_img_sizes, new_img_size):
    """(u,v) * (W_new/W_orig, H_new/H_orig); scales computed in fp32 (:117-122), -1 coords scaled too."""
    if isinstance(new_img_size, int):
        h_new = w_new = float(new_img_size)
    else:
        h_new, w_new = map(float, new_img_size)
    hw = np.asarray(origin_img_sizes, dtype=F32)
    if hw.ndim == 1:
        hw = hw[None]
    if hw.shape[-1] != 2:
        raise ValueError("origin_img_sizes must have shape (B, 2) in (H, W) order.")
    h0 = np.maximum(hw[:, 0], F32(1e-6))
    w0 = np.maximum(hw[:, 1], F32(1e-6))
    scale = np.stack([(F32(w_new) / w0).astype(F32), (F32(h_new) / h0).astype(F32)], axis=-1)
    shape = (scale.shape[0],) + (1,) * (pixel_coords.ndim - 2) + (2,)
    return (np.asarray(pixel_coords, dtype=F32) * scale.reshape(shape)).astype(F32)


# ------------------------------------------------------------------------------------------------
# a3  patch index                                               utils/fusion_model.py:172-180,188
# ------------------------------------------------------------------------------------------------
def patch_index(scaled_coords, patch_size, resize_h, resize_w):
    """trunc(pix / float(patch)) -> clamp x [0,gw-1], y [0,gh-1] -> y*gw+x -> clamp [0,M-1]; int64."""
    gh = resize_h // patch_size
    gw = resize_w // patch_size
    q = (np.asarray(scaled_coords, dtype=F32) / F32(float(patch_size))).astype(F32)
    pxy = np.trunc(q).astype(np.int64)
    px = np.clip(pxy[..., 0], 0, gw - 1)
    py = np.clip(pxy[..., 1], 0, gh - 1)
    return np.clip(py * gw + px, 0, gh * gw - 1)


# ------------------------------------------------------------------------------------------------
# a4+a5  gather + masked mean over cameras                      utils/fusion_model.py:184-201
# ------------------------------------------------------------------------------------------------
def gather_mean(tokens, flat_idx, valid):
    """tokens (B,NC,M,Ci); flat_idx (B,NC,V) int; valid (B,NC,V) bool -> (B,V,Ci) in the token dtype.
    Every (point, camera) pair gathers (invalid ones read patch 0.. and are multiplied by 0, F4);
    sum over cameras in order 0..NC-1, then / clamp(count, 1)  (:198-201)."""
    tokens = np.asarray(tokens)
    dt = tokens.dtype
    B, NC, M, Ci = tokens.shape
    V = flat_idx.shape[-1]
    acc = np.zeros((B, V, Ci), dtype=np.float32)
    for c in range(NC):
        g = np.take_along_axis(tokens[:, c], flat_idx[:, c, :, None].astype(np.int64), axis=1)  # (B,V,Ci)
        m = valid[:, c, :, None].astype(dt)
        acc = acc + (g * m).astype(np.float32)  # fp32 accumulation (exact order for fp32 tokens)
    cnt = np.maximum(valid.astype(np.float32).sum(axis=1), 1.0)[..., None]
    if dt == np.float32:
        return (acc / cnt).astype(np.float32)
    s = acc.astype(dt)  # same-dtype reduction result, then same-dtype quotient
    return (s.astype(np.float32) / cnt).astype(dt)


# ------------------------------------------------------------------------------------------------
# a6  concat + fusion MLP                                       utils/fusion_model.py:144-150,203-204
# ------------------------------------------------------------------------------------------------
def mlp_forward(x_pt, x_img, w, dtype=np.float32, return_hidden=False):
    """cat([point, image]) -> Linear(448,256)+ReLU -> Linear(256,64)+ReLU -> Linear(64,16).
    w: dict with the reference's state_dict names mlp.{0,2,4}.{weight,bias} (weight = (out,in))."""
    x = np.concatenate([np.asarray(x_pt, dtype=dtype), np.asarray(x_img).astype(dtype)], axis=-1)
    g = lambda k: np.asarray(w[k], dtype=dtype)
    h1 = np.maximum(x @ g("mlp.0.weight").T + g("mlp.0.bias"), 0)
    h2 = np.maximum(h1 @ g("mlp.2.weight").T + g("mlp.2.bias"), 0)
    out = h2 @ g("mlp.4.weight").T + g("mlp.4.bias")
    if return_hidden:
        return out, h1, h2
    return out


def mlp_backward(x_pt, x_img, w, dlogits, dtype=np.float64):
    """Gradients of the MLP w.r.t. its parameters and the point feature (first 64 input columns; the
    image branch carries no gradient).  Returns (grads dict, dx_pt)."""
    x = np.concatenate([np.asarray(x_pt, dtype=dtype), np.asarray(x_img).astype(dtype)], axis=-1)
    g = lambda k: np.asarray(w[k], dtype=dtype)
    _, h1, h2 = mlp_forward(x_pt, x_img, w, dtype=dtype, return_hidden=True)
    dl = np.asarray(dlogits, dtype=dtype)
    x2, h1f, h2f, dlf = x.reshape(-1, x.shape[-1]), h1.reshape(-1, 256), h2.reshape(-1, 64), dl.reshape(-1, dl.shape[-1])
    grads = {"mlp.4.weight": dlf.T @ h2f, "mlp.4.bias": dlf.sum(0)}
    dh2 = (dlf @ g("mlp.4.weight")) * (h2f > 0)
    grads["mlp.2.weight"] = dh2.T @ h1f
    grads["mlp.2.bias"] = dh2.sum(0)
    dh1 = (dh2 @ g("mlp.2.weight")) * (h1f > 0)
    grads["mlp.0.weight"] = dh1.T @ x2
    grads["mlp.0.bias"] = dh1.sum(0)
    cp = np.asarray(x_pt).shape[-1]
    dx_pt = (dh1 @ g("mlp.0.weight")[:, :cp]).reshape(np.asarray(x_pt).shape)
    return grads, dx_pt


# ------------------------------------------------------------------------------------------------
# a1..a6  FeatureFusionModel.forward up to voxel_scores         utils/fusion_model.py:152-204
# ------------------------------------------------------------------------------------------------
def fusion_forward(tokens, voxel_features, voxel_raw, image_sizes, K, Rt, w, patch_size, resize_h, resize_w):
    pix, _, valid = project(np.asarray(voxel_raw)[:, :, :3], K, Rt, image_sizes)
    pix = scale_pixel_coords(pix, np.asarray(image_sizes, dtype=F32), (resize_h, resize_w))
    flat = patch_index(pix, patch_size, resize_h, resize_w)
    img = gather_mean(tokens, flat, valid)
    return mlp_forward(voxel_features, img.astype(np.float32), w), img, flat, valid


# ------------------------------------------------------------------------------------------------
# a7  devoxelize                                                utils/lidar.py:120-172
# ------------------------------------------------------------------------------------------------
def devox_nn(pts_xyz, vox_xyz, chunk=2048):
    """argmin_v ||p - v|| per point over ALL V rows (zero padding included, F7), first index on ties.
    Distances in float64 from direct differences: this is the exact nearest neighbour.  The reference
    (torch.cdist, :147) evaluates ||x||^2+||y||^2-2x.y in fp32, so where the two best candidates are
    closer than its rounding error it may pick the other one; `devox_gap` measures that margin."""
    p = np.asarray(pts_xyz, dtype=np.float64)
    v = np.asarray(vox_xyz, dtype=np.float64)
    out = np.empty(p.shape[0], dtype=np.int64)
    for s in range(0, p.shape[0], chunk):
        d = ((p[s:s + chunk, None, :] - v[None, :, :]) ** 2).sum(-1)
        out[s:s + chunk] = d.argmin(axis=1)
    return out


def devox_gap(pts_xyz, vox_xyz, chunk=2048):
    """(best, second best) distance per point, float64 — for the H7 ambiguity margin."""
    p = np.asarray(pts_xyz, dtype=np.float64)
    v = np.asarray(vox_xyz, dtype=np.float64)
    best = np.empty(p.shape[0])
    second = np.empty(p.shape[0])
    for s in range(0, p.shape[0], chunk):
        d = np.sqrt(((p[s:s + chunk, None, :] - v[None, :, :]) ** 2).sum(-1))
        part = np.partition(d, 1, axis=1) if d.shape[1] > 1 else np.concatenate([d, d + np.inf], 1)
        best[s:s + chunk] = part[:, 0]
        second[s:s + chunk] = part[:, 1]
    return best, second


def devoxelize(voxel_scores, lidar_points_raw, voxel_raw):
    """(B,V,C) scores -> (B,P,C) point scores + (B,P) int64 indices (:126-172, equal P per sweep)."""
    B, V, C = voxel_scores.shape
    P = lidar_points_raw.shape[1]
    out = np.zeros((B, P, C), dtype=np.float32)
    idx = np.zeros((B, P), dtype=np.int64)
    for b in range(B):
        idx[b] = devox_nn(lidar_points_raw[b][:, :3], voxel_raw[b][:, :3])
        out[b] = voxel_scores[b][idx[b]]
    return out, idx


# ------------------------------------------------------------------------------------------------
# a10  lovasz_grad                             LovaszSoftmax/pytorch/lovasz_losses.py:18-30
# ------------------------------------------------------------------------------------------------
def lovasz_grad(gt_sorted):
    """gts - cumsum(fg), gts + cumsum(1-fg), jaccard = 1 - inter/union, first difference (:23-29).
    fp32; the cumulative sums of 0/1 values are exact integers below 2^24."""
    gt = np.asarray(gt_sorted, dtype=np.float32)
    p = gt.shape[0]
    gts = gt.sum(dtype=np.float32)
    inter = gts - np.cumsum(gt, dtype=np.float32)
    union = gts + np.cumsum(F32(1) - gt, dtype=np.float32)
    jac = (F32(1.0) - inter / union).astype(np.float32)
    if p > 1:
        jac[1:p] = jac[1:p] - jac[0:-1]
    return jac


# ------------------------------------------------------------------------------------------------
# a9  lovasz_softmax_flat(classes='present')   LovaszSoftmax/pytorch/lovasz_losses.py:171-199
# ------------------------------------------------------------------------------------------------
def lovasz_softmax_flat(probas, labels, return_grad=False):
    """probas (N,C) f32, labels (N,) in 0..C-1.  Per present class: errors=|fg-p_c| (:194), sort
    descending (:195) — stable, ties by ascending index — dot with lovasz_grad (:198),
    mean over present classes (:199, helper :232-250).  Empty input returns probas*0 (:178-180)."""
    probas = np.asarray(probas, dtype=np.float32)
    labels = np.asarray(labels)
    if probas.size == 0:
        return (probas * F32(0.0), np.zeros_like(probas)) if return_grad else probas * F32(0.0)
    N, C = probas.shape
    losses = []
    dprob = np.zeros((N, C), dtype=np.float64)
    for c in range(C):
        fg = (labels == c).astype(np.float32)
        if fg.sum() == 0:
            continue
        err = np.abs(fg - probas[:, c]).astype(np.float32)
        perm = np.argsort(-err, kind="stable")
        g = lovasz_grad(fg[perm])
        losses.append(np.dot(err[perm].astype(np.float64), g.astype(np.float64)))
        if return_grad:
            # d|fg - p|/dp = -sign(fg - p); sign(0) = 0 like torch.abs' backward
            dprob[perm, c] = g.astype(np.float64) * (-np.sign(fg[perm] - probas[perm, c]))
    n = len(losses)
    loss = F32(sum(losses) / n) if n else F32(0.0)
    if return_grad:
        return loss, (dprob / max(n, 1))
    return loss


# ------------------------------------------------------------------------------------------------
# a8  CELSLoss.forward                                          utils/losses.py:19-70
# ------------------------------------------------------------------------------------------------
def _softmax(x):
    x = np.asarray(x, dtype=np.float32)
    m = x.max(axis=-1, keepdims=True)
    e = np.exp(x - m).astype(np.float32)
    return (e / e.sum(axis=-1, keepdims=True, dtype=np.float32)).astype(np.float32)


def cels_loss(pred_scores, gt_labels, mask=None, weight=None, ignore_index=0, lamda_lovasz=1.0,
              return_grad=False):
    """Returns (total, ce, lovasz, predictions (N,) int64 in 1..C, gt_valid (N,) int64) and, with
    return_grad, d total / d pred_scores (same shape as pred_scores, float64).
    Steps: align mask/pred lengths (:32-43); compact by mask (:46-48) then label != ignore (:54-57);
    softmax (:56); weighted CE on label-1, mean = sum(w*nll)/sum(w) (:15,60); Lovasz on the probs
    (:63); predictions = argmax(probs)+1 (:66); total = ce + lambda*lovasz (:68)."""
    pred_scores = np.asarray(pred_scores, dtype=np.float32)
    gt_labels = np.asarray(gt_labels)
    B, P, C = pred_scores.shape
    if mask is not None:
        mask = np.asarray(mask)
        if mask.shape[1] > P:
            mask = mask[:, :P]
        elif mask.shape[1] < P:
            pred_scores = pred_scores[:, :mask.shape[1]]
            gt_labels = gt_labels[:, :mask.shape[1]]
            P = mask.shape[1]
        keep = mask.reshape(-1).astype(bool)
    else:
        keep = np.ones(B * P, dtype=bool)
    flat_scores = pred_scores.reshape(-1, C)
    flat_labels = gt_labels.reshape(-1)
    keep = keep & (flat_labels != ignore_index)
    rows = np.nonzero(keep)[0]
    x = flat_scores[rows]
    y = flat_labels[rows].astype(np.int64)
    probs = _softmax(x)
    w = np.ones(C, dtype=np.float32) if weight is None else np.asarray(weight, dtype=np.float32)
    if rows.size:
        x64 = x.astype(np.float64)
        lse = np.log(np.exp(x64 - x64.max(-1, keepdims=True)).sum(-1)) + x64.max(-1)
        nll = lse - x64[np.arange(rows.size), y - 1]
        wy = w[y - 1].astype(np.float64)
        ce = F32((wy * nll).sum() / wy.sum())
    else:
        ce = F32(np.nan)  # nn.CrossEntropyLoss over an empty batch
    if return_grad:
        lov, dprob = lovasz_softmax_flat(probs, y - 1, return_grad=True)
    else:
        lov = lovasz_softmax_flat(probs, y - 1)
    preds = probs.argmax(axis=-1).astype(np.int64) + 1 if rows.size else np.zeros(0, dtype=np.int64)
    total = ce + F32(lamda_lovasz) * lov
    if not return_grad:
        return total, ce, lov, preds, y
    grad = np.zeros((pred_scores.shape[0] * P, C), dtype=np.float64)
    if rows.size:
        p64 = probs.astype(np.float64)
        onehot = np.zeros_like(p64)
        onehot[np.arange(rows.size), y - 1] = 1.0
        g_ce = (wy / wy.sum())[:, None] * (p64 - onehot)
        dp = dprob * float(lamda_lovasz)
        g_lov = p64 * (dp - (dp * p64).sum(-1, keepdims=True))
        grad[rows] = g_ce + g_lov
    return total, ce, lov, preds, y, grad.reshape(pred_scores.shape[0], P, C)


# ------------------------------------------------------------------------------------------------
# a11  confusion matrix and metrics                             utils/evaluation.py:35-73
# ------------------------------------------------------------------------------------------------
def confusion_matrix(preds, labels, num_classes):
    """bincount(C*(label-1) + (pred-1), minlength=C^2).reshape(C,C); rows = truth (:36-42)."""
    p = np.asarray(preds, dtype=np.int64) - 1
    l = np.asarray(labels, dtype=np.int64) - 1
    return np.bincount(num_classes * l + p, minlength=num_classes ** 2).reshape(num_classes, num_classes)


def compute_iou(cm):
    inter = np.diag(cm)
    union = cm.sum(1) + cm.sum(0) - inter
    iou = inter.astype(np.float32) / np.maximum(union.astype(np.float32), F32(1))
    return iou, iou.mean(dtype=np.float32)


def per_class_accuracy(cm):
    acc = np.diag(cm).astype(np.float32) / np.maximum(cm.sum(1).astype(np.float32), F32(1))
    return acc, acc.mean(dtype=np.float32)


def overall_accuracy(cm):
    return F32(np.diag(cm).sum()) / max(F32(cm.sum()), F32(1))


def precision_recall_f1(cm):
    tp = np.diag(cm).astype(np.float32)
    fp = cm.sum(0).astype(np.float32) - tp
    fn = cm.sum(1).astype(np.float32) - tp
    precision = (tp / np.maximum(tp + fp, F32(1))).mean(dtype=np.float32)
    recall = (tp / np.maximum(tp + fn, F32(1))).mean(dtype=np.float32)
    f1 = F32(2) * precision * recall / max(precision + recall, F32(1e-6))
    return precision, recall, f1


# ------------------------------------------------------------------------------------------------
# f1  voxelize_open3d                                           utils/lidar.py:83-118
#
# PARITY UNPINNED for the down-sampling step: it lives in Open3D (pinned 0.18.0, environment.yml:278), which is
# neither under the reference tree nor installed here.  Restated from its published algorithm
# (geometry::PointCloud::VoxelDownSample, cpp/open3d/geometry/PointCloud.cpp): all arithmetic in double on the
# fp32 coordinates; voxel_min_bound = min_bound - voxel_size/2; voxel index = int(floor((p - voxel_min_bound) /
# voxel_size)) per axis; per occupied voxel the points are accumulated in point order and divided by their count.
# Open3D emits the voxels in the iteration order of a std::unordered_map (implementation-defined); here they come
# in the order of the first point of each voxel, so comparisons with Open3D itself would be as sets.
# The 5-NN intensity step is scikit-learn (installed here: `voxelize(..., knn="sklearn")` runs the reference's own
# call) and numpy; `knn="brute"` restates it (exact double squared distances, ties to the lower point index).
# ------------------------------------------------------------------------------------------------
KNN_WEIGHTS = np.array([0.4, 0.3, 0.2, 0.05, 0.05])


def voxel_down_sample(xyz, voxel_size):
    """xyz (P,3) fp32 -> centroids (V,3) float64 in order of first appearance, plus per-point voxel id (P,)."""
    pts = np.asarray(xyz, dtype=F32).astype(np.float64)
    vmin = pts.min(axis=0) - np.float64(voxel_size) * 0.5
    vidx = np.floor((pts - vmin) / np.float64(voxel_size)).astype(np.int64)
    _, first, inv = np.unique(vidx, axis=0, return_index=True, return_inverse=True)
    inv = inv.reshape(-1)
    order = np.argsort(first, kind="stable")            # voxels by their first point
    rank = np.empty_like(order)
    rank[order] = np.arange(order.size)
    vid = rank[inv]
    sums = np.zeros((order.size, 3), dtype=np.float64)
    np.add.at(sums, vid, pts)                            # unbuffered: accumulates in point order, like AddPoint
    cnt = np.bincount(vid, minlength=order.size).astype(np.float64)
    return sums / cnt[:, None], vid


def knn5_brute(xyz, queries, chunk=512):
    """Indices (V,5) of the five nearest points per query, nearest first; double squared distances
    ((dx*dx + dy*dy) + dz*dz), ties to the lower index; also the gap d2[5] - d2[4] to the sixth neighbour."""
    pts = np.asarray(xyz, dtype=F32).astype(np.float64)
    out = np.empty((queries.shape[0], 5), dtype=np.int64)
    gap = np.full(queries.shape[0], np.inf)
    mingap = np.full(queries.shape[0], np.inf)
    for s in range(0, queries.shape[0], chunk):
        q = queries[s:s + chunk]
        d = q[:, None, :] - pts[None, :, :]
        d2 = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]
        k = min(6, pts.shape[0])
        part = np.argpartition(d2, k - 1, axis=1)[:, :k]
        pd = np.take_along_axis(d2, part, axis=1)
        o = np.lexsort((part, pd), axis=1)               # by distance, then by index
        part = np.take_along_axis(part, o, axis=1)
        pd = np.take_along_axis(pd, o, axis=1)
        out[s:s + chunk] = part[:, :5]
        if k == 6:
            gap[s:s + chunk] = pd[:, 5] - pd[:, 4]
        mingap[s:s + chunk] = np.diff(pd[:, :k], axis=1).min(axis=1)
    return out, gap, mingap


def voxelize(points, voxel_size, knn="brute"):
    """points (P,4) fp32 [x y z intensity] -> voxel_raw (V,4) fp32, voxel_coords (V,3) int32, following
    utils/lidar.py:83-118 step by step (Step 2 per the note above).  Extra returns for the tests: per-point voxel
    id and the smallest distance gap among each voxel's six nearest points (0 = a tie: neighbour order undefined)."""
    pts = np.asarray(points, dtype=F32)
    centres, vid = voxel_down_sample(pts[:, :3], voxel_size)             # :92-93
    if pts.shape[1] < 4:                                                  # :104-105: no intensity column
        return centres.astype(F32), np.floor(centres / voxel_size).astype(np.int32), vid, None
    if knn == "sklearn":
        from sklearn.neighbors import NearestNeighbors
        _, idx = NearestNeighbors(n_neighbors=5, algorithm="auto").fit(pts[:, :3]).kneighbors(centres)   # :98-99
        mingap = None
    else:
        idx, _, mingap = knn5_brute(pts[:, :3], centres)
    inten = (pts[idx, 3] * KNN_WEIGHTS.reshape(1, 5)).sum(axis=1).reshape(-1, 1)                        # :101-102
    raw = np.hstack([centres, inten])                                                                   # :103
    coords = np.floor(centres / voxel_size).astype(np.int32)                                            # :108-110
    return raw.astype(F32), coords, vid, mingap                                                         # :113-114


# ------------------------------------------------------------------------------------------------
# f4  calibration composition                                     utils/dataloader.py:123-152
#     multi-sweep accumulation: nuscenes-devkit LidarPointCloud.from_file_multisweep (dependency of the reference,
#     environment.yml; source not under the reference tree -> that part is a restatement, parity unpinned)
# ------------------------------------------------------------------------------------------------
def pose_matrix(pose):
    """pose = (qw,qx,qy,qz,tx,ty,tz) float64 -> 4x4.  utils/dataloader.py:127-131,139-143:
    `T = eye(4); T[:3,:3] = Quaternion(rotation).rotation_matrix; T[:3,3] = translation`.  pyquaternion normalises the
    quaternion and returns the 3x3 block of Q(q) @ Qbar(q)^T, i.e. the w²+x²−y²−z² form used here."""
    w, x, y, z = (float(v) for v in pose[:4])
    n = np.sqrt(w * w + x * x + y * y + z * z)
    if n > 0:
        w, x, y, z = w / n, x / n, y / n, z / n
    T = np.eye(4)
    T[:3, :3] = [[w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
                 [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
                 [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z]]
    T[:3, 3] = pose[4:7]
    return T


def compose_lidar2cam(lidar_sensor, cam_sensor, lidar_ego=None, cam_ego=None):
    """lidar_sensor (B,7), cam_sensor (B,NC,7) [, lidar_ego (B,7), cam_ego (B,NC,7)] -> (B,NC,4,4) float64.
    Without ego poses: utils/dataloader.py:146-149 verbatim, `np.linalg.inv(T_cam_ego) @ T_lidar_ego` (the reference fetches
    pose_record_lidar / pose_record at :125,138 and never uses them).  With them: the four-stage chain
    inv(T_cam_ego) @ inv(T_egocam_global) @ T_egolidar_global @ T_lidar_ego."""
    B, NC = cam_sensor.shape[:2]
    out = np.zeros((B, NC, 4, 4))
    for b in range(B):
        T_lidar_ego = pose_matrix(lidar_sensor[b])
        for c in range(NC):
            T = T_lidar_ego
            if lidar_ego is not None:
                T = np.linalg.inv(pose_matrix(cam_ego[b, c])) @ (pose_matrix(lidar_ego[b]) @ T)
            out[b, c] = np.linalg.inv(pose_matrix(cam_sensor[b, c])) @ T
    return out


def compose_sweep2ref(sweep_sensor, sweep_ego, ref_sensor, ref_ego):
    """(S,7),(S,7),(7,),(7,) -> (S,4,4) float64: ref_from_car @ car_from_global @ global_from_car @ car_from_current."""
    ref_from_car = np.linalg.inv(pose_matrix(ref_sensor))
    car_from_global = np.linalg.inv(pose_matrix(ref_ego))
    return np.stack([ref_from_car @ car_from_global @ pose_matrix(sweep_ego[s]) @ pose_matrix(sweep_sensor[s])
                     for s in range(len(sweep_sensor))])


def accumulate_sweeps(pts, counts, T):
    """pts (S,P,4) f32, counts (S,), T (S,4,4) f64 -> (sum counts, 4) f32: xyz of sweep s mapped by T[s] in float64
    (`T.dot(vstack((xyz, ones)))`), stored as fp32; intensity copied; sweeps concatenated in order."""
    out = []
    for s in range(len(counts)):
        p = pts[s, :counts[s]]
        h = np.vstack((p[:, :3].T.astype(np.float64), np.ones(len(p))))
        q = p.copy()
        q[:, :3] = (T[s] @ h)[:3].T.astype(F32)
        out.append(q)
    return np.concatenate(out) if out else np.zeros((0, 4), F32)
