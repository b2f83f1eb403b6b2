"""torch-facing wrappers of the C ABI (include/tdssf.h).

PyTorch is plumbing here: device memory, streams, autograd bookkeeping.  Every op launches the
hand-written sm_100a kernels of libtdssf.so through ctypes and raises if the tensors are not on a
CUDA device — there is no CPU or eager fallback.
"""
from __future__ import annotations

import ctypes as C
import functools
import os
from typing import Optional, Sequence

import torch

from . import _lib
from ._lib import lib, check

_PACK_BYTES = lib.tdssf_mlp_pack_bytes()
_TOK_DTYPES = {torch.float32: _lib.F32, torch.float16: _lib.F16, torch.bfloat16: _lib.BF16}


def _ptr(t: Optional[torch.Tensor]):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _stream(t: torch.Tensor):
    # the raw handle of torch's current stream on the tensor's device (the Stream-object route costs ~7 us per call)
    return C.c_void_p(torch._C._cuda_getCurrentRawStream(t.device.index))


def _need_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("tdssf ops run on CUDA tensors only (no CPU fallback); got a tensor on " + str(t.device))


def _f32c(t: torch.Tensor) -> torch.Tensor:
    return t if (t.dtype == torch.float32 and t.is_contiguous()) else t.to(torch.float32).contiguous()


def _on_device(fn):
    """The library launches on the CURRENT device (and caches per-device state by it), while the stream and the
    memory belong to the tensors' device: run `fn` with the device of its first CUDA tensor argument current, so a
    process that drives several GPUs (or a model on cuda:1 without set_device) launches where its data lives."""
    @functools.wraps(fn)
    def guarded(*args, **kwargs):
        for a in args:
            if torch.is_tensor(a) and a.is_cuda:
                if a.device.index == torch.cuda.current_device():
                    break
                with torch.cuda.device(a.device):
                    return fn(*args, **kwargs)
        return fn(*args, **kwargs)
    return guarded


# ------------------------------------------------------------------------------------------------
# calibration stacking: mirrors multi_camera_projector's `_to_tensor` (utils/fusion_model.py:36-49)
# ------------------------------------------------------------------------------------------------
def stack_calibration(cam_intrinsics: Sequence, lidar2cam_extrinsics: Sequence, device, image_sizes=None):
    """(K (B,NC,3,3), Rt (B,NC,4,4)) fp32 on `device`; with `image_sizes` also hw (B,2) fp32 — host inputs go up in ONE copy."""
    n_cam = cam_intrinsics[0].shape[0]

    def as_f32(data):
        if all(torch.is_tensor(item) and item.dtype == torch.float32 for item in data):
            return list(data)
        return [torch.as_tensor(item, dtype=torch.float32) if not torch.is_tensor(item) else item.to(dtype=torch.float32)
                for item in data]

    def check(t, tail):
        if tuple(t.shape[1:]) != tail:
            raise ValueError(f"Expected shape (*, {tail}) but got {tuple(t.shape)}.")

    ks, rts = as_f32(cam_intrinsics), as_f32(lidar2cam_extrinsics)
    hw = None
    if not any(t.is_cuda for t in ks) and not any(t.is_cuda for t in rts):
        # the loops hand over lists of host tensors (utils/train.py:83): one stack and ONE upload for everything
        K, Rt = torch.stack(ks, dim=0), torch.stack(rts, dim=0)
        check(K, (n_cam, 3, 3))
        check(Rt, (n_cam, 4, 4))
        B = K.shape[0]
        parts = [K.reshape(-1), Rt.reshape(-1)]
        hw_host = image_sizes is not None and not (torch.is_tensor(image_sizes) and image_sizes.is_cuda)
        if hw_host:
            h = torch.as_tensor(image_sizes, dtype=torch.float32)
            if tuple(h.shape) != (B, 2):
                raise ValueError("image_sizes must be of shape (B, 2) where each entry is (H, W).")
            parts.append(h.reshape(-1))
        flat = torch.cat(parts).to(device, non_blocking=True)
        nk, nr = K.numel(), Rt.numel()
        K, Rt = flat[:nk].view(B, n_cam, 3, 3), flat[nk:nk + nr].view(B, n_cam, 4, 4)   # contiguous views of one buffer
        if hw_host:
            hw = flat[nk + nr:].view(B, 2)
    else:
        K = torch.stack([t.to(device) for t in ks], dim=0).contiguous()
        Rt = torch.stack([t.to(device) for t in rts], dim=0).contiguous()
        check(K, (n_cam, 3, 3))
        check(Rt, (n_cam, 4, 4))
    if image_sizes is None:
        return K, Rt
    return K, Rt, (hw if hw is not None else image_sizes_f32(image_sizes, K.shape[0], device))


def image_sizes_f32(image_sizes, B: int, device) -> torch.Tensor:
    hw = torch.as_tensor(image_sizes, dtype=torch.float32, device=device)
    if tuple(hw.shape) != (B, 2):
        raise ValueError("image_sizes must be of shape (B, 2) where each entry is (H, W).")
    return hw.contiguous()


def _points_layout(points: torch.Tensor):
    """(B,V,>=3) fp32 view with unit inner stride -> (tensor kept alive, sweep_stride, point_stride)."""
    if points.dtype != torch.float32 or points.stride(-1) != 1 or points.dim() != 3:
        points = points.to(torch.float32).contiguous()
    B, V, _ = points.shape
    ss = points.stride(0) if B > 1 else points.stride(1) * V
    return points, ss, points.stride(1)


# ------------------------------------------------------------------------------------------------
# a1-a3
# ------------------------------------------------------------------------------------------------
@_on_device
def project(points, K, Rt, img_hw, encoder=None, *, want_pix=True, want_depth=True, want_valid=True,
            want_flat=False, want_rec=False):
    """points (B,V,>=3) CUDA fp32 (may be a strided view of (B,V,4)); K (B,NC,3,3); Rt (B,NC,4,4);
    img_hw (B,2) fp32.  Returns dict with the requested outputs."""
    _need_cuda(points, K, Rt, img_hw)
    pts, ss, ps = _points_layout(points)
    B, V, _ = pts.shape
    NC = K.shape[1]
    dev = pts.device
    out = {}
    pix = torch.empty((B, NC, V, 2), dtype=torch.float32, device=dev) if want_pix else None
    depth = torch.empty((B, NC, V), dtype=torch.float32, device=dev) if want_depth else None
    valid = torch.empty((B, NC, V), dtype=torch.uint8, device=dev) if want_valid else None
    flat = torch.empty((B, NC, V), dtype=torch.int32, device=dev) if want_flat else None
    rec = torch.empty((B, V, 4), dtype=torch.int32, device=dev) if want_rec else None
    rh = rw = ps_ = 0
    if want_flat or want_rec:
        if encoder is None:
            raise ValueError("patch indices need the encoder's patch_size / resize_height / resize_width")
        rh, rw, ps_ = int(encoder.resize_height), int(encoder.resize_width), int(encoder.patch_size)
    check(lib.tdssf_project(_ptr(pts), ss, ps, _ptr(_f32c(K)), _ptr(_f32c(Rt)), _ptr(_f32c(img_hw)), B, V, NC,
                            rh, rw, ps_, _ptr(pix), _ptr(depth), _ptr(valid), _ptr(flat), _ptr(rec), _stream(pts)),
          "tdssf_project")
    out.update(pix=pix, depth=depth, valid=valid, flat_idx=flat, rec=rec)
    return out


@_on_device
def scale_pixels(pixel_coords: torch.Tensor, origin_hw: torch.Tensor, new_h: float, new_w: float) -> torch.Tensor:
    """pixel_coords (B,...,2) CUDA fp32; origin_hw (B,2) or (1,2) fp32 on the same device."""
    _need_cuda(pixel_coords, origin_hw)
    pix = _f32c(pixel_coords)
    hw = _f32c(origin_hw)
    B = pix.shape[0]
    per_sweep = pix.numel() // (2 * B) if B else 0
    out = torch.empty_like(pix)
    check(lib.tdssf_scale_pixels(_ptr(pix), _ptr(hw), hw.shape[0], B, per_sweep, float(new_h), float(new_w), _ptr(out),
                                 _stream(pix)), "tdssf_scale_pixels")
    return out


# ------------------------------------------------------------------------------------------------
# a4-a5
# ------------------------------------------------------------------------------------------------
@_on_device
def gather_mean(tokens: torch.Tensor, rec: torch.Tensor) -> torch.Tensor:
    """tokens (B,NC,M,Ci) CUDA fp32/fp16/bf16; rec (B,V,4) int32 records -> (B,V,Ci) fp32."""
    _need_cuda(tokens, rec)
    if tokens.dtype not in _TOK_DTYPES:
        tokens = tokens.float()
    tokens = tokens.contiguous()
    B, NC, M, Ci = tokens.shape
    V = rec.shape[1]
    feat = torch.empty((B, V, Ci), dtype=torch.float32, device=tokens.device)
    check(lib.tdssf_gather_mean(_ptr(tokens), _TOK_DTYPES[tokens.dtype], _ptr(rec), B, V, NC, M, Ci, _ptr(feat),
                                _stream(tokens)), "tdssf_gather_mean")
    return feat


# ------------------------------------------------------------------------------------------------
# a6: fp32 CUDA-core MLP with autograd, and the tensor-core paths
# ------------------------------------------------------------------------------------------------
def _weights_struct(params: Sequence[torch.Tensor]) -> _lib.MlpWeights:
    W1, b1, W2, b2, W3, b3 = params
    if tuple(W1.shape) != (256, 448) or tuple(W2.shape) != (64, 256) or tuple(W3.shape) != (16, 64):
        raise ValueError("tdssf MLP kernels are built for the reference shape 448->256->64->16 "
                         f"(got {tuple(W1.shape)}, {tuple(W2.shape)}, {tuple(W3.shape)})")
    return _lib.MlpWeights(*[t.data_ptr() for t in (W1, b1, W2, b2, W3, b3)])


def _flat_zero_grads(params: Sequence[torch.Tensor]) -> list:
    """The six parameter gradients as views of ONE zero-filled fp32 buffer (one fill launch instead of six; the bias
    sums are accumulated, so they start from zero).  Every view starts on a 16-byte boundary: the reference shapes
    (checked by _weights_struct) all have a multiple of four elements."""
    sizes = [t.numel() for t in params]
    if any(n & 3 for n in sizes):
        raise ValueError("parameter sizes must be multiples of 4 elements")
    flat = torch.zeros(sum(sizes), dtype=torch.float32, device=params[0].device)
    return [piece.view(t.shape) for piece, t in zip(flat.split_with_sizes(sizes), params)]


class _MlpF32(torch.autograd.Function):
    @staticmethod
    @_on_device
    def forward(ctx, x_pt, x_img, W1, b1, W2, b2, W3, b3):
        _need_cuda(x_pt, x_img, W1)
        shape = x_pt.shape[:-1]
        xp = _f32c(x_pt).reshape(-1, 64)
        xi = _f32c(x_img).reshape(-1, 384)
        params = [_f32c(t.detach()) for t in (W1, b1, W2, b2, W3, b3)]
        N = xp.shape[0]
        dev = xp.device
        h1 = torch.empty((N, 256), dtype=torch.float32, device=dev)
        h2 = torch.empty((N, 64), dtype=torch.float32, device=dev)
        logits = torch.empty((N, 16), dtype=torch.float32, device=dev)
        ws = _weights_struct(params)
        check(lib.tdssf_mlp_fwd_f32(_ptr(xp), _ptr(xi), N, C.byref(ws), _ptr(h1), _ptr(h2), _ptr(logits), _stream(xp)),
              "tdssf_mlp_fwd_f32")
        ctx.save_for_backward(xp, xi, h1, h2, *params)
        ctx.in_shape = tuple(x_pt.shape)
        return logits.reshape(*shape, 16)

    @staticmethod
    @_on_device
    def backward(ctx, dlogits):
        xp, xi, h1, h2, W1, b1, W2, b2, W3, b3 = ctx.saved_tensors
        N = xp.shape[0]
        dev = xp.device
        dl = _f32c(dlogits).reshape(-1, 16)
        grads = _flat_zero_grads((W1, b1, W2, b2, W3, b3))
        dx_pt = torch.empty((N, 64), dtype=torch.float32, device=dev) if ctx.needs_input_grad[0] else None
        ws_bytes = lib.tdssf_mlp_bwd_workspace_bytes(N)
        work = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        w = _weights_struct([W1, b1, W2, b2, W3, b3])
        g = _lib.MlpGrads(*[t.data_ptr() for t in grads])
        check(lib.tdssf_mlp_bwd_f32(_ptr(xp), _ptr(xi), _ptr(h1), _ptr(h2), _ptr(dl), N, C.byref(w), C.byref(g),
                                    _ptr(dx_pt), _ptr(work), ws_bytes, _stream(xp)), "tdssf_mlp_bwd_f32")
        return (dx_pt.reshape(ctx.in_shape) if dx_pt is not None else None, None, *grads)


def mlp_f32(x_pt, x_img, W1, b1, W2, b2, W3, b3):
    """fp32 parity mode of the fusion MLP, differentiable w.r.t. x_pt and the six parameters."""
    return _MlpF32.apply(x_pt, x_img, W1, b1, W2, b2, W3, b3)


@_on_device
def pack_weights(W1, b1, W2, b2, W3, b3) -> torch.Tensor:
    """bf16 tensor-core operand image of the MLP weights (both CTA ranks + biases)."""
    _need_cuda(W1)
    params = [t if (not t.requires_grad and t.dtype == torch.float32 and t.is_contiguous()) else _f32c(t.detach())
              for t in (W1, b1, W2, b2, W3, b3)]
    packed = torch.empty(_PACK_BYTES, dtype=torch.uint8, device=W1.device)
    ws = _weights_struct(params)
    check(lib.tdssf_mlp_pack_weights(C.byref(ws), _ptr(packed), _stream(W1)), "tdssf_mlp_pack_weights")
    return packed


@_on_device
def mlp_tc(x_pt, x_img, packed) -> torch.Tensor:
    """tcgen05 MLP on pre-gathered fp32 features: (N,64),(N,384) -> (N,16)."""
    _need_cuda(x_pt, x_img, packed)
    shape = x_pt.shape[:-1]
    xp = _f32c(x_pt).reshape(-1, 64)
    xi = _f32c(x_img).reshape(-1, 384)
    N = xp.shape[0]
    logits = torch.empty((N, 16), dtype=torch.float32, device=xp.device)
    check(lib.tdssf_mlp_fwd_tc(_ptr(xp), _ptr(xi), N, _ptr(packed), _ptr(logits), None, None, _stream(xp)),
          "tdssf_mlp_fwd_tc")
    return logits.reshape(*shape, 16)


class _MlpTC(torch.autograd.Function):
    """Training through the tensor cores: tcgen05 forward that saves its bf16 activations, bf16 backward
    (tdssf_mlp_bwd_tc).  Tolerance 1e-2 (north_star's bf16 bound); `mlp_f32` is the 1e-5 parity mode."""

    @staticmethod
    @_on_device
    def forward(ctx, x_pt, x_img, W1, b1, W2, b2, W3, b3):
        _need_cuda(x_pt, x_img, W1)
        shape = x_pt.shape[:-1]
        xp = _f32c(x_pt).reshape(-1, 64)
        xi = _f32c(x_img).reshape(-1, 384)
        params = [_f32c(t.detach()) for t in (W1, b1, W2, b2, W3, b3)]
        _weights_struct(params)  # shape check
        N = xp.shape[0]
        dev = xp.device
        packed = pack_weights(*params)
        h1 = torch.empty((N, 256), dtype=torch.bfloat16, device=dev)
        h2 = torch.empty((N, 64), dtype=torch.bfloat16, device=dev)
        logits = torch.empty((N, 16), dtype=torch.float32, device=dev)
        check(lib.tdssf_mlp_fwd_tc(_ptr(xp), _ptr(xi), N, _ptr(packed), _ptr(logits), _ptr(h1), _ptr(h2), _stream(xp)),
              "tdssf_mlp_fwd_tc")
        ctx.save_for_backward(xp, xi, h1, h2, *params)
        ctx.in_shape = tuple(x_pt.shape)
        return logits.reshape(*shape, 16)

    @staticmethod
    @_on_device
    def backward(ctx, dlogits):
        xp, xi, h1, h2, W1, b1, W2, b2, W3, b3 = ctx.saved_tensors
        N = xp.shape[0]
        dev = xp.device
        dl = _f32c(dlogits).reshape(-1, 16)
        grads = _flat_zero_grads((W1, b1, W2, b2, W3, b3))
        dx_pt = torch.empty((N, 64), dtype=torch.float32, device=dev) if ctx.needs_input_grad[0] else None
        ws_bytes = lib.tdssf_mlp_bwd_tc_workspace_bytes(N)
        work = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        w = _weights_struct([W1, b1, W2, b2, W3, b3])
        g = _lib.MlpGrads(*[t.data_ptr() for t in grads])
        check(lib.tdssf_mlp_bwd_tc(_ptr(xp), _ptr(xi), None, _ptr(h1), _ptr(h2), _ptr(dl), N, C.byref(w), C.byref(g),
                                   _ptr(dx_pt), _ptr(work), ws_bytes, _stream(xp)), "tdssf_mlp_bwd_tc")
        return (dx_pt.reshape(ctx.in_shape) if dx_pt is not None else None, None, *grads)


def mlp_tc_train(x_pt, x_img, W1, b1, W2, b2, W3, b3):
    """bf16 tensor-core MLP, differentiable w.r.t. x_pt and the six parameters (1e-2 tolerance)."""
    return _MlpTC.apply(x_pt, x_img, W1, b1, W2, b2, W3, b3)


_fusion_ws = {}  # (device index, stream) -> workspace tensor: scratch of the launches queued on that stream, grown on demand


def _fusion_workspace(device, B, V, tok_code):
    need = lib.tdssf_fusion_workspace_bytes(B, V, tok_code)
    key = (device.index, torch._C._cuda_getCurrentRawStream(device.index))
    ws = _fusion_ws.get(key)
    if ws is None or ws.numel() < need:
        ws = torch.empty(need, dtype=torch.uint8, device=device)
        _fusion_ws[key] = ws
    return ws


@_on_device
def fusion_fwd(points, K, Rt, img_hw, encoder, tokens, x_pt, packed, want_rec=False, out=None, preds_out=None,
               save=None):
    """Fully fused inference: points (B,V,>=3), tokens (B,NC,M,384), x_pt (B,V,64) -> logits (B,V,16).
    preds_out: optional CUDA int64 (B,V) that receives argmax(logits)+1 (class ids 1..16) from the same launch.
    save: optional (x, h1, h2) contiguous CUDA bf16 tensors of B*V rows x 448 / 256 / 64 that receive the MLP
    input rows [x_pt | masked camera mean] and the two hidden activations as the tensor cores saw them (for the
    backward pass, `fusion_train`)."""
    _need_cuda(points, K, Rt, img_hw, tokens, x_pt, packed)
    pts, ss, ps = _points_layout(points)
    B, V, _ = pts.shape
    if tokens.dtype not in _TOK_DTYPES:
        tokens = tokens.float()
    tokens = tokens.contiguous()
    _, NC, M, Ci = tokens.shape
    xp = _f32c(x_pt)
    logits = out if out is not None else torch.empty((B, V, 16), dtype=torch.float32, device=pts.device)
    rec = torch.empty((B, V, 4), dtype=torch.int32, device=pts.device) if want_rec else None
    if preds_out is not None and (preds_out.dtype != torch.int64 or not preds_out.is_cuda or preds_out.numel() != B * V
                                  or not preds_out.is_contiguous()):
        raise ValueError("preds_out must be a contiguous CUDA int64 tensor with B*V elements")
    sx, sh1, sh2 = save if save is not None else (None, None, None)
    for t, c in ((sx, 448), (sh1, 256), (sh2, 64)):
        if t is not None and (t.dtype != torch.bfloat16 or not t.is_cuda or not t.is_contiguous()
                              or t.numel() != B * V * c):
            raise ValueError(f"save tensors must be contiguous CUDA bf16 with B*V x {c} elements")
    ws = _fusion_workspace(pts.device, B, V, _TOK_DTYPES[tokens.dtype])
    check(lib.tdssf_fusion_fwd(_ptr(pts), ss, ps, _ptr(_f32c(K)), _ptr(_f32c(Rt)), _ptr(_f32c(img_hw)), B, V, NC,
                               int(encoder.resize_height), int(encoder.resize_width), int(encoder.patch_size),
                               _ptr(tokens), _TOK_DTYPES[tokens.dtype], M, Ci, _ptr(xp), _ptr(packed), _ptr(logits),
                               _ptr(rec), _ptr(preds_out), _ptr(sx), _ptr(sh1), _ptr(sh2), _ptr(ws), ws.numel(),
                               _stream(pts)), "tdssf_fusion_fwd")
    return (logits, rec) if want_rec else logits


class _FusedTrain(torch.autograd.Function):
    """Training through the fused launch: projection, token gather, camera mean and the three layers in one
    kernel that also writes the bf16 MLP input rows and hidden activations; backward = tdssf_mlp_bwd_tc on those
    (no fp32 image feature, no concat pass).  Differentiable w.r.t. x_pt and the six parameters; the tokens come
    from the frozen encoder (reference utils/fusion_model.py:68, torch.no_grad) and get no gradient."""

    @staticmethod
    @_on_device
    def forward(ctx, x_pt, W1, b1, W2, b2, W3, b3, points, K, Rt, img_hw, encoder, tokens):
        params = [_f32c(t.detach()) for t in (W1, b1, W2, b2, W3, b3)]
        packed = pack_weights(*params)   # (checks the shapes)
        B, V = x_pt.shape[0], x_pt.shape[1]
        N = B * V
        dev = x_pt.device
        xb = torch.empty((N, 448), dtype=torch.bfloat16, device=dev)
        h1 = torch.empty((N, 256), dtype=torch.bfloat16, device=dev)
        h2 = torch.empty((N, 64), dtype=torch.bfloat16, device=dev)
        logits = fusion_fwd(points, K, Rt, img_hw, encoder, tokens, x_pt.detach(), packed, save=(xb, h1, h2))
        ctx.save_for_backward(xb, h1, h2, *params)
        ctx.in_shape = tuple(x_pt.shape)
        return logits

    @staticmethod
    @_on_device
    def backward(ctx, dlogits):
        xb, h1, h2, W1, b1, W2, b2, W3, b3 = ctx.saved_tensors
        N = xb.shape[0]
        dev = xb.device
        dl = _f32c(dlogits).reshape(-1, 16)
        grads = _flat_zero_grads((W1, b1, W2, b2, W3, b3))
        dx_pt = torch.empty((N, 64), dtype=torch.float32, device=dev) if ctx.needs_input_grad[0] else None
        ws_bytes = lib.tdssf_mlp_bwd_tc_workspace_bytes(N)
        work = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        w = _weights_struct([W1, b1, W2, b2, W3, b3])
        g = _lib.MlpGrads(*[t.data_ptr() for t in grads])
        check(lib.tdssf_mlp_bwd_tc(None, None, _ptr(xb), _ptr(h1), _ptr(h2), _ptr(dl), N, C.byref(w), C.byref(g),
                                   _ptr(dx_pt), _ptr(work), ws_bytes, _stream(xb)), "tdssf_mlp_bwd_tc")
        return (dx_pt.reshape(ctx.in_shape) if dx_pt is not None else None, *grads, None, None, None, None, None, None)


def fusion_train(points, K, Rt, img_hw, encoder, tokens, x_pt, W1, b1, W2, b2, W3, b3):
    """Fused forward with a bf16 tensor-core backward: points (B,V,>=3), tokens (B,NC,M,384), x_pt (B,V,64) ->
    logits (B,V,16), differentiable w.r.t. x_pt and the six MLP parameters (1e-2 tolerance)."""
    return _FusedTrain.apply(x_pt, W1, b1, W2, b2, W3, b3, points, K, Rt, img_hw, encoder, tokens)


# ------------------------------------------------------------------------------------------------
# f1
# ------------------------------------------------------------------------------------------------
_voxel_ws: dict = {}


@_on_device
def voxelize(points: torch.Tensor, voxel_size: float, knn_cell: Optional[float] = None):
    """points (B,P,3|4) CUDA fp32 -> voxel_raw (B,P,4) fp32 [centroid xyz, 5-NN weighted intensity] with the rows past
    each cloud's voxel count zeroed, voxel_coords (B,P,3) int32, n_voxels (B,) int32 (utils/lidar.py:83-118 for B
    clouds in one call; voxels in the order of their first point).  Nothing is read back to the host."""
    _need_cuda(points)
    pts = _f32c(points)
    if pts.dim() != 3 or pts.shape[-1] < 3:
        raise ValueError("points must have shape (B, P, >=3)")
    B, P, st = pts.shape
    dev = pts.device
    raw = torch.empty((B, P, 4), dtype=torch.float32, device=dev)
    coords = torch.empty((B, P, 3), dtype=torch.int32, device=dev)
    n_vox = torch.zeros((B,), dtype=torch.int32, device=dev)
    status = torch.zeros((1,), dtype=torch.int32, device=dev)
    need = lib.tdssf_voxelize_workspace_bytes(B, P)
    key = (dev.index, torch._C._cuda_getCurrentRawStream(dev.index))
    ws = _voxel_ws.get(key)
    if ws is None or ws.numel() < need:
        ws = torch.empty(need, dtype=torch.uint8, device=dev)
        _voxel_ws[key] = ws
    check(lib.tdssf_voxelize(_ptr(pts), st, B, P, int(st >= 4), float(voxel_size), float(knn_cell or 0.0), _ptr(raw),
                             _ptr(coords), _ptr(n_vox), _ptr(status), _ptr(ws), ws.numel(), _stream(pts)), "tdssf_voxelize")
    return raw, coords, n_vox, status


# ------------------------------------------------------------------------------------------------
# a7
# ------------------------------------------------------------------------------------------------
_devox_ws: dict = {}


@_on_device
def devox_nn(pts: torch.Tensor, vox: torch.Tensor, cell: Optional[float] = 0.25) -> torch.Tensor:
    """pts (B,P,>=3), vox (B,V,>=3) CUDA fp32 -> (B,P) int64 nearest-voxel indices (first index on ties).
    `cell`: edge of the search grid's cells in coordinate units (any positive value gives the same indices);
    None = brute force over all P x V pairs."""
    _need_cuda(pts, vox)
    pts = _f32c(pts)
    vox = _f32c(vox)
    B, P, sp = pts.shape
    V, sv = vox.shape[1], vox.shape[2]
    idx = torch.empty((B, P), dtype=torch.int64, device=pts.device)
    ws, ws_bytes = None, 0
    if cell is not None:
        ws_bytes = lib.tdssf_devox_nn_workspace_bytes(B, P, V)
        key = (pts.device.index, torch._C._cuda_getCurrentRawStream(pts.device.index))
        ws = _devox_ws.get(key)
        if ws is None or ws.numel() < ws_bytes:
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=pts.device)
            _devox_ws[key] = ws
    check(lib.tdssf_devox_nn(_ptr(pts), sp, _ptr(vox), sv, B, P, V, _ptr(idx), float(cell or 0.0), _ptr(ws),
                             ws.numel() if ws is not None else 0, _stream(pts)), "tdssf_devox_nn")
    return idx


class _DevoxGather(torch.autograd.Function):
    @staticmethod
    @_on_device
    def forward(ctx, scores, idx):
        _need_cuda(scores, idx)
        s = _f32c(scores)
        B, V, Cn = s.shape
        P = idx.shape[1]
        out = torch.empty((B, P, Cn), dtype=torch.float32, device=s.device)
        check(lib.tdssf_devox_gather(_ptr(s), _ptr(idx), B, P, V, Cn, _ptr(out), _stream(s)), "tdssf_devox_gather")
        ctx.save_for_backward(idx)
        ctx.shape = (B, P, V, Cn)
        return out

    @staticmethod
    @_on_device
    def backward(ctx, dout):
        (idx,) = ctx.saved_tensors
        B, P, V, Cn = ctx.shape
        d = _f32c(dout)
        ds = torch.zeros((B, V, Cn), dtype=torch.float32, device=d.device)
        check(lib.tdssf_devox_scatter_add(_ptr(d), _ptr(idx), B, P, V, Cn, _ptr(ds), _stream(d)),
              "tdssf_devox_scatter_add")
        return ds, None


def devox_gather(scores: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    return _DevoxGather.apply(scores, idx)


# ------------------------------------------------------------------------------------------------
# a8-a10
# ------------------------------------------------------------------------------------------------
_pinned = {}


def _pinned_int(dev) -> torch.Tensor:
    """One pinned int32 per device: the landing place of the valid-row count (read right after its event)."""
    t = _pinned.get(dev.index)
    if t is None:
        t = torch.zeros(1, dtype=torch.int32).pin_memory()
        _pinned[dev.index] = t
    return t


class _Cels(torch.autograd.Function):
    @staticmethod
    @_on_device
    def forward(ctx, logits, labels, mask, class_w, lam, ignore_index):
        _need_cuda(logits, labels)
        x = _f32c(logits).reshape(-1, 16)
        n_rows = x.shape[0]
        dev = x.device
        lab = labels.reshape(-1).to(torch.int64).contiguous()
        m = None if mask is None else mask.reshape(-1).to(torch.uint8).contiguous()
        w = _f32c(class_w.to(dev))
        scal = torch.empty(4, dtype=torch.float32, device=dev)
        preds = torch.empty(max(n_rows, 1), dtype=torch.int64, device=dev)
        gt = torch.empty(max(n_rows, 1), dtype=torch.int64, device=dev)
        need_grad = ctx.needs_input_grad[0]
        dlog = torch.empty_like(x) if need_grad else None
        ws_bytes = lib.tdssf_cels_workspace_bytes(n_rows)
        work = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        if os.environ.get("TDSSF_CELS_SYNC") == "1":
            # one call with a stream synchronisation in the middle (where the reference's boolean indexing synchronises)
            n_valid = C.c_int64(0)
            check(lib.tdssf_cels_fwd_bwd(_ptr(x), _ptr(lab), _ptr(m), n_rows, _ptr(w), float(lam), int(ignore_index),
                                         _ptr(scal), C.byref(n_valid), _ptr(preds), _ptr(gt), _ptr(dlog), _ptr(work),
                                         ws_bytes, _stream(x)), "tdssf_cels_fwd_bwd")
            n = int(n_valid.value)
        else:
            # two stages: the count (two small kernels) with its host copy started at once, then everything else as one
            # launch sequence (a CUDA graph when the arguments repeat) sized for n_rows; the host then waits for the COUNT only,
            # i.e. it returns to the caller while the loss kernels are still running and the backward pass can be queued behind them
            n_dev = torch.empty(1, dtype=torch.int32, device=dev)
            check(lib.tdssf_cels_count(_ptr(lab), _ptr(m), n_rows, int(ignore_index), _ptr(n_dev), _ptr(work), ws_bytes, _stream(x)),
                  "tdssf_cels_count")
            n_host = _pinned_int(dev)
            n_host.copy_(n_dev, non_blocking=True)
            counted = torch.cuda.Event()
            counted.record(torch.cuda.current_stream(dev))
            check(lib.tdssf_cels_fwd_bwd_counted(_ptr(x), _ptr(lab), _ptr(m), n_rows, _ptr(w), float(lam), int(ignore_index),
                                                 _ptr(scal), _ptr(preds), _ptr(gt), _ptr(dlog), _ptr(work), ws_bytes, _stream(x)),
                  "tdssf_cels_fwd_bwd_counted")
            counted.synchronize()
            n = int(n_host.item())
        ctx.dlog = dlog
        # kept for a backward through ce / lovasz themselves (tdssf_cels_backward reads the workspace)
        ctx.terms = (work, w, n_rows, n, float(lam)) if need_grad else None
        ctx.in_shape = tuple(logits.shape)
        ctx.mark_non_differentiable(preds, gt)
        ctx.set_materialize_grads(False)     # outputs nobody differentiates arrive as None, not as zero tensors
        total, ce, lov, _ = scal.unbind(0)   # `scal` is this call's own buffer: no copies needed
        return total, ce, lov, preds[:n], gt[:n]

    @staticmethod
    @_on_device
    def backward(ctx, g_total, g_ce, g_lov, _gp, _gg):
        d = ctx.dlog
        if d is None or (g_total is None and g_ce is None and g_lov is None):
            return None, None, None, None, None, None
        if g_ce is None and g_lov is None:
            # the loops call backward on total_loss (utils/train.py:91): the forward call already produced d total
            return (d * g_total).reshape(ctx.in_shape), None, None, None, None, None
        # ce_loss / lovasz_loss are differentiable outputs of the reference too (utils/losses.py:60-70):
        # d = (g_total + g_ce) * d ce + (g_total * lambda + g_lov) * d lovasz, from the kept workspace
        work, w, n_rows, n, lam = ctx.terms
        zero = torch.zeros((), dtype=torch.float32, device=d.device)
        gt_, gc_, gl_ = (zero if g is None else g.to(torch.float32) for g in (g_total, g_ce, g_lov))
        coef = torch.stack([gt_ + gc_, gt_ * lam + gl_]).contiguous()
        out = torch.empty_like(d)
        check(lib.tdssf_cels_backward(_ptr(work), work.numel(), n_rows, n, _ptr(w), _ptr(coef), _ptr(out), _stream(out)),
              "tdssf_cels_backward")
        return out.reshape(ctx.in_shape), None, None, None, None, None


def cels(logits, labels, mask, class_w, lam=1.0, ignore_index=0):
    """(total, ce, lovasz, predictions, gt_valid); `total.backward()` reaches `logits`."""
    return _Cels.apply(logits, labels, mask, class_w, lam, ignore_index)


# ------------------------------------------------------------------------------------------------
# a11
# ------------------------------------------------------------------------------------------------
@_on_device
def confusion(preds: torch.Tensor, labels: torch.Tensor, num_classes: int, out: Optional[torch.Tensor] = None):
    """int64 (C,C) confusion matrix, rows = truth; accumulates into `out` when given."""
    _need_cuda(preds, labels)
    p = preds.reshape(-1).to(torch.int64).contiguous()
    l = labels.reshape(-1).to(torch.int64).contiguous()
    if p.numel() != l.numel():
        raise ValueError("preds and labels must have the same number of elements")
    cm = out if out is not None else torch.zeros((num_classes, num_classes), dtype=torch.int64, device=p.device)
    check(lib.tdssf_confusion(_ptr(p), _ptr(l), p.numel(), num_classes, _ptr(cm), _stream(p)), "tdssf_confusion")
    return cm


# ------------------------------------------------------------------------------------------------
# f4  calibration composition, multi-sweep accumulation
# ------------------------------------------------------------------------------------------------
def _f64c(t, device) -> torch.Tensor:
    return torch.as_tensor(t, dtype=torch.float64).to(device).contiguous()


def compose_lidar2cam(lidar_sensor, cam_sensor, lidar_ego=None, cam_ego=None, device="cuda", want_f64=False):
    """Poses (qw,qx,qy,qz,tx,ty,tz): lidar_sensor (B,7), cam_sensor (B,NC,7) [, lidar_ego (B,7), cam_ego (B,NC,7)] ->
    lidar2cam (B,NC,4,4) fp32 on the device (the `lidar2cam_extrinsics` of utils/dataloader.py:150,226)."""
    ls, cs = _f64c(lidar_sensor, device), _f64c(cam_sensor, device)
    if ls.dim() != 2 or ls.shape[1] != 7 or cs.dim() != 3 or cs.shape[2] != 7 or cs.shape[0] != ls.shape[0]:
        raise ValueError(f"compose_lidar2cam: expected (B,7) and (B,NC,7) poses, got {tuple(ls.shape)} and {tuple(cs.shape)}")
    if (lidar_ego is None) != (cam_ego is None):
        raise ValueError("compose_lidar2cam: lidar_ego and cam_ego go together")
    le = ce = None
    if lidar_ego is not None:
        le, ce = _f64c(lidar_ego, device), _f64c(cam_ego, device)
        if le.shape != ls.shape or ce.shape != cs.shape:
            raise ValueError("compose_lidar2cam: ego poses must have the shapes of the sensor poses")
    _need_cuda(ls)
    B, NC = cs.shape[0], cs.shape[1]
    out32 = torch.empty((B, NC, 4, 4), dtype=torch.float32, device=ls.device)
    out64 = torch.empty((B, NC, 4, 4), dtype=torch.float64, device=ls.device) if want_f64 else None
    with torch.cuda.device(ls.device):
        check(lib.tdssf_compose_lidar2cam(_ptr(ls), _ptr(cs), _ptr(le), _ptr(ce), B, NC, _ptr(out32), _ptr(out64),
                                          _stream(ls)), "tdssf_compose_lidar2cam")
    return (out32, out64) if want_f64 else out32


@_on_device
def accumulate_sweeps(points, counts, sweep_sensor, sweep_ego, ref_sensor, ref_ego):
    """points (S,P,4) fp32 CUDA (padded sweeps), counts (S,) valid rows; poses as in compose_lidar2cam ->
    (cloud (S*P,4) fp32 with the first `total` rows filled in sweep order, total (1,) int32 on the device,
    T (S,4,4) float64).  No host synchronisation: slice with `cloud[:int(total)]` when the length is needed."""
    _need_cuda(points)
    if points.dim() != 3 or points.shape[2] != 4 or points.dtype != torch.float32:
        raise ValueError(f"accumulate_sweeps: points must be (S,P,4) fp32, got {tuple(points.shape)} {points.dtype}")
    pts = points.contiguous()
    S, P = pts.shape[0], pts.shape[1]
    dev = pts.device
    cnt = torch.as_tensor(counts, dtype=torch.int32).to(dev).contiguous()
    if cnt.shape != (S,):
        raise ValueError("accumulate_sweeps: counts must have one entry per sweep")
    ss, se = _f64c(sweep_sensor, dev), _f64c(sweep_ego, dev)
    rs, re = _f64c(ref_sensor, dev).reshape(-1), _f64c(ref_ego, dev).reshape(-1)
    if ss.shape != (S, 7) or se.shape != (S, 7) or rs.shape != (7,) or re.shape != (7,):
        raise ValueError("accumulate_sweeps: poses must be (S,7), (S,7), (7,), (7,)")
    T = torch.empty((S, 4, 4), dtype=torch.float64, device=dev)
    check(lib.tdssf_compose_sweep2ref(_ptr(ss), _ptr(se), _ptr(rs), _ptr(re), S, _ptr(T), _stream(pts)), "tdssf_compose_sweep2ref")
    cloud = torch.empty((S * P, 4), dtype=torch.float32, device=dev)
    total = torch.zeros(1, dtype=torch.int32, device=dev)
    check(lib.tdssf_accumulate_sweeps(_ptr(pts), 4 * P, _ptr(cnt), _ptr(T), S, P, _ptr(cloud), _ptr(total), _stream(pts)),
          "tdssf_accumulate_sweeps")
    return cloud, total, T
