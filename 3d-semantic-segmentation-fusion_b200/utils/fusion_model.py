"""Drop-in for the reference's utils/fusion_model.py: same public names, argument order, return
values and error behaviour; the internals run in libtdssf.so (hand-written sm_100a kernels).

Reference interface mirrored (file:line in the reference tree):
    multi_camera_projector(lidar_points, cam_intrinsics, lidar2cam_extrinsics, image_sizes)   utils/fusion_model.py:14
    scale_pixel_coords(pixel_coords, origin_img_sizes, new_img_size)                          utils/fusion_model.py:89
    FeatureFusionModel(image_encoder, pcd_encoder, point_feat_dim, patch_tok_dim, output_dim, device)   :133-150
        .forward(patch_tokens, voxel_features, voxel_raw, voxel_coords, image_sizes,
                 cam_intrinsics, lidar2cam_extrinsics) -> point_scores                        :152-207

MLP arithmetic is selected with the environment variable TDSSF_MLP:
    bf16 (default)  the fully fused tcgen05 kernel; when a gradient is required the same launch also saves its bf16
                    input rows and activations and the backward runs in bf16 (1e-2 tolerance, north_star)
    fp32            fp32 CUDA-core kernels everywhere (1e-5 tolerance)
"""
from __future__ import annotations

import os
from typing import Sequence

import torch
import torch.nn as nn

from tdssf import ops


def _device_of(t: torch.Tensor, fallback=None) -> torch.device:
    if t.is_cuda:
        return t.device
    if fallback is not None and torch.device(fallback).type == "cuda":
        return torch.device(fallback)
    raise RuntimeError("tdssf runs on CUDA tensors only (no CPU fallback)")


# -------------------------
# 1. Projector
# -------------------------
def multi_camera_projector(lidar_points, cam_intrinsics, lidar2cam_extrinsics, image_sizes):
    """Project batched LiDAR points into the camera views.

    lidar_points (B,V,3) on a CUDA device; cam_intrinsics / lidar2cam_extrinsics: iterables of length B
    with (N_cam,3,3) / (N_cam,4,4) tensors or arrays on any device; image_sizes (B,2) in (H,W).
    Returns pixel_coords (B,N_cam,V,2) fp32, depth (B,N_cam,V) fp32, valid_mask (B,N_cam,V) bool.
    Raises ValueError on the shape mismatches the reference checks (utils/fusion_model.py:45,53).
    """
    device = _device_of(lidar_points)
    B = lidar_points.shape[0]
    K, Rt = ops.stack_calibration(cam_intrinsics, lidar2cam_extrinsics, device)
    hw = ops.image_sizes_f32(image_sizes, B, device)
    out = ops.project(lidar_points, K, Rt, hw)
    return out["pix"], out["depth"], out["valid"].view(torch.bool)


def scale_pixel_coords(pixel_coords: torch.Tensor, origin_img_sizes, new_img_size) -> torch.Tensor:
    """Scale (u,v) from the original resolution to `new_img_size` ((H_new,W_new) or one int)."""
    if isinstance(new_img_size, int):
        h_new = w_new = float(new_img_size)
    else:
        h_new, w_new = map(float, new_img_size)
    device = _device_of(pixel_coords)
    hw = torch.as_tensor(origin_img_sizes, dtype=torch.float32, device=device)
    if hw.dim() == 1:
        hw = hw.unsqueeze(0)
    if hw.shape[-1] != 2:
        raise ValueError("origin_img_sizes must have shape (B, 2) in (H, W) order.")
    return ops.scale_pixels(pixel_coords, hw, h_new, w_new)


# -------------------------
# 2. Fusion Model
# -------------------------
class FeatureFusionModel(nn.Module):
    def __init__(self, image_encoder=None, pcd_encoder=None, point_feat_dim=64, patch_tok_dim=384, output_dim=16,
                 device="cuda"):
        super().__init__()
        self.device = device
        self.image_encoder = image_encoder
        self.pcd_encoder = pcd_encoder
        # identical module tree => identical parameter names/shapes/initialisation and checkpoint keys
        self.mlp = nn.Sequential(
            nn.Linear(point_feat_dim + patch_tok_dim, 256),
            nn.ReLU(),
            nn.Linear(256, 64),
            nn.ReLU(),
            nn.Linear(64, output_dim),
        )
        self._packed = None
        self._packed_key = None

    # fp32 master weights stay nn.Parameters; the bf16 tensor-core image is a derived cache
    def _mlp_params(self):
        return [self.mlp[0].weight, self.mlp[0].bias, self.mlp[2].weight, self.mlp[2].bias, self.mlp[4].weight,
                self.mlp[4].bias]

    def _packed_weights(self):
        params = self._mlp_params()
        key = tuple((p.data_ptr(), p._version) for p in params)
        if self._packed is None or key != self._packed_key:
            self._packed = ops.pack_weights(*params)
            self._packed_key = key
        return self._packed

    def forward(self, patch_tokens, voxel_features, voxel_raw, voxel_coords, image_sizes, cam_intrinsics,
                lidar2cam_extrinsics):
        B, V, _ = voxel_features.shape                  # (B, V, 64)
        B2, num_cams, M, dim = patch_tokens.shape       # (B, 6, M, 384)
        if os.environ.get("TDSSF_VERBOSE"):
            print("B2:", B2, "num_cams:", num_cams, "M:", M, "dim:", dim)

        device = _device_of(voxel_raw, self.device)
        enc = self.image_encoder
        K, Rt, hw = ops.stack_calibration(cam_intrinsics, lidar2cam_extrinsics, device, image_sizes)
        tokens = patch_tokens.to(device, non_blocking=True)  # the reference moves them per camera (F11)
        voxel_raw = voxel_raw.to(device)
        voxel_features = voxel_features.to(device)

        params = self._mlp_params()
        needs_grad = torch.is_grad_enabled() and (voxel_features.requires_grad or any(p.requires_grad for p in params))
        mode = os.environ.get("TDSSF_MLP", "bf16").lower()
        if mode not in ("bf16", "fp32"):
            raise ValueError(f"TDSSF_MLP must be bf16 or fp32, got {mode!r}")

        fused_ok = mode == "bf16" and dim == 384 and self.mlp[0].in_features == 448
        if fused_ok and not needs_grad:
            voxel_scores = ops.fusion_fwd(voxel_raw, K, Rt, hw, enc, tokens, voxel_features, self._packed_weights())
        elif fused_ok:
            # training: the same fused launch, saving its bf16 input rows and activations for the backward pass
            voxel_scores = ops.fusion_train(voxel_raw, K, Rt, hw, enc, tokens, voxel_features, *params)
        else:
            rec = ops.project(voxel_raw, K, Rt, hw, enc, want_pix=False, want_depth=False, want_valid=False,
                              want_rec=True)["rec"]
            fused_img_feat = ops.gather_mean(tokens, rec)
            voxel_scores = ops.mlp_f32(voxel_features, fused_img_feat, *params)

        point_scores, _ = self.pcd_encoder.devoxelize(voxel_scores)
        return point_scores
