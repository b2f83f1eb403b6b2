"""Seeded nuScenes-shaped synthetic inputs for the fusion hot path.

Everything the path consumes is generated here with `torch.Generator().manual_seed(seed)` on the
CPU, in the layout `fusion_collate_fn` produces (reference utils/dataloader.py:194-242):
points (B,V,4) fp32, per-sweep lists of K (6,3,3) and lidar2cam (6,4,4) fp32 composed in float64
then cast (utils/dataloader.py:150,226), image_sizes (B,2) int64 (H,W), patch tokens (B,6,M,Ci),
point features (B,V,64), labels in 0..16 with `mask = labels != 0` (utils/dataloader.py:239-240).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np
import torch

CAM_YAW_DEG = (0.0, -55.0, -110.0, 180.0, 110.0, 55.0)  # nuScenes camera order
IMG_H, IMG_W = 900, 1600
# soft-log-compressed class weights used by the reference's config 12 (config/nuscenesTestConfig12.yaml:29-31)
CLASS_WEIGHTS = (2.7324509, 4.7631689, 3.3921255, 2.3901170, 3.4012512, 3.9967013, 3.3400925, 4.3070731,
                 3.2343460, 2.5893233, 1.0000000, 3.3888311, 1.7566124, 1.4314391, 1.3798489, 1.3449555)


@dataclass
class Encoder:
    """Attribute contract of the image encoder the fusion model reads (utils/camera.py:24-34)."""
    patch_size: int
    resize_height: int
    resize_width: int

    @property
    def grid(self):
        return self.resize_height // self.patch_size, self.resize_width // self.patch_size

    @property
    def n_patches(self):
        gh, gw = self.grid
        return gh * gw


ENCODERS = {
    "dinov3": Encoder(16, 640, 1138),        # 40 x 71 = 2 840 patches
    "dinov3_hires": Encoder(16, 896, 1600),  # 56 x 100 = 5 600
    "dinov2": Encoder(14, 224, 224),         # 16 x 16 = 256
}


def make_cameras(B: int, gen: torch.Generator):
    """Six-camera rig: returns lists (len B) of K (6,3,3) and lidar2cam (6,4,4) fp32 tensors."""
    Ks, Rts = [], []
    # lidar (x fwd, y left, z up) -> camera (x right, y down, z fwd)
    base = np.array([[0.0, -1.0, 0.0], [0.0, 0.0, -1.0], [1.0, 0.0, 0.0]])
    for _ in range(B):
        K = np.zeros((6, 3, 3), dtype=np.float64)
        Rt = np.zeros((6, 4, 4), dtype=np.float64)
        jitter = torch.randn(6, 3, generator=gen, dtype=torch.float64).numpy() * 0.05
        for c, yaw in enumerate(CAM_YAW_DEG):
            a = math.radians(yaw)
            rz = np.array([[math.cos(a), math.sin(a), 0.0], [-math.sin(a), math.cos(a), 0.0], [0.0, 0.0, 1.0]])
            R = base @ rz
            t = np.array([0.0, -0.3, -0.5]) + jitter[c]
            Rt[c, :3, :3] = R
            Rt[c, :3, 3] = t
            Rt[c, 3, 3] = 1.0
            f = 809.2 if c == 3 else 1266.4
            K[c] = [[f, 0.0, 816.3], [0.0, f, 491.5], [0.0, 0.0, 1.0]]
        Ks.append(torch.from_numpy(K).float())
        Rts.append(torch.from_numpy(Rt).float())
    return Ks, Rts


def make_points(B: int, V: int, gen: torch.Generator) -> torch.Tensor:
    """32-beam spinning-LiDAR-like cloud, (B,V,4) fp32 [x,y,z,intensity]."""
    az = torch.rand(B, V, generator=gen) * (2 * math.pi)
    el = torch.deg2rad(torch.rand(B, V, generator=gen) * 40.0 - 30.0)
    rng = 2.0 + 78.0 * torch.rand(B, V, generator=gen) ** 2
    x = rng * torch.cos(el) * torch.cos(az)
    y = rng * torch.cos(el) * torch.sin(az)
    z = rng * torch.sin(el)
    inten = torch.rand(B, V, generator=gen)
    return torch.stack([x, y, z, inten], dim=-1).float().contiguous()


def make_labels(B: int, P: int, gen: torch.Generator) -> torch.Tensor:
    """int64 labels in 0..16 with a nuScenes-like skew; 0 = noise (ignored)."""
    p = torch.tensor([0.05, 0.002, 0.001, 0.004, 0.04, 0.002, 0.001, 0.003, 0.001, 0.006, 0.015,
                      0.28, 0.01, 0.06, 0.06, 0.25, 0.215], dtype=torch.float64)
    p = p / p.sum()
    return torch.multinomial(p, B * P, replacement=True, generator=gen).view(B, P).long()


@dataclass
class Sweeps:
    points: torch.Tensor          # (B,V,4) fp32
    point_feat: torch.Tensor      # (B,V,64) fp32
    tokens: torch.Tensor          # (B,6,M,Ci)
    K: list                       # B x (6,3,3) fp32
    Rt: list                      # B x (6,4,4) fp32
    image_sizes: torch.Tensor     # (B,2) int64 (H,W)
    labels: torch.Tensor          # (B,V) int64
    mask: torch.Tensor            # (B,V) bool
    encoder: Encoder


def make_sweeps(B: int, V: int, encoder: str | Encoder = "dinov3", seed: int = 0, Ci: int = 384,
                tok_dtype: torch.dtype = torch.float32, Cp: int = 64) -> Sweeps:
    enc = ENCODERS[encoder] if isinstance(encoder, str) else encoder
    gen = torch.Generator().manual_seed(seed)
    pts = make_points(B, V, gen)
    Ks, Rts = make_cameras(B, gen)
    tokens = torch.randn(B, 6, enc.n_patches, Ci, generator=gen).to(tok_dtype)
    feat = torch.randn(B, V, Cp, generator=gen)
    labels = make_labels(B, V, gen)
    sizes = torch.tensor([[IMG_H, IMG_W]] * B, dtype=torch.int64)
    return Sweeps(pts, feat, tokens, Ks, Rts, sizes, labels, labels != 0, enc)


def make_mlp_weights(seed: int = 0, Cp: int = 64, Ci: int = 384, C: int = 16):
    """nn.Linear default init of the 448->256->64->16 MLP (utils/fusion_model.py:144-150)."""
    gen = torch.Generator().manual_seed(10_000 + seed)

    def linear(out_f, in_f):
        bound = 1.0 / math.sqrt(in_f)
        w = (torch.rand(out_f, in_f, generator=gen) * 2 - 1) * bound
        b = (torch.rand(out_f, generator=gen) * 2 - 1) * bound
        return w.contiguous(), b.contiguous()

    W1, b1 = linear(256, Cp + Ci)
    W2, b2 = linear(64, 256)
    W3, b3 = linear(C, 64)
    return {"mlp.0.weight": W1, "mlp.0.bias": b1, "mlp.2.weight": W2, "mlp.2.bias": b2,
            "mlp.4.weight": W3, "mlp.4.bias": b3}
