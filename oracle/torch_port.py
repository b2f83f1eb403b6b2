"""CPU port of the reference's fusion forward in PyTorch ops — TEST / BASELINE INFRASTRUCTURE, NOT PRODUCT.

Purpose: the timed CPU baseline (`bench.py` cpu_baseline leg and `--impl reference`).  The reference's
path *is* a sequence of ATen ops (utils/fusion_model.py:14-204), so its CPU cost is only represented
fairly by the same op sequence run by the same multithreaded ATen kernels; the numpy oracle
(np_oracle.py) is the checker, this file is the stopwatch's subject.  It follows the reference op by
op — per-camera matmuls, clamp, divide, 5-way mask, where; broadcast rescale; `.long()` patch index;
int64 `expand` + `torch.gather` per camera; stack, mask multiply, sum, clamp, divide; cat; three
Linear layers — including the temporaries that make the reference slow.
tests/test_oracle_golden.py pins its outputs to the golden vectors produced by the reference itself.
Only tests/, __graft_entry__.smoke() and bench.py may import this module.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def project_cameras(xyz, K_list, Rt_list, image_sizes):
    """utils/fusion_model.py:14-86.  xyz (B,V,3); returns (pix (B,NC,V,2), depth (B,NC,V), valid (B,NC,V))."""
    dev, dt = xyz.device, xyz.dtype
    B, V, _ = xyz.shape
    n_cam = K_list[0].shape[0]

    def stacked(seq, tail):
        t = torch.stack([torch.as_tensor(s, dtype=torch.float32, device=dev) for s in seq], dim=0)
        if t.shape[1:] != tail:
            raise ValueError(f"Expected shape (*, {tail}) but got {t.shape}.")
        return t

    K = stacked(K_list, (n_cam, 3, 3))
    Rt = stacked(Rt_list, (n_cam, 4, 4))
    hw = torch.as_tensor(image_sizes, dtype=dt, device=dev)
    if hw.shape != (B, 2):
        raise ValueError("image_sizes must be of shape (B, 2) where each entry is (H, W).")
    h_img, w_img = hw[:, 0].view(B, 1), hw[:, 1].view(B, 1)
    homog = torch.cat([xyz, torch.ones(B, V, 1, device=dev, dtype=torch.float32)], dim=-1)
    pix_all = torch.full((B, n_cam, V, 2), -1.0, dtype=dt, device=dev)
    depth_all = torch.zeros((B, n_cam, V), dtype=dt, device=dev)
    valid_all = torch.zeros((B, n_cam, V), dtype=torch.bool, device=dev)
    for c in range(n_cam):
        cam = torch.matmul(Rt[:, c].float().unsqueeze(1), homog.unsqueeze(-1)).squeeze(-1)[:, :, :3]  # :68-69
        z = cam[:, :, 2]
        img = torch.matmul(K[:, c], cam.transpose(1, 2)).transpose(1, 2)                             # :72
        den = img[:, :, 2].clamp(min=1e-12)
        u, v = img[:, :, 0] / den, img[:, :, 1] / den
        ok = (z > 0) & (u >= 0) & (u < w_img) & (v >= 0) & (v < h_img)                                # :77
        uv = torch.stack([u, v], dim=-1)
        pix_all[:, c] = torch.where(ok.unsqueeze(-1), uv, torch.full_like(uv, -1.0))
        depth_all[:, c] = torch.where(ok, z, torch.full_like(z, -1.0))
        valid_all[:, c] = ok
    return pix_all, depth_all, valid_all


def rescale(pix, origin_hw, new_hw):
    """utils/fusion_model.py:89-127."""
    h_new, w_new = (float(new_hw),) * 2 if isinstance(new_hw, int) else map(float, new_hw)
    hw = torch.as_tensor(origin_hw, dtype=pix.dtype, device=pix.device)
    if hw.dim() == 1:
        hw = hw.unsqueeze(0)
    if hw.shape[-1] != 2:
        raise ValueError("origin_img_sizes must have shape (B, 2) in (H, W) order.")
    scale = torch.stack([w_new / hw[:, 1].clamp(min=1e-6), h_new / hw[:, 0].clamp(min=1e-6)], dim=-1)
    return pix * scale.view(scale.shape[0], *([1] * (pix.dim() - 2)), 2)


def fusion_forward(tokens, voxel_features, voxel_raw, image_sizes, K_list, Rt_list, weights, patch_size, resize_h,
                   resize_w, return_parts=False):
    """utils/fusion_model.py:152-204 (everything before devoxelize) on CPU tensors."""
    B, V, _ = voxel_features.shape
    n_cam, dim = tokens.shape[1], tokens.shape[3]
    pix, _, valid = project_cameras(voxel_raw[:, :, :3], K_list, Rt_list, image_sizes)
    pix = rescale(pix, torch.as_tensor(image_sizes, dtype=pix.dtype), (resize_h, resize_w))
    gh, gw = resize_h // patch_size, resize_w // patch_size
    cell = (pix / float(patch_size)).long()
    cell[..., 0] = cell[..., 0].clamp(0, gw - 1)
    cell[..., 1] = cell[..., 1].clamp(0, gh - 1)
    per_cam = []
    flats = []
    for c in range(n_cam):
        flat = (cell[:, c, :, 1] * gw + cell[:, c, :, 0]).clamp(0, gh * gw - 1)
        index = flat.unsqueeze(-1).expand(-1, -1, dim)                 # int64 (B,V,dim) like :189
        per_cam.append(torch.gather(tokens[:, c], dim=1, index=index))
        flats.append(flat)
    stack = torch.stack(per_cam, dim=1)
    m = valid.unsqueeze(-1).to(stack.dtype)
    img_feat = (stack * m).sum(dim=1) / m.sum(dim=1).clamp(min=1.0)
    x = torch.cat([voxel_features, img_feat], dim=-1)
    h = F.relu(F.linear(x, weights["mlp.0.weight"], weights["mlp.0.bias"]))
    h = F.relu(F.linear(h, weights["mlp.2.weight"], weights["mlp.2.bias"]))
    out = F.linear(h, weights["mlp.4.weight"], weights["mlp.4.bias"])
    if return_parts:
        return out, img_feat, torch.stack(flats, dim=1), valid, pix
    return out
