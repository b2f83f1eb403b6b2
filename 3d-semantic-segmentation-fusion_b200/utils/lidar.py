"""Drop-in for the reference's utils/lidar.py (LiDARFeatureEncoder, utils/lidar.py:14-172).

In scope here: the padding / caching contract of `forward` and `devoxelize` (nearest voxel per raw
point, gather, differentiable through the gathered scores), which runs on libtdssf.so
(csrc/devox.cu), and the voxelizer (csrc/voxel.cu: the Open3D down-sampling + scikit-learn 5-NN of
utils/lidar.py:83-118 without leaving the device, all B clouds of a batch in one call and one host
read).  There is no host voxelizer in this module; a caller who wants another one passes a callable
`voxelizer(points, voxel_size) -> (voxel_raw, voxel_coords)`.
The PTv3 backbone stays stock third-party code exactly as in the reference: it is imported lazily,
and a caller may inject a replacement (`backbone=`) — the synthetic harness does, because spconv
does not exist in this image.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from tdssf import ops


class LiDARFeatureEncoder(nn.Module):
    """LiDAR -> voxelize -> PTv3 -> per-voxel feature (V x 64); devoxelize maps voxel scores to points."""

    def __init__(self, config, backbone: nn.Module | None = None, voxelizer=None):
        super().__init__()
        self.voxel_size = config['dataset_params']['lidar']['voxel_size']
        if backbone is None:
            try:  # the reference's stock backbone (utils/lidar.py:12,25)
                from PointTransformerV3.model import PointTransformerV3
            except Exception as exc:  # spconv / torch_scatter / flash_attn / addict / timm
                raise ImportError("PointTransformerV3 (and its spconv/torch_scatter dependencies) is required for the "
                                  "stock backbone; pass backbone= to use another module") from exc
            backbone = PointTransformerV3(in_channels=4)
        self.ptv3 = backbone
        self._voxelizer = voxelizer

    def forward(self, lidar_points):
        """lidar_points (B,P,4) -> voxel_features (B,V_max,64), voxel_raw (B,V_max,4), voxel_coords (B,V_max,3) int32,
        voxel_mask (B,V_max) bool.  Caches `lidar_points_raw` and `voxel_raw` for devoxelize (:37,79)."""
        self.lidar_points_raw = lidar_points
        dev = lidar_points.device
        B = lidar_points.shape[0]
        if callable(self._voxelizer):
            per_cloud = [self._voxelizer(cloud, self.voxel_size) for cloud in lidar_points]
        else:
            per_cloud = self._voxelize_batch(lidar_points)
        feats = []
        for voxel_raw, _ in per_cloud:
            sample = {"coord": voxel_raw[:, :3], "feat": voxel_raw,
                      "grid_size": torch.tensor(self.voxel_size, device=dev),
                      "batch": torch.zeros(voxel_raw.shape[0], dtype=torch.long, device=dev)}
            out = self.ptv3(sample)
            feats.append(out.feat if hasattr(out, "feat") else out)
        v_max = max(f.shape[0] for f in feats)
        voxel_features = torch.zeros((B, v_max, feats[0].shape[1]), device=dev)
        voxel_raw_t = torch.zeros((B, v_max, 4), device=dev)
        voxel_coords_t = torch.zeros((B, v_max, 3), device=dev, dtype=torch.int)
        voxel_mask = torch.zeros((B, v_max), device=dev, dtype=torch.bool)
        for i, (f, (r, c)) in enumerate(zip(feats, per_cloud)):
            n = f.shape[0]
            voxel_features[i, :n] = f
            voxel_raw_t[i, :n, :r.shape[1]] = r
            voxel_coords_t[i, :n] = c
            voxel_mask[i, :n] = True
        self.voxel_raw = voxel_raw_t  # zero padding included: devoxelize searches all V_max rows
        return voxel_features, voxel_raw_t, voxel_coords_t, voxel_mask

    def _voxelize_batch(self, lidar_points):
        """All B clouds through one tdssf_voxelize call; one host read (B voxel counts + the status word) where the
        reference makes a device -> host -> device round trip per cloud (utils/lidar.py:87,116-117)."""
        B, P, ch = lidar_points.shape
        if ch == 4 and P < 5:
            raise ValueError("Expected n_neighbors <= n_samples_fit")   # what scikit-learn raises at :99
        raw, coords, n_vox, status = ops.voxelize(lidar_points.detach(), self.voxel_size)
        host = torch.cat([n_vox, status]).tolist()
        if host[-1]:
            raise RuntimeError("voxelize: non-finite coordinates or more than 2^21 voxels along an axis")
        return [(raw[b, :host[b], :ch], coords[b, :host[b]]) for b in range(B)]

    def voxelize_open3d(self, lidar_points):
        """(P,4) -> voxel_raw (V,4) [centroid xyz, 5-NN weighted intensity], voxel_coords (V,3) int32
        (utils/lidar.py:83-118), on the device (csrc/voxel.cu), voxels in the order of their first point."""
        if callable(self._voxelizer):
            return self._voxelizer(lidar_points, self.voxel_size)
        return self._voxelize_batch(lidar_points[None])[0]

    def devoxelize(self, voxel_scores):
        """voxel_scores (B,V_max,C) -> (point_scores (B,P,C), voxel_indices (B,P) int64); gradients flow to
        voxel_scores through the gather (backward = scatter-add)."""
        pts = self.lidar_points_raw
        vox = self.voxel_raw
        if pts.dim() != 3 or vox.shape[0] != voxel_scores.shape[0]:
            raise RuntimeError("devoxelize must follow forward() on the same batch (it reads the cached clouds)")
        # search-grid cells of about two voxel edges: the nearest centre of an occupied neighbourhood is then within
        # 0.99 cell and the point never falls back to the full scan (any positive value gives the same indices)
        idx = ops.devox_nn(pts.to(voxel_scores.device), vox.to(voxel_scores.device),
                           cell=max(0.25, 2.0 * float(self.voxel_size)))
        return ops.devox_gather(voxel_scores, idx), idx
