"""Synthetic stand-ins for the parts either side of the hot path ("a synthetic dataloaders dict in front of
the unchanged loops"): batches in the layout of fusion_collate_fn (utils/dataloader.py:194-242), an image encoder with
the attribute contract of utils/camera.py:24-34 that returns pre-computed token maps, and a point-cloud encoder with the
forward / devoxelize contract of utils/lidar.py:27,120 whose "voxels" are the points themselves."""
import torch

from tdssf import synth


class TokenLookupEncoder:
    """Stands in for DINO: the 'image' of (batch i, sample b, view v) carries those three numbers and the encoder returns
    the pre-computed token map sweeps[i].tokens[b, v] — on the host, like the reference's encoder (utils/camera.py:59)."""

    def __init__(self, sweeps, device=None):
        self.sweeps = sweeps
        self.device = device
        enc = sweeps[0].encoder
        self.patch_size, self.resize_height, self.resize_width = enc.patch_size, enc.resize_height, enc.resize_width

    def __call__(self, images):
        out = []
        for img in images:
            i, b, v = (int(x) for x in img.flatten()[:3])
            t = self.sweeps[i].tokens[b, v]
            out.append(t.to(self.device) if self.device else t)
        return {"patch_features": torch.stack(out), "global_features": torch.zeros(len(out), 384)}


class IdentityVoxels(torch.nn.Module):
    """forward: every point is its own voxel, feature = a fixed function of the coordinates followed by one trainable
    scale (so that the loops' optimizer sees a pcd_encoder parameter); devoxelize: identity."""

    def __init__(self):
        super().__init__()
        self.scale = torch.nn.Parameter(torch.ones(64))

    def forward(self, pts):
        base = torch.cat([pts, pts.flip(-1)] * 8, -1)[..., :64].contiguous()
        return base * self.scale, pts, None, None

    def devoxelize(self, s):
        return s, None


def make_batches(n_batches, B, V, encoder="dinov2", seed=50):
    sweeps = [synth.make_sweeps(B, V, encoder, seed=seed + i) for i in range(n_batches)]

    def batch(i):
        sw = sweeps[i]
        images = torch.zeros(B, 6, 3, 1, 1)
        for b in range(B):
            for v in range(6):
                images[b, v, :, 0, 0] = torch.tensor([i, b, v], dtype=torch.float32)
        return images, sw.image_sizes, sw.points, sw.labels, sw.mask, sw.K, sw.Rt

    return sweeps, [batch(i) for i in range(n_batches)]
