"""Times the grid-accelerated devoxelize search on a voxelised synthetic sweep (for an ncu launch list)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "3d-semantic-segmentation-fusion_b200")]
import torch
from tdssf import ops, synth
gen = torch.Generator().manual_seed(5)
Bd, P = 2, 34720
raw = synth.make_points(Bd, P, gen)[..., :3].contiguous().cuda()
vs = 0.1
cent = [torch.unique(torch.floor(raw[b] / vs), dim=0) * vs + vs / 2 for b in range(Bd)]
Vv = min(c.shape[0] for c in cent)
vox_v = torch.stack([c[:Vv] for c in cent]).contiguous()
for cell in (0.25, 0.5, 1.0):
    for _ in range(3):
        idx = ops.devox_nn(raw, vox_v, cell=cell)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(10):
        idx = ops.devox_nn(raw, vox_v, cell=cell)
    e1.record(); torch.cuda.synchronize()
    d = (raw - torch.gather(vox_v, 1, idx[..., None].expand(-1, -1, 3))).norm(dim=-1)
    print(f"cell {cell}: {e0.elapsed_time(e1) / 10:.4f} ms  V={Vv}  share of points farther than 0.99*cell from their voxel: {(d > 0.99 * cell).float().mean().item():.4f}")
