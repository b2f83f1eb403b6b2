"""Times the device voxelizer (csrc/voxel.cu) on synthetic sweeps: B x 34 720 points, voxel 0.1 / 0.5 m, several
neighbour-grid cells; prints ms per call, voxels per cloud and the share of voxels that needed the full scan."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "3d-semantic-segmentation-fusion_b200")]
import torch
from tdssf import ops, synth

QUICK = "--quick" in sys.argv   # one configuration, few calls: for an ncu launch list
for B, P in (((8, 34720),) if QUICK else ((8, 34720), (1, 350000))):
    pts = synth.make_points(B, P, torch.Generator().manual_seed(5)).cuda()
    for vs in ((0.1,) if QUICK else (0.1, 0.5)):
        for cell in ((None,) if QUICK else (None, 0.25, 1.0, 2.0)):
            for _ in range(1 if QUICK else 3):
                out = ops.voxelize(pts, vs, cell)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); e0.record()
            for _ in range(2 if QUICK else 10):
                out = ops.voxelize(pts, vs, cell)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / (2 if QUICK else 10)
            print(json.dumps({"row": "f1", "kernel": "tdssf_voxelize", "B": B, "P": P, "voxel_size": vs, "knn_cell": cell or max(2 * vs, 0.25),
                              "ms": round(ms, 4), "points_per_s": round(B * P / ms * 1e3), "voxels_per_cloud": int(out[2].float().mean()),
                              "status": int(out[3][0])}))
