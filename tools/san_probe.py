"""Small end-to-end pass for compute-sanitizer (memcheck / racecheck / synccheck): fused inference, one training step
(fused forward with saves -> CE + Lovasz -> tcgen05 backward), devoxelize, confusion matrix, at sizes that stay quick
under the tool.  Prints `san_probe ok` when every result is finite."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "3d-semantic-segmentation-fusion_b200")]
import torch
from tdssf import ops, synth
from utils.fusion_model import FeatureFusionModel
from utils.losses import CELSLoss

B, V = 2, int(sys.argv[1]) if len(sys.argv) > 1 else 1500
sw = synth.make_sweeps(B, V, "dinov3", seed=9)
dev = torch.device("cuda")


class Pcd(torch.nn.Module):
    def devoxelize(self, s):
        idx = ops.devox_nn(self.pts, self.pts)
        return ops.devox_gather(s, idx), idx


pcd = Pcd()
pcd.pts = sw.points.to(dev)
model = FeatureFusionModel(sw.encoder, pcd, 64, 384, 16, device=dev).to(dev)
model.load_state_dict(synth.make_mlp_weights(9), strict=True)
crit = CELSLoss(weight=torch.tensor(synth.CLASS_WEIGHTS, device=dev))
tok, feat, pts = sw.tokens.to(dev), sw.point_feat.to(dev).requires_grad_(True), sw.points.to(dev)
with torch.no_grad():
    out = model(tok, feat, pts, None, sw.image_sizes, sw.K, sw.Rt)
assert torch.isfinite(out).all()
for _ in range(3):   # the third call replays the loss sequence as a CUDA graph
    model.zero_grad(set_to_none=True)
    feat.grad = None
    scores = model(tok, feat, pts, None, sw.image_sizes, sw.K, sw.Rt)
    total, ce, lov, preds, gt = crit(scores, sw.labels.to(dev), sw.mask.to(dev))
    total.backward()
assert torch.isfinite(total) and all(torch.isfinite(p.grad).all() for p in model.mlp.parameters()) and torch.isfinite(feat.grad).all()
cm = ops.confusion(preds, gt, 16)
assert int(cm.sum()) == preds.numel()
raw, coords, n_vox, status = ops.voxelize(pts, 0.5)
assert int(status[0]) == 0
torch.cuda.synchronize()
print("san_probe ok", float(total))
