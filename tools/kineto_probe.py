"""Warm-cache per-kernel times of one training step (configs[2] per-GPU share: 4 sweeps x 34 720 points) from torch.profiler
(CUPTI activity records: no replay, no cache flush, kernels back to back as in a normal run)."""
import os, sys, json, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "3d-semantic-segmentation-fusion_b200")]
import torch
from torch.profiler import profile, ProfilerActivity
from tdssf import synth
from utils.fusion_model import FeatureFusionModel
from utils.losses import CELSLoss

class Pcd(torch.nn.Module):
    def devoxelize(self, s):
        return s, None

B, V = 4, 34720
sw = synth.make_sweeps(B, V, "dinov3", seed=3000)
dev = torch.device("cuda")
model = FeatureFusionModel(sw.encoder, Pcd(), 64, 384, 16, device=dev).to(dev)
model.load_state_dict(synth.make_mlp_weights(0), strict=True)
crit = CELSLoss()
tok, feat, pts = sw.tokens.to(dev), sw.point_feat.to(dev).requires_grad_(True), sw.points.to(dev)
lab, mask = sw.labels.to(dev), sw.mask.to(dev)
def step():
    scores = model(tok, feat, pts, None, sw.image_sizes, sw.K, sw.Rt)
    total = crit(scores, lab, mask)[0]
    total.backward()
    model.zero_grad(set_to_none=True)
    feat.grad = None
for _ in range(5): step()
torch.cuda.synchronize()
n = 10
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(n): step()
e1.record(); torch.cuda.synchronize()
print(f"step (no profiler): {e0.elapsed_time(e1) / n * 1e3:.1f} us")
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(n): step()
    torch.cuda.synchronize()
agg = collections.OrderedDict()
for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CUDA:
        d = agg.setdefault(ev.name, [0.0, 0])
        d[0] += ev.device_time; d[1] += 1
tot = sum(v[0] for v in agg.values()) / n
print(f"kernel time per step: {tot:.1f} us over {sum(v[1] for v in agg.values()) // n} launches")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{v[0] / n:9.1f} us  x{v[1] // n:3d}  {k[:120]}")
