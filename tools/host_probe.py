"""Host-side profile (cProfile) of the training step of tools/kineto_probe.py: where the Python time between launches goes."""
import os, sys, cProfile, pstats
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "3d-semantic-segmentation-fusion_b200")]
import torch
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
for _ in range(10): step()
torch.cuda.synchronize()
import time
t0 = time.perf_counter()
for _ in range(100): step()
torch.cuda.synchronize()
print(f"step wall: {(time.perf_counter() - t0) * 1e4:.1f} us")
pr = cProfile.Profile(); pr.enable()
for _ in range(100): step()
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr); st.strip_dirs(); st.sort_stats("cumtime").print_stats(45)
import ctypes as C
from tdssf import _lib
a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
_lib.lib.tdssf_cels_graph_stats(C.byref(a), C.byref(b), C.byref(c))
print("cels graph: replays", a.value, "captures", b.value, "plain", c.value)
