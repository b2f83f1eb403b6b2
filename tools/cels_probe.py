"""Times the CE + Lovasz loss (forward + backward, tdssf_cels_fwd_bwd) at configs[2] size: 4 sweeps x 34 720 points."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "3d-semantic-segmentation-fusion_b200")]
import torch
from tdssf import ops, synth
B, V = 4, 34720
sw = synth.make_sweeps(B, V, "dinov3", seed=3000)
g = torch.Generator().manual_seed(1)
mode = sys.argv[1] if len(sys.argv) > 1 else "randn"
lg = torch.randn(B, V, 16, generator=g) * (4.0 if mode == "peaked" else 1.0)
lg = lg.cuda().requires_grad_(True)
lab, mask, cw = sw.labels.cuda(), sw.mask.cuda(), torch.tensor(synth.CLASS_WEIGHTS).cuda()
def step():
    total = ops.cels(lg, lab, mask, cw, 1.0, 0)[0]
    total.backward()
    lg.grad = None
for _ in range(3): step()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(20): step()
e1.record(); torch.cuda.synchronize()
print(json.dumps({"tag": os.environ.get("TDSSF_LIB", "default"), "logits": mode, "cels_fwd_bwd_ms": e0.elapsed_time(e1) / 20, "points": B * V}))
