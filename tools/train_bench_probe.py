"""Times bench.py's train sub-record loop alone and prints the CELS graph counters (debugging aid)."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "3d-semantic-segmentation-fusion_b200")]
import torch, ctypes as C
from tdssf import synth, parallel, _lib
from utils.fusion_model import FeatureFusionModel
from utils.losses import CELSLoss
import bench
dev = torch.device("cuda")
st = synth.make_sweeps(4, 34720, "dinov3", seed=3000)
model = FeatureFusionModel(st.encoder, bench._IdentityPcd(), 64, 384, 16, device=dev).to(dev)
model.load_state_dict(synth.make_mlp_weights(0), strict=True)
crit = CELSLoss(weight=torch.tensor(synth.CLASS_WEIGHTS, device=dev), ignore_index=0, lamda_lovasz=1.0)
d_tok, d_pts = st.tokens.to(dev), st.points.to(dev)
d_feat = st.point_feat.to(dev).requires_grad_(True)
d_lab, d_mask = st.labels.to(dev), st.mask.to(dev)
params = list(model.mlp.parameters())
def stats():
    gs = [C.c_int64(), C.c_int64(), C.c_int64()]
    _lib.lib.tdssf_cels_graph_stats(*[C.byref(g) for g in gs])
    return [g.value for g in gs]
def train_step(reduce=True):
    for p in params: p.grad = None
    d_feat.grad = None
    scores = model(d_tok, d_feat, d_pts, None, st.image_sizes, st.K, st.Rt)
    total = crit(scores, d_lab, d_mask)[0]
    total.backward()
    if reduce: parallel.allreduce_mean_grads(params)
    return total
for mode in (True, False):
    for _ in range(6): train_step(mode)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(50): train_step(mode)
    torch.cuda.synchronize()
    print(json.dumps({"allreduce_call": mode, "step_us": (time.perf_counter() - t0) / 50 * 1e6, "graph_stats": stats()}))

# host time per phase of the step (no synchronisation inside: the step is host-bound, the GPU queue stays short)
import collections
acc = collections.OrderedDict((k, 0.0) for k in ("zero_grads", "forward", "loss", "backward"))
for _ in range(10): train_step(False)
torch.cuda.synchronize()
N_IT = 200
for _ in range(N_IT):
    t0 = time.perf_counter()
    for p in params: p.grad = None
    d_feat.grad = None
    t1 = time.perf_counter()
    scores = model(d_tok, d_feat, d_pts, None, st.image_sizes, st.K, st.Rt)
    t2 = time.perf_counter()
    total = crit(scores, d_lab, d_mask)[0]
    t3 = time.perf_counter()
    total.backward()
    t4 = time.perf_counter()
    acc["zero_grads"] += t1 - t0; acc["forward"] += t2 - t1; acc["loss"] += t3 - t2; acc["backward"] += t4 - t3
torch.cuda.synchronize()
print(json.dumps({"host_us_per_phase": {k: v / N_IT * 1e6 for k, v in acc.items()}}))

# finer: wall time inside selected host functions (wrappers add ~0.3 us per call)
from tdssf import ops as _ops
import utils.fusion_model as _fm
tacc = collections.OrderedDict()
def wrap(obj, name, label=None):
    fn = getattr(obj, name); label = label or name
    tacc[label] = 0.0
    def timed(*a, **k):
        t = time.perf_counter()
        try: return fn(*a, **k)
        finally: tacc[label] += time.perf_counter() - t
    setattr(obj, name, timed)
for n in ("stack_calibration", "image_sizes_f32", "pack_weights", "fusion_fwd", "_flat_zero_grads", "_fusion_workspace", "_weights_struct"):
    wrap(_ops, n)
for n in ("tdssf_fusion_fwd", "tdssf_mlp_pack_weights", "tdssf_mlp_bwd_tc", "tdssf_cels_count", "tdssf_cels_fwd_bwd_counted"):
    wrap(_ops.lib, n, "C:" + n)
for _ in range(5): train_step(False)
for k in tacc: tacc[k] = 0.0
torch.cuda.synchronize()
for _ in range(N_IT): train_step(False)
torch.cuda.synchronize()
print(json.dumps({"host_us_inside": {k: round(v / N_IT * 1e6, 1) for k, v in tacc.items()}}))
