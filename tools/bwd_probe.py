"""Numerics and timing of tdssf_mlp_bwd_tc (hand-written tcgen05 backward) against an fp64 model of its arithmetic."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "3d-semantic-segmentation-fusion_b200")]
import torch
from tdssf import ops, synth
W_KEYS = ("mlp.0.weight", "mlp.0.bias", "mlp.2.weight", "mlp.2.bias", "mlp.4.weight", "mlp.4.bias")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
gen = torch.Generator().manual_seed(4)
x_pt = torch.randn(n, 64, generator=gen).cuda(); x_img = torch.randn(n, 384, generator=gen).cuda(); dl = torch.randn(n, 16, generator=gen).cuda()
w = synth.make_mlp_weights(4)
params = [w[k].cuda().clone().requires_grad_(True) for k in W_KEYS]
xp = x_pt.clone().requires_grad_(True)
out = ops.mlp_tc_train(xp, x_img, *params)
out.backward(dl)
torch.cuda.synchronize()
rel = lambda a, b: float((a.double() - b.double()).norm() / b.double().norm())
q = lambda t: t.bfloat16().double()
W1, b1, W2, b2, W3, b3 = [w[k].cuda() for k in W_KEYS]
x = q(torch.cat([x_pt, x_img], -1))
h1 = q(torch.relu(x @ q(W1).T + b1.double())); h2 = q(torch.relu(h1 @ q(W2).T + b2.double()))
dh2 = (dl.double() @ W3.double()) * (h2 > 0); dh1 = (q(dh2) @ q(W2)) * (h1 > 0)
model = {"mlp.0.weight": q(dh1).T @ x, "mlp.0.bias": q(dh1).sum(0), "mlp.2.weight": q(dh2).T @ h1, "mlp.2.bias": q(dh2).sum(0),
         "mlp.4.weight": dl.double().T @ h2, "mlp.4.bias": dl.double().sum(0)}
res = {k: rel(p.grad, model[k]) for k, p in zip(W_KEYS, params)}
res["dx_pt"] = rel(xp.grad, q(dh1) @ q(W1)[:, :64])
print(json.dumps({"n": n, "rel_err": res}))
# timing of the backward alone
N = 4 * 34720
xb = torch.randn(N, 448, device="cuda").bfloat16(); h1b = torch.randn(N, 256, device="cuda").bfloat16(); h2b = torch.randn(N, 64, device="cuda").bfloat16()
dlN = torch.randn(N, 16, device="cuda")
import ctypes as C
from tdssf import _lib
from tdssf.ops import _ptr, _stream, _weights_struct, _flat_zero_grads
ws_bytes = _lib.lib.tdssf_mlp_bwd_tc_workspace_bytes(N); work = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
wp = [t.detach() for t in params]; grads = _flat_zero_grads(wp); dx = torch.empty(N, 64, device="cuda")
wst = _weights_struct(wp); g = _lib.MlpGrads(*[t.data_ptr() for t in grads])
def run():
    _lib.check(_lib.lib.tdssf_mlp_bwd_tc(None, None, _ptr(xb), _ptr(h1b), _ptr(h2b), _ptr(dlN), N, C.byref(wst), C.byref(g), _ptr(dx), _ptr(work), ws_bytes, _stream(xb)))
for _ in range(3): run()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(20): run()
e1.record(); torch.cuda.synchronize()
print(json.dumps({"bwd_ms_4sweeps": e0.elapsed_time(e1) / 20}))
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(5): run()
    torch.cuda.synchronize()
agg = {}
for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CUDA:
        d = agg.setdefault(ev.name[:70], [0.0, 0]); d[0] += ev.device_time; d[1] += 1
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0]): print(f"{v[0]/5:9.1f} us x{v[1]//5} {k}")
