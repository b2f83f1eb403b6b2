"""Per-row timings of the hot path on one GPU: every stand-alone kernel with its algorithmic bytes or
FLOPs against the measured peaks, and the two composite steps BASELINE's configs[2] (training fwd/bwd per GPU: 4 sweeps)
and configs[4] (evaluation step) describe.  CUDA events, 3 warm-up + `--iters` timed calls, inputs larger than L2 where
the row's working set allows it.  One JSON object per line; `python tools/bench_rows.py > profiles/rN_rows.jsonl`."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "3d-semantic-segmentation-fusion_b200")]
import torch  # noqa: E402

import bench  # noqa: E402
from tdssf import ops, synth  # noqa: E402
from utils.fusion_model import FeatureFusionModel  # noqa: E402
from utils.losses import CELSLoss  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--iters", type=int, default=20)
a = ap.parse_args()
peaks, peak_kind = bench.measured_peaks()
HBM = peaks["hbm_gbs"]
FP32_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12   # CUDA-core FMA peak at the maximum SM clock (no measured figure)
dev = torch.device("cuda")


def timeit(fn, n=a.iters):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def emit(row, kernel, ms, n_units, unit, bytes_per_unit=None, flops_per_unit=None, note=""):
    out = {"row": row, "kernel": kernel, "ms": round(ms, 5), "units": n_units, "unit": unit,
           "throughput": n_units / (ms * 1e-3), "note": note}
    if bytes_per_unit is not None:
        gbs = bytes_per_unit * n_units / (ms * 1e-3) / 1e9
        out.update(bound="hbm", algorithmic_bytes_per_unit=bytes_per_unit, achieved_gbs=gbs, peak_gbs=HBM, frac=gbs / HBM)
    if flops_per_unit is not None:
        tf = flops_per_unit * n_units / (ms * 1e-3) / 1e12
        out.update(bound="fp32 fma", flops_per_unit=flops_per_unit, achieved_tflops=tf, peak_tflops=FP32_TFLOPS,
                   frac=tf / FP32_TFLOPS)
    print(json.dumps(out), flush=True)


# ---------------------------------------------------------------- config 2 shapes: 8 sweeps x 34 720 points -----
B, V = 8, 34720
sw = synth.make_sweeps(B, V, "dinov3", seed=2000)
enc = sw.encoder
M = enc.n_patches
w = synth.make_mlp_weights(0)
wd = [w[k].to(dev) for k in bench.W_KEYS]
pts, K, Rt = sw.points.to(dev), torch.stack(sw.K).to(dev), torch.stack(sw.Rt).to(dev)
hw, tok, feat = sw.image_sizes.float().to(dev), sw.tokens.to(dev), sw.point_feat.to(dev)
N = B * V

# a1-a3 projection
ms = timeit(lambda: ops.project(pts, K, Rt, hw, enc, want_pix=False, want_depth=False, want_valid=False, want_rec=True))
emit("a1-a3", "project_kernel (records only)", ms, N, "points", bytes_per_unit=16 + 16,
     note="16 B point in, 16 B camera record out; 6 cameras x (27 flops + 4 IEEE divisions) per point")
ms = timeit(lambda: ops.project(pts, K, Rt, hw, enc, want_flat=True, want_rec=True))
emit("a1-a3", "project_kernel (all reference outputs)", ms, N, "points", bytes_per_unit=16 + 6 * (8 + 4 + 1 + 4) + 16,
     note="pix (B,6,V,2) + depth + valid + flat index + record")
rec = ops.project(pts, K, Rt, hw, enc, want_pix=False, want_depth=False, want_valid=False, want_rec=True)["rec"]

# a4-a5 gather + masked camera mean (stand-alone)
ms = timeit(lambda: ops.gather_mean(tok, rec))
emit("a4-a5", "gather_mean_kernel<float>", ms, N, "points", bytes_per_unit=16 + 6 * M * 1536 / V + 1536,
     note="record in, token maps read once, 1536 B fp32 feature out (the fused kernel never writes this)")
img = ops.gather_mean(tok, rec)

# a6 MLP, fp32 CUDA-core kernels (1e-5 parity mode, training)
flops = 2 * (448 * 256 + 256 * 64 + 64 * 16)
ms = timeit(lambda: ops.mlp_f32(feat, img, *wd), n=5)
emit("a6", "mlp_fwd_f32", ms, N, "points", flops_per_unit=flops, note="fp32 FMA, inference call (no saved activations)")
wg = [t.clone().requires_grad_(True) for t in wd]
featg = feat.clone().requires_grad_(True)


def mlp_train():
    out = ops.mlp_f32(featg, img, *wg)
    out.backward(torch.ones_like(out))
    for t in wg:
        t.grad = None
    featg.grad = None


ms = timeit(mlp_train, n=5)
emit("a6", "mlp_fwd_f32 + mlp_bwd_f32", ms, N, "points", flops_per_unit=3 * flops - 2 * 384 * 256,
     note="forward with saved h1/h2 + backward (dW, db, dx_pt; no gradient to the image feature)")

# a6 on the tensor cores (dense features) and a1-a6 fused
packed = ops.pack_weights(*wd)
ms = timeit(lambda: ops.mlp_tc(feat, img, packed))
emit("a6", "fused_mlp_kernel<float,false> (dense features)", ms, N, "points", bytes_per_unit=4 * 448 + 64,
     note="tcgen05 bf16 MLP fed by a dense fp32 [N,384] feature")
out = torch.empty((B, V, 16), device=dev)
ms = timeit(lambda: ops.fusion_fwd(pts, K, Rt, hw, enc, tok, feat, packed, out=out))
emit("a1-a6", "fused_mlp_kernel<float,true>", ms, N, "points", bytes_per_unit=bench.algorithmic_bytes_per_point(V, M),
     note="the bench.py kernel, for reference")

# ---------------------------------------------------------------- a7 devoxelize ---------------------------------
Bd, P, Vd = 2, 34720, 30000
gen = torch.Generator().manual_seed(5)
raw = synth.make_points(Bd, P, gen)[..., :3].contiguous().to(dev)
vox = (raw[:, torch.randperm(P, generator=gen)[:Vd]] + 0.05).contiguous()
ms = timeit(lambda: ops.devox_nn(raw, vox, cell=None), n=5)
emit("a7", "devox_nn_kernel (brute force)", ms, Bd * P * Vd, "point-voxel pairs", flops_per_unit=8,
     note=f"{Bd} sweeps x {P} points x {Vd} voxels, direct-difference distance + running arg-min (~8 flops per pair); "
          "the reference materialises a P x V_max matrix")
ms = timeit(lambda: ops.devox_nn(raw, vox, cell=0.5))
emit("a7", "devox grid build + query + fallback scan (cell 0.5 m)", ms, Bd * P, "points", bytes_per_unit=16 + 8 + 16 * Vd / P,
     note="same indices as the brute force; uniformly random synthetic cloud, so the share of fallback points is "
          "far above a voxelised sweep's")
# voxel centres of an actual voxelisation of the same points (0.1 m voxels): the reference's situation
vs = 0.1
cent = [torch.unique(torch.floor(raw[b] / vs), dim=0) * vs + vs / 2 for b in range(Bd)]
Vv = min(c.shape[0] for c in cent)
vox_v = torch.stack([c[:Vv] for c in cent]).contiguous()
ms = timeit(lambda: ops.devox_nn(raw, vox_v, cell=None), n=5)
emit("a7", "devox_nn_kernel (brute force), voxelised sweep", ms, Bd * P * Vv, "point-voxel pairs", flops_per_unit=8,
     note=f"{Vv} occupied 0.1 m voxels per sweep")
ms = timeit(lambda: ops.devox_nn(raw, vox_v, cell=0.5))
emit("a7", "devox grid search, voxelised sweep (cell 0.5 m)", ms, Bd * P, "points", bytes_per_unit=16 + 8 + 16 * Vv / P,
     note="every point lies inside an occupied voxel: no fallback")
idx = ops.devox_nn(raw, vox)
scores = torch.randn(Bd, Vd, 16, device=dev)
ms = timeit(lambda: ops.devox_gather(scores, idx))
emit("a7", "devox_gather_kernel", ms, Bd * P, "points", bytes_per_unit=8 + 64 + 64, note="index + 16 logits in (L2) + 16 out")

# ---------------------------------------------------------------- a8-a10 CE + Lovasz ----------------------------
Bl = 4
logits = torch.randn(Bl, V, 16, device=dev)
labels = synth.make_labels(Bl, V, gen).to(dev)
mask = torch.ones(Bl, V, dtype=torch.bool, device=dev)
cw = torch.ones(16, device=dev)
lg = logits.clone().requires_grad_(True)


def cels_step():
    total = ops.cels(lg, labels, mask, cw)[0]
    total.backward()
    lg.grad = None


ms = timeit(cels_step, n=10)
emit("a8-a10", "cels_fwd_bwd (all kernels, incl. the host read of n_valid)", ms, Bl * V, "points",
     bytes_per_unit=64 + 8 + 1 + 64 + 16 * (4 * 12 * 2),
     note="logits + labels + mask in, dlogits out, and per class 4 radix passes reading and writing a 12-byte key/payload")

# ---------------------------------------------------------------- a11 confusion matrix --------------------------
n_cm = 8 * 34720
pr = torch.randint(1, 17, (n_cm,), device=dev)
lb = torch.randint(1, 17, (n_cm,), device=dev)
cm = torch.zeros(16, 16, dtype=torch.int64, device=dev)
ms = timeit(lambda: ops.confusion(pr, lb, 16, out=cm))
emit("a11", "confusion_kernel", ms, n_cm, "points", bytes_per_unit=16, note="two int64 per point; 4.4 MB: L2-resident after warm-up")

# ---------------------------------------------------------------- composite steps -------------------------------
class Pcd(torch.nn.Module):
    def devoxelize(self, s):
        return s, None


# configs[2]: training fwd/bwd, 4 sweeps per GPU (batch 32 over 8 GPUs), fp32 kernels (a gradient is required)
Bt = 4
model = FeatureFusionModel(enc, Pcd(), 64, 384, 16, device=dev).to(dev)
model.load_state_dict(w, strict=True)
crit = CELSLoss()
t_tok, t_feat, t_pts = tok[:Bt], feat[:Bt].clone().requires_grad_(True), pts[:Bt]
t_hw, t_K, t_Rt = sw.image_sizes[:Bt], sw.K[:Bt], sw.Rt[:Bt]
t_lab, t_mask = labels[:Bt], mask[:Bt]


def train_step():
    scores = model(t_tok, t_feat, t_pts, None, t_hw, t_K, t_Rt)
    total = crit(scores, t_lab, t_mask)[0]
    total.backward()
    model.zero_grad(set_to_none=True)
    t_feat.grad = None


os.environ["TDSSF_MLP"] = "fp32"
ms = timeit(train_step, n=5)
emit("config3", "training step per GPU, TDSSF_MLP=fp32 (drop-in modules: forward + CE/Lovasz + backward)", ms, Bt * V, "points",
     note="4 sweeps x 34 720 points; project + gather + fp32 MLP fwd, cels fwd/bwd, fp32 MLP bwd; no optimizer, no all-reduce")
os.environ["TDSSF_MLP"] = "bf16"
ms = timeit(train_step, n=5)
emit("config3", "training step per GPU, default bf16 (fused tcgen05 forward, tcgen05 backward: data-gradient chain + weight-gradient kernels)", ms, Bt * V, "points",
     note="4 sweeps x 34 720 points; fused launch saving its bf16 input rows + activations, cels fwd/bwd, bf16 backward")
wt = [t.clone().requires_grad_(True) for t in wd]
featt = feat.clone().requires_grad_(True)


def mlp_tc_train_step():
    o = ops.mlp_tc_train(featt, img, *wt)
    o.backward(torch.ones_like(o))
    for t in wt:
        t.grad = None
    featt.grad = None


ms = timeit(mlp_tc_train_step, n=5)
emit("a6", "mlp_fwd_tc (saving h1/h2) + mlp_bwd_tc", ms, N, "points", flops_per_unit=3 * flops - 2 * 384 * 256,
     note="8 sweeps; bf16 tensor-core forward + tcgen05 backward (hand-written chain + weight-gradient kernels); frac is against the fp32 FMA peak for comparison with the fp32 row")


def fused_train_step():
    o = ops.fusion_train(pts, K, Rt, hw, enc, tok, featt, *wt)
    o.backward(torch.ones_like(o))
    for t in wt:
        t.grad = None
    featt.grad = None


ms = timeit(fused_train_step, n=5)
emit("a1-a6", "fusion_fwd (saving x/h1/h2) + mlp_bwd_tc", ms, N, "points", flops_per_unit=3 * flops - 2 * 384 * 256,
     note="8 sweeps; the fused launch also writes its bf16 input rows; backward reads them (no fp32 image feature, no concat)")

# configs[4]: evaluation step: fused inference + argmax + confusion accumulation (8 sweeps)
cm.zero_()
lab8 = torch.randint(1, 17, (B, V), device=dev)


pred8 = torch.empty(B, V, dtype=torch.int64, device=dev)


def eval_step():
    with torch.no_grad():
        ops.fusion_fwd(pts, K, Rt, hw, enc, tok, feat, packed, out=out, preds_out=pred8)
        ops.confusion(pred8, lab8, 16, out=cm)


ms = timeit(eval_step)
emit("config5", "evaluation step per GPU (fused forward with class ids + confusion)", ms, N, "points",
     note="8 sweeps x 34 720 points; two launches, the arg-max comes out of the fused kernel's last epilogue")

# f1: voxelizer on the device (8 sweeps, voxel 0.1 and 0.5 m).  Beside it, as a LOWER bound of the reference's host path
# (utils/lidar.py:83-118: D2H copy, Open3D down-sampling, scikit-learn 5-NN, H2D copy), the reference's own scikit-learn
# call alone on ONE sweep on this box's cores (queries = the device voxelizer's centroids).
import time
from sklearn.neighbors import NearestNeighbors
for vs in (0.1, 0.5):
    ms = timeit(lambda: ops.voxelize(pts, vs))
    raw_v, _, n_v, _ = ops.voxelize(pts, vs)
    one = pts[0].cpu().numpy()
    cen = raw_v[0, :int(n_v[0]), :3].double().cpu().numpy()
    t0 = time.perf_counter()
    NearestNeighbors(n_neighbors=5, algorithm="auto").fit(one[:, :3]).kneighbors(cen)
    cpu_ms = (time.perf_counter() - t0) * 1e3
    emit("f1", f"tdssf_voxelize, voxel {vs} m", ms, N, "points",
         note=f"8 sweeps x 34 720 points in one call (down-sample + 5-NN intensity + coords), {int(n_v.float().mean())} voxels per sweep; "
              f"scikit-learn 5-NN alone (part of the reference's host path) on one sweep: {cpu_ms:.1f} ms = {V / cpu_ms * 1e3:.3g} points/s")

# f4: calibration composition (one launch for 8 sweeps x 6 cameras) and the 10-sweep cloud of configs[3]
import numpy as np
rng = np.random.default_rng(0)
def _poses(*shape):
    q = rng.normal(size=shape + (4,)); q /= np.linalg.norm(q, axis=-1, keepdims=True)
    return np.concatenate([q, rng.normal(size=shape + (3,))], -1)
ls_, cs_, le_, ce_ = (torch.from_numpy(x).to(dev) for x in (_poses(B), _poses(B, 6), _poses(B), _poses(B, 6)))
ms = timeit(lambda: ops.compose_lidar2cam(ls_, cs_, le_, ce_))
emit("f4", "tdssf_compose_lidar2cam (four-stage chain, float64)", ms, B * 6, "cameras",
     note="launch-latency bound: 48 threads; the reference composes these on the host with numpy per sample (utils/dataloader.py:123-152)")
S10 = 10
p10 = synth.make_points(S10, V, torch.Generator().manual_seed(9)).to(dev)
cnt10 = torch.full((S10,), V, dtype=torch.int32, device=dev)
ss_, se_ = (torch.from_numpy(x).to(dev) for x in (_poses(S10), _poses(S10)))
ms = timeit(lambda: ops.accumulate_sweeps(p10, cnt10, ss_, se_, ss_[0], se_[0]))
emit("f4", "tdssf_compose_sweep2ref + tdssf_accumulate_sweeps (10 sweeps x 34 720 points)", ms, S10 * V, "points", bytes_per_unit=32,
     note="16 B in + 16 B out per point, float64 transform; includes the wrapper's two small allocations and the pose kernel")
