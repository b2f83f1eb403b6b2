"""Experiment harness: times the fused kernel, the stand-alone gather and projector for whatever
libtdssf.so variant TDSSF_LIB points at (CUDA events, config 2 shapes unless --workload)."""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "3d-semantic-segmentation-fusion_b200")]
import torch
from tdssf import ops, synth
import bench

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="config2")
ap.add_argument("--tok", default="float32")
ap.add_argument("--iters", type=int, default=20)
ap.add_argument("--points", default="normal", choices=["normal", "invalid", "hot", "sorted"],
                help="invalid: no camera sees any point (zero rows only); hot: 64 distinct points repeated (L1/L2 hits); "
                     "sorted: points of a sweep ordered by azimuth (spatially coherent gather)")
ap.add_argument("--fused-only", action="store_true")
ap.add_argument("--tag", default=os.environ.get("TDSSF_LIB", "default"))
a = ap.parse_args()
B, V, enc_key, _ = bench.WORKLOADS[a.workload]
sw = synth.make_sweeps(B, V, enc_key, seed=2000, tok_dtype=getattr(torch, a.tok))
w = synth.make_mlp_weights(0)
wd = [w[k].cuda() for k in bench.W_KEYS]
packed = ops.pack_weights(*wd)
if a.points == "hot":
    sw.points = sw.points[:, :64].repeat(1, (V + 63) // 64, 1)[:, :V].contiguous()
elif a.points == "sorted":
    order = torch.atan2(sw.points[..., 1], sw.points[..., 0]).argsort(dim=1)
    sw.points = torch.gather(sw.points, 1, order[..., None].expand(-1, -1, 4)).contiguous()
elif a.points == "invalid":
    sw.image_sizes = torch.zeros_like(sw.image_sizes)
pts, K, Rt = sw.points.cuda(), torch.stack(sw.K).cuda(), torch.stack(sw.Rt).cuda()
hw, tok, feat = sw.image_sizes.float().cuda(), sw.tokens.cuda(), sw.point_feat.cuda()
out = torch.empty((B, V, 16), device="cuda")

def timeit(fn, n=a.iters):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

res = {"tag": a.tag, "workload": a.workload, "tok": a.tok, "points": a.points}
res["fused_ms"] = timeit(lambda: ops.fusion_fwd(pts, K, Rt, hw, sw.encoder, tok, feat, packed, out=out))
if a.fused_only:
    res["fused_Gpts"] = B * V / res["fused_ms"] / 1e6
    print(json.dumps(res))
    sys.exit(0)
rec = ops.project(pts, K, Rt, hw, sw.encoder, want_pix=False, want_depth=False, want_valid=False, want_rec=True)["rec"]
res["project_rec_ms"] = timeit(lambda: ops.project(pts, K, Rt, hw, sw.encoder, want_pix=False, want_depth=False, want_valid=False, want_rec=True))
res["project_full_ms"] = timeit(lambda: ops.project(pts, K, Rt, hw, sw.encoder, want_flat=True, want_rec=True))
res["gather_ms"] = timeit(lambda: ops.gather_mean(tok, rec))
img = ops.gather_mean(tok, rec)
res["mlp_tc_dense_ms"] = timeit(lambda: ops.mlp_tc(feat, img, packed))
res["mlp_f32_ms"] = timeit(lambda: ops.mlp_f32(feat, img, *wd), n=3)
n = B * V
res["fused_Gpts"] = n / res["fused_ms"] / 1e6
print(json.dumps(res))
