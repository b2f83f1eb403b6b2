"""Multi-GPU probe of the two places where ranks exchange anything, over NCCL, one process per GPU:

  configs[2]  training step: every rank runs forward + CE/Lovasz + backward on its 4 sweeps of a 4*N-sweep batch, then
              ONE all-reduce of the flat MLP gradient (tdssf.parallel.allreduce_mean_grads).  Rank 0 also runs every
              shard itself and checks that the all-reduced gradient is the mean of the per-shard gradients.
  configs[4]  evaluation: every rank accumulates the int64 confusion matrix of its sweeps on the device, one all-reduce
              at the end (allreduce_confusion); rank 0 checks it against its own pass over all shards, bit for bit.

Launch:  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/dist_probe.py
Times are CUDA events on each rank, max over ranks (tdssf.parallel.max_over_ranks).  One JSON line per config on rank 0."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "3d-semantic-segmentation-fusion_b200")]
import torch
import torch.distributed as dist

from tdssf import ops, parallel, synth
from utils.fusion_model import FeatureFusionModel
from utils.losses import CELSLoss

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


class Pcd(torch.nn.Module):
    def devoxelize(self, s):
        return s, None


PER_RANK, V = 4, 34720
shards = [synth.make_sweeps(PER_RANK, V, "dinov3", seed=3000 + r) for r in range(world)]   # the same on every rank
weights = synth.make_mlp_weights(0)
crit = CELSLoss()


def make_model():
    m = FeatureFusionModel(shards[0].encoder, Pcd(), 64, 384, 16, device=dev).to(dev)
    m.load_state_dict(weights, strict=True)
    return m


def to_dev(sw):
    return dict(tok=sw.tokens.to(dev), feat=sw.point_feat.to(dev), pts=sw.points.to(dev), lab=sw.labels.to(dev),
                mask=sw.mask.to(dev), sizes=sw.image_sizes, K=sw.K, Rt=sw.Rt)


def train_step(model, d, reduce):
    model.zero_grad(set_to_none=True)
    scores = model(d["tok"], d["feat"], d["pts"], None, d["sizes"], d["K"], d["Rt"])
    total = crit(scores, d["lab"], d["mask"])[0]
    total.backward()
    if reduce:
        parallel.allreduce_mean_grads(model.mlp.parameters())
    return total.detach()


mine = to_dev(shards[rank])
model = make_model()
for _ in range(3):
    train_step(model, mine, True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
STEPS = 20
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
e0.record()
for _ in range(STEPS):
    loss = train_step(model, mine, True)
e1.record()
torch.cuda.synchronize()
ms = parallel.max_over_ranks(e0.elapsed_time(e1) / STEPS, dev)
got = [p.grad.clone() for p in model.mlp.parameters()]
if rank == 0:
    ref_model = make_model()
    acc = None
    for r in range(world):
        train_step(ref_model, to_dev(shards[r]), False)
        g = [p.grad.double() for p in ref_model.mlp.parameters()]
        acc = g if acc is None else [a + b for a, b in zip(acc, g)]
    err = max(float((a.double() - b / world).norm() / (b / world).norm()) for a, b in zip(got, acc))
    print(json.dumps({"row": "config3", "n_gpus": world, "sweeps_per_gpu": PER_RANK, "points_per_sweep": V,
                      "ms_per_step": round(ms, 4), "points_per_s": round(world * PER_RANK * V / ms * 1e3),
                      "allreduce": "one flat fp32 bucket of the 132 432 MLP gradients, mean",
                   "bucket_in_place": parallel._shared_flat_grad([p for p in model.mlp.parameters()]) is not None,
                      "grad_vs_mean_of_shards_rel_err": err, "loss_rank0": float(loss)}), flush=True)
    assert err < 1e-5, err

# evaluation: fused forward writing class ids + confusion on the device, one all-reduce at the end
packed = ops.pack_weights(*[weights[k].to(dev) for k in ("mlp.0.weight", "mlp.0.bias", "mlp.2.weight", "mlp.2.bias", "mlp.4.weight", "mlp.4.bias")])


def eval_pass(d, cm, sw):
    preds = torch.empty(PER_RANK, V, dtype=torch.int64, device=dev)
    with torch.no_grad():
        ops.fusion_fwd(d["pts"], torch.stack(d["K"]).to(dev), torch.stack(d["Rt"]).to(dev), d["sizes"].float().to(dev), sw.encoder,
                       d["tok"], d["feat"], packed, preds_out=preds)
        lab = d["lab"].clamp_min(1)   # the evaluation loop only sees labelled points; ids 1..16
        ops.confusion(preds, lab, 16, out=cm)


cm = torch.zeros(16, 16, dtype=torch.int64, device=dev)
for _ in range(2):
    eval_pass(mine, torch.zeros_like(cm), shards[rank])
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
e0.record()
eval_pass(mine, cm, shards[rank])
parallel.allreduce_confusion(cm)
e1.record()
torch.cuda.synchronize()
ms = parallel.max_over_ranks(e0.elapsed_time(e1), dev)
if rank == 0:
    want = torch.zeros_like(cm)
    for r in range(world):
        eval_pass(to_dev(shards[r]), want, shards[r])
    torch.cuda.synchronize()
    same = bool(torch.equal(cm, want))
    print(json.dumps({"row": "config5", "n_gpus": world, "ms_eval_pass_plus_allreduce": round(ms, 4),
                      "points": world * PER_RANK * V, "confusion_equals_single_process": same,
                      "confusion_total": int(cm.sum())}), flush=True)
    assert same and int(cm.sum()) == world * PER_RANK * V
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
