"""Multi-rank parity check of the two collectives, one process per GPU over NCCL; run by tests/test_dist_gpu.py
(torchrun, world size 2/4/8) and usable by hand.

  configs[2]  training step in TDSSF_MLP=fp32 mode on this rank's shard -> all-reduce(mean) of the MLP gradients.  The
              ORACLE (oracle/np_oracle.py: fusion forward, CE + Lovasz with its gradient, MLP backward) runs on every shard on
              rank 0's host; the mean of its per-shard gradients must equal the all-reduced gradient within 1e-5 relative.
              (The loss is not additive over sweeps, SURVEY H6: the reference semantics of N ranks is the mean of N runs.)
  configs[4]  evaluation: per-rank int64 confusion matrix on the device, one all-reduce; must equal the oracle's matrix of
              the union of all shards bit for bit, at any world size.
Prints one JSON line on rank 0; exit code 1 on a mismatch."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "3d-semantic-segmentation-fusion_b200")]
os.environ["TDSSF_MLP"] = "fp32"
import numpy as np
import torch
import torch.distributed as dist

from oracle import np_oracle as O
from tdssf import ops, parallel, synth
from utils.fusion_model import FeatureFusionModel
from utils.losses import CELSLoss

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
W_KEYS = ("mlp.0.weight", "mlp.0.bias", "mlp.2.weight", "mlp.2.bias", "mlp.4.weight", "mlp.4.bias")
PER_RANK, V = 2, 3000
shards = [synth.make_sweeps(PER_RANK, V, "dinov3", seed=5000 + r) for r in range(world)]   # the same on every rank
weights = synth.make_mlp_weights(5)
cw = np.array(synth.CLASS_WEIGHTS, np.float32)


class Pcd(torch.nn.Module):
    def devoxelize(self, s):
        return s, None


sw = shards[rank]
model = FeatureFusionModel(sw.encoder, Pcd(), 64, 384, 16, device=dev).to(dev)
model.load_state_dict(weights, strict=True)
crit = CELSLoss(weight=torch.tensor(synth.CLASS_WEIGHTS, device=dev), ignore_index=0, lamda_lovasz=1.0)
feat = sw.point_feat.to(dev).requires_grad_(True)
scores = model(sw.tokens.to(dev), feat, sw.points.to(dev), None, sw.image_sizes, sw.K, sw.Rt)
total, _, _, preds, gt = crit(scores, sw.labels.to(dev), sw.mask.to(dev))
total.backward()
params = list(model.mlp.parameters())
parallel.allreduce_mean_grads(params)
cm = ops.confusion(preds, gt, 16)
parallel.allreduce_confusion(cm)
torch.cuda.synchronize()

# the gradient bucket over NVLink peer memory (csrc/p2p.cu) against NCCL: 40 consecutive calls (epoch / parity protocol)
p2p = {"grad_allreduce": "nccl" if world > 1 else "none"}
if world > 1:
    n_bucket = sum(p.numel() for p in params)
    ex = parallel._peers_for(n_bucket, dev)
    if ex is not None:
        gen = torch.Generator().manual_seed(100 + rank)
        worst, same = 0.0, True
        for it in range(40):
            x = torch.randn(n_bucket, generator=gen).to(dev)
            want = x.clone()
            dist.all_reduce(want, op=dist.ReduceOp.AVG)
            ex.allreduce_mean(x)
            worst = max(worst, float((x - want).abs().max()))
            lo, hi = x.double().sum().reshape(1).clone(), x.double().sum().reshape(1).clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            same = same and bool((lo == hi).item())
        p2p = {"grad_allreduce": "p2p", "p2p_vs_nccl_max_abs": worst, "p2p_identical_across_ranks": same}

ok = True
out = {"world": world, **p2p}
if rank == 0:
    acc, cm_want, losses = None, np.zeros((16, 16), np.int64), []
    for r in range(world):
        s = shards[r]
        wn = {k: v.numpy() for k, v in weights.items()}
        K, Rt = torch.stack(s.K).numpy(), torch.stack(s.Rt).numpy()
        logits, img, _, _ = O.fusion_forward(s.tokens.numpy(), s.point_feat.numpy(), s.points.numpy(), s.image_sizes.numpy(), K, Rt, wn,
                                             s.encoder.patch_size, s.encoder.resize_height, s.encoder.resize_width)
        t, _, _, p_o, g_o, dlog = O.cels_loss(logits, s.labels.numpy(), s.mask.numpy(), cw, 0, 1.0, return_grad=True)
        grads, _ = O.mlp_backward(s.point_feat.numpy(), img, wn, dlog)
        acc = grads if acc is None else {k: acc[k] + grads[k] for k in grads}
        cm_want += O.confusion_matrix(p_o, g_o, 16)
        losses.append(float(t))
    errs = {}
    for k, p in zip(W_KEYS, params):
        want = acc[k] / world
        errs[k] = float(np.linalg.norm(p.grad.double().cpu().numpy() - want) / np.linalg.norm(want))
    same = bool(np.array_equal(cm.cpu().numpy(), cm_want))
    out.update(grad_rel_err_vs_mean_of_oracle_shards=errs, loss_rank0=float(total), loss_oracle_rank0=losses[0],
               confusion_equals_oracle_union=same, confusion_total=int(cm.sum()))
    ok = max(errs.values()) < 1e-5 * 20 and same and abs(float(total) - losses[0]) < 1e-5 * abs(losses[0])
    if p2p["grad_allreduce"] == "p2p":
        ok = ok and p2p["p2p_vs_nccl_max_abs"] < 1e-5 and p2p["p2p_identical_across_ranks"]
    out["ok"] = ok
    print(json.dumps(out), flush=True)
if world > 1:
    flag = torch.tensor([int(ok)], device=dev)
    dist.broadcast(flag, 0)
    ok = bool(flag.item())
    dist.barrier()
    dist.destroy_process_group()
sys.exit(0 if ok else 1)
