"""BASELINE configs[2] / configs[4] across ranks, one process per GPU over NCCL: the all-reduced MLP gradient
against the mean of the ORACLE's per-shard gradients (TDSSF_MLP=fp32, 1e-5 class tolerance), the all-reduced int64
confusion matrix against the oracle's matrix of the union, bit for bit.  World size 1 always runs (no collective: the same
check of the single-rank path); 2 / 4 / 8 need that many visible GPUs and are skipped otherwise."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_collectives_match_the_oracle_per_shard(world):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs, {torch.cuda.device_count()} visible")
    script = os.path.join(ROOT, "tools", "dist_check.py")
    if world == 1:
        cmd = [sys.executable, script]
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
               "--master-port", str(_free_port()), script]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert r.returncode == 0 and lines, r.stdout[-2000:] + r.stderr[-2000:]
    out = json.loads(lines[-1])
    assert out["ok"] and out["world"] == world and out["confusion_equals_oracle_union"], out
    assert max(out["grad_rel_err_vs_mean_of_oracle_shards"].values()) < 2e-4, out
    if world > 1 and out.get("grad_allreduce") == "p2p":   # the peer-memory kernel: same mean as NCCL, identical on every rank
        assert out["p2p_vs_nccl_max_abs"] < 1e-5 and out["p2p_identical_across_ranks"], out
