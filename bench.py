"""bench.py — fused LiDAR points/s through project -> gather -> camera mean -> concat -> MLP -> logits.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload config2|config4|config1]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
           bench.py --gpus N --steps K --warmup W

One "step" = one pass of the hot path over one batch of synthetic nuScenes-shaped sweeps (BASELINE.json
configs[1]: 8 sweeps x 34 720 points, 6 cameras, DINOv3 ViT-S/16 40x71x384 fp32 patch tokens, inference).
Every rank owns its own batch (sweep-sharded, weak scaling, no data-path collective); `value` is the
whole-job points/s with inputs resident in HBM, timed with CUDA events, max over ranks.  `e2e` is the same
metric through the drop-in `FeatureFusionModel.forward` with HOST (pinned) inputs and a host read of the
logits, copies inside the timed region.  `--impl reference` times the CPU port of the reference's PyTorch
path (oracle/torch_port.py) on the host cores.  Prints ONE JSON line on rank 0.
"""
import argparse
import contextlib
import io
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "3d-semantic-segmentation-fusion_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

METRIC = "fused LiDAR points/sec (project+gather+MLP)"
UNIT = "points/s"
W_KEYS = ("mlp.0.weight", "mlp.0.bias", "mlp.2.weight", "mlp.2.bias", "mlp.4.weight", "mlp.4.bias")
WORKLOADS = {
    # name: (sweeps per GPU, points per sweep, encoder key, description)
    "config2": (8, 34720, "dinov3", "BASELINE configs[1]: 8 sweeps x 34720 pts, 6 cams, DINOv3 ViT-S/16 40x71x384 fp32 tokens, fusion inference"),
    "config4": (1, 350000, "dinov3_hires", "BASELINE configs[3]: 10-sweep cloud 350000 pts, 6 cams, 56x100x384 fp32 tokens"),
    "config1": (1, 34720, "dinov2", "BASELINE configs[0]: 1 sweep x 34720 pts, DINOv2 ViT-S/14 16x16x384 tokens"),
}


def algorithmic_bytes_per_point(V, M, tok_bytes=4, NC=6, Ci=384, Cp=64, C=16):
    """xyz+intensity 16 B + point feature 4*Cp + logits 4*C + the sweep's token maps read once."""
    return 16 + 4 * Cp + 4 * C + NC * M * Ci * tok_bytes / V


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index, period=0.004):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.window = [], None
        self._stop_evt = threading.Event()
        self.max_mhz = None
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self._stop_evt.is_set():
            try:
                self.samples.append((time.perf_counter(), nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM),
                                     int(get_reasons(self.h))))
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()

    def summary(self, t0, t1):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        inside = [s for s in self.samples if t0 <= s[0] <= t1]
        scope = "timed regions (device-resident steps + end-to-end steps)"
        if len(inside) < 3:
            inside, scope = self.samples, "whole run (timed region shorter than 3 samples)"
        mhz = sorted(s[1] for s in inside)
        bits = 0
        for s in inside:
            bits |= s[2]
        names = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
                 0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
                 0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}
        reasons = [n for b, n in names.items() if bits & b and n != "gpu_idle"]
        return {"sm_mhz": mhz[len(mhz) // 2], "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(inside),
                "scope": scope}


# --------------------------------------------------------------------------------------------------
# CPU side: the reference's PyTorch path, ported op for op (oracle/torch_port.py), on the host cores
# --------------------------------------------------------------------------------------------------
def cpu_pass(sw, weights, n_sweeps, n_pts=None):
    """One forward of the reference-equivalent CPU path over the first n_sweeps sweeps (first n_pts points of
    each); returns seconds."""
    from oracle import torch_port as T
    enc = sw.encoder
    n_pts = n_pts or sw.points.shape[1]
    t0 = time.perf_counter()
    with torch.no_grad():
        T.fusion_forward(sw.tokens[:n_sweeps], sw.point_feat[:n_sweeps, :n_pts], sw.points[:n_sweeps, :n_pts],
                         sw.image_sizes[:n_sweeps], sw.K[:n_sweeps], sw.Rt[:n_sweeps], weights, enc.patch_size,
                         enc.resize_height, enc.resize_width)
    return time.perf_counter() - t0


def cpu_baseline(sw, weights, budget_s=20.0):
    """Bounded sample for the own-arm line: 1 warm-up + 3..24 timed passes of the batch (fewer sweeps if slow)."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    B, V = sw.points.shape[0], sw.points.shape[1]
    t1 = cpu_pass(sw, weights, 1)  # warm-up on one sweep, also sizes the sample
    n = max(1, min(B, int(budget_s / 4 / max(t1, 1e-3))))
    times = []
    while len(times) < 3 or (sum(times) < 0.6 * budget_s and len(times) < 40):   # about 12 s of CPU work
        times.append(cpu_pass(sw, weights, n))
    passes = len(times)
    best = min(times)
    return {"value": n * V / best, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"oracle/torch_port.py (CPU port of the reference's ATen op sequence), 1 warm-up sweep + best of {passes} "
                      f"passes over {n} of {B} sweeps x {V} pts ({sum(times):.1f} s), {cores} threads, fp32"}


def run_reference_arm(args, rank):
    """--impl reference: the CPU port on the host cores, same metric/config; rank 0 only."""
    if rank != 0:
        return
    from tdssf import synth
    B, V, enc_key, desc = WORKLOADS[args.workload]
    sw = synth.make_sweeps(B, V, enc_key, seed=1000 * 2)
    weights = synth.make_mlp_weights(0)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    t1 = cpu_pass(sw, weights, 1)
    total = args.steps + args.warmup
    per_step = 150.0 / max(total, 1)                                # whole run within a few minutes
    n = max(1, min(B, int(per_step / max(t1, 1e-3))))
    vs = V if per_step >= t1 else max(1024, int(V * per_step / t1))  # many steps: a slice of one sweep
    for _ in range(args.warmup):
        cpu_pass(sw, weights, n, vs)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_pass(sw, weights, n, vs)
    dt = time.perf_counter() - t0
    value = args.steps * n * vs / dt
    sample = (f"each step = oracle/torch_port.py forward over {n} of {B} sweeps x {vs} of {V} pts (bounded so that "
              f"{total} steps end within minutes), {cores} threads, fp32")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / max(args.steps, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "sweeps_per_gpu": B, "points_per_sweep": V, "tokens": "fp32",
                   "mlp": "fp32 ATen ops on the host cores", "sample_per_step": f"{n} sweeps x {vs} pts",
                   "parallelism": "host CPU, rank 0 only"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# --------------------------------------------------------------------------------------------------
# GPU side
# --------------------------------------------------------------------------------------------------
def run_own_arm(args, rank, world, local_rank):
    from tdssf import ops, synth, _lib
    from utils.fusion_model import FeatureFusionModel

    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    _lib.check(_lib.lib.tdssf_check_device(), "device")

    B, V, enc_key, desc = WORKLOADS[args.workload]
    sw = synth.make_sweeps(B, V, enc_key, seed=1000 * 2 + rank)   # every rank: its own shard of sweeps
    enc = sw.encoder
    weights = synth.make_mlp_weights(0)
    wd = [weights[k].to(dev) for k in W_KEYS]
    packed = ops.pack_weights(*wd)
    pts, K, Rt = sw.points.to(dev), torch.stack(sw.K).to(dev), torch.stack(sw.Rt).to(dev)
    hw, tok, feat = sw.image_sizes.float().to(dev), sw.tokens.to(dev), sw.point_feat.to(dev)
    logits = torch.empty((B, V, 16), dtype=torch.float32, device=dev)
    n_pts = B * V

    def step():
        ops.fusion_fwd(pts, K, Rt, hw, enc, tok, feat, packed, out=logits)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    sampler.start()
    for _ in range(max(args.warmup, 3)):
        step()
    # ---- device-resident throughput: K steps, one CUDA-event pair per launch + one around the region ----
    stream = torch.cuda.current_stream()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.perf_counter()
    r0.record(stream)
    for a, b in ev:
        a.record(stream)
        step()
        b.record(stream)
    r1.record(stream)
    barrier()
    t_wall1 = time.perf_counter()
    region_ms = r0.elapsed_time(r1)
    kernel_ms = sorted(a.elapsed_time(b) for a, b in ev)
    kernel_avg_ms = sum(kernel_ms) / len(kernel_ms)
    if dist is not None:
        t = torch.tensor([region_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        region_ms = float(t.item())
    value = world * n_pts * args.steps / (region_ms * 1e-3)

    # ---- end to end through the drop-in module: pinned host inputs -> logits back on the host ----------
    class Pcd(torch.nn.Module):
        def devoxelize(self, s):
            return s, None

    model = FeatureFusionModel(enc, Pcd(), 64, 384, 16, device=dev).to(dev)
    model.load_state_dict(weights, strict=True)
    h_tok, h_feat, h_pts = sw.tokens.pin_memory(), sw.point_feat.pin_memory(), sw.points.pin_memory()
    h_sizes = sw.image_sizes.pin_memory()
    h_out = torch.empty((B, V, 16), dtype=torch.float32).pin_memory()
    h2d = sum(t.numel() * t.element_size() for t in (h_tok, h_feat, h_pts, h_sizes)) + sum(
        t.numel() * 4 for t in sw.K + sw.Rt)
    d2h = h_out.numel() * 4

    # Two streams in alternation: a step's device->host read of the logits overlaps the next step's host->device copies
    # (PCIe is full duplex); every step still uploads all of its inputs and downloads its result.
    e2e_streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
    h_outs = [h_out, torch.empty_like(h_out).pin_memory()]

    def e2e_step(k):
        with torch.cuda.stream(e2e_streams[k % 2]), torch.no_grad():
            out = model(h_tok, h_feat, h_pts, None, h_sizes, sw.K, sw.Rt)
            h_outs[k % 2].copy_(out, non_blocking=True)

    e2e_steps = max(3, min(args.steps, 50))
    for k in range(4):
        e2e_step(k)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for st_ in e2e_streams:
        st_.wait_stream(stream)
    for k in range(e2e_steps):
        e2e_step(k)
    for st_ in e2e_streams:
        stream.wait_stream(st_)
    e1.record(stream)
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    if dist is not None:
        t = torch.tensor([e2e_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    e2e_value = world * n_pts * e2e_steps / (e2e_ms * 1e-3)
    t_wall2 = time.perf_counter()
    sampler.stop()
    clocks = sampler.summary(t_wall0, t_wall2)   # both timed regions: device-resident steps and end-to-end steps

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    peaks, peak_kind = measured_peaks()
    bpp = algorithmic_bytes_per_point(V, enc.n_patches)
    achieved = bpp * n_pts / (kernel_avg_ms * 1e-3) / 1e9
    flops_pt = 2 * (448 * 256 + 256 * 64 + 64 * 16)
    roofline = {
        "bound": "hbm", "kernel": "tdssf::tc::fused_mlp_kernel<float,true>", "achieved": achieved,
        "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"], "traffic": TRAFFIC_BYTES.get(args.workload),
        "peak_source": f"{peak_kind} MEASURED_PEAKS.json hbm_gbs",
        "algorithmic_bytes_per_point": bpp, "points_per_launch": n_pts, "kernel_avg_ms": kernel_avg_ms,
        "kernel_min_ms": kernel_ms[0],
        "tensor": {"achieved_tflops": flops_pt * n_pts / (kernel_avg_ms * 1e-3) / 1e12, "peak_tflops": peaks["bf16_tflops"],
                   "frac": flops_pt * n_pts / (kernel_avg_ms * 1e-3) / 1e12 / peaks["bf16_tflops"]},
    }
    cpu = cpu_baseline(sw, weights) if world == 1 and not args.no_cpu_baseline else None
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": region_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": desc, "sweeps_per_gpu": B, "points_per_sweep": V, "tokens": "fp32", "mlp": "bf16 operands, fp32 accumulate (tcgen05)",
                   "l2": f"inputs per step {int(bpp * n_pts / 1e6)} MB > 126 MB L2 (no flush needed)", "parallelism": f"sweep-sharded x{world}"},
        "roofline": roofline, "cpu_baseline": cpu,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                "api": "utils.fusion_model.FeatureFusionModel.forward (drop-in), pinned host inputs, logits copied to host; steps alternate between two streams"},
        "gpu_launches": args.steps, "clocks": clocks,
    }
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


# dram__bytes_read.sum + dram__bytes_write.sum per launch of the fused kernel, from the ncu --set full capture
# under profiles/ (None until captured)
TRAFFIC_BYTES = {"config2": 291_800_000}  # profiles/r1_ncu_fused_v7_summary.txt: 249.9 MB read + 41.9 MB written


_REAL_STDOUT = None


def _quiet_stdout():
    """Libraries (NCCL's version banner, for one) write to fd 1; the contract is ONE JSON line on stdout.  Everything
    written to fd 1 from here on goes to stderr, and emit() writes the line to the real stdout."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="tdssf", choices=["tdssf", "reference"])
    ap.add_argument("--workload", default="config2", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    _quiet_stdout()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the tdssf arm has no CPU fallback (use --impl reference for the CPU port)")
    if world == 1 and args.gpus > 1:
        raise SystemExit("launch multi-GPU runs with torch.distributed.run (one rank per GPU)")
    run_own_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
