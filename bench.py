"""bench.py — fused LiDAR points/s through project -> gather -> camera mean -> concat -> MLP -> logits.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload config2|config4|config1]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
           bench.py --gpus N --steps K --warmup W

Headline (`value`, `roofline`, `e2e`): one "step" = one pass of the hot path over one batch of synthetic
nuScenes-shaped sweeps (BASELINE.json configs[1]: 8 sweeps x 34 720 points, 6 cameras, DINOv3 ViT-S/16 40x71x384 fp32
patch tokens, inference).  Every rank owns its own batch (sweep-sharded, weak scaling, no data-path collective);
`value` is the whole-job points/s with inputs resident in HBM, timed with CUDA events, max over ranks.  `e2e` is the
same metric through the drop-in `FeatureFusionModel.forward` with HOST (pinned) inputs and a host read of the logits,
copies inside the timed region.

Sub-records of the same JSON line (each with its own timing, launches per step and roofline entry):
  sustained  the headline launch back to back for >= 2 s (clocks sampled): the figure to hold against the sustained peaks
  tok16      the same batch with bf16 patch tokens (SURVEY f2 hand-off, 712.9 algorithmic B/pt)
  config4    BASELINE configs[3]: one 350 000-point cloud, 56x100 token maps (tensor-bound regime)
  train      BASELINE configs[2]: 4 sweeps per rank, fused forward saving its bf16 rows -> CE + Lovasz-Softmax -> backward ->
             ONE mean all-reduce of the 132 432 MLP gradients when N > 1 (a tdssf kernel over NVLink peer memory; NCCL when the
             peer set-up is not possible)
  eval       BASELINE configs[4]: per-rank shard of sweeps, fused forward writing class ids -> int64 confusion matrix on
             the device, ONE int64 all-reduce at the end, mIoU

`--impl reference` times the reference's own fusion forward (the unmodified utils/fusion_model.py from oracle/_ref/,
else the port oracle/torch_port.py) on the host cores.  Prints ONE JSON line on rank 0.
"""
import argparse
import contextlib
import glob
import io
import json
import os
import re
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
FLOPS_FWD = 2 * (448 * 256 + 256 * 64 + 64 * 16)                 # 264 192 per point
FLOPS_TRAIN = FLOPS_FWD + FLOPS_FWD + 2 * (16 * 64 + 64 * 256 + 256 * 64)  # + wgrad + dgrad (first layer: 64 columns) = 595 968
LOSS_BYTES_PER_POINT = 137                                       # logits 64 + label 8 + mask 1 in, dlogits 64 out


def algorithmic_bytes_per_point(V, M, tok_bytes=4, NC=6, Ci=384, Cp=64, C=16):
    """xyz+intensity 16 B + point feature 4*Cp + logits 4*C + the sweep's token maps read once."""
    return 16 + 4 * Cp + 4 * C + NC * M * Ci * tok_bytes / V


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def ncu_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the fused kernel, read from the newest ncu --set full
    summary of this workload under profiles/ (file name returned next to the value), or (None, None)."""
    pats = {"config2": "profiles/r*_ncu_fused_v*_summary.txt", "config4": "profiles/r*_ncu_fused_config4_v*_summary.txt"}
    files = glob.glob(os.path.join(ROOT, pats.get(workload, "profiles/none")))
    if workload == "config2":
        files = [f for f in files if "config4" not in f]

    def order(f):
        m = re.search(r"r(\d+)_ncu_fused_(?:config4_)?v(\d+)_", os.path.basename(f))
        return (int(m.group(1)), int(m.group(2))) if m else (0, 0)

    for f in sorted(files, key=order, reverse=True):
        text = open(f).read()
        rd = re.search(r"dram__bytes_read\.sum \[Mbyte\]: \['([\d.]+)'\]", text)
        wr = re.search(r"dram__bytes_write\.sum \[Mbyte\]: \['([\d.]+)'\]", text)
        if rd and wr:
            return int((float(rd.group(1)) + float(wr.group(1))) * 1e6), os.path.relpath(f, ROOT)
    return None, None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed regions run."""

    def __init__(self, index, period=0.004):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples = []
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

    def summary(self, t0, t1, scope):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        inside = [s for s in self.samples if t0 <= s[0] <= t1]
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
# CPU side: the reference's own fusion forward on the host cores
# --------------------------------------------------------------------------------------------------
class _IdentityPcd(torch.nn.Module):
    """devoxelize = identity: the metric counts voxel rows, devoxelize is timed separately."""

    def devoxelize(self, s):
        return s, None


class CpuPath:
    """One forward of the reference's CPU path.  kind 'reference': the UNMODIFIED utils/fusion_model.py
    (FeatureFusionModel.forward, utils/fusion_model.py:152-206) loaded by file path from oracle/_ref/ (a verbatim copy
    made by oracle/make_ref.py in the build container) or from the mounted tree; kind 'port': oracle/torch_port.py, the
    same ATen op sequence restated, when neither exists."""

    def __init__(self, sw, weights):
        from oracle import ref_loader
        self.sw, self.weights = sw, weights
        root = ref_loader.ref_root()
        self.kind = ref_loader.kind(root)
        if root is not None:
            ref = ref_loader.load_hot_path(root)
            self.what = ("the reference's own FeatureFusionModel.forward (utils/fusion_model.py, unmodified, from "
                         + ("oracle/_ref" if "_ref" in root else root) + ")")
            self.model = ref.fusion_model.FeatureFusionModel(sw.encoder, _IdentityPcd(), 64, 384, 16, device="cpu")
            self.model.load_state_dict(weights, strict=True)
            self.model.eval()
        else:
            self.what = "oracle/torch_port.py (CPU port of the reference's ATen op sequence)"
            self.model = None

    def run(self, n_sweeps, n_pts=None):
        """Forward over the first n_sweeps sweeps (first n_pts points of each); returns seconds."""
        sw = self.sw
        n_pts = n_pts or sw.points.shape[1]
        args = (sw.tokens[:n_sweeps], sw.point_feat[:n_sweeps, :n_pts], sw.points[:n_sweeps, :n_pts])
        t0 = time.perf_counter()
        with torch.no_grad():
            if self.model is not None:
                with contextlib.redirect_stdout(io.StringIO()):   # utils/fusion_model.py:155 prints on every call
                    self.model(*args, None, sw.image_sizes[:n_sweeps], sw.K[:n_sweeps], sw.Rt[:n_sweeps])
            else:
                from oracle import torch_port as T
                enc = sw.encoder
                T.fusion_forward(*args, sw.image_sizes[:n_sweeps], sw.K[:n_sweeps], sw.Rt[:n_sweeps], self.weights,
                                 enc.patch_size, enc.resize_height, enc.resize_width)
        return time.perf_counter() - t0


def cpu_baseline(sw, weights, budget_s=16.0):
    """Bounded sample for the own-arm line: 1 warm-up + 3.. timed passes of the batch (fewer sweeps if slow) on all
    host cores, then a 1-thread figure on one sweep."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    path = CpuPath(sw, weights)
    B, V = sw.points.shape[0], sw.points.shape[1]
    t1 = path.run(1)  # warm-up on one sweep, also sizes the sample
    n = max(1, min(B, int(budget_s / 4 / max(t1, 1e-3))))
    times = []
    while len(times) < 3 or (sum(times) < 0.6 * budget_s and len(times) < 40):
        times.append(path.run(n))
    best = min(times)
    torch.set_num_threads(1)
    path.run(1, V // 4)
    one = min(path.run(1, V // 4) for _ in range(2))
    torch.set_num_threads(cores)
    return {"value": n * V / best, "unit": UNIT, "cores": cores, "kind": path.kind,
            "one_thread_value": (V // 4) / one,
            "sample": f"{path.what}, 1 warm-up sweep + best of {len(times)} passes over {n} of {B} sweeps x {V} pts "
                      f"({sum(times):.1f} s), {cores} threads, fp32; one_thread_value: best of 2 passes over {V // 4} pts of one sweep"}


def run_reference_arm(args, rank):
    """--impl reference: the reference's CPU forward on the host cores, same metric/config; rank 0 only."""
    if rank != 0:
        return
    from tdssf import synth
    B, V, enc_key, desc = WORKLOADS[args.workload]
    sw = synth.make_sweeps(B, V, enc_key, seed=1000 * 2)
    weights = synth.make_mlp_weights(0)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    path = CpuPath(sw, weights)
    t1 = path.run(1)
    total = args.steps + args.warmup
    per_step = 150.0 / max(total, 1)                                # whole run within a few minutes
    n = max(1, min(B, int(per_step / max(t1, 1e-3))))
    vs = V if per_step >= t1 else max(1024, int(V * per_step / t1))  # many steps: a slice of one sweep
    for _ in range(args.warmup):
        path.run(n, vs)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        path.run(n, vs)
    dt = time.perf_counter() - t0
    value = args.steps * n * vs / dt
    sample = (f"each step = {path.what} over {n} of {B} sweeps x {vs} of {V} pts (bounded so that "
              f"{total} steps end within minutes), {cores} threads, fp32")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / max(args.steps, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "sweeps_per_gpu": B, "points_per_sweep": V, "tokens": "fp32",
                   "parallelism": "host CPU, rank 0 only"},
        "sample_per_step": f"{n} sweeps x {vs} pts",
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": path.kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# --------------------------------------------------------------------------------------------------
# GPU side
# --------------------------------------------------------------------------------------------------
class Timer:
    """CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks."""

    def __init__(self, dist, dev):
        self.dist, self.dev = dist, dev

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        torch.cuda.synchronize()

    def max_ranks(self, ms):
        if self.dist is None:
            return ms
        t = torch.tensor([ms], device=self.dev, dtype=torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def region(self, fn, steps, per_launch=False):
        """(region ms max over ranks, sorted per-launch ms on this rank or None, wall t0, wall t1)."""
        stream = torch.cuda.current_stream()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)] if per_launch else None
        r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        w0 = time.perf_counter()
        r0.record(stream)
        for k in range(steps):
            if ev:
                ev[k][0].record(stream)
            fn(k)
            if ev:
                ev[k][1].record(stream)
        r1.record(stream)
        self.barrier()
        w1 = time.perf_counter()
        ms = self.max_ranks(r0.elapsed_time(r1))
        each = sorted(a.elapsed_time(b) for a, b in ev) if ev else None
        return ms, each, w0, w1


def count_kernels(fn, n=3):
    """Kernel launches per call of `fn` from CUPTI activity records (torch.profiler), None if unavailable."""
    done = 0
    try:
        from torch.profiler import profile, ProfilerActivity
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for k in range(n):
                fn(k)
                done += 1
            torch.cuda.synchronize()
    except Exception as exc:  # profiler not usable on this box: the claim is then the static count
        for k in range(done, n):   # `fn` may hold a collective: every rank runs it exactly n times whatever happens
            fn(k)
        return {"error": repr(exc)[:120]}
    try:
        names = [e.name for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
        kernels = [x for x in names if "memcpy" not in x.lower() and "memset" not in x.lower()]
        return {"kernels_per_step": len(kernels) // n, "memcpy_memset_per_step": (len(names) - len(kernels)) // n,
                "own_kernels_per_step": sum("tdssf" in x for x in kernels) // n,
                "library_kernels": sorted({x[:60] for x in kernels if "tdssf" not in x})[:12]}
    except Exception as exc:  # profiler not usable on this box: the claim is then the static count
        return {"error": repr(exc)[:120]}


def fused_roofline(kernel_name, bpp, n_pts, kernel_ms, peaks, peak_kind, traffic, sustained=False):
    achieved = bpp * n_pts / (kernel_ms * 1e-3) / 1e9
    tf = FLOPS_FWD * n_pts / (kernel_ms * 1e-3) / 1e12
    tf_peak = peaks["bf16_tflops_sustained" if sustained else "bf16_tflops"]
    return {"bound": "hbm", "kernel": kernel_name, "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
            "frac": achieved / peaks["hbm_gbs"], "traffic": traffic[0], "traffic_source": traffic[1],
            "peak_source": f"{peak_kind} MEASURED_PEAKS.json hbm_gbs",
            "algorithmic_bytes_per_point": bpp, "points_per_launch": n_pts, "kernel_avg_ms": kernel_ms,
            "tensor": {"achieved_tflops": tf, "peak_tflops": tf_peak, "frac": tf / tf_peak,
                       "peak_source": "bf16_tflops_sustained" if sustained else "bf16_tflops"}}


def run_own_arm(args, rank, world, local_rank):
    from tdssf import ops, parallel, synth, _lib
    from utils.fusion_model import FeatureFusionModel
    from utils.losses import CELSLoss
    from utils.evaluation import compute_iou

    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    _lib.check(_lib.lib.tdssf_check_device(), "device")
    timer = Timer(dist, dev)
    peaks, peak_kind = measured_peaks()

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
    warm = max(args.warmup, 3)

    def step(_k=0):
        ops.fusion_fwd(pts, K, Rt, hw, enc, tok, feat, packed, out=logits)

    sampler = ClockSampler(local_rank)
    sampler.start()
    for _ in range(warm):
        step()
    # untimed rehearsal of the timed region (part of the warm-up): the first barrier / all-reduce of a process group sets up
    # the NCCL communicator, whose helper threads otherwise compete with the launch loop of the first timed region for the
    # host cores (seen at 8 ranks: a GPU-idle first region while every later region scaled)
    timer.region(step, min(args.steps, 20), per_launch=True)
    launches_rehearsal = min(args.steps, 20)
    # ---- headline: device-resident throughput, K steps, one CUDA-event pair per launch + one around the region ----
    region_ms, kernel_ms, t_head0, t_head1 = timer.region(step, args.steps, per_launch=True)
    kernel_avg_ms = sum(kernel_ms) / len(kernel_ms)
    value = world * n_pts * args.steps / (region_ms * 1e-3)
    launches = args.steps   # (timed regions only; the rehearsal's `launches_rehearsal` launches are warm-up)

    # ---- sustained: the same launch back to back for >= 2 s ------------------------------------------------------------
    sus_chunk = max(200, int(0.25 / (kernel_avg_ms * 1e-3)))
    sus_ms, sus_n = 0.0, 0
    t_sus0 = time.perf_counter()
    while sus_ms < 2000.0 and sus_n < 400 * sus_chunk:
        ms, _, _, _ = timer.region(step, sus_chunk)
        sus_ms += ms
        sus_n += sus_chunk
    t_sus1 = time.perf_counter()
    launches += sus_n
    sus_kernel_ms = sus_ms / sus_n

    # ---- end to end through the drop-in module: pinned host inputs -> logits back on the host ----------
    model = FeatureFusionModel(enc, _IdentityPcd(), 64, 384, 16, device=dev).to(dev)
    model.load_state_dict(weights, strict=True)
    h_tok, h_feat, h_pts = sw.tokens.pin_memory(), sw.point_feat.pin_memory(), sw.points.pin_memory()
    h_sizes = sw.image_sizes.pin_memory()
    h_out = torch.empty((B, V, 16), dtype=torch.float32).pin_memory()
    h2d = sum(t.numel() * t.element_size() for t in (h_tok, h_feat, h_pts, h_sizes)) + sum(
        t.numel() * 4 for t in sw.K + sw.Rt)
    d2h = h_out.numel() * 4

    # Two streams in alternation: a step's device->host read of the logits overlaps the next step's host->device copies
    # (PCIe is full duplex); every step still uploads all of its inputs and downloads its result.
    stream = torch.cuda.current_stream()
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
    timer.barrier()
    t_e2e0 = time.perf_counter()
    e0.record(stream)
    for st_ in e2e_streams:
        st_.wait_stream(stream)
    for k in range(e2e_steps):
        e2e_step(k)
    for st_ in e2e_streams:
        stream.wait_stream(st_)
    e1.record(stream)
    timer.barrier()
    t_e2e1 = time.perf_counter()
    e2e_ms = timer.max_ranks(e0.elapsed_time(e1))
    e2e_value = world * n_pts * e2e_steps / (e2e_ms * 1e-3)
    launches += e2e_steps + 4
    # the ceiling of that step: nothing but the host->device copies of the same pinned buffers, all ranks at once
    d_tok, d_feat, d_pts = torch.empty_like(tok), torch.empty_like(feat), torch.empty_like(pts)

    def copy_only(_k):
        d_tok.copy_(h_tok, non_blocking=True)
        d_feat.copy_(h_feat, non_blocking=True)
        d_pts.copy_(h_pts, non_blocking=True)

    copy_only(0)
    copy_ms, _, _, _ = timer.region(copy_only, 5)
    h2d_gbs = 5 * h2d / (copy_ms * 1e-3) / 1e9
    del d_tok, d_feat, d_pts, model

    sub = {}
    if args.workload == "config2" and not args.headline_only:
        sub = sub_records(args, rank, world, dev, timer, peaks, peak_kind, sw, wd, packed, weights, pts, K, Rt, hw, tok, feat)
        launches += sub.pop("_launches", 0)
    sampler.stop()

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    bpp = algorithmic_bytes_per_point(V, enc.n_patches)
    traffic = ncu_traffic(args.workload)
    kname = "tdssf::tc::fused_mlp_kernel<float,true,false>"
    roofline = fused_roofline(kname, bpp, n_pts, kernel_avg_ms, peaks, peak_kind, traffic)
    roofline["kernel_min_ms"] = kernel_ms[0]
    sustained = {
        "launches": sus_n, "seconds": sus_ms * 1e-3, "kernel_avg_ms": sus_kernel_ms,
        "value": world * n_pts / (sus_kernel_ms * 1e-3), "unit": UNIT,
        "roofline": fused_roofline(kname, bpp, n_pts, sus_kernel_ms, peaks, peak_kind, traffic, sustained=True),
        "clocks": sampler.summary(t_sus0, t_sus1, "sustained region"),
    }
    cpu = cpu_baseline(sw, weights) if world == 1 and not args.no_cpu_baseline else None
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
        "ms_per_step": region_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": desc, "sweeps_per_gpu": B, "points_per_sweep": V, "tokens": "fp32", "mlp": "bf16 operands, fp32 accumulate (tcgen05)",
                   "l2": f"inputs per step {int(bpp * n_pts / 1e6)} MB > 126 MB L2 (no flush needed)", "parallelism": f"sweep-sharded x{world}",
                   "warmup_detail": f"{warm} launches + an untimed rehearsal of {launches_rehearsal} launches through the timed code path"},
        "roofline": roofline, "sustained": sustained, "cpu_baseline": cpu,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                "ms_per_step": e2e_ms / e2e_steps,
                "h2d_copy_only_gbs_per_rank": h2d_gbs,
                "h2d_ceiling_points_per_s": world * n_pts / (copy_ms / 5 * 1e-3),
                "api": "utils.fusion_model.FeatureFusionModel.forward (drop-in), pinned host inputs, logits copied to host; steps alternate between two streams; "
                       "h2d_copy_only = the same pinned buffers copied host->device by all ranks at once with no kernel (the ceiling of this step)"},
        "gpu_launches": launches,
        "clocks": sampler.summary(t_head0, t_e2e1, "headline + sustained + end-to-end regions"),
        "clocks_headline": sampler.summary(t_head0, t_head1, "headline region"),
    }
    line.update(sub)
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


def sub_records(args, rank, world, dev, timer, peaks, peak_kind, sw, wd, packed, weights, pts, K, Rt, hw, tok, feat):
    """tok16 / config4 / train / eval sub-records (see the module docstring); returns them keyed by name."""
    from tdssf import ops, parallel, synth
    from utils.fusion_model import FeatureFusionModel
    from utils.losses import CELSLoss
    from utils.evaluation import compute_iou
    enc = sw.encoder
    B, V = sw.points.shape[0], sw.points.shape[1]
    steps = max(5, min(args.steps, 50))
    out = {}
    launches = 0

    # ---- tok16: bf16 patch tokens (the f2 hand-off) -------------------------------------------------------------------
    tok16 = sw.tokens.to(torch.bfloat16).to(dev)
    logits = torch.empty((B, V, 16), dtype=torch.float32, device=dev)

    def step16(_k=0):
        ops.fusion_fwd(pts, K, Rt, hw, enc, tok16, feat, packed, out=logits)

    for _ in range(3):
        step16()
    ms, each, _, _ = timer.region(step16, steps, per_launch=True)
    launches += steps + 3
    k16 = sum(each) / len(each)
    bpp16 = algorithmic_bytes_per_point(V, enc.n_patches, tok_bytes=2)
    out["tok16"] = {"workload": "configs[1] with bf16 patch tokens (SURVEY f2 hand-off)", "value": world * B * V * steps / (ms * 1e-3),
                    "unit": UNIT, "steps": steps,
                    "roofline": fused_roofline("tdssf::tc::fused_mlp_kernel<__nv_bfloat16,true,false>", bpp16, B * V, k16, peaks,
                                               peak_kind, (None, None))}
    del tok16

    # ---- config4: one 350 000-point cloud, 56 x 100 token maps ---------------------------------------------------------
    B4, V4, key4, desc4 = WORKLOADS["config4"]
    s4 = synth.make_sweeps(B4, V4, key4, seed=4000 + rank)
    p4, K4, R4 = s4.points.to(dev), torch.stack(s4.K).to(dev), torch.stack(s4.Rt).to(dev)
    h4, t4, f4 = s4.image_sizes.float().to(dev), s4.tokens.to(dev), s4.point_feat.to(dev)
    l4 = torch.empty((B4, V4, 16), dtype=torch.float32, device=dev)

    def step4(_k=0):
        ops.fusion_fwd(p4, K4, R4, h4, s4.encoder, t4, f4, packed, out=l4)

    for _ in range(3):
        step4()
    ms, each, _, _ = timer.region(step4, steps, per_launch=True)
    launches += steps + 3
    k4 = sum(each) / len(each)
    r4 = fused_roofline("tdssf::tc::fused_mlp_kernel<float,true,false>", algorithmic_bytes_per_point(V4, s4.encoder.n_patches),
                        B4 * V4, k4, peaks, peak_kind, ncu_traffic("config4"))
    r4["bound"] = "tensor"   # 483.5 B/pt against 264 192 FLOP/pt: the tensor floor is above the HBM floor
    out["config4"] = {"workload": desc4, "value": world * B4 * V4 * steps / (ms * 1e-3), "unit": UNIT, "steps": steps,
                      "roofline": r4}
    del p4, t4, f4, l4, s4

    # ---- train: configs[2], 4 sweeps per rank, forward + CE/Lovasz + backward + gradient all-reduce --------------------
    PER = 4
    st = synth.make_sweeps(PER, V, "dinov3", seed=3000 + rank)
    model = FeatureFusionModel(st.encoder, _IdentityPcd(), 64, 384, 16, device=dev).to(dev)
    model.load_state_dict(weights, strict=True)
    crit = CELSLoss(weight=torch.tensor(synth.CLASS_WEIGHTS, device=dev), ignore_index=0, lamda_lovasz=1.0)
    d_tok, d_pts = st.tokens.to(dev), st.points.to(dev)
    d_feat = st.point_feat.to(dev).requires_grad_(True)       # PTv3 output: needs its gradient
    d_lab, d_mask = st.labels.to(dev), st.mask.to(dev)
    params = list(model.mlp.parameters())

    def train_step(_k=0):
        for p in params:
            p.grad = None
        d_feat.grad = None
        scores = model(d_tok, d_feat, d_pts, None, st.image_sizes, st.K, st.Rt)
        total = crit(scores, d_lab, d_mask)[0]
        total.backward()
        parallel.allreduce_mean_grads(params)
        return total

    for _ in range(10):   # the loss sequence is captured into a CUDA graph once its buffers repeat
        loss = train_step()
    ms, _, _, _ = timer.region(train_step, steps)
    step_ms = ms / steps
    counts = count_kernels(train_step)   # on EVERY rank: the step holds a collective, so all ranks must run it equally often
    kps = (counts or {}).get("kernels_per_step")
    launches += (steps + 10 + 3) * (kps or 45)
    import ctypes as C
    from tdssf import _lib
    gs = [C.c_int64(), C.c_int64(), C.c_int64()]
    _lib.lib.tdssf_cels_graph_stats(*[C.byref(g) for g in gs])
    grad_path = "none"
    if world > 1:   # which branch tdssf.parallel.allreduce_mean_grads took for this bucket (decided collectively at its first call)
        ex = parallel._peer_exchange.get((dev.index, sum(p.numel() for p in params)))
        grad_path = "one tdssf kernel over NVLink peer memory, csrc/p2p.cu" if ex is not None and ex.ok else "ncclAllReduce"
    tf = FLOPS_TRAIN * PER * V / (step_ms * 1e-3) / 1e12
    out["train"] = {
        "workload": f"BASELINE configs[2]: {PER} sweeps x {V} pts per GPU, fused forward (saving bf16 rows) + CE/Lovasz-Softmax + backward"
                    + (f" + all-reduce(mean) of the 132432 MLP gradients x{world} ({grad_path})" if world > 1 else " (no all-reduce at N=1)"),
        "value": world * PER * V / (step_ms * 1e-3), "unit": UNIT, "steps": steps, "ms_per_step": step_ms,
        "loss": float(loss.detach()), "launches": counts, "grad_allreduce": grad_path,
        "cels_graph": {"replays": gs[0].value, "captures": gs[1].value, "plain_sequences": gs[2].value},
        "roofline": {"bound": "tensor", "achieved": tf, "peak": peaks["bf16_tflops"], "unit": "TFLOP/s", "frac": tf / peaks["bf16_tflops"],
                     "flops_per_point": FLOPS_TRAIN,
                     "loss_hbm": {"algorithmic_bytes_per_point": LOSS_BYTES_PER_POINT,
                                  "note": "CE/Lovasz compulsory traffic; the sort passes are latency-bound integer work (DESIGN.md §4.2)"}},
    }
    del model, d_tok, d_feat

    # ---- eval: configs[4], per-rank shard, class ids out of the fused launch -> confusion on the device -> one all-reduce ----
    EVAL_BATCHES = 12
    labels = sw.labels.to(dev)
    preds = torch.empty((B, V), dtype=torch.int64, device=dev)
    cm = torch.zeros((16, 16), dtype=torch.int64, device=dev)

    def eval_batch(_k=0):
        ops.fusion_fwd(pts, K, Rt, hw, enc, tok, feat, packed, out=logits, preds_out=preds)
        ops.confusion(preds, labels, 16, out=cm)   # label 0 (noise) lies outside 1..16 and is dropped by the kernel

    for _ in range(2):
        eval_batch()
    cm.zero_()

    def eval_pass(_k=0):
        for b in range(EVAL_BATCHES):
            eval_batch(b)
        parallel.allreduce_confusion(cm)

    ms, _, _, _ = timer.region(eval_pass, 1)
    launches += 2 * (EVAL_BATCHES + 2)
    _, miou = compute_iou(cm)
    labelled = int((sw.labels != 0).sum())
    out["eval"] = {
        "workload": f"BASELINE configs[4]: {EVAL_BATCHES} batches x {B} sweeps x {V} pts per GPU, fused forward writing class ids + "
                    f"int64 confusion on the device, one int64 all-reduce over {world} rank(s) at the end",
        "value": world * EVAL_BATCHES * B * V / (ms * 1e-3), "unit": UNIT, "ms_total": ms,
        "ms_per_batch": ms / EVAL_BATCHES, "launches_per_batch": 2, "mean_iou": float(miou),
        "confusion_total": int(cm.sum()),
        "confusion_total_expected_rank0_share": EVAL_BATCHES * labelled,
    }
    out["_launches"] = launches
    return out


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
    ap.add_argument("--headline-only", action="store_true", help="skip the tok16 / config4 / train / eval sub-records")
    args = ap.parse_args()
    _quiet_stdout()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the tdssf arm has no CPU fallback (use --impl reference for the CPU path)")
    if world == 1 and args.gpus > 1:
        raise SystemExit("launch multi-GPU runs with torch.distributed.run (one rank per GPU)")
    run_own_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
