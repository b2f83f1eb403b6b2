"""GPU parity tests of the tcgen05 (bf16 operands, fp32 accumulate) kernels.  Tolerance 1e-2 relative to the
largest |logit| (north_star's bf16-MLP bound); integer by-products (camera records) are bit-exact."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import np_oracle as O  # noqa: E402
from tdssf import synth  # noqa: E402

BF16_TOL = 1e-2
W_KEYS = ("mlp.0.weight", "mlp.0.bias", "mlp.2.weight", "mlp.2.bias", "mlp.4.weight", "mlp.4.bias")


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from tdssf import ops as _ops, _lib
    _lib.check(_lib.lib.tdssf_check_device(), "device")
    return _ops


def _rel(a, b):
    return float(np.abs(a - b).max() / np.abs(b).max())


def _bf16_round(x):
    return torch.from_numpy(np.asarray(x, dtype=np.float32)).bfloat16().float().numpy()


def _oracle_bf16_mlp(x_pt, x_img, w):
    """fp64 MLP on bf16-rounded operands with bf16-rounded hidden activations: what the kernel computes."""
    wq = {k: (_bf16_round(v) if k.endswith("weight") else np.asarray(v, np.float32)) for k, v in w.items()}
    x = np.concatenate([_bf16_round(x_pt), _bf16_round(x_img)], -1).astype(np.float64)
    h1 = _bf16_round(np.maximum(x @ wq["mlp.0.weight"].T.astype(np.float64) + wq["mlp.0.bias"], 0)).astype(np.float64)
    h2 = _bf16_round(np.maximum(h1 @ wq["mlp.2.weight"].T.astype(np.float64) + wq["mlp.2.bias"], 0)).astype(np.float64)
    return h2 @ wq["mlp.4.weight"].T.astype(np.float64) + wq["mlp.4.bias"]


@pytest.mark.parametrize("n", [1, 100, 256, 257, 5000, 40_000])
def test_mlp_tc_dense(ops, n):
    gen = torch.Generator().manual_seed(n)
    x_pt = torch.randn(n, 64, generator=gen)
    x_img = torch.randn(n, 384, generator=gen)
    w = synth.make_mlp_weights(3)
    packed = ops.pack_weights(*[w[k].cuda() for k in W_KEYS])
    out = ops.mlp_tc(x_pt.cuda(), x_img.cuda(), packed)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    wn = {k: v.numpy() for k, v in w.items()}
    exact = _oracle_bf16_mlp(x_pt.numpy(), x_img.numpy(), wn)
    assert _rel(got, exact) < 2e-3, "kernel disagrees with a bf16-operand model of itself"
    assert _rel(got, O.mlp_forward(x_pt.numpy(), x_img.numpy(), wn)) < BF16_TOL


@pytest.mark.parametrize("name", ["fusion_dinov3", "fusion_dinov2", "fusion_hires", "fusion_fp16tok"])
def test_fused_forward_golden(ops, golden_dir, name):
    g = np.load(os.path.join(golden_dir, name + ".npz"), allow_pickle=False)
    sw = synth.make_sweeps(int(g["B"]), int(g["V"]), str(g["encoder"]), seed=int(g["seed"]),
                           tok_dtype=getattr(torch, str(g["tok_dtype"])))
    w = synth.make_mlp_weights(int(g["seed"]))
    packed = ops.pack_weights(*[w[k].cuda() for k in W_KEYS])
    logits, rec = ops.fusion_fwd(sw.points.cuda(), torch.stack(sw.K).cuda(), torch.stack(sw.Rt).cuda(),
                                 sw.image_sizes.float().cuda(), sw.encoder, sw.tokens.cuda(), sw.point_feat.cuda(),
                                 packed, want_rec=True)
    torch.cuda.synchronize()
    assert _rel(logits.cpu().numpy(), g["voxel_scores"]) < BF16_TOL
    # the records the fused kernel computes inline are the stand-alone projector's, bit for bit
    ref = ops.project(sw.points.cuda(), torch.stack(sw.K).cuda(), torch.stack(sw.Rt).cuda(),
                      sw.image_sizes.float().cuda(), sw.encoder, want_pix=False, want_depth=False, want_valid=False,
                      want_rec=True)["rec"]
    assert torch.equal(rec, ref)


def test_dropin_forward_bf16_default(ops, golden_dir, monkeypatch):
    from utils.fusion_model import FeatureFusionModel
    monkeypatch.delenv("TDSSF_MLP", raising=False)
    g = np.load(os.path.join(golden_dir, "fusion_dinov3.npz"), allow_pickle=False)
    sw = synth.make_sweeps(2, 4096, "dinov3", seed=1)

    class Pcd(torch.nn.Module):
        def devoxelize(self, s):
            return s, None

    model = FeatureFusionModel(sw.encoder, Pcd(), device="cuda").cuda()
    model.load_state_dict(synth.make_mlp_weights(1), strict=True)
    with torch.no_grad():
        out = model(sw.tokens, sw.point_feat.cuda(), sw.points.cuda(), None, sw.image_sizes.cuda(), sw.K, sw.Rt)
    assert _rel(out.cpu().numpy(), g["voxel_scores"]) < BF16_TOL
    # an optimizer step invalidates the packed bf16 image
    with torch.no_grad():
        model.mlp[4].bias.add_(1.0)
        out2 = model(sw.tokens, sw.point_feat.cuda(), sw.points.cuda(), None, sw.image_sizes.cuda(), sw.K, sw.Rt)
    assert abs(float((out2 - out).mean()) - 1.0) < 1e-3


def _cpu_reference(sw, w):
    """Logits, flat patch indices and validity of the CPU path at full size: the reference's own FeatureFusionModel.forward
    where the reference tree or its verbatim copy oracle/_ref/ is present (the GPU box has the copy), and always
    oracle/torch_port.py (pinned bit-exactly to the reference-generated goldens), which also returns the parts."""
    import contextlib, io
    from oracle import ref_loader, torch_port as T
    enc = sw.encoder
    with torch.no_grad():
        logits, _, flat, valid, _ = T.fusion_forward(sw.tokens.float(), sw.point_feat, sw.points, sw.image_sizes, sw.K, sw.Rt, w,
                                                     enc.patch_size, enc.resize_height, enc.resize_width, return_parts=True)
        if ref_loader.ref_root() is not None:
            class Pcd(torch.nn.Module):
                def devoxelize(self, s):
                    return s, None
            ref = ref_loader.load_hot_path()
            model = ref.fusion_model.FeatureFusionModel(enc, Pcd(), 64, 384, 16, device="cpu")
            model.load_state_dict(w, strict=True)
            with contextlib.redirect_stdout(io.StringIO()):
                own = model(sw.tokens.float(), sw.point_feat, sw.points, None, sw.image_sizes, sw.K, sw.Rt)
            assert torch.equal(own, logits), "oracle/torch_port.py no longer reproduces the reference's forward bit for bit"
    return logits.numpy(), flat.numpy(), valid.numpy()


def _check_records(rec, flat, valid):
    """(B,V,4) int32 camera records against the CPU path's flat indices (B,NC,V) and validity (B,NC,V): bit-exact."""
    r = rec.cpu().numpy().view(np.uint16).reshape(rec.shape[0], rec.shape[1], 8)
    nc = flat.shape[1]
    assert np.array_equal(r[..., :nc].astype(np.int64), flat.transpose(0, 2, 1))
    mask = (r[..., 6] & 0xff).astype(np.int64)
    want = sum(valid[:, c].astype(np.int64) << c for c in range(nc))
    assert np.array_equal(mask, want)
    assert np.array_equal((r[..., 6] >> 8).astype(np.int64), valid.sum(1))


@pytest.mark.parametrize("enc_key,B", [("dinov2", 1), ("dinov3", 8)])
def test_fused_full_size_vs_cpu_reference(ops, enc_key, B):
    """BASELINE configs[0] (1 sweep, DINOv2 16x16 tokens) and configs[1] (8 sweeps, DINOv3 40x71) at V = 34 720: fused
    logits against the reference's CPU forward at north_star's bf16 tolerance, camera records bit-exact."""
    sw = synth.make_sweeps(B, 34720, enc_key, seed=2100 + B)
    w = synth.make_mlp_weights(2100 + B)
    ref_logits, flat, valid = _cpu_reference(sw, w)
    packed = ops.pack_weights(*[w[k].cuda() for k in W_KEYS])
    fused, rec = ops.fusion_fwd(sw.points.cuda(), torch.stack(sw.K).cuda(), torch.stack(sw.Rt).cuda(), sw.image_sizes.float().cuda(),
                                sw.encoder, sw.tokens.cuda(), sw.point_feat.cuda(), packed, want_rec=True)
    assert _rel(fused.cpu().numpy(), ref_logits) < BF16_TOL
    _check_records(rec, flat, valid)


def test_fused_full_size_config2(ops):
    """BASELINE config 2 at full size: fused tcgen05 kernel vs the unfused fp32 kernels (already parity-checked)."""
    sw = synth.make_sweeps(8, 34720, "dinov3", seed=2002)
    w = synth.make_mlp_weights(2002)
    wd = [w[k].cuda() for k in W_KEYS]
    packed = ops.pack_weights(*wd)
    pts, K, Rt, hw = sw.points.cuda(), torch.stack(sw.K).cuda(), torch.stack(sw.Rt).cuda(), sw.image_sizes.float().cuda()
    tok, feat = sw.tokens.cuda(), sw.point_feat.cuda()
    fused = ops.fusion_fwd(pts, K, Rt, hw, sw.encoder, tok, feat, packed)
    rec = ops.project(pts, K, Rt, hw, sw.encoder, want_pix=False, want_depth=False, want_valid=False, want_rec=True)["rec"]
    ref = ops.mlp_f32(feat, ops.gather_mean(tok, rec), *wd)
    assert _rel(fused.cpu().numpy(), ref.cpu().numpy()) < BF16_TOL
    # idempotence: a second launch into the same buffers gives identical bits (no stale TMEM / barrier state)
    again = ops.fusion_fwd(pts, K, Rt, hw, sw.encoder, tok, feat, packed)
    assert torch.equal(fused, again)


def test_fused_full_size_config4_accumulated_cloud(ops):
    """BASELINE configs[3] at full size: ten sweeps (each with its own sensor / ego pose) accumulated into one 350 000-point
    cloud by tdssf_accumulate_sweeps, hi-res 56 x 100 token maps; fused kernel vs the unfused fp32 kernels on that cloud,
    and the cloud itself against the oracle's float64 transform."""
    from oracle import np_oracle as O
    S, P = 10, 35000
    sw = synth.make_sweeps(1, S * P, "dinov3_hires", seed=4004)
    rng = np.random.default_rng(4004)
    raw = sw.points[0].reshape(S, P, 4)
    q = rng.normal(size=(S, 4)) * 0.02 + np.array([1.0, 0, 0, 0]); q /= np.linalg.norm(q, axis=1, keepdims=True)
    ego = np.concatenate([q, np.cumsum(rng.normal(size=(S, 3)) * np.array([0.5, 0.1, 0.01]), 0) + np.array([600.0, 900.0, 0])], 1)
    sensor = np.tile(np.array([1.0, 0, 0, 0, 0.9, 0.0, 1.8]), (S, 1))
    cloud, total, T = ops.accumulate_sweeps(raw.cuda(), [P] * S, sensor, ego, sensor[0], ego[0])
    assert int(total.item()) == S * P
    ref_cloud = O.accumulate_sweeps(raw.numpy(), [P] * S, O.compose_sweep2ref(sensor, ego, sensor[0], ego[0]))
    got = cloud.cpu().numpy()
    def _ord(a):   # fp32 bits -> integers in value order (sign-magnitude to two's complement)
        i = np.ascontiguousarray(a).view(np.int32).astype(np.int64)
        return np.where(i < 0, -(i & 0x7fffffff), i)
    ulp = np.abs(_ord(got[:, :3]) - _ord(ref_cloud[:, :3]))
    assert np.array_equal(got[:, 3], ref_cloud[:, 3]) and ulp.max() <= 2 and (ulp == 0).mean() >= 0.99
    w = synth.make_mlp_weights(4004)
    wd = [w[k].cuda() for k in W_KEYS]
    packed = ops.pack_weights(*wd)
    pts = cloud.reshape(1, S * P, 4)
    K, Rt, hw = torch.stack(sw.K).cuda(), torch.stack(sw.Rt).cuda(), sw.image_sizes.float().cuda()
    tok, feat = sw.tokens.cuda(), sw.point_feat.cuda()
    fused = ops.fusion_fwd(pts, K, Rt, hw, sw.encoder, tok, feat, packed)
    rec = ops.project(pts, K, Rt, hw, sw.encoder, want_pix=False, want_depth=False, want_valid=False, want_rec=True)["rec"]
    ref = ops.mlp_f32(feat, ops.gather_mean(tok, rec), *wd)
    assert _rel(fused.cpu().numpy(), ref.cpu().numpy()) < BF16_TOL
    assert int((rec.view(torch.uint8).reshape(1, S * P, 16)[..., 13] > 0).sum()) > S * P // 4   # a real share of the cloud is seen
    # ... and against the reference's CPU forward on the same accumulated cloud (350 000 points, 56 x 100 maps)
    import copy
    sw_cpu = copy.copy(sw)
    sw_cpu.points = pts.cpu()
    ref_logits, flat, valid = _cpu_reference(sw_cpu, w)
    assert _rel(fused.cpu().numpy(), ref_logits) < BF16_TOL
    _check_records(rec, flat, valid)


@pytest.mark.parametrize("same", [6, 3])
def test_fused_overlapping_cameras_and_edge_points(ops, same):
    """Calibrations in which `same` cameras coincide (points seen by up to six cameras: the general camera mean and the
    all-cameras projection path of the fused kernel), plus points on image borders, on / behind the image plane and
    non-finite points.  Records bit-exact against the stand-alone projector; logits against the unfused fp32 route."""
    B, V = 2, 3001
    sw = synth.make_sweeps(B, V, "dinov3", seed=77)
    K, Rt = torch.stack(sw.K).clone(), torch.stack(sw.Rt).clone()
    K[:, :same] = K[:, :1]
    Rt[:, :same] = Rt[:, :1]
    pts = sw.points.clone()
    hw = sw.image_sizes.float()
    # points constructed in camera 0's frame and mapped back to the lidar frame: exact borders, the image plane, behind
    Rinv = torch.linalg.inv(Rt[0, 0].double())
    k = K[0, 0].double()
    H, W = float(hw[0, 0]), float(hw[0, 1])
    special = []
    for u, v, zc in [(0.0, 0.0, 5.0), (W, 10.0, 5.0), (W - 1e-3, H - 1e-3, 7.0), (10.0, H, 3.0), (-1e-4, 5.0, 2.0),
                     (W / 2, H / 2, 0.0), (W / 2, H / 2, -4.0), (W / 2, H / 2, 1e-9), (W * (1 + 1e-7), H / 2, 9.0),
                     (1e-30, 1e-30, 1.0), (W / 2, H * (1 - 1e-7), 60.0)]:
        cam = torch.tensor([(u - k[0, 2]) / k[0, 0] * zc, (v - k[1, 2]) / k[1, 1] * zc, zc, 1.0], dtype=torch.float64)
        special.append((Rinv @ cam)[:3].float())
    n = len(special)
    pts[0, :n, :3] = torch.stack(special)
    pts[0, n, :3] = torch.tensor([float("nan"), 1.0, 2.0])
    pts[0, n + 1, :3] = torch.tensor([float("inf"), 1.0, 2.0])
    pts[0, n + 2, :3] = torch.tensor([1e30, -1e30, 1e30])
    pts[0, n + 3, :3] = 0.0
    w = synth.make_mlp_weights(5)
    wd = [w[kk].cuda() for kk in W_KEYS]
    packed = ops.pack_weights(*wd)
    args = (pts.cuda(), K.cuda(), Rt.cuda(), hw.cuda(), sw.encoder)
    logits, rec = ops.fusion_fwd(*args, sw.tokens.cuda(), sw.point_feat.cuda(), packed, want_rec=True)
    ref_rec = ops.project(*args, want_pix=False, want_depth=False, want_valid=False, want_rec=True)["rec"]
    torch.cuda.synchronize()
    assert torch.equal(rec, ref_rec)
    counts = rec.view(torch.uint8).reshape(B, V, 16)[..., 13]
    assert int(counts.max()) >= same          # the multi-camera paths were really taken
    ref = ops.mlp_f32(sw.point_feat.cuda(), ops.gather_mean(sw.tokens.cuda(), ref_rec), *wd)
    assert _rel(logits.cpu().numpy(), ref.cpu().numpy()) < BF16_TOL
    # and the stand-alone projector agrees with the oracle on these points
    pix, depth, valid = O.project(pts[..., :3].numpy(), K.numpy(), Rt.numpy(), hw.numpy())
    assert np.array_equal(valid.astype(np.uint8).sum(1), counts.cpu().numpy())


def test_fused_many_sweeps_chunked(ops):
    """More sweeps than one launch stages calibration for (84 at six cameras): consecutive launches over sweep chunks."""
    B, V = 100, 333
    sw = synth.make_sweeps(B, V, "dinov2", seed=9)
    w = synth.make_mlp_weights(9)
    wd = [w[k].cuda() for k in W_KEYS]
    packed = ops.pack_weights(*wd)
    args = (sw.points.cuda(), torch.stack(sw.K).cuda(), torch.stack(sw.Rt).cuda(), sw.image_sizes.float().cuda(), sw.encoder)
    tok, feat = sw.tokens.cuda(), sw.point_feat.cuda()
    logits, rec = ops.fusion_fwd(*args, tok, feat, packed, want_rec=True)
    ref_rec = ops.project(*args, want_pix=False, want_depth=False, want_valid=False, want_rec=True)["rec"]
    assert torch.equal(rec, ref_rec)
    ref = ops.mlp_f32(feat, ops.gather_mean(tok, ref_rec), *wd)
    assert _rel(logits.cpu().numpy(), ref.cpu().numpy()) < BF16_TOL


@pytest.mark.parametrize("tok_dtype", [torch.float32, torch.bfloat16, torch.float16])
def test_fused_saved_rows_chunked_launches(ops, tok_dtype):
    """The training launch's saved rows (x by TMA out of the A ring, h1 / h2 from the epilogues) when the sweeps need
    several launches (100 > 84) and the tile count is ragged, for every token type: x = bf16([x_pt | camera mean]) of the
    stand-alone kernels (one bf16 ulp on multi-camera points), nothing outside the N rows is written."""
    B, V = 100, 333
    sw = synth.make_sweeps(B, V, "dinov2", seed=19, tok_dtype=tok_dtype)
    w = synth.make_mlp_weights(19)
    wd = [w[k].cuda() for k in W_KEYS]
    packed = ops.pack_weights(*wd)
    args = (sw.points.cuda(), torch.stack(sw.K).cuda(), torch.stack(sw.Rt).cuda(), sw.image_sizes.float().cuda(), sw.encoder)
    tok, xp = sw.tokens.cuda(), sw.point_feat.cuda()
    N, PAD = B * V, 300
    bufs = [torch.full((N + PAD, c), float("nan"), dtype=torch.bfloat16, device="cuda") for c in (448, 256, 64)]
    xb, h1, h2 = (b[:N] for b in bufs)
    logits, rec = ops.fusion_fwd(*args, tok, xp, packed, want_rec=True, save=(xb, h1, h2))
    assert torch.equal(logits, ops.fusion_fwd(*args, tok, xp, packed))
    for b in bufs:
        assert torch.isnan(b[N:].float()).all() and not torch.isnan(b[:N].float()).any()
    feat = ops.gather_mean(tok, rec)
    want = torch.cat([xp.reshape(N, 64), feat.reshape(N, 384)], -1).bfloat16()
    assert torch.equal(xb[:, :64], want[:, :64])
    d = (xb.float() - want.float()).abs()
    assert float((d / want.float().abs().clamp_min(1e-3)).max()) < 2 ** -7
    q = lambda t: t.bfloat16().double()
    m1 = torch.relu(xb.double() @ q(wd[0]).T + wd[1].double())
    assert float((h1.double() - m1).abs().max() / m1.abs().max()) < 1e-2


def test_fused_large_biases(ops):
    """The biases enter through a K = 16 MMA as bf16 hi + lo parts: biases far larger than the products still arrive to
    fp32 accuracy (2^-17 relative), so the logits track an fp64 model of the kernel's bf16 operand roundings."""
    B, V = 2, 3000
    sw = synth.make_sweeps(B, V, "dinov3", seed=23)
    w = {k: v.clone() for k, v in synth.make_mlp_weights(23).items()}
    gen = torch.Generator().manual_seed(23)
    for k, scale in (("mlp.0.bias", 300.0), ("mlp.2.bias", 50.0), ("mlp.4.bias", 1000.0)):
        w[k] = (torch.rand(w[k].shape, generator=gen) - 0.3) * scale
    wd = [w[k].cuda() for k in W_KEYS]
    packed = ops.pack_weights(*wd)
    args = (sw.points.cuda(), torch.stack(sw.K).cuda(), torch.stack(sw.Rt).cuda(), sw.image_sizes.float().cuda(), sw.encoder)
    tok, xp = sw.tokens.cuda(), sw.point_feat.cuda()
    logits, rec = ops.fusion_fwd(*args, tok, xp, packed, want_rec=True)
    x = torch.cat([xp.reshape(-1, 64), ops.gather_mean(tok, rec).reshape(-1, 384)], -1)
    q = lambda t: t.bfloat16().double()
    h1 = torch.relu(q(x) @ q(wd[0]).T + wd[1].double())
    h2 = torch.relu(q(h1.float()) @ q(wd[2]).T + wd[3].double())
    ref = q(h2.float()) @ q(wd[4]).T + wd[5].double()
    err = float((logits.reshape(-1, 16).double() - ref).abs().max() / ref.abs().max())
    assert err < 2e-3, err
    # and the bias alone (zero weights) is reproduced to fp32 rounding
    zw = [torch.zeros_like(t) if i % 2 == 0 else t for i, t in enumerate(wd)]
    only_b = ops.fusion_fwd(*args, tok, xp, ops.pack_weights(*zw)).reshape(-1, 16)
    assert float((only_b - wd[5]).abs().max() / wd[5].abs().max()) < 1e-5


@pytest.mark.parametrize("B,V,NC", [(1, 1, 6), (1, 255, 1), (3, 257, 3), (2, 1000, 5), (1, 40000, 6)])
def test_fused_odd_shapes(ops, B, V, NC):
    """Tile remainders, a single point, fewer than six cameras: fused kernel against the unfused fp32 route."""
    sw = synth.make_sweeps(B, V, "dinov2", seed=B * 1000 + V)
    K, Rt = torch.stack(sw.K)[:, :NC].contiguous(), torch.stack(sw.Rt)[:, :NC].contiguous()
    tok = sw.tokens[:, :NC].contiguous()
    w = synth.make_mlp_weights(3)
    wd = [w[k].cuda() for k in W_KEYS]
    packed = ops.pack_weights(*wd)
    args = (sw.points.cuda(), K.cuda(), Rt.cuda(), sw.image_sizes.float().cuda(), sw.encoder)
    logits, rec = ops.fusion_fwd(*args, tok.cuda(), sw.point_feat.cuda(), packed, want_rec=True)
    ref_rec = ops.project(*args, want_pix=False, want_depth=False, want_valid=False, want_rec=True)["rec"]
    assert torch.equal(rec, ref_rec)
    ref = ops.mlp_f32(sw.point_feat.cuda(), ops.gather_mean(tok.cuda(), ref_rec), *wd)
    assert _rel(logits.cpu().numpy(), ref.cpu().numpy()) < BF16_TOL


def test_fused_forward_under_cuda_graph_capture(ops):
    """After one warm-up call (workspace allocated, kernel attributes set) the fused launch is a pure stream operation:
    it can be captured into a CUDA graph and replayed - new point features in the same buffers give the eager result."""
    sw = synth.make_sweeps(2, 6000, "dinov3", seed=77)
    w = synth.make_mlp_weights(77)
    packed = ops.pack_weights(*[w[k].cuda() for k in W_KEYS])
    args = (sw.points.cuda(), torch.stack(sw.K).cuda(), torch.stack(sw.Rt).cuda(), sw.image_sizes.float().cuda(), sw.encoder)
    tok, xp = sw.tokens.cuda(), sw.point_feat.cuda().clone()
    out = torch.empty(2, 6000, 16, device="cuda")
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        ops.fusion_fwd(*args, tok, xp, packed, out=out)          # warm-up on the capture stream (its workspace)
        side.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            ops.fusion_fwd(*args, tok, xp, packed, out=out)
    torch.cuda.current_stream().wait_stream(side)
    xp.copy_(torch.randn(xp.shape, generator=torch.Generator().manual_seed(78)).cuda())
    out.fill_(float("nan"))
    g.replay()
    torch.cuda.synchronize()
    want = ops.fusion_fwd(*args, tok, xp, packed)
    assert torch.equal(out, want)


def test_fused_preds_out(ops):
    """The class ids written by the fused launch are torch.argmax(logits) + 1 of the logits it wrote (utils/losses.py:66)."""
    sw = synth.make_sweeps(2, 5000, "dinov3", seed=21)
    w = synth.make_mlp_weights(21)
    packed = ops.pack_weights(*[w[k].cuda() for k in W_KEYS])
    preds = torch.zeros(2, 5000, dtype=torch.int64, device="cuda")
    logits = ops.fusion_fwd(sw.points.cuda(), torch.stack(sw.K).cuda(), torch.stack(sw.Rt).cuda(),
                            sw.image_sizes.float().cuda(), sw.encoder, sw.tokens.cuda(), sw.point_feat.cuda(), packed,
                            preds_out=preds)
    assert torch.equal(preds, logits.argmax(-1) + 1)
    assert int(preds.min()) >= 1 and int(preds.max()) <= 16
    with pytest.raises(ValueError):
        ops.fusion_fwd(sw.points.cuda(), torch.stack(sw.K).cuda(), torch.stack(sw.Rt).cuda(), sw.image_sizes.float().cuda(),
                       sw.encoder, sw.tokens.cuda(), sw.point_feat.cuda(), packed, preds_out=preds.int())


def test_mlp_tc_training_gradients(ops):
    """bf16 tensor-core training path (tcgen05 forward with saved activations + bf16 backward).
    (1) Against an fp64 model of exactly what it computes (bf16-rounded operands and activations, the same ReLU masks):
    every gradient within 1e-2 (a handful of ReLU masks differ between fp32 and fp64 accumulation).  (2) Against the fp32 kernels, which are parity-checked against the reference at 1e-5:
    logits within north_star's 1e-2; gradients within 1e-1 here - about 0.3 % of the pre-activations round to the other
    side of zero in bf16, each switching a whole ReLU path, and with a random upstream gradient that is a 6 % change in
    norm (the drop-in test below, through the CE + Lovasz loss, agrees to 4 % with a loss equal within 1e-2)."""
    n = 20_000
    gen = torch.Generator().manual_seed(4)
    x_pt = torch.randn(n, 64, generator=gen).cuda()
    x_img = torch.randn(n, 384, generator=gen).cuda()
    dl = torch.randn(n, 16, generator=gen).cuda()
    w = synth.make_mlp_weights(4)

    def run(fn):
        params = [w[k].cuda().clone().requires_grad_(True) for k in W_KEYS]
        xp = x_pt.clone().requires_grad_(True)
        out = fn(xp, x_img, *params)
        out.backward(dl)
        return out.detach(), xp.grad, [p.grad for p in params]

    out_a, dx_a, g_a = run(ops.mlp_tc_train)
    rel = lambda a, b: float((a.double() - b.double()).norm() / b.double().norm())
    # (1) fp64 model of the kernels' arithmetic
    q = lambda t: t.bfloat16().double()
    W1, b1, W2, b2, W3, b3 = [w[k].cuda() for k in W_KEYS]
    x = q(torch.cat([x_pt, x_img], -1))
    h1 = q(torch.relu(x @ q(W1).T + b1.double()))
    h2 = q(torch.relu(h1 @ q(W2).T + b2.double()))
    logits = h2 @ q(W3).T + b3.double()
    dlq = q(dl)
    dh2 = (dlq @ q(W3)) * (h2 > 0)
    dh1 = (q(dh2) @ q(W2)) * (h1 > 0)
    model = {"mlp.0.weight": q(dh1).T @ x, "mlp.0.bias": dh1.sum(0), "mlp.2.weight": q(dh2).T @ h1, "mlp.2.bias": dh2.sum(0),
             "mlp.4.weight": dlq.T @ h2, "mlp.4.bias": dl.double().sum(0)}
    assert rel(out_a, logits) < 2e-3
    assert rel(dx_a, q(dh1) @ q(W1)[:, :64]) < 1e-2   # a handful of masks differ between fp32 and fp64 accumulation
    errs = {k: rel(g, model[k]) for k, g in zip(W_KEYS, g_a)}
    print("bf16 backward vs its fp64 model:", {k: f"{v:.2e}" for k, v in errs.items()})
    assert max(errs.values()) < 1e-2, errs
    # (2) the fp32 kernels
    out_b, dx_b, g_b = run(ops.mlp_f32)
    assert _rel(out_a.cpu().numpy(), out_b.cpu().numpy()) < BF16_TOL
    assert rel(dx_a, dx_b) < 0.1
    for k, a, b in zip(W_KEYS, g_a, g_b):
        assert rel(a, b) < 0.1, (k, rel(a, b))   # random upstream gradient: no mean component to average the flips out


def test_fused_training_saved_rows_and_gradients(ops):
    """Training through the fused launch (ops.fusion_train).  The rows it saves for the backward pass are bit-exactly
    bf16([x_pt | masked camera mean]) of the stand-alone projection + gather kernels for single-camera points (a plain
    copy of one token row) and within one bf16 ulp for points seen by several cameras (the fused kernel averages in a
    different order); logits equal the inference launch bit for bit; gradients match the unfused tensor-core path."""
    B, V = 3, 7001   # not a multiple of the 256-row tile: the last tile is ragged
    sw = synth.make_sweeps(B, V, "dinov3", seed=41)
    w = synth.make_mlp_weights(41)
    K, Rt, hw = torch.stack(sw.K).cuda(), torch.stack(sw.Rt).cuda(), sw.image_sizes.float().cuda()
    pts, tok, xp = sw.points.cuda(), sw.tokens.cuda(), sw.point_feat.cuda()
    wd = [w[k].cuda() for k in W_KEYS]
    packed = ops.pack_weights(*wd)
    N = B * V
    xb = torch.full((N, 448), float("nan"), dtype=torch.bfloat16, device="cuda")
    h1 = torch.full((N, 256), float("nan"), dtype=torch.bfloat16, device="cuda")
    h2 = torch.full((N, 64), float("nan"), dtype=torch.bfloat16, device="cuda")
    logits, rec = ops.fusion_fwd(pts, K, Rt, hw, sw.encoder, tok, xp, packed, want_rec=True, save=(xb, h1, h2))
    plain = ops.fusion_fwd(pts, K, Rt, hw, sw.encoder, tok, xp, packed)
    assert torch.equal(logits, plain)
    feat = ops.gather_mean(tok, rec)
    want = torch.cat([xp.reshape(N, 64), feat.reshape(N, 384)], -1).bfloat16()
    assert not torch.isnan(xb.float()).any() and not torch.isnan(h1.float()).any() and not torch.isnan(h2.float()).any()
    assert torch.equal(xb[:, :64], want[:, :64])
    single = ops.project(pts, K, Rt, hw, sw.encoder, want_pix=False, want_depth=False, want_rec=False)["valid"].sum(1).reshape(-1) <= 1
    assert int(single.sum()) > N // 4
    assert torch.equal(xb[single], want[single])
    d = (xb.float() - want.float()).abs()
    assert float((d / want.float().abs().clamp_min(1e-3)).max()) < 2 ** -7   # one bf16 ulp
    # hidden activations against a model of the kernel fed with the saved rows
    q = lambda t: t.bfloat16().double()
    m1 = torch.relu(xb.double() @ q(wd[0]).T + wd[1].double())
    assert float((h1.double() - m1).abs().max() / m1.abs().max()) < 1e-2
    m2 = torch.relu(h1.double() @ q(wd[2]).T + wd[3].double())
    assert float((h2.double() - m2).abs().max() / m2.abs().max()) < 1e-2

    # gradients: fused training path against the unfused tensor-core path (same arithmetic up to the mean's rounding)
    dl = torch.randn(B, V, 16, generator=torch.Generator().manual_seed(5)).cuda()

    def run(fused):
        params = [t.clone().requires_grad_(True) for t in wd]
        x = xp.clone().requires_grad_(True)
        if fused:
            out = ops.fusion_train(pts, K, Rt, hw, sw.encoder, tok, x, *params)
        else:
            out = ops.mlp_tc_train(x, feat, *params)
        out.backward(dl)
        return out.detach(), x.grad, [p.grad for p in params]

    out_a, dx_a, g_a = run(True)
    out_b, dx_b, g_b = run(False)
    rel = lambda a, b: float((a.double() - b.double()).norm() / b.double().norm())
    assert torch.equal(out_a, logits)
    assert rel(out_a, out_b) < 2e-3
    assert dx_a.shape == xp.shape and rel(dx_a, dx_b) < 1e-2
    for k, a, b in zip(W_KEYS, g_a, g_b):
        assert rel(a, b) < 1e-2, (k, rel(a, b))


def test_dropin_training_step_bf16(ops, monkeypatch):
    """FeatureFusionModel in training mode with the default (bf16) arithmetic: loss and parameter gradients against
    TDSSF_MLP=fp32 on the same inputs."""
    from utils.fusion_model import FeatureFusionModel
    from utils.losses import CELSLoss
    sw = synth.make_sweeps(2, 6000, "dinov3", seed=31)

    class Pcd(torch.nn.Module):
        def devoxelize(self, s):
            return s, None

    def step(mode):
        monkeypatch.setenv("TDSSF_MLP", mode)
        model = FeatureFusionModel(sw.encoder, Pcd(), device="cuda").cuda()
        model.load_state_dict(synth.make_mlp_weights(31), strict=True)
        scores = model(sw.tokens, sw.point_feat.cuda(), sw.points.cuda(), None, sw.image_sizes.cuda(), sw.K, sw.Rt)
        total = CELSLoss()(scores, sw.labels.cuda(), sw.mask.cuda())[0]
        total.backward()
        return float(total.detach()), [p.grad.clone() for p in model.mlp.parameters()]

    loss_a, g_a = step("bf16")
    loss_b, g_b = step("fp32")
    assert abs(loss_a - loss_b) / abs(loss_b) < BF16_TOL
    for a, b in zip(g_a, g_b):
        assert float((a - b).norm() / b.norm()) < 0.1   # measured 4 % on W1 (random labels: see the test above)


def test_training_trajectory_bf16_tracks_fp32(ops, monkeypatch):
    """50 AdamW steps of the drop-in model + CELSLoss on four alternating synthetic batches, default bf16 tensor-core mode
    against TDSSF_MLP=fp32 (the 1e-5 parity mode) from the same initial weights: the two loss curves stay within 2 % of
    each other at every step (north_star's 1e-2 bound is per evaluation; the per-step gradients differ by a few per cent
    through ReLU sign flips, test above) and both go down - the bf16 mode trains like the reference arithmetic."""
    from utils.fusion_model import FeatureFusionModel
    from utils.losses import CELSLoss
    batches = [synth.make_sweeps(2, 4096, "dinov3", seed=600 + i) for i in range(4)]
    # learnable structure: labels follow the sign pattern of two point-feature channels, so the loss can really fall
    for sw in batches:
        code = (sw.point_feat[..., 0] > 0).long() * 2 + (sw.point_feat[..., 1] > 0).long()
        sw.labels = torch.where(sw.labels == 0, sw.labels, 11 + code)
        sw.mask = sw.labels != 0

    class Pcd(torch.nn.Module):
        def devoxelize(self, s):
            return s, None

    def run(mode):
        monkeypatch.setenv("TDSSF_MLP", mode)
        model = FeatureFusionModel(batches[0].encoder, Pcd(), device="cuda").cuda()
        model.load_state_dict(synth.make_mlp_weights(600), strict=True)
        crit = CELSLoss(weight=torch.tensor(synth.CLASS_WEIGHTS).cuda(), ignore_index=0, lamda_lovasz=1.0)
        opt = torch.optim.AdamW(model.parameters(), lr=1e-3)
        dev = [(sw.tokens.cuda(), sw.point_feat.cuda(), sw.points.cuda(), sw.labels.cuda(), sw.mask.cuda()) for sw in batches]
        curve = []
        for step in range(50):
            sw = batches[step % 4]
            tok, feat, pts, lab, mask = dev[step % 4]
            opt.zero_grad()
            total = crit(model(tok, feat, pts, None, sw.image_sizes.cuda(), sw.K, sw.Rt), lab, mask)[0]
            total.backward()
            opt.step()
            curve.append(float(total.detach()))
        return np.array(curve)

    a, b = run("bf16"), run("fp32")
    print("loss bf16:", np.round(a[::7], 4), "fp32:", np.round(b[::7], 4))
    assert b[-1] < 0.7 * b[0] and a[-1] < 0.7 * a[0], (a[0], a[-1], b[0], b[-1])
    assert np.max(np.abs(a - b) / b) < 2e-2, float(np.max(np.abs(a - b) / b))
