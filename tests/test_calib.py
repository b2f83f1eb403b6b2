"""f4: calibration composition (utils/dataloader.py:123-152) and multi-sweep accumulation.  CPU part: the oracle against
hand-checkable poses and against the float64 geometry of the projector; GPU part (`-m gpu`): csrc/calib.cu through the
C ABI against the oracle.

Tolerances: the composition is float64 on both sides but the device uses the closed-form rigid inverse where the
reference calls np.linalg.inv (LU), so float64 results agree to 1e-12 and the fp32 casts to 1 ulp (asserted: at most one
ulp, and at least 99 % of the entries identical).  Accumulated points: float64 transform, fp32 store, same rule.
The multi-sweep chain restates nuscenes-devkit (not vendored in the reference): parity unpinned for that part."""
import numpy as np
import pytest
import torch

from oracle import np_oracle as O


def _poses(rng, *shape):
    q = rng.normal(size=shape + (4,))
    q /= np.linalg.norm(q, axis=-1, keepdims=True)
    t = rng.normal(size=shape + (3,)) * np.array([2.0, 2.0, 0.5])
    return np.concatenate([q, t], -1)


def _ulps(a, b):
    ai = a.astype(np.float32).view(np.int32).astype(np.int64)
    bi = b.astype(np.float32).view(np.int32).astype(np.int64)
    ai = np.where(ai < 0, -(ai & 0x7fffffff), ai)
    bi = np.where(bi < 0, -(bi & 0x7fffffff), bi)
    return np.abs(ai - bi)


# ------------------------------------------------------------------------------------------------ oracle (CPU)
def test_oracle_pose_matrix_kat():
    T = O.pose_matrix(np.array([1.0, 0, 0, 0, 1, 2, 3]))
    assert np.array_equal(T, np.array([[1, 0, 0, 1], [0, 1, 0, 2], [0, 0, 1, 3], [0, 0, 0, 1.0]]))
    s = np.sqrt(0.5)   # 90 degrees about z: x -> y
    T = O.pose_matrix(np.array([s, 0, 0, s, 0, 0, 0]))
    assert np.allclose(T[:3, :3] @ np.array([1.0, 0, 0]), [0, 1, 0], atol=1e-15)
    T2 = O.pose_matrix(np.array([2 * s, 0, 0, 2 * s, 0, 0, 0]))   # non-unit quaternions are normalised (pyquaternion)
    assert np.allclose(T, T2, atol=1e-15)


def test_oracle_compose_matches_reference_expression():
    """utils/dataloader.py:146-149: T_lidar_cam = inv(T_cam_ego) @ T_lidar_ego; rotation blocks stay orthonormal and a
    lidar point mapped to the camera, then back through T_cam_ego, lands where T_lidar_ego puts it."""
    rng = np.random.default_rng(0)
    ls, cs = _poses(rng, 3), _poses(rng, 3, 6)
    T = O.compose_lidar2cam(ls, cs)
    p = np.array([4.0, -1.0, 0.5, 1.0])
    for b in range(3):
        for c in range(6):
            assert np.allclose(T[b, c, :3, :3] @ T[b, c, :3, :3].T, np.eye(3), atol=1e-12)
            assert np.allclose(O.pose_matrix(cs[b, c]) @ (T[b, c] @ p), O.pose_matrix(ls[b]) @ p, atol=1e-12)
    # equal ego poses at the lidar and camera timestamps cancel: the four-stage chain reduces to the reference's product
    ego = _poses(rng, 3)
    T4 = O.compose_lidar2cam(ls, cs, ego, np.repeat(ego[:, None], 6, 1))
    assert np.allclose(T4, T, atol=1e-12)


def test_oracle_accumulate_identity_and_order():
    rng = np.random.default_rng(1)
    pts = rng.normal(size=(3, 5, 4)).astype(np.float32)
    ident = np.array([1.0, 0, 0, 0, 0, 0, 0])
    T = O.compose_sweep2ref(np.tile(ident, (3, 1)), np.tile(ident, (3, 1)), ident, ident)
    out = O.accumulate_sweeps(pts, [5, 2, 0], T)
    assert out.shape == (7, 4) and np.array_equal(out[:5], pts[0]) and np.array_equal(out[5:], pts[1, :2])
    # a sweep taken 1 m further along x in the global frame appears shifted by +1 m in the reference frame
    ego = np.tile(ident, (3, 1)); ego[1, 4] = 1.0
    T = O.compose_sweep2ref(np.tile(ident, (3, 1)), ego, ident, ident)
    out = O.accumulate_sweeps(pts, [1, 1, 0], T)
    assert np.allclose(out[1, :3], pts[1, 0, :3] + np.array([1, 0, 0], np.float32), atol=1e-6) and out[1, 3] == pts[1, 0, 3]


# ------------------------------------------------------------------------------------------------ device (GPU)
@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from tdssf import ops as _ops, _lib
    _lib.check(_lib.lib.tdssf_check_device(), "device")
    return _ops


@pytest.mark.gpu
@pytest.mark.parametrize("with_ego", [False, True])
def test_gpu_compose_lidar2cam(ops, with_ego):
    rng = np.random.default_rng(7)
    B, NC = 8, 6
    ls, cs = _poses(rng, B), _poses(rng, B, NC)
    le, ce = (_poses(rng, B), _poses(rng, B, NC)) if with_ego else (None, None)
    if with_ego:   # ego poses live in a global frame hundreds of metres from the origin
        le[:, 4:6] += 600.0
        ce[:, :, 4:6] += 600.0
    ref = O.compose_lidar2cam(ls, cs, le, ce)
    out32, out64 = ops.compose_lidar2cam(ls, cs, le, ce, want_f64=True)
    got64, got32 = out64.cpu().numpy(), out32.cpu().numpy()
    assert np.abs(got64 - ref).max() < 1e-9 if with_ego else np.abs(got64 - ref).max() < 1e-12
    u = _ulps(got32, ref.astype(np.float32))
    assert u.max() <= 1 and (u == 0).mean() >= 0.99
    assert np.array_equal(got32[..., 3, :], np.broadcast_to(np.array([0, 0, 0, 1], np.float32), (B, NC, 4)))


@pytest.mark.gpu
def test_gpu_compose_feeds_projector(ops):
    """The composed matrices are a valid `Rt` input of tdssf_project: same pixels as the oracle projector on the oracle's
    composition wherever the two fp32 matrices are identical."""
    from tdssf import synth
    rng = np.random.default_rng(11)
    ls, cs = _poses(rng, 2), _poses(rng, 2, 6)
    Rt = ops.compose_lidar2cam(ls, cs)
    ref = O.compose_lidar2cam(ls, cs).astype(np.float32)
    same = (Rt.cpu().numpy() == ref).all(axis=(2, 3))
    assert same.mean() > 0.5
    sw = synth.make_sweeps(2, 2000, "dinov3", seed=5)
    K = torch.stack(sw.K).cuda()
    out = ops.project(sw.points.cuda(), K, Rt, sw.image_sizes.float().cuda(), sw.encoder)
    pix, depth, valid = O.project(sw.points.numpy()[..., :3], K.cpu().numpy(), ref, sw.image_sizes.numpy())
    for b in range(2):
        for c in range(6):
            if same[b, c]:
                assert np.array_equal(out["valid"][b, c].cpu().numpy().astype(bool), valid[b, c])
                assert np.array_equal(out["pix"][b, c].cpu().numpy(), pix[b, c])


@pytest.mark.gpu
@pytest.mark.parametrize("S,P", [(10, 34720), (3, 1000), (1, 1)])
def test_gpu_accumulate_sweeps(ops, S, P):
    rng = np.random.default_rng(S * 1000 + P)
    pts = (rng.normal(size=(S, P, 4)) * np.array([30, 30, 2, 1])).astype(np.float32)
    counts = rng.integers(max(P - 500, 0), P + 1, size=S).astype(np.int32)
    if S > 2:
        counts[1] = 0   # an empty sweep in the middle
    counts[0] = max(counts[0], 1)
    ss, se = _poses(rng, S), _poses(rng, S)
    se[:, 4:6] += 600.0
    rs, re = ss[0], se[0]
    T_ref = O.compose_sweep2ref(ss, se, rs, re)
    ref = O.accumulate_sweeps(pts, counts, T_ref)
    cloud, total, T = ops.accumulate_sweeps(torch.from_numpy(pts).cuda(), counts, ss, se, rs, re)
    n = int(total.item())
    assert n == int(counts.sum()) == len(ref)
    assert np.abs(T.cpu().numpy() - T_ref).max() < 1e-9
    assert np.allclose(T.cpu().numpy()[0], np.eye(4), atol=1e-12)          # the reference sweep maps onto itself
    got = cloud[:n].cpu().numpy()
    assert np.array_equal(got[:, 3], ref[:, 3])
    u = _ulps(got[:, :3], ref[:, :3])
    # |x| up to ~100 m: one fp32 ulp is 8e-6 m; allow the last float64 bit of the transform to flip a rounding
    assert u.max() <= 2 and (u == 0).mean() >= 0.99


@pytest.mark.gpu
def test_gpu_calib_errors(ops):
    with pytest.raises(ValueError):
        ops.compose_lidar2cam(np.zeros((2, 6)), np.zeros((2, 6, 7)))
    with pytest.raises(ValueError):
        ops.compose_lidar2cam(np.zeros((2, 7)), np.zeros((2, 6, 7)), lidar_ego=np.zeros((2, 7)))
    with pytest.raises(ValueError):
        ops.accumulate_sweeps(torch.zeros((2, 4, 3), device="cuda"), [1, 1], np.zeros((2, 7)), np.zeros((2, 7)), np.zeros(7), np.zeros(7))
