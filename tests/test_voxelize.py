"""f1: the voxelizer (utils/lidar.py:83-118).  CPU part: the oracle's restatement against hand-checkable cases and
against scikit-learn (the reference's own neighbour search) where the neighbour order is defined.  GPU part
(`-m gpu`): csrc/voxel.cu through the C ABI against the oracle, bit for bit.

Parity of the down-sampling step is UNPINNED: Open3D (0.18.0, environment.yml:278) is not in this image and its source
is not under the reference tree; the oracle restates its published algorithm (see oracle/np_oracle.py)."""
import numpy as np
import pytest
import torch

from oracle import np_oracle as O
from tdssf import synth


def _cloud(P, seed):
    return synth.make_points(1, P, torch.Generator().manual_seed(seed))[0].numpy()


# ------------------------------------------------------------------------------------------------ oracle (CPU)
def test_oracle_kat_two_voxels():
    """Hand-checkable: voxel 1.0, min corner (0,0,0) -> voxel_min_bound -0.5; points 0,1,3 share the voxel [−0.5,0.5)³…"""
    pts = np.array([[0.0, 0.0, 0.0, 1.0],
                    [0.25, 0.25, 0.0, 2.0],
                    [2.0, 0.0, 0.0, 3.0],
                    [0.25, 0.0, 0.25, 4.0],
                    [2.25, 0.25, 0.25, 5.0],
                    [9.0, 9.0, 9.0, 6.0]], dtype=np.float32)
    raw, coords, vid, _ = O.voxelize(pts, 1.0)
    assert vid.tolist() == [0, 0, 1, 0, 1, 2]                       # numbered by first point
    assert np.array_equal(raw[:, :3], np.array([[0.5 / 3, 0.25 / 3, 0.25 / 3], [2.125, 0.125, 0.125], [9, 9, 9]], dtype=np.float64).astype(np.float32))
    assert coords.tolist() == [[0, 0, 0], [2, 0, 0], [9, 9, 9]]
    # voxel 2 (one far point): neighbours by distance = 5 (itself), then 4, 2, 1|3 ... ; weights 0.4 .3 .2 .05 .05
    d = ((pts[:, :3].astype(np.float64) - np.array([9.0, 9.0, 9.0])) ** 2).sum(1)
    order = np.lexsort((np.arange(6), d))[:5]
    want = (pts[order, 3].astype(np.float64) * O.KNN_WEIGHTS).sum()
    assert raw[2, 3] == np.float32(want)


@pytest.mark.parametrize("voxel_size", [0.5, 0.1])
def test_oracle_matches_sklearn_where_order_is_defined(voxel_size):
    pts = _cloud(6000, 7)
    raw, coords, vid, mingap = O.voxelize(pts, voxel_size)
    raw_sk, coords_sk, _, _ = O.voxelize(pts, voxel_size, knn="sklearn")
    assert np.array_equal(raw[:, :3], raw_sk[:, :3]) and np.array_equal(coords, coords_sk)
    defined = mingap > 0        # no two of the six nearest points equidistant (2-point voxels tie at their midpoint)
    assert defined.mean() > 0.9
    assert np.array_equal(raw[defined, 3], raw_sk[defined, 3])
    # every voxel is the mean of its points, every point lies in the cube of its voxel
    v = 17
    assert np.allclose(raw[v, :3], pts[vid == v, :3].astype(np.float64).mean(0), rtol=0, atol=1e-6)
    lo = pts[:, :3].astype(np.float64).min(0) - voxel_size / 2
    cell = np.floor((pts[:, :3].astype(np.float64) - lo) / voxel_size)
    assert len(np.unique(cell, axis=0)) == raw.shape[0]


# ------------------------------------------------------------------------------------------------ device
@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from tdssf import ops as _ops, _lib
    _lib.check(_lib.lib.tdssf_check_device(), "device")
    return _ops


def _check_against_oracle(ops, clouds, voxel_size, knn_cell=None):
    pts = torch.from_numpy(np.stack(clouds)).cuda()
    raw, coords, n_vox, status = ops.voxelize(pts, voxel_size, knn_cell)
    assert int(status[0]) == 0
    raw, coords, n_vox = raw.cpu().numpy(), coords.cpu().numpy(), n_vox.cpu().numpy()
    for b, cloud in enumerate(clouds):
        want_raw, want_coords, _, _ = O.voxelize(cloud, voxel_size)
        n = want_raw.shape[0]
        assert n_vox[b] == n
        assert np.array_equal(raw[b, :n, :3], want_raw[:, :3]), "centroids"
        assert np.array_equal(coords[b, :n], want_coords), "grid coordinates"
        if cloud.shape[1] == 4:
            assert np.array_equal(raw[b, :n, 3], want_raw[:, 3]), "weighted intensity"
        assert not raw[b, n:].any() and not coords[b, n:].any(), "padding rows"
    return raw, n_vox


@pytest.mark.gpu
@pytest.mark.parametrize("voxel_size", [0.5, 0.1])
def test_voxelize_bit_exact(ops, voxel_size):
    _check_against_oracle(ops, [_cloud(7000, 11), _cloud(7000, 12)], voxel_size)


@pytest.mark.gpu
def test_voxelize_search_grid_does_not_change_results(ops):
    clouds = [_cloud(5000, 13)]
    base, n = _check_against_oracle(ops, clouds, 0.5)
    for cell in (0.05, 0.3, 3.0, 50.0):     # tiny cells: nearly every voxel takes the full scan; huge: one bucket walk
        got, n2 = _check_against_oracle(ops, clouds, 0.5, knn_cell=cell)
        assert np.array_equal(got, base) and np.array_equal(n, n2)


@pytest.mark.gpu
def test_voxelize_edge_cases(ops):
    base = _cloud(3000, 14)
    padded = np.concatenate([base, np.zeros((500, 4), np.float32)])          # zero padding rows of a collated batch: one voxel
    dup = np.concatenate([base[:1500], base[:1500]])                           # every point twice: ties everywhere
    _check_against_oracle(ops, [padded], 0.5)
    _check_against_oracle(ops, [dup], 0.5)
    _check_against_oracle(ops, [base], 500.0)                                  # the whole cloud in one voxel: 3000 ordered adds
    _check_against_oracle(ops, [base[:5]], 0.1)                                # the smallest cloud scikit-learn accepts
    _check_against_oracle(ops, [base[:, :3].copy()], 0.5)                      # xyz only: no intensity column (utils/lidar.py:104-105)
    neg = base.copy(); neg[:, :3] -= 1000.0                                    # far from the origin, negative coordinates
    _check_against_oracle(ops, [neg], 0.25)
    with pytest.raises(ValueError):
        ops.voxelize(torch.from_numpy(base[None, :4]).cuda(), 0.5)             # 5 neighbours of 4 points
    with pytest.raises(ValueError):
        ops.voxelize(torch.from_numpy(base[None]).cuda(), 0.0)                 # Open3D: voxel_size <= 0
    bad = base.copy(); bad[7, 1] = np.nan
    assert int(ops.voxelize(torch.from_numpy(bad[None]).cuda(), 0.5)[3][0]) == 1


@pytest.mark.gpu
def test_voxelize_full_size_properties(ops):
    """BASELINE sizes (8 x 34 720 points, voxel 0.1 m): size-independent properties instead of the O(P*V) oracle."""
    pts = synth.make_points(8, 34_720, torch.Generator().manual_seed(15)).cuda()
    raw, coords, n_vox, status = ops.voxelize(pts, 0.1)
    raw2, coords2, n_vox2, _ = ops.voxelize(pts, 0.1)
    assert int(status[0]) == 0
    assert torch.equal(raw, raw2) and torch.equal(coords, coords2) and torch.equal(n_vox, n_vox2)   # run-to-run identical
    for b in range(8):
        n = int(n_vox[b])
        p64 = pts[b, :, :3].double()
        lo = p64.min(0).values - 0.05
        cell = torch.floor((p64 - lo) / 0.1).long()
        uniq, inv, cnt = torch.unique(cell, dim=0, return_inverse=True, return_counts=True)
        assert n == uniq.shape[0]
        # count-weighted mean of the centroids == mean of the points; every centroid inside its voxel
        sums = torch.zeros(uniq.shape[0], 3, dtype=torch.float64, device="cuda").index_add_(0, inv, p64)
        cen = raw[b, :n, :3].double()
        ccell = torch.floor((cen - lo) / 0.1).long()
        key = lambda c: (c[:, 0] * 4096 + c[:, 1]) * 4096 + c[:, 2]
        order_got, order_want = torch.argsort(key(ccell)), torch.argsort(key(uniq))
        assert torch.equal(key(ccell)[order_got], key(uniq)[order_want])
        assert torch.allclose(cen[order_got], (sums / cnt[:, None])[order_want], rtol=0, atol=1e-5)
        assert torch.equal(coords[b, :n], torch.floor(cen / 0.1).int()) or \
            (coords[b, :n] - torch.floor(cen / 0.1).int()).abs().max() <= 1   # fp32-rounded centroid on a cell face
        inten = raw[b, :n, 3]
        assert float(inten.min()) >= 0.0 and float(inten.max()) <= 1.0 + 1e-6   # convex combination of intensities in [0,1]


@pytest.mark.gpu
def test_lidar_encoder_default_voxelizer(ops):
    """LiDARFeatureEncoder.forward with the device voxelizer + devoxelize on its output."""
    from utils.lidar import LiDARFeatureEncoder

    class Backbone(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.lin = torch.nn.Linear(4, 64)

        def forward(self, d):
            return self.lin(d["feat"])

    pts = synth.make_points(2, 4000, torch.Generator().manual_seed(16)).cuda()
    pts[1, 2000:] = pts[1, :2000].clone()          # fewer voxels in the second cloud -> padding
    enc = LiDARFeatureEncoder({"dataset_params": {"lidar": {"voxel_size": 0.5}}}, backbone=Backbone()).cuda()
    feats, raw, coords, mask = enc(pts)
    for b in range(2):
        want_raw, want_coords, _, _ = O.voxelize(pts[b].cpu().numpy(), 0.5)
        n = want_raw.shape[0]
        assert int(mask[b].sum()) == n
        assert np.array_equal(raw[b, :n].cpu().numpy(), want_raw) and np.array_equal(coords[b, :n].cpu().numpy(), want_coords)
    assert mask[0].sum() != mask[1].sum() and float(raw[~mask].abs().sum()) == 0.0
    out, idx = enc.devoxelize(torch.randn(2, raw.shape[1], 16, device="cuda"))
    assert out.shape == (2, 4000, 16)
