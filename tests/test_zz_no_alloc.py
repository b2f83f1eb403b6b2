"""tdssf_devox_nn_static: the nearest-voxel search with its hash-grid workspace in a static device array (include/tdssf.h),
for callers that cannot allocate device memory.  Called through the C-ABI, checked bit for bit (int64 indices: the
tolerance of every devoxelize test here is equality) against tdssf_devox_nn with a caller-owned workspace, on two calls in
a row that share the array - the second one smaller, so stale buckets of the first would show.  Collected last on purpose
(file name)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

STATIC_BYTES = 16 << 20   # TDSSF_DEVOX_NN_STATIC_BYTES


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _cloud(B, P, V, seed):
    g = torch.Generator().manual_seed(seed)
    pts = (torch.rand(B, P, 4, generator=g) - 0.5) * torch.tensor([80.0, 80.0, 6.0, 1.0])
    vs = 0.25
    vox = torch.zeros(B, V, 3)          # voxel centres of a subset of the points; the tail stays zero padding
    n = int(V * 0.8)
    vox[:, :n] = torch.floor(pts[:, torch.randperm(P, generator=g)[:n], :3] / vs) * vs + vs / 2
    return pts.cuda().contiguous(), vox.cuda().contiguous()


def _both(lib, check, pts, vox, cell):
    B, P, sp = pts.shape
    V, sv = vox.shape[1], vox.shape[2]
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    need = lib.tdssf_devox_nn_workspace_bytes(B, P, V)
    assert need <= STATIC_BYTES
    ws = torch.empty(need, dtype=torch.uint8, device=pts.device)
    want = torch.full((B, P), -1, dtype=torch.int64, device=pts.device)
    got = torch.full((B, P), -2, dtype=torch.int64, device=pts.device)
    check(lib.tdssf_devox_nn(_ptr(pts), sp, _ptr(vox), sv, B, P, V, _ptr(want), cell, _ptr(ws), need, stream), "tdssf_devox_nn")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    check(lib.tdssf_devox_nn_static(_ptr(pts), sp, _ptr(vox), sv, B, P, V, _ptr(got), cell, stream), "tdssf_devox_nn_static")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return want, got


def test_static_workspace_matches_caller_workspace_twice():
    from tdssf import _lib
    lib, check = _lib.lib, _lib.check
    torch.cuda.set_device(0)
    # first call: 4 sweeps x 30 000 points / 24 000 voxel rows (about 7 MB of the 16 MiB array)
    pts, vox = _cloud(4, 30000, 24000, seed=5)
    want, got = _both(lib, check, pts, vox, 0.5)
    assert torch.equal(got, want)
    assert int(got.min()) >= 0 and int(got.max()) < 24000
    # second call right after, other data, smaller table: nothing of the first call may leak through the shared array
    pts2, vox2 = _cloud(2, 9000, 3000, seed=6)
    want2, got2 = _both(lib, check, pts2, vox2, 1.0)
    assert torch.equal(got2, want2)
    # and the first inputs again give the first answer
    _, got3 = _both(lib, check, pts, vox, 0.5)
    assert torch.equal(got3, want)


def test_static_workspace_rejects_sizes_beyond_its_limit():
    from tdssf import _lib
    lib = _lib.lib
    B, P, V = 8, 200000, 200000
    assert lib.tdssf_devox_nn_workspace_bytes(B, P, V) > STATIC_BYTES
    pts = torch.zeros(1, 8, 3, device="cuda")   # never read: the size check comes first
    idx = torch.zeros(1, 8, dtype=torch.int64, device="cuda")
    rc = lib.tdssf_devox_nn_static(_ptr(pts), 3, _ptr(pts), 3, B, P, V, _ptr(idx), 0.5, None)
    assert rc == -1   # TDSSF_EINVAL
    msg = lib.tdssf_last_error().decode()
    assert "static workspace" in msg and str(STATIC_BYTES) in msg
