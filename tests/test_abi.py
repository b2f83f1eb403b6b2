"""CPU-side checks of the drop-in boundary: libtdssf.so loads, exports every symbol include/tdssf.h declares,
the ctypes table matches the header, and the product has no path into oracle/ or a CPU fallback."""
import ctypes
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "3d-semantic-segmentation-fusion_b200")
HEADER = os.path.join(ROOT, "include", "tdssf.h")


@pytest.fixture(scope="module")
def lib_path():
    path = os.path.join(PKG, "libtdssf.so")
    if not os.path.exists(path):
        sys.path.insert(0, ROOT)
        import __graft_entry__
        __graft_entry__.build()
    return path


def _declared():
    text = open(HEADER).read()
    return sorted(set(re.findall(r"TDSSF_API\s+[\w\s\*]+?\b(tdssf_\w+)\s*\(", text)))


def test_header_symbols_exported(lib_path):
    names = _declared()
    assert len(names) >= 19
    lib = ctypes.CDLL(lib_path)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/tdssf.h but not exported"
    out = subprocess.run(["nm", "-D", "--defined-only", lib_path], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    assert set(names) <= exported
    assert not [s for s in exported if not s.startswith("tdssf_")], "only the C ABI is visible"


def test_ctypes_table_matches_header(lib_path):
    from tdssf import _lib
    assert sorted(_lib.SIGNATURES) == _declared()
    text = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    for name, (_, args) in _lib.SIGNATURES.items():
        m = re.search(r"\b%s\s*\(([^;]*?)\)\s*;" % name, text, flags=re.S)
        params = m.group(1).strip()
        n = 0 if params in ("", "void") else params.count(",") + 1
        assert n == len(args), f"{name}: header has {n} parameters, ctypes table {len(args)}"
    assert _lib.lib.tdssf_version() == 100
    assert ctypes.sizeof(_lib.MlpWeights) == 48 and _lib.REC_BYTES == 16


def test_no_cpu_fallback_and_no_oracle_in_product(lib_path):
    from tdssf import ops
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.gather_mean(torch.zeros(1, 6, 4, 384), torch.zeros(1, 2, 4, dtype=torch.int32))
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.confusion(torch.ones(3, dtype=torch.long), torch.ones(3, dtype=torch.long), 16)
    for dirpath, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f"{f} imports the oracle"


def test_error_convention(lib_path):
    from tdssf import _lib
    rc = _lib.lib.tdssf_confusion(None, None, 5, 1000, None, None)
    assert rc == _lib.EINVAL and "tdssf_confusion" in _lib.last_error()
    with pytest.raises(ValueError):
        _lib.check(rc)
    assert _lib.lib.tdssf_mlp_pack_bytes() == 2 * 139264 + 64   # two CTA images: weights 133 120 B + bias MMA operands, 1 KB-aligned
    assert _lib.lib.tdssf_cels_workspace_bytes(1000) > 1000 * 16 * 4 * 6
