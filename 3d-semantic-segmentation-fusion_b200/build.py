"""Builds libtdssf.so (sm_100a only) from csrc/*.cu with nvcc, in-tree, next to this file.

    python 3d-semantic-segmentation-fusion_b200/build.py [--force] [--verbose]

nvcc cross-compiles without a GPU.  Objects are cached under build/ by source mtime; the .so is
git-ignored but travels with the repo snapshot to the GPU box.
"""
import argparse
import concurrent.futures as cf
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libtdssf.so")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "-Xcompiler", "-fvisibility=hidden",
    "-I", INCLUDE,
]


def nvcc_path():
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found (needed to build libtdssf.so for sm_100a)")
    return p


def _newer(src, dst, extra=()):
    if not os.path.exists(dst):
        return True
    t = os.path.getmtime(dst)
    return any(os.path.getmtime(s) > t for s in (src, *extra))


def build(force=False, verbose=False, defines=(), out=None):
    """defines/out build an experiment variant (e.g. TDSSF_LDMODE=1) into its own object dir and .so."""
    global OBJ, LIB
    nvcc = nvcc_path()
    if defines or out:
        tag = "_".join(d.replace("=", "") for d in defines) or "variant"
        OBJ = os.path.join(HERE, "build", "variants", tag)
        LIB = out or os.path.join(OBJ, "libtdssf.so")
    os.makedirs(OBJ, exist_ok=True)
    sources = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(INCLUDE, "tdssf.h"))
    jobs = []
    for f in sources:
        src = os.path.join(CSRC, f)
        obj = os.path.join(OBJ, f[:-3] + ".o")
        if force or _newer(src, obj, headers):
            cmd = [nvcc, *NVCC_FLAGS, *["-D" + d for d in defines], "-c", src, "-o", obj]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            jobs.append((f, cmd))

    def run(job):
        name, cmd = job
        r = subprocess.run(cmd, capture_output=True, text=True)
        return name, r.returncode, r.stdout + r.stderr

    failed = False
    with cf.ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for name, rc, out in ex.map(run, jobs):
            if verbose or rc != 0:
                sys.stderr.write(f"--- nvcc {name} (rc={rc})\n{out}\n")
            failed |= rc != 0
    if failed:
        raise RuntimeError("nvcc failed; see messages above")
    objs = [os.path.join(OBJ, f[:-3] + ".o") for f in sources]
    if force or jobs or not os.path.exists(LIB):
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB, *objs, "-lcudart", "-ldl"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--define", action="append", default=[], help="experiment variant, e.g. TDSSF_LDMODE=1")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    print(build(force=a.force, verbose=a.verbose, defines=tuple(a.define), out=a.out))
