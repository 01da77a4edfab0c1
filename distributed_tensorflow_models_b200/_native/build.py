"""In-tree build of the native sm_100a kernel library (``libdtfb200.so``).

Plain ``nvcc`` (no torch headers) so a full rebuild takes seconds and cross-compiles on a
box without a GPU.  The shared object is written next to this file: it is git-ignored but
travels with the source snapshot to the GPU box, and it is what ``ctypes`` loads.

Run as ``python -m distributed_tensorflow_models_b200._native.build [--force] [--ptxas-v]``.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"
OBJ_DIR = HERE / "build"
LIB_PATH = HERE / "libdtfb200.so"

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-lineinfo", "--std=c++17", "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall",
          "--expt-relaxed-constexpr", "-I", str(INCLUDE), "-I", str(CSRC)]


def nvcc_path() -> str:
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found; the native library cannot be built")
    return p


def sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _digest(src: Path) -> str:
    h = hashlib.sha256()
    h.update(src.read_bytes())
    for hdr in sorted(list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + list(INCLUDE.glob("*.h"))):
        h.update(hdr.read_bytes())
    h.update(" ".join(ARCH_FLAGS + COMMON).encode())
    return h.hexdigest()


def _compile_one(src: Path, force: bool, ptxas_v: bool) -> tuple[Path, str]:
    obj = OBJ_DIR / (src.stem + ".o")
    stamp = OBJ_DIR / (src.stem + ".sha")
    dig = _digest(src)
    if not force and obj.exists() and stamp.exists() and stamp.read_text() == dig:
        return obj, ""
    cmd = [nvcc_path(), *ARCH_FLAGS, *COMMON, "-c", str(src), "-o", str(obj)]
    if ptxas_v:
        cmd += ["-Xptxas", "-v"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(dig)
    return obj, (r.stdout + r.stderr)


def build(force: bool = False, ptxas_v: bool = False, verbose: bool = False) -> Path:
    """Compile every ``csrc/*.cu`` for sm_100a and link ``libdtfb200.so``. Returns its path."""
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    srcs = sources()
    if not srcs:
        raise RuntimeError(f"no CUDA sources under {CSRC}")
    logs = []
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        futs = [ex.submit(_compile_one, s, force, ptxas_v) for s in srcs]
        objs = []
        for f in futs:
            o, log = f.result()
            objs.append(o)
            if log:
                logs.append(log)
    newest = max(o.stat().st_mtime for o in objs)
    if force or not LIB_PATH.exists() or LIB_PATH.stat().st_mtime < newest:
        tmp = LIB_PATH.with_suffix(".so.tmp")
        cmd = [nvcc_path(), *ARCH_FLAGS, "-shared", "-o", str(tmp), *map(str, objs), "-lcudart", "-lpthread"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        os.replace(tmp, LIB_PATH)
    if verbose or ptxas_v:
        for log in logs:
            sys.stderr.write(log)
    return LIB_PATH


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, ptxas_v="--ptxas-v" in sys.argv, verbose=True)
    print(path)
