"""In-tree build of the native library ``_lib/libdapb200.so`` for sm_100a.

``python -m distributed_algorithm_pyspark_b200.build`` (or ``__graft_entry__.build()``) compiles every
``csrc/*.cu`` with ``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo`` and links one shared
library next to the package, so the binary travels with the source tree (no JIT cache).  nvcc
cross-compiles without a GPU.  The reference has no build system at all.
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from pathlib import Path
from typing import List

PKG_DIR = Path(__file__).resolve().parent
REPO = PKG_DIR.parent
CSRC = REPO / "csrc"
INCLUDE = REPO / "include"
LIB_DIR = PKG_DIR / "_lib"
OBJ_DIR = REPO / "build" / "obj"
LIB_PATH = LIB_DIR / "libdapb200.so"

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
                "--expt-relaxed-constexpr", "-I", str(INCLUDE), "-I", str(CSRC)]
# per-file extras: the fp64 parity kernel must not contract a*b+c differently from NumPy
PER_FILE = {"lr.cu": ["-fmad=false"]}


def nvcc_path() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found; the native library cannot be built")
    return exe


def sources() -> List[Path]:
    return sorted(CSRC.glob("*.cu"))


def _headers_mtime() -> float:
    hs = list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + list(INCLUDE.glob("*.h"))
    return max((h.stat().st_mtime for h in hs), default=0.0)


def _compile(src: Path, verbose: bool, ptxas_v: bool) -> Path:
    obj = OBJ_DIR / (src.stem + ".o")
    newest = max(src.stat().st_mtime, _headers_mtime())
    if obj.exists() and obj.stat().st_mtime >= newest:
        return obj
    extra = os.environ.get("DAPB_EXTRA_NVCC_FLAGS", "").split()      # experiments only
    cmd = [nvcc_path(), *ARCH_FLAGS, *COMMON_FLAGS, *PER_FILE.get(src.name, []), *extra, "-c", str(src), "-o", str(obj)]
    if ptxas_v:
        cmd += ["-Xptxas", "-v"]
    if verbose:
        print(" ".join(cmd), flush=True)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{res.stdout}\n{res.stderr}")
    if ptxas_v or verbose:
        sys.stderr.write(res.stderr)
    return obj


def build(force: bool = False, verbose: bool = False, ptxas_v: bool = False) -> Path:
    """Compile (incrementally) and link the library; returns its path."""
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    if force:
        for o in OBJ_DIR.glob("*.o"):
            o.unlink()
    srcs = sources()
    if not srcs:
        raise RuntimeError(f"no CUDA sources under {CSRC}")
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose, ptxas_v), srcs))
    newest_obj = max(o.stat().st_mtime for o in objs)
    if force or not LIB_PATH.exists() or LIB_PATH.stat().st_mtime < newest_obj:
        cmd = [nvcc_path(), *ARCH_FLAGS, "-shared", "-o", str(LIB_PATH), *map(str, objs)]
        if verbose:
            print(" ".join(cmd), flush=True)
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    return LIB_PATH


def is_built() -> bool:
    return LIB_PATH.exists()


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--ptxas-v", action="store_true", help="print registers / spills / smem per kernel")
    a = ap.parse_args()
    print(build(force=a.force, verbose=a.verbose, ptxas_v=a.ptxas_v))
