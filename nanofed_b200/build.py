"""Build ``libnanofed_b200.so`` in-tree with nvcc for sm_100a.

    python -m nanofed_b200.build [--force] [--verbose]

Every ``csrc/*.cu`` is compiled to an object (in parallel) with
``-gencode arch=compute_100a,code=sm_100a -lineinfo`` and linked into
``nanofed_b200/_native/libnanofed_b200.so``.  No libtorch dependency, so a full
rebuild is seconds (tcgen05 files a little longer).  Objects are cached by
source mtime under ``build/obj``.  nvcc cross-compiles without a GPU.
"""

from __future__ import annotations

import argparse
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
OUT_DIR = ROOT / "nanofed_b200" / "_native"
OBJ_DIR = ROOT / "build" / "obj"
LIB = OUT_DIR / "libnanofed_b200.so"

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "--expt-relaxed-constexpr", "--expt-extended-lambda",
          "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default", "-I", str(ROOT / "include")]


def nvcc() -> str:
    found = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(found).exists():
        raise RuntimeError("nvcc not found (expected on PATH or /usr/local/cuda/bin)")
    return found


def _stale(src: Path, obj: Path) -> bool:
    if not obj.exists():
        return True
    newest = max([src.stat().st_mtime]
                 + [h.stat().st_mtime for h in list(CSRC.glob("*.cuh")) + list((ROOT / "include").glob("*.h"))])
    return obj.stat().st_mtime < newest


def _compile(src: Path, verbose: bool) -> Path:
    obj = OBJ_DIR / (src.stem + ".o")
    if _stale(src, obj):
        cmd = [nvcc(), *ARCH, *COMMON, "-c", str(src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
            print(" ".join(cmd), flush=True)
        proc = subprocess.run(cmd, capture_output=True, text=True)
        if proc.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{proc.stdout}\n{proc.stderr}")
        if verbose and proc.stderr:
            print(proc.stderr)
    return obj


def build(force: bool = False, verbose: bool = False) -> Path:
    OUT_DIR.mkdir(parents=True, exist_ok=True)
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    if force:
        for o in OBJ_DIR.glob("*.o"):
            o.unlink()
    sources = sorted(CSRC.glob("*.cu"))
    if not sources:
        raise RuntimeError(f"no CUDA sources under {CSRC}")
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 4)) as pool:
        objects = list(pool.map(lambda s: _compile(s, verbose), sources))
    if force or not LIB.exists() or any(o.stat().st_mtime > LIB.stat().st_mtime for o in objects):
        cmd = [nvcc(), *ARCH, "-shared", "-o", str(LIB), *map(str, objects), "-lcuda"]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        if proc.returncode != 0:
            raise RuntimeError(f"link failed:\n{proc.stdout}\n{proc.stderr}")
    return LIB


def main() -> None:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    args = ap.parse_args()
    print(build(args.force, args.verbose))


if __name__ == "__main__":
    sys.exit(main())
