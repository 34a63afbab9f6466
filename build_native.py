#!/usr/bin/env python
"""Builds ``libfps_b200.so`` (every .cu under csrc/) and ``libfps_cpu.so`` (csrc/cpu/*.cpp) in-tree.

    python build_native.py [--force] [--verbose]

nvcc cross-compiles for sm_100a without a GPU; objects are cached under build/ by source mtime.
"""
from __future__ import annotations

import argparse
import os
import pathlib
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = pathlib.Path(__file__).resolve().parent
CSRC = ROOT / "csrc"
BUILD = ROOT / "build"
PKG_LIB = ROOT / "flink-parameter-server-1_b200" / "ops" / "_lib"

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--expt-relaxed-constexpr"]
CXX_FLAGS = ["-O3", "-march=x86-64-v3", "-std=c++17", "-fPIC", "-pthread"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _newer(src: pathlib.Path, obj: pathlib.Path, deps) -> bool:
    if not obj.exists():
        return True
    t = obj.stat().st_mtime
    return any(p.stat().st_mtime > t for p in [src, *deps])


def build(force: bool = False, verbose: bool = False) -> dict:
    BUILD.mkdir(exist_ok=True)
    PKG_LIB.mkdir(parents=True, exist_ok=True)
    headers = list(CSRC.glob("*.cuh")) + list((ROOT / "include").glob("*.h"))
    cu = sorted(CSRC.glob("*.cu"))
    nvcc = _nvcc()
    logs = {}

    def compile_one(src: pathlib.Path):
        obj = BUILD / (src.stem + ".o")
        if not force and not _newer(src, obj, headers):
            return obj, ""
        cmd = [nvcc, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
        (BUILD / (src.stem + ".ptxas.log")).write_text(r.stderr)
        return obj, r.stderr

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(cu)))) as ex:
        results = list(ex.map(compile_one, cu))
    objs = [str(o) for o, _ in results]
    for (o, log), src in zip(results, cu):
        logs[src.name] = log
        if verbose and log:
            print(log)
    lib = PKG_LIB / "libfps_b200.so"
    if force or not lib.exists() or any(pathlib.Path(o).stat().st_mtime > lib.stat().st_mtime for o in objs):
        r = subprocess.run([nvcc, "-shared", "-o", str(lib), *objs, "-lcudart"], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    # CPU baseline library
    cpu_src = sorted((CSRC / "cpu").glob("*.cpp")) if (CSRC / "cpu").exists() else []
    if cpu_src:
        cpu_lib = PKG_LIB / "libfps_cpu.so"
        if force or not cpu_lib.exists() or any(s.stat().st_mtime > cpu_lib.stat().st_mtime for s in cpu_src):
            r = subprocess.run(["g++", *CXX_FLAGS, "-shared", "-o", str(cpu_lib), *map(str, cpu_src)],
                               capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"g++ failed:\n{r.stdout}\n{r.stderr}")
    return logs


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    build(a.force, a.verbose)
    print("built", PKG_LIB / "libfps_b200.so")
