"""In-tree build of the native libraries.

Two shared objects are produced under ``b200flow/lib/``:

* ``libb200flow_cuda.so`` — every sm_100a kernel under ``csrc/kernels`` and
  ``csrc/comm`` (nvcc, ``-gencode arch=compute_100a,code=sm_100a -lineinfo``).
* ``libb200flow_cuda_fp16.so`` — the same sources with ``-DB2F_HALF_FP16``: IEEE fp16 as the 16-bit storage /
  tensor-core operand type instead of bf16 (``B2F_HALF=fp16``; the reference's GPT-2 examples train with ``--fp16``).
* ``libb200flow_host.so`` — pure C++ runtime pieces under ``csrc/host`` (dataset
  index builders, shared-memory mailbox, bucket planner); loads without a GPU.

The build is incremental (mtime based) and parallel, and needs no GPU: nvcc
cross-compiles.  ``python -m b200flow.build`` builds everything.
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
LIB_DIR = Path(__file__).resolve().parent / "lib"
OBJ_DIR = ROOT / "build" / "obj"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
    "-I", str(ROOT / "include"),
    *os.environ.get("B2F_NVCC_EXTRA", "").split(),  # e.g. -DB2F_ATTN_DEBUG for the timing counters
]
CXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-Wall", "-pthread", "-I", str(ROOT / "include")]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; set NVCC=/path/to/nvcc")


def _newer(src: Path, dst: Path, extra_deps: list[Path]) -> bool:
    if not dst.exists():
        return True
    t = dst.stat().st_mtime
    return any(p.stat().st_mtime > t for p in [src, *extra_deps])


def _headers() -> list[Path]:
    return [*CSRC.rglob("*.cuh"), *CSRC.rglob("*.h"), *(ROOT / "include").glob("*.h")]


def _compile(cmd: list[str], log: Path) -> None:
    res = subprocess.run(cmd, capture_output=True, text=True)
    log.write_text(" ".join(cmd) + "\n" + res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError(f"compile failed: {' '.join(cmd)}\n{res.stdout}\n{res.stderr}")


def build_cuda(verbose: bool = False, half: str = "bf16", variant: str = "", variant_flags=()) -> Path:
    """`variant`: an instrumented copy next to the product libraries (e.g. "attndbg" with -DB2F_ATTN_DEBUG for the
    cycle counters); a process picks it with B2F_CUDA_LIB=libb200flow_cuda_<variant>.so."""
    nvcc = _nvcc()
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    obj_dir = OBJ_DIR if half == "bf16" else OBJ_DIR.parent / "obj_fp16"
    if variant:
        obj_dir = OBJ_DIR.parent / f"obj_{variant}"
    obj_dir.mkdir(parents=True, exist_ok=True)
    extra = ([] if half == "bf16" else ["-DB2F_HALF_FP16"]) + list(variant_flags)
    hdrs = _headers()
    srcs = sorted([*CSRC.glob("kernels/*.cu"), *CSRC.glob("comm/*.cu"), *CSRC.glob("common/*.cpp")])
    jobs, objs = [], []
    for s in srcs:
        o = obj_dir / (s.parent.name + "_" + s.stem + ".o")
        objs.append(o)
        if _newer(s, o, hdrs):
            cmd = [nvcc, *NVCC_FLAGS, *extra, "-x", "cu", "-c", str(s), "-o", str(o)]
            jobs.append((cmd, o.with_suffix(".log")))
    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(lambda j: _compile(*j), jobs))
    out = LIB_DIR / ("libb200flow_cuda.so" if half == "bf16" else "libb200flow_cuda_fp16.so")
    if variant:
        out = LIB_DIR / f"libb200flow_cuda_{variant}.so"
    if jobs or not out.exists():
        _compile([nvcc, "-shared", "-o", str(out), *map(str, objs), "-lcudart"],
                 obj_dir / "link_cuda.log")
    if verbose:
        print(f"built {out} ({len(jobs)} recompiled / {len(srcs)} sources)")
    return out


def build_host(verbose: bool = False) -> Path:
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    cxx = os.environ.get("CXX", "g++")
    srcs = sorted(CSRC.glob("host/*.cpp"))
    hdrs = _headers()
    out = LIB_DIR / "libb200flow_host.so"
    if srcs and (not out.exists() or any(_newer(s, out, hdrs) for s in srcs)):
        _compile([cxx, *CXX_FLAGS, "-shared", "-o", str(out), *map(str, srcs), "-lrt"],
                 OBJ_DIR / "link_host.log")
    if verbose:
        print(f"built {out} ({len(srcs)} sources)")
    return out


def build_all(verbose: bool = False) -> None:
    build_host(verbose)
    build_cuda(verbose)
    build_cuda(verbose, half="fp16")


if __name__ == "__main__":
    build_all(verbose=True)
    if "--ptxas" in sys.argv:
        for log in sorted(OBJ_DIR.glob("*.log")):
            print("==", log.name)
            print(log.read_text())
