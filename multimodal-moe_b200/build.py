"""Build libb200moe.so (the C-ABI CUDA library) in-tree for sm_100a with nvcc.

    python multimodal-moe_b200/build.py [--force] [--verbose]

nvcc cross-compiles without a GPU; the .so is git-ignored but travels to the GPU box with the snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT_DIR = os.path.join(HERE, "b200moe")
# B200MOE_BUILD_DIR / B200MOE_LIB_OUT redirect the objects / the library (compile checks that must not replace the
# library a queued GPU run is about to pick up)
LIB = os.environ.get("B200MOE_LIB_OUT", os.path.join(OUT_DIR, "libb200moe.so"))
OBJ_DIR = os.environ.get("B200MOE_BUILD_DIR", os.path.join(HERE, "build"))
SOURCES = ["api.cu", "router.cu", "router_bwd.cu", "plan.cu", "movement.cu", "gemm.cu", "ep.cu", "ep_p2p.cu", "layer.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]


def _deps(src: str) -> list[str]:
    deps = [os.path.join(CSRC, src), os.path.join(HERE, "..", "include", "b200moe.h")]
    deps += [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".inc"))]
    return deps


def _stale(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_lib(force: bool = False, verbose: bool = False, report: list | None = None) -> str:
    """Compile every stale (or, with `force`, every) translation unit and link.  `report` receives the names of the
    objects that were actually rebuilt."""
    os.makedirs(OBJ_DIR, exist_ok=True)
    extra = ["-Xptxas", "-v"] if verbose else []

    def compile_one(src: str) -> str:
        obj = os.path.join(OBJ_DIR, src.replace(".cu", ".o"))
        if force or _stale(obj, _deps(src)):
            cmd = [NVCC, *FLAGS, *extra, "-c", os.path.join(CSRC, src), "-o", obj]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
            if verbose:
                sys.stderr.write(r.stderr)
            if report is not None:
                report.append(os.path.basename(obj))
        return obj

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    if force or _stale(LIB, objs):
        cmd = [NVCC, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        if report is not None:
            report.append(os.path.basename(LIB))
    return LIB


if __name__ == "__main__":
    print(build_lib(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
