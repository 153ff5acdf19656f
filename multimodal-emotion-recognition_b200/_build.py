"""In-tree build of libm2f_b200.so with nvcc for sm_100a (no torch headers: the library is a plain C ABI)."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
BUILD = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libm2f_b200.so")
SOURCES = ["abi.cu", "gemm_sm100.cu", "gemm_persist.cu", "layernorm.cu", "attention.cu", "attention_mma_p1.cu", "attention_mma_p2.cu", "attention_mma_p3.cu", "attention_mma_long.cu", "attention_tc.cu", "loss.cu", "loss_static.cu", "optim.cu", "pipeline.cu", "exchange.cu", "plan.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--use_fast_math=false"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libm2f_b200.so cannot be built (there is no CPU fallback)")


def _deps():
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hdrs.append(os.path.join(os.path.dirname(HERE), "include", "m2f_b200.h"))
    return hdrs


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every CUDA source for sm_100a and link the shared library.  Returns its path."""
    os.makedirs(BUILD, exist_ok=True)
    nvcc = _nvcc()
    hdrs = _deps()
    srcs = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    jobs = []
    for s in srcs:
        src, obj = os.path.join(CSRC, s), os.path.join(BUILD, s.replace(".cu", ".o"))
        if force or _stale(obj, [src] + hdrs):
            jobs.append((src, obj))

    # build-time experiments: extra -D definitions (e.g. M2F_NVCC_DEFS="-DM2F_PAIR_RING_KB=96 -DM2F_PAIR_EPI_WARPS=4");
    # when set, everything is recompiled, and the next build without it recompiles again (stamp file)
    extra = os.environ.get("M2F_NVCC_DEFS", "").split()
    stamp = os.path.join(BUILD, "defs.stamp")
    prev = open(stamp).read() if os.path.exists(stamp) else ""
    if " ".join(extra) != prev:
        jobs = [(os.path.join(CSRC, s), os.path.join(BUILD, s.replace(".cu", ".o"))) for s in srcs]
        with open(stamp, "w") as f:
            f.write(" ".join(extra))

    def compile_one(job):
        src, obj = job
        cmd = [nvcc] + [f for f in NVCC_FLAGS if not f.startswith("--use_fast_math")] + extra + ["-c", src, "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)

    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(compile_one, jobs))
    objs = [os.path.join(BUILD, s.replace(".cu", ".o")) for s in srcs]
    if force or jobs or _stale(LIB, objs):
        cmd = [nvcc, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
