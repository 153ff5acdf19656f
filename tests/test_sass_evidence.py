"""What the built library contains, read from its SASS (B200_PROFILING.md, "What proves a Blackwell-native kernel"):
the GEMM issues tcgen05 MMAs with TMEM reads and TMA loads / stores, the gradient exchange uses multimem loads with
in-switch reduction.  Runs on the CPU-only box: cuobjdump disassembles the objects `build()` left in-tree."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_DIR = os.path.join(ROOT, "multimodal-emotion-recognition_b200")


def _sass(obj_name):
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    obj = os.path.join(PKG_DIR, "build", obj_name)
    if not os.path.exists(obj):
        pytest.skip(f"{obj} not built on this box (the library is; objects are not shipped)")
    r = subprocess.run([cuobjdump, "-sass", obj], capture_output=True, text=True, check=True)
    return r.stdout


def _count(sass, pattern):
    return len(re.findall(pattern, sass))


def test_gemm_is_tcgen05_tmem_tma(pkg):
    sass = _sass("gemm_sm100.o")
    assert "sm_100a" in sass
    assert _count(sass, r"\bUTCHMMA") > 0        # tcgen05.mma (bf16)
    assert _count(sass, r"\bLDTM") > 0           # tcgen05.ld: accumulators out of TMEM
    assert _count(sass, r"\bUTMALDG") > 0        # cp.async.bulk.tensor loads (operands, epilogue tiles)
    assert _count(sass, r"\bUTMASTG") > 0        # TMA tile stores of the results
    assert _count(sass, r"\bHMMA") == 0          # no legacy mma.sync path in the GEMM
    assert _count(sass, r"\bHGMMA") == 0


def test_exchange_reduces_in_the_switch(pkg):
    sass = _sass("exchange.o")
    assert _count(sass, r"\bLDGMC\.E\.ADD\.F32x4") > 0   # multimem.ld_reduce.add.v4.f32
    assert _count(sass, r"\bATOMG\.E\.CAS\.STRONG\.SYS") > 0  # barrier words: system-scope CAS
