"""The C-ABI library loads on a CPU-only box and exports every symbol include/m2f_b200.h declares; the ctypes binding
lists a prototype for each of them (no compute calls here)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared():
    src = open(os.path.join(ROOT, "include", "m2f_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(m2f_[a-z0-9_]+)\s*\(", src)) - {"m2f_bucket_cb"})


def test_header_declares_the_expected_surface():
    names = declared()
    for must in ("m2f_gemm", "m2f_attn_fwd", "m2f_attn_bwd", "m2f_ln_fwd", "m2f_ln_bwd", "m2f_ce_fwd_bwd", "m2f_plan_forward",
                 "m2f_plan_backward", "m2f_plan_backward_bucketed", "m2f_adam_step", "m2f_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol(pkg):
    lib = ctypes.CDLL(pkg._lib.LIB_PATH)
    for n in declared():
        assert hasattr(lib, n), n


def test_binding_covers_every_declared_symbol(pkg):
    assert sorted(pkg._lib.PROTOTYPES) == declared()
    L = pkg._lib.lib()
    assert L.m2f_abi_version() == 2
    assert L.m2f_launch_count() == 0


def test_errors_are_codes_with_messages(pkg):
    L = pkg._lib.lib()
    cfg = pkg._lib.ModelCfg()  # everything disabled
    h = ctypes.c_void_p()
    rc = L.m2f_model_create(ctypes.byref(cfg), ctypes.byref(h))
    assert rc == -1 and b"At least one of audio and text must be enabled!" in L.m2f_last_error()


def test_struct_layouts_match_the_header(pkg, tmp_path):
    """ctypes mirrors of the header's structs: sizes and field offsets as gcc lays them out."""
    import subprocess
    L = pkg._lib
    pairs = [("m2f_planes_t", L.Planes), ("m2f_kseg_t", L.KSeg), ("m2f_epilogue_t", L.Epilogue), ("m2f_modality_cfg_t", L.ModalityCfg),
             ("m2f_model_cfg_t", L.ModelCfg), ("m2f_peer_group_t", L.PeerGroup)]
    prog = ['#include <stdio.h>', '#include <stddef.h>', '#include "m2f_b200.h"', 'int main(void) {']
    for cname, cls in pairs:
        prog.append(f'printf("{cname} %zu\\n", sizeof({cname}));')
        for f, *_ in cls._fields_:
            prog.append(f'printf("{cname}.{f} %zu\\n", offsetof({cname}, {f}));')
    prog += ['return 0; }']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(prog))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = dict(line.split() for line in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    for cname, cls in pairs:
        assert int(got[cname]) == ctypes.sizeof(cls), cname
        for f, *_ in cls._fields_:
            assert int(got[f"{cname}.{f}"]) == getattr(cls, f).offset, (cname, f)
    assert L.XCHG_SIGNAL_BYTES == 64 * 16 * 4


def test_exchange_argument_errors_without_a_gpu(pkg):
    L = pkg._lib
    lib = L.lib()
    assert lib.m2f_grad_allreduce_peer(None, 0, 16, 1.0, 0, 0, None) == L.M2F_ERR_ARG
    pg = L.PeerGroup()
    pg.rank, pg.world = 0, 17
    assert lib.m2f_grad_allreduce_peer(ctypes.byref(pg), 0, 16, 1.0, 0, 0, None) == L.M2F_ERR_ARG
    assert b"world" in lib.m2f_last_error()
