"""CPU: the C-ABI library builds, loads, and exports every symbol include/ofx.h declares; the
product never touches the oracle; without a GPU the product fails loudly (no fallback)."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "ofx.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ofx_[a-z_0-9]+)\s*\(", text)))


def test_header_symbols_exported():
    from occlusionfusion_b200 import _capi, build
    lib_path = build.build()
    lib = ctypes.CDLL(lib_path)
    syms = _header_symbols()
    assert len(syms) >= 16
    for s in syms:
        assert hasattr(lib, s), f"libofx.so does not export {s}"
    assert sorted(_capi.SYMBOLS) == syms, "occlusionfusion_b200/_capi.py SYMBOLS out of sync with include/ofx.h"
    lib.ofx_abi_version.restype = ctypes.c_int
    assert lib.ofx_abi_version() == 1
    nm = subprocess.run(["nm", "-D", "--defined-only", lib_path], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (ofx_[a-z_0-9]+)", nm))
    assert set(syms) <= exported


def test_library_is_sm100a_with_tma():
    from occlusionfusion_b200 import build
    lib_path = build.build()
    out = subprocess.run(["cuobjdump", "-lelf", lib_path], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    sass = subprocess.run(["cuobjdump", "-sass", lib_path], capture_output=True, text=True).stdout
    assert "UBLKCP" in sass, "the brick kernels should stage node records with TMA bulk copies"
    assert "SYNCS" in sass, "mbarrier pipeline missing"
    # the running averages of the persistent kernel work on voxel pairs with the packed float32 instructions of sm_100
    persistent = sass[sass.index("fused_warped_persistentILi4ELb1"):]
    persistent = persistent[:persistent.index("Function :", 10)] if "Function :" in persistent[10:] else persistent
    for op in ("FFMA2", "FADD2", "FMUL2"):
        assert op in persistent, f"{op} missing from fused_warped_persistent<4, true>"
    # ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 even under --fmad=false; the colour blend
    # w*old + obs*new must keep the rounding of its products (ofx_fused.cuh: mul2_pos = FFMA2 with a zero addend)
    assert re.search(r"FFMA2 R\d+, R\d+\.F32x2\.HI_LO, R\d+\.F32x2\.HI_LO, RZ", persistent), \
        "the unfused product (FFMA2 ..., RZ) of the colour blend is gone: check for a contracted mul+add"


def test_product_does_not_touch_oracle():
    pkg = os.path.join(ROOT, "occlusionfusion_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dp, f)).read()
                assert "import oracle" not in text and "libofx_oracle" not in text and "ofx_oracle_" not in text, f


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from occlusionfusion_b200 import TSDFVolume, _capi
    from occlusionfusion_b200.synthetic import Workload
    w = Workload("tiny")
    with pytest.raises(_capi.OfxError):
        TSDFVolume(w.bbox, w.max_depth, w.intr, w.fopt, None)
    # the C entry point itself refuses too
    lib = _capi.load()
    h = ctypes.c_void_p()
    desc = _capi.VolumeDesc()
    desc.dims[:] = [8, 8, 8]
    desc.n_shards, desc.shard, desc.x_block = 1, 0, 8
    desc.voxel_size, desc.trunc_margin = 0.01, 0.04
    desc.tsdf = desc.weight = desc.colour = 16
    assert lib.ofx_volume_create(ctypes.byref(h), ctypes.byref(desc)) != 0
    assert b"no CUDA device" in lib.ofx_last_error()
