"""CPU: the C-ABI library builds, loads, and exports exactly what include/pb_b200.h declares."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pb_b200.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pb_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def libpath():
    import __graft_entry__ as ge
    ge.build()
    import proteinbert_b200 as pb
    assert os.path.exists(pb._lib.LIB_PATH)
    return pb._lib.LIB_PATH


def test_header_declares_the_hot_path_ops():
    syms = declared_symbols()
    for need in ("pb_conv_pair_fwd", "pb_conv_pair_bwd", "pb_ln_joint_apply", "pb_ln_joint_bwd", "pb_local_linear_fwd",
                 "pb_attn_pool_fwd", "pb_token_head_fwd", "pb_pretrain_loss", "pb_corrupt_pad", "pb_embedding_fwd",
                 "pb_linear_fwd", "pb_res_ln_rows_fwd"):
        assert need in syms


def test_library_exports_every_declared_symbol(libpath):
    h = ctypes.CDLL(libpath)
    for s in declared_symbols():
        assert hasattr(h, s), f"{s} declared in pb_b200.h but not exported"
    out = subprocess.run(["nm", "-D", "--defined-only", libpath], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (pb_[a-z0-9_]+)$", out, flags=re.M))
    assert set(declared_symbols()) <= exported


def test_python_prototypes_cover_the_header(libpath):
    import proteinbert_b200 as pb
    assert set(pb._lib.PROTOTYPES) == set(declared_symbols())
    assert pb._lib.lib().pb_version() >= 100
    assert pb._lib.lib().pb_workspace_bytes(b"linear_fwd", 64, 512, 8943) == 32 * 64 * 512 * 4


def test_native_code_is_sm100a(libpath):
    out = subprocess.run(["cuobjdump", "-lelf", libpath], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_stage_desc_layout_matches_python_table():
    """The numpy record the host builds must be byte-compatible with `struct pb_stage_desc`."""
    import numpy as np
    import proteinbert_b200 as pb

    class Desc(ctypes.Structure):
        _fields_ = [("src", ctypes.c_void_p), ("dst0", ctypes.c_void_p), ("dst1", ctypes.c_void_p)] + \
                   [(n, ctypes.c_int) for n in ("kind", "d0", "d1", "d2", "ld0", "ld1", "reserved")]
    dt = np.dtype(pb.ops._STAGE_DTYPE)
    assert dt.itemsize == ctypes.sizeof(Desc) == 56
    for name in ("src", "dst0", "dst1", "kind", "d0", "d1", "d2", "ld0", "ld1", "reserved"):
        assert dt.fields[name][1] == getattr(Desc, name).offset, name
