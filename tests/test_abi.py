"""CPU: the C-ABI library loads and exports exactly the symbols include/s2p_b200.h
declares; the Python binding covers each of them; no compute is launched."""
import ctypes
import os
import re
import subprocess

import pytest


def _header_symbols(root):
    txt = open(os.path.join(root, "include", "s2p_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(s2p_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(repo_root):
    from s2p_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    syms = _header_symbols(repo_root)
    assert len(syms) >= 13
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/s2p_b200.h but not exported"
    assert set(syms) == set(_lib.SIGNATURES), "ctypes binding and header disagree"
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (s2p_[a-z0-9_]+)", out))
    assert exported == set(syms), f"undeclared exports: {exported ^ set(syms)}"


def test_host_only_entry_points(repo_root):
    from s2p_b200 import _lib
    lib = _lib.load()
    assert lib.s2p_version() == 1
    # 100x100 -> 25x5 tiles (125 of 128 MMA rows used): 80 tiles, one statistics partial row each
    assert lib.s2p_conv_stats_rows(100, 100, 256, 3, 1) == 80
    assert lib.s2p_conv_stats_rows(50, 50, 256, 3, 1) == 20
    assert lib.s2p_conv_stats_rows(25, 25, 512, 3, 1) == 5
    # Cout <= 64, 3x3 stride 1: horizontal-halo tiles (25 x 4 pixels in 27-row pitch): 100 tiles
    assert lib.s2p_conv_stats_rows(100, 100, 64, 3, 1) == 100
    assert lib.s2p_conv_stats_rows(100, 100, 64, 1, 1) == 80
    assert lib.s2p_stem_stats_rows(100, 100) == 13   # 8 full-width rows per CTA
    assert lib.s2p_mat_fused_ct(512) == 128 and lib.s2p_mat_fused_ct(64) == 64
    assert lib.s2p_launch_count() == 0


def test_bad_arguments_are_reported_not_launched():
    from s2p_b200 import _lib
    lib = _lib.load()
    rc = lib.s2p_conv_igemm_fwd(16, 64, 16, None, None, 0, 16, 64, None, 1, 25, 25, 48, 64, 3, 1, 1, 0, None)
    assert rc == -1 and b"Cin" in lib.s2p_last_error_string()
    rc = lib.s2p_mat_norm_fwd(None, None, 64, 1, None, None, None, 0, None, 0, None, None, 64, 1, 25, 25, 64, 2, 0, None)
    assert rc == -1
    assert lib.s2p_launch_count() == 0


def test_sass_is_blackwell_native(repo_root):
    """the conv kernel must contain tcgen05 MMA, TMEM loads and TMA loads (B200_PROFILING.md)."""
    from s2p_b200 import _lib
    r = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True)
    if r.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    for mnemonic in ("UTCHMMA", "LDTM", "UTMALDG"):
        assert mnemonic in r.stdout, f"{mnemonic} missing from SASS"
