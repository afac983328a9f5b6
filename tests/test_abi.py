"""CPU tests of the drop-in boundary: the C-ABI library builds for sm_100a, loads, and exports every symbol
that include/wxddp.h declares (no compute calls -- there is no GPU here)."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "wxddp.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(wx_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def wx():
    import wxddp
    wxddp.build_library()          # no-op when the in-tree .so is current
    return wxddp


def test_header_declares_the_expected_surface():
    syms = header_symbols()
    for must in ["wx_create", "wx_destroy", "wx_last_error", "wx_label_encode", "wx_col_stats", "wx_stats_merge",
                 "wx_normalize_pack", "wx_impute_index_list", "wx_mlp_grad", "wx_adam_step", "wx_train_step",
                 "wx_train_step_raw", "wx_train_steps_small", "wx_mlp_eval", "wx_mlp_predict", "wx_build_arch"]:
        assert must in syms


def test_library_loads_and_exports_every_declared_symbol(wx):
    lib = wx.load()
    from wxddp import _lib
    declared = header_symbols()
    assert sorted(_lib.SIGNATURES) == declared, set(declared) ^ set(_lib.SIGNATURES)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.wx_build_arch() == b"sm_100a"
    assert lib.wx_version() >= 100
    out = subprocess.run(["nm", "-D", "--defined-only", wx.lib_path()], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (wx_[a-z0-9_]+)", out))
    assert set(declared) <= exported


def test_library_contains_sm100a_code_only(wx):
    out = subprocess.run(["cuobjdump", "--list-elf", wx.lib_path()], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_argument_errors_need_no_gpu(wx):
    lib = wx.load()
    # NULL ctx is rejected before any CUDA call, with a message
    assert lib.wx_mlp_grad(None, None, None, None, None, 0, 1, 0, None, 0, 0, 1.0, None, None) == -1
    assert b"ctx is NULL" in lib.wx_last_error()
    assert lib.wx_n_params(None) == -1


def test_context_without_gpu_fails_loudly(wx):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(wx.WxError):
        wx.Context(5)


def _sass_of(lib_path: str, mangled: str) -> str:
    r = subprocess.run(["cuobjdump", "-sass", "-fun", mangled, lib_path], capture_output=True, text=True, timeout=120)
    return "\n".join(ln for ln in r.stdout.splitlines() if re.match(r"\s+/\*[0-9a-f]+\*/", ln))


@pytest.mark.skipif(subprocess.run(["which", "cuobjdump"], capture_output=True).returncode != 0, reason="no cuobjdump")
def test_the_dominant_kernels_are_tcgen05_code(wx):
    """What the bench lines name as the dominant kernels must be Blackwell tensor-core code in the built library: UTC*MMA
    (tcgen05.mma), LDTM / STTM (tcgen05.ld / st) and, for the wide GEMM, UTMALDG (TMA tensor loads) -- and no legacy
    HMMA (mma.sync) anywhere in them.  (SASS mnemonics: /opt/skills/guides/B200_PROFILING.md; the whole library is
    listed in profiles/r2_sass_mnemonics.txt.)"""
    lib = wx.lib_path() if hasattr(wx, "lib_path") else None
    if lib is None:
        from wxddp import _lib
        lib = _lib.lib_path()
    step = _sass_of(lib, "_ZN8wxtcstep14tc_step_kernelILb1ELi2ELi20EEEvN3wxk8GridArgsEiNS_6TcGeomEPi")
    assert len(step.splitlines()) > 1000, "tc_step_kernel<true, PHILOX, 20> not found in the library"
    assert len(re.findall(r"\bUTC[A-Z]*MMA\b", step)) >= 25          # 9 GEMM1 + 16 GEMM2 MMAs per tile
    assert "LDTM" in step and "STTM" in step and "LDGSTS" in step    # TMEM operands / accumulators, cp.async producer
    assert "FFMA2" in step and not re.search(r"\bHMMA\b", step)
    gemm = _sass_of(lib, "_ZN6wxwide17wide_gemm2_kernelILb1ELb1ELi0ELi2ELi8EEEv14CUtensorMap_stS1_NS_8GemmArgsE")
    assert len(gemm.splitlines()) > 500, "wide_gemm2_kernel<K, K, FWD, 256x512, 8> not found in the library"
    assert re.search(r"\bUTC[A-Z]*MMA\b", gemm) and "UTMALDG" in gemm and "LDTM" in gemm
    assert not re.search(r"\bHMMA\b", gemm)
