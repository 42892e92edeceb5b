"""The C-ABI library loads and exports every symbol include/duckhts_cuda.h declares (no compute calls: no GPU here)."""
import ctypes
import os
import re

from duckhts_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "duckhts_cuda.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dh_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound():
    syms = declared_symbols()
    assert len(syms) >= 20
    lib = ctypes.CDLL(_lib.lib_path)
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/duckhts_cuda.h but not exported"
        assert s in _lib.SIGNATURES, f"{s} has no ctypes signature in duckhts_b200/_lib.py"
    for s in _lib.SIGNATURES:
        assert s in syms, f"{s} bound in _lib.py but not declared in the header"


def test_no_gpu_means_loud_failure_not_fallback():
    import duckhts_b200 as dh

    if dh.lib().dh_device_count() > 0:
        return
    try:
        dh.Context(0)
    except dh.DuckhtsError as e:
        assert "no CUDA device" in str(e)
    else:
        raise AssertionError("Context() must fail without a GPU: there is no CPU fallback")


def test_product_does_not_reference_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "duckhts_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".c", ".cpp", "Makefile")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle" not in txt.lower(), f"{f} mentions oracle"


def test_host_framing_matches_oracle():
    """dh_bgzf_walk is host code: compare with the oracle's walk on a python-made BGZF image."""
    import numpy as np

    import duckhts_b200 as dh
    from oracle import dh_oracle_py as orc
    from tests import util

    rng = np.random.default_rng(3)
    data = rng.integers(0, 4, 300000, dtype=np.uint8).tobytes()
    img = util.make_bgzf(data)
    a, enda = dh.bgzf_walk(img)
    b, endb = orc.bgzf_walk(img)
    assert a == b and enda == endb == len(img)
    # truncated tail and corrupt header
    try:
        dh.bgzf_walk(img[:-5])
        raise AssertionError("strict walk must reject a truncated block")
    except dh.DuckhtsError:
        pass
    bad = bytearray(img)
    bad[a[2][0] + 12] = ord("X")
    try:
        dh.bgzf_walk(bytes(bad))
        raise AssertionError("bad header accepted")
    except dh.DuckhtsError:
        pass
    assert orc.bgzf_walk(bytes(bad))[0] is None


def test_duckdb_extension_exports_only_the_entry_point():
    """Like the reference's build (CMAKE_C_VISIBILITY_PRESET hidden): one default-visibility symbol, duckhts_init_c_api."""
    import subprocess

    ext = os.path.join(ROOT, "duckhts_b200", "lib", "duckhts.duckdb_extension.so")
    assert os.path.exists(ext), "run __graft_entry__.build()"
    out = subprocess.check_output(["nm", "-D", "--defined-only", ext], text=True)
    names = [ln.split()[-1] for ln in out.splitlines() if ln.strip() and ln.split()[-2] in ("T", "t") or " T " in ln]
    assert "duckhts_init_c_api" in names
    assert [n for n in names if not n.startswith("_")] == ["duckhts_init_c_api"], names
