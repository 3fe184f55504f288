"""CPU: the C-ABI library builds, loads and exports every symbol include/ecolab_b200.h declares
(no compute calls -- there is no GPU here), and the product refuses to run without a device."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ecolab_b200.h")
LIB = os.path.join(ROOT, "ecolab_b200", "libecolab_b200.so")


def declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(eco_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def built():
    if not os.path.exists(LIB):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "ecolab_b200", "csrc"), "-s"])
    return LIB


def test_header_declares_the_hot_path():
    names = declared()
    for n in ("eco_generate", "eco_migrate", "eco_condense", "eco_mutate", "eco_set_grid", "eco_set_density", "eco_step"):
        assert n in names
    assert len(names) >= 50


def test_library_exports_every_declared_symbol(built):
    L = ctypes.CDLL(built)
    missing = [n for n in declared() if not hasattr(L, n)]
    assert not missing, missing


def test_python_binding_matches_header(built):
    from ecolab_b200 import _cabi
    bound = set(_cabi.SIGNATURES) | set(_cabi._SPECIAL)
    assert bound == set(declared())
    _cabi.lib()


def test_every_entry_point_cites_the_reference():
    src = open(HEADER).read()
    assert len(re.findall(r":\d+(-\d+)?", src)) >= 30 and "models/ecolab_model" in src


def test_no_cpu_fallback(built):
    """Without a CUDA device the product must fail loudly, not fall back (the oracle is never reachable from it)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import ecolab_b200
    with pytest.raises(ecolab_b200.EcolabError):
        ecolab_b200.PanmicticModel(0)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "ecolab_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cpp", ".cu", ".h", "Makefile")):
                txt = open(os.path.join(dp, f), errors="replace").read()
                for pat in ("import oracle", "from oracle", "libecolab_oracle", "orc_", "oracle/"):
                    assert pat not in txt, (os.path.join(dp, f), pat)


def test_sass_is_sm100a_with_bulk_copy(built):
    out = subprocess.run(["cuobjdump", "-sass", built], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    assert "UBLKCP" in out          # cp.async.bulk (TMA unit) staging of the generate tiles
    # no FP64 contraction in the reference-order arithmetic: bit-exactness vs the reference's x86-64 build.  DFMA may
    # appear in exactly two places: the tensor-core path, whose Horner recombination of integer accumulators is exact
    # either way (k_dense.cuh), and the generate kernels' exact product fma(val, 2^52+n, -val*2^52) = fl(val*n)
    # (k_generate.cuh, template flag FMA = true, the last template argument).  The FMA = false twins of the same source
    # prove that nothing else in generate contracts; migrate, mutate, condense and the diagnostics hold none.
    funcs = out.split("Function : ")
    assert any("generate_dense_kernel" in f.split("\n", 1)[0] for f in funcs)
    twins = 0
    for f in funcs[1:]:
        name = f.split("\n", 1)[0]
        fma_generate = re.search(r"generate_(pipe|multi)_kernelI.*Lb1EEEvNS", name) is not None
        if re.search(r"generate_pipe_kernelI.*Lb0EEEvNS", name):
            twins += 1
        if "generate_dense_kernel" not in name and not fma_generate:
            assert "DFMA" not in f, name
    assert twins >= 4
    assert "UTCIMMA" in out and "LDTM" in out     # tcgen05.mma kind::i8 and tcgen05.ld of the dense generate kernel
