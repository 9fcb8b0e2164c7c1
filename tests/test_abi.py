"""The C-ABI library loads and exports every symbol include/phylo_b200.h declares (no compute calls)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "phylo_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(phb_[a-z0-9_]+)\s*\(", txt)))


def test_header_and_binding_agree():
    from phylo_rs_b200 import capi

    assert sorted(capi.EXPORTS) == declared_symbols()


def test_product_library_exports_every_symbol():
    import __graft_entry__ as g

    path = g.build_cuda()
    lib = ctypes.CDLL(path)
    for name in declared_symbols():
        assert hasattr(lib, name), name
    lib.phb_version.restype = ctypes.c_char_p
    assert b"sm_100a" in lib.phb_version()


def test_product_library_has_only_sm100a_code():
    import subprocess

    import __graft_entry__ as g

    out = subprocess.run(["cuobjdump", "-lelf", g.build_cuda()], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_no_gpu_means_loud_failure():
    """Without a CUDA device the product library refuses to create a context (no CPU fallback)."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from phylo_rs_b200 import capi

    with pytest.raises(capi.PhbError) as e:
        capi.Context()
    assert e.value.variant == "Cuda"


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "phylo-rs_b200")
    for dp, _dn, fn in os.walk(pkg):
        for f in fn:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), f
                assert "phylo_oracle" not in src, f
