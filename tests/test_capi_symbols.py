"""CPU checks of the drop-in boundary: liblgs_b200.so loads, exports every symbol include/lgs_b200.h declares, the
ctypes table mirrors the header, and without a CUDA device the library refuses to work (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT

HEADER = os.path.join(ROOT, "include", "lgs_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"LGS_API\s+[\w\s\*]+?\b(lgs_\w+)\s*\(", src)))


@pytest.fixture(scope="module")
def built_lib():
    import __graft_entry__ as g
    g.build()
    from lattice_gaussian_mcmc_b200 import _capi
    return _capi


def test_header_declares_expected_entry_points():
    syms = declared_symbols()
    for must in ("lgs_create", "lgs_destroy", "lgs_klein_sample", "lgs_klein_replay", "lgs_imhk_run",
                 "lgs_imhk_replay", "lgs_klein_sample_host", "lgs_imhk_counts_host", "lgs_histogram"):
        assert must in syms
    assert len(syms) >= 20


def test_library_exports_every_declared_symbol(built_lib):
    L = ctypes.CDLL(built_lib.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(L, name), f"{name} declared in include/lgs_b200.h but not exported"
    assert L.lgs_abi_version() == 1


def test_ctypes_table_mirrors_header(built_lib):
    assert sorted(built_lib.PROTOTYPES) == declared_symbols()


def test_no_cpu_fallback(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    import lattice_gaussian_mcmc_b200 as lgs
    from lattice_gaussian_mcmc_b200.lattices import identity_lattice
    with pytest.raises(built_lib.LgsError) as ei:
        lgs.KleinSampler(identity_lattice(4), 2.0)
    assert ei.value.code == built_lib.LGS_E_CUDA
    with pytest.raises(built_lib.LgsError):
        lgs.IMHKSampler(identity_lattice(4), 2.0, burn_in=0)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "lattice-gaussian-mcmc_b200")
    bad = re.compile(r"(import\s+oracle|from\s+oracle|liblgs_oracle|#include\s+[\"<].*oracle)")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not bad.search(txt), f"{f} links or imports the oracle"
