"""The C-ABI library builds, loads, and exports exactly what include/tb_b200.h declares; without
a GPU the compute entry points fail loudly instead of falling back. CPU only."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "tb_b200.h")


@pytest.fixture(scope="module")
def lib():
    from taboo_brittleness_b200 import _lib, build
    build.build()
    return _lib.load()


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tb_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported(lib):
    names = _declared()
    assert "tb_lens_partials" in names and "tb_sae_encode" in names
    for n in names:
        assert hasattr(lib, n), f"{n} declared in tb_b200.h but not exported"
    from taboo_brittleness_b200._lib import SIGNATURES
    assert sorted(SIGNATURES) == names, "ctypes prototypes and header disagree"


def test_argument_counts_match_header():
    src = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    from taboo_brittleness_b200._lib import SIGNATURES
    for name, (_, argtypes) in SIGNATURES.items():
        m = re.search(name + r"\s*\((.*?)\)\s*;", src, flags=re.S)
        assert m, name
        args = m.group(1).strip()
        n = 0 if args in ("", "void") else args.count(",") + 1
        assert n == len(argtypes), f"{name}: header has {n} parameters, ctypes {len(argtypes)}"


def test_sass_has_blackwell_tensor_and_tma_instructions(lib):
    from taboo_brittleness_b200._lib import LIB_PATH
    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("no cuobjdump")
    sass = subprocess.run([cuobjdump, "-sass", LIB_PATH], capture_output=True, text=True).stdout
    for mnemonic in ("UTCHMMA", "UTMALDG", "LDTM"):
        assert mnemonic in sass, f"{mnemonic} missing: not a tcgen05/TMA build"
    assert "HMMA.16816" not in sass


def test_no_gpu_means_loud_failure(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    assert lib.tb_version() >= 100
    rc = lib.tb_device_check()
    assert rc != 0
    assert lib.tb_last_error()
    rc = lib.tb_rmsnorm(None, None, ctypes.c_float(1e-6), 1, 8, None, None)
    assert rc != 0


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "taboo_brittleness_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
