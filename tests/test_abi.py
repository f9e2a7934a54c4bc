"""CPU tests of the drop-in boundary: the C-ABI library builds, loads and exports every symbol
include/b2mlp.h declares, and fails loudly (never silently falls back) without a GPU."""
import ctypes
import os
import re

import pytest
import torch

import tensorflow_examples_b200 as pkg
from tensorflow_examples_b200 import _ffi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = set()
    for fn in os.listdir(os.path.join(ROOT, "include")):
        if fn.endswith(".h"):
            src = open(os.path.join(ROOT, "include", fn)).read()
            src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
            names |= set(re.findall(r"\b(b2_[a-z0-9_]+)\s*\(", src))
    return names


def test_library_exports_every_declared_symbol(built_library):
    lib = ctypes.CDLL(built_library)
    declared = _declared_symbols()
    assert declared, "no declarations found in include/*.h"
    for name in sorted(declared):
        assert hasattr(lib, name), "libb2mlp.so does not export %s" % name
    assert declared >= set(_ffi.EXPORTS), set(_ffi.EXPORTS) - declared


def test_version_and_sass_target(built_library):
    lib = pkg.load_library()
    assert b"sm_100a" in lib.b2_version()


def test_bad_arguments_are_rejected_with_a_message(built_library):
    lib = pkg.load_library()
    h = ctypes.c_void_p()
    dims = (ctypes.c_int * 2)(784, 10)
    rc = lib.b2_create(ctypes.byref(h), 0, 0, dims, 0, 1, 0, 128)       # n_layers = 0
    assert rc == 1 and b"n_layers" in lib.b2_last_error()
    rc = lib.b2_create(ctypes.byref(h), 0, 1, dims, 7, 1, 0, 128)       # unknown activation
    assert rc == 1 and b"activation" in lib.b2_last_error()
    rc = lib.b2_train_steps(None, None, None, 0, 1, 1, 1, 0, 0.1, None, None, None)
    assert rc == 1 and b"ctx is NULL" in lib.b2_last_error()
    assert lib.b2_param_count(None) == 0


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_gpu_means_error_not_fallback(built_library):
    lib = pkg.load_library()
    h = ctypes.c_void_p()
    dims = (ctypes.c_int * 2)(784, 10)
    rc = lib.b2_create(ctypes.byref(h), 0, 1, dims, 0, 1, 0, 128)
    assert rc == 2, lib.b2_last_error()          # B2_ERR_CUDA
    assert not h.value
    with pytest.raises(RuntimeError):
        pkg.DenseTrainer([784, 10], pkg.ACT_IDENTITY, pkg.LOSS_SOFTMAX_CE, 128)


def test_product_code_never_imports_the_oracle():
    bad = []
    for base in ("tensorflow-examples_b200", "simple", "distributed"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            for fn in files:
                if fn.endswith((".py", ".cu", ".cuh", ".h")):
                    src = open(os.path.join(dp, fn)).read()
                    if re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M):
                        bad.append(os.path.join(dp, fn))
    assert not bad, bad
