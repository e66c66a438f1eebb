"""CPU tests of the boundary: the library builds, loads and exports every symbol that
include/b200lsap.h declares; no compute call is made (there is no GPU here)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def ab():
    from auctionalgorithm_b200 import build
    build.build()
    import auctionalgorithm_b200 as ab
    return ab


def test_header_symbols_are_exported(ab):
    header = open(os.path.join(ROOT, "include", "b200lsap.h")).read()
    declared = set(re.findall(r"\b(b200lsap_[a-z0-9_]+)\s*\(", header))
    declared -= {"b200lsap_shard_"}
    assert declared, "no declarations parsed"
    lib = ab.lib()
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in b200lsap.h but not exported"
    assert declared == set(ab.SIGNATURES), declared ^ set(ab.SIGNATURES)


def test_struct_sizes(ab):
    import ctypes as C
    from auctionalgorithm_b200._capi import Options, Stats
    assert C.sizeof(Options) == 72 and C.sizeof(Stats) == 9 * 8 + 48 + 24
    o = Options()
    ab.lib().b200lsap_options_default(C.byref(o))
    assert o.eps_multiplier == 1e-5 and o.eps_scaling == -1 and o.driver == 0


def test_status_strings(ab):
    lib = ab.lib()
    assert lib.b200lsap_status_string(0) == b"ok"
    assert b"rows" in lib.b200lsap_status_string(ab.ERR_SHAPE)


def test_fails_loudly_without_device(ab):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(ab.B200LsapError) as e:
        ab.Handle()
    assert e.value.status == ab.ERR_NO_DEVICE


def test_missing_library_is_an_import_error(ab, monkeypatch):
    from auctionalgorithm_b200 import _capi
    monkeypatch.setattr(_capi, "_lib", None)
    monkeypatch.setattr(_capi, "LIB_PATH", "/nonexistent/libb200lsap.so")
    with pytest.raises(ImportError):
        _capi.lib()


def test_product_does_not_touch_the_oracle():
    pkg = os.path.join(ROOT, "auctionalgorithm_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("oracle_generate_f32", ""), f"{f} mentions the oracle"
