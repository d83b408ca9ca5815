"""The C-ABI library loads and exports every symbol include/avrecode_b200.h declares; without a
GPU the compute entry points fail loudly (no CPU fallback)."""
import ctypes
import os
import re

import pytest

from conftest import GOLDEN, PKG, ROOT


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "avrecode_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(avr_[a-z0-9_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(os.path.join(PKG, "libavrecode_b200.so"))
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), n


def test_ctypes_stub_lists_the_same_symbols():
    import avr_ctypes
    assert sorted(avr_ctypes.EXPORTS) == declared_symbols()


def test_host_recorder_runs_without_gpu():
    import avr_ctypes as A
    rec = A.Record(open(os.path.join(GOLDEN, "multislice_multiref.h264"), "rb").read())
    assert rec.rc == 0 and rec.slices == 12 and rec.bins > 10000 and rec.out_of_domain == 0
    spans = rec.spans()
    assert spans[0]["flip"] == 2 and all(s["size"] >= 8 for s in spans)
    rec.close()


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import avr_ctypes as A
    with pytest.raises(A.AvrError):
        A.Context(0)
    rc = os.system(f"{PKG}/bin/recode roundtrip {GOLDEN}/ip_basic.h264 2>/dev/null")
    assert rc != 0
