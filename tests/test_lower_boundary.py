"""The lower boundary driven from OUTSIDE the library (VERDICT r1 row b): the AVCodecHooks-shaped sinks
avr_rec_frame_spec / slice_begin / mb_xy / begin_sub_mb / begin_coding_type / bin fed with the events the REFERENCE's own
av_decoder<> delivers (coding_hooks.h:27-55, recode.cpp:145-235):
  * replayed through ctypes from the hook traces `oracle/_ref/ref_harness trace` printed for the golden clips,
  * and live, by oracle/gpu_driver.cpp: reference libavformat + libavcodec-hooks -> avr_rec_* -> GPU -> reference container code."""
import gzip
import os

import pytest

from avrec_wire import parse_avrec
from conftest import GOLDEN, GOLDEN_NAMES, REF, run

DRIVER = os.path.join(REF, "gpu_driver")


def golden_events(name):
    return gzip.open(os.path.join(GOLDEN, name + ".trace.gz"), "rt").read().splitlines()


def golden_payloads(name):
    return [b["cabac"] for b in parse_avrec(open(os.path.join(GOLDEN, name + ".avrec"), "rb").read()) if "cabac" in b]


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_hook_sinks_fill_the_same_record_as_the_native_front_end(name):
    import avr_ctypes as A
    hooked = A.HookRecord(golden_events(name))
    native = A.Record(open(os.path.join(GOLDEN, name + ".h264"), "rb").read())
    assert native.rc == 0 and hooked.out_of_domain == 0
    assert (hooked.bins, hooked.blocks, hooked.slices, hooked.nspans, hooked.pictures) == \
           (native.bins, native.blocks, native.slices, native.nspans, native.pictures)
    hs, ns = hooked.spans(), native.spans()
    assert [(s["payload_off"], s["size"], s["slice_index"], s["flip"]) for s in hs] == \
           [(s["payload_off"], s["size"], s["slice_index"], s["flip"]) for s in ns]
    hooked.close(); native.close()


@pytest.mark.skipif(not os.path.exists(DRIVER), reason="oracle/_ref/gpu_driver not built (needs /root/reference at build time)")
@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_compiled_driver_records_what_the_native_front_end_records(name):
    """reference libavformat + libavcodec-hooks -> avr_rec_* (no GPU needed for this half)"""
    import avr_ctypes as A
    rc, out, err = run([DRIVER, "records", os.path.join(GOLDEN, name + ".h264")])
    assert rc == 0, err
    got = dict(zip(out.split()[0::2], [int(x) for x in out.split()[1::2]]))
    native = A.Record(open(os.path.join(GOLDEN, name + ".h264"), "rb").read())
    assert (got[b"bins"], got[b"blocks"], got[b"slices"], got[b"spans"], got[b"pictures"], got[b"out_of_domain"]) == \
           (native.bins, native.blocks, native.slices, native.nspans, native.pictures, 0)
    native.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_replayed_hook_events_compress_to_the_reference_payloads(name):
    import avr_ctypes as A
    with A.Context(0) as ctx:
        rec = A.HookRecord(golden_events(name))
        for mode in (A.MODE_PIPELINE, A.MODE_CHAIN):
            b = A.CBatch(ctx, [rec])
            b.run(mode)
            b.fetch()
            assert b.payloads()[0] == golden_payloads(name), mode
            b.close()
        rec.close()


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(DRIVER), reason="oracle/_ref/gpu_driver not built (needs /root/reference at build time)")
@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_compiled_driver_writes_the_reference_container(name, tmpdir_path):
    """reference demux + hooks -> avr_rec_* -> avr_cbatch on the GPU -> reference Recoded container == the reference compressor's bytes"""
    out = os.path.join(tmpdir_path, "d.avrec")
    rc, _, err = run([DRIVER, "compress", os.path.join(GOLDEN, name + ".h264"), out], timeout=300)
    assert rc == 0, err
    assert open(out, "rb").read() == open(os.path.join(GOLDEN, name + ".avrec"), "rb").read()
