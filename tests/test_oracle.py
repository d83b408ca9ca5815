"""Pins the oracle (oracle/avr_oracle.cpp) against the reference: golden .avrec files minted by
the unmodified reference compressor (tests/golden/make_golden.py), and -- when oracle/_ref/recode
is present -- fresh reference runs on new clips."""
import os

import pytest

from avrec_wire import parse_avrec
from conftest import GOLDEN, GOLDEN_NAMES, REF, gen_clip, golden_trace, have_ref, oracle_encode, ref_trace, run


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_oracle_matches_golden_avrec(name, tmpdir_path):
    blocks = parse_avrec(open(os.path.join(GOLDEN, name + ".avrec"), "rb").read())
    want = [b["cabac"] for b in blocks if "cabac" in b]
    got = oracle_encode(golden_trace(name, tmpdir_path))
    assert len(want) > 0 and got == want


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_oracle_decode_inverts_and_reencodes_cabac(name, tmpdir_path):
    """decode(encode(trace)) reproduces every bin, and the CABAC re-encode reproduces the original
    slice bytes up to the reference's length_parity/last_byte patch (recode.cpp:1352-1362)."""
    tr = golden_trace(name, tmpdir_path)
    pay = os.path.join(tmpdir_path, "p")
    rc, out, _ = run([os.path.join(REF, "avr_oracle"), "encode", tr])
    open(pay, "wb").write(out)
    rc, out, _ = run([os.path.join(REF, "avr_oracle"), "decode", tr, pay])
    assert rc == 0 and out.splitlines()[-1].startswith(b"OK")
    cab = [bytes.fromhex(l[2:].decode()) for l in out.splitlines() if l.startswith(b"C ")]
    data = open(os.path.join(GOLDEN, name + ".h264"), "rb").read()
    spans = [l.split() for l in open(tr, "rb").read().splitlines() if l.startswith(b"S ")]
    coded = [(int(s[1]), int(s[2])) for s in spans if int(s[2]) >= 0]
    assert len(coded) == len(cab)
    for (size, off), c in zip(coded, cab):
        orig = data[off:off + size]
        c = bytearray(c)
        if (size & 1) != (len(c) & 1):
            c.append(orig[-1])
        else:
            c[-1] = orig[-1]
        assert bytes(c) == orig


def test_reference_decompressor_quirk_mode_on_indomain_clip(tmpdir_path):
    """Statement-faithful mode of the oracle decoder also inverts a clip inside the reference's own
    valid domain (no 8x8 transform)."""
    tr = golden_trace("ip_basic", tmpdir_path)
    pay = os.path.join(tmpdir_path, "p")
    open(pay, "wb").write(run([os.path.join(REF, "avr_oracle"), "encode", tr])[1])
    rc, out, _ = run([os.path.join(REF, "avr_oracle"), "decode", tr, pay, "--ref-quirks"])
    assert rc == 0


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (no /root/reference here)")
@pytest.mark.parametrize("params", [
    dict(seed=21, mb_w=12, mb_h=8, frames=4, slices_per_frame=2),
    dict(seed=22, mb_w=9, mb_h=7, frames=3, transform_8x8=1, num_ref=2),
    dict(seed=23, mb_w=8, mb_h=8, frames=3, p_sig=240, p_last=4, p_cbf=250, p_cbp=250),
])
def test_oracle_matches_live_reference(params, tmpdir_path):
    clip = os.path.join(tmpdir_path, "c.h264")
    gen_clip(clip, **params)
    avrec = os.path.join(tmpdir_path, "c.avrec")
    rc, _, err = run([os.path.join(REF, "recode"), "compress", clip, avrec])
    assert rc == 0, err
    want = [b["cabac"] for b in parse_avrec(open(avrec, "rb").read()) if "cabac" in b]
    tr = os.path.join(tmpdir_path, "c.trace")
    ref_trace(clip, tr)
    assert oracle_encode(tr) == want


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built")
def test_reference_arith_test_plumbing():
    """BASELINE config 1: the reference's own test/arithmetic_code round trip (p = 1/2)."""
    rc, out, _ = run([os.path.join(REF, "arith_test"), "100000"])
    assert rc == 0 and b"compressed size: 12501" in out
