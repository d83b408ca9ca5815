"""BASELINE configs[0]: independent bin streams through the generic range coder (reference test/arithmetic_code.cpp:93-110).
Golden = bytes produced by the reference's own arithmetic_code.h (tests/golden/make_streams_golden.py)."""
import json
import os
import subprocess

import numpy as np
import pytest
from conftest import GOLDEN, REF, ROOT, have_ref

CASES = json.load(open(os.path.join(GOLDEN, "streams.json")))["cases"]
IDS = [f"d{c['digit_bits']}m{c['model']}_{c['nstreams']}x{c['nbins']}" for c in CASES]


def case_bits(c):
    n = c["nstreams"] * c["nbins"]
    return np.unpackbits(np.frombuffer(bytes.fromhex(c["bits"]), np.uint8))[:n].astype(np.uint8).tobytes()


def tool_streams(exe, c, td):
    p = os.path.join(td, "bits.bin")
    open(p, "wb").write(case_bits(c))
    r = subprocess.run([exe, "streams", str(c["digit_bits"]), str(c["model"]), str(c["nstreams"]), str(c["nbins"]), p],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr
    return [l[2:].strip().decode() for l in r.stdout.splitlines() if l.startswith(b"A")]


@pytest.mark.parametrize("c", CASES, ids=IDS)
def test_oracle_streams_match_reference_golden(c, tmpdir_path):
    assert tool_streams(os.path.join(REF, "avr_oracle"), c, tmpdir_path) == c["coded"]


@pytest.mark.parametrize("c", CASES, ids=IDS)
def test_stream_bodies_match_reference_golden(c, tmpdir_path):
    """the kernels' bodies compiled for the host (emulated grid): encode == golden, decode restores the bits"""
    assert tool_streams(os.path.join(ROOT, "build", "avr_emul"), c, tmpdir_path) == c["coded"]


@pytest.mark.parametrize("c", CASES[::3], ids=IDS[::3])
def test_stream_bodies_under_asan(c, tmpdir_path):
    """same bodies in the ASan/UBSan build of the scaffold: no access outside the per-stream scratch, inputs or outputs"""
    assert tool_streams(os.path.join(ROOT, "build", "avr_emul_asan"), c, tmpdir_path) == c["coded"]


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built")
def test_oracle_streams_match_live_reference(tmpdir_path):
    rng = np.random.RandomState(77)
    for digit_bits, model in ((16, 0), (8, 1)):
        bits = (rng.random_sample((6, 5003)) < 0.03).astype(np.uint8)
        p = os.path.join(tmpdir_path, "b.bin")
        open(p, "wb").write(bits.tobytes())
        args = [str(digit_bits), str(model), "6", "5003", p]
        a = subprocess.run([os.path.join(REF, "ref_harness"), "arith"] + args, stdout=subprocess.PIPE, check=True).stdout
        b = subprocess.run([os.path.join(REF, "avr_oracle"), "streams"] + args, stdout=subprocess.PIPE, check=True).stdout
        assert a == b and a.count(b"A ") == 6


# ---- GPU ----------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def ctx():
    import avr_ctypes as A
    with A.Context(0) as c:
        yield c


@pytest.mark.gpu
@pytest.mark.parametrize("c", CASES, ids=IDS)
def test_gpu_streams_match_reference_golden(ctx, c):
    import avr_ctypes as A
    bits = case_bits(c)
    coded, t = A.streams_encode(ctx, bits, c["nstreams"], c["nbins"], c["digit_bits"], c["model"])
    assert [x.hex() for x in coded] == c["coded"]
    back, _ = A.streams_decode(ctx, coded, c["nstreams"], c["nbins"], c["digit_bits"], c["model"])
    assert back == bits


@pytest.mark.gpu
@pytest.mark.parametrize("digit_bits,model", [(16, 0), (8, 0), (16, 1), (8, 1)])
def test_gpu_streams_large_roundtrip_and_oracle_sample(ctx, digit_bits, model, tmpdir_path):
    """4096 streams x 40 000 bins (p(1) cycling through 0.5 / 0.1 / 0.01 / 0.001): decode(encode(x)) == x for all of them,
    FIXED_HALF streams cost one bit per bin, and a sample of streams is byte-identical to the oracle."""
    import avr_ctypes as A
    S, N = 4096, 40000
    rng = np.random.RandomState(1234 + digit_bits + model)
    p = np.array([0.5, 0.1, 0.01, 0.001])[np.arange(S) % 4][:, None]
    bits = (rng.random_sample((S, N)) < p).astype(np.uint8)
    coded, t = A.streams_encode(ctx, bits.tobytes(), S, N, digit_bits, model)
    assert t["bins"] == S * N and t["kernel_ms"] > 0
    if model == 0:
        assert all(N // 8 <= len(x) <= N // 8 + 10 for x in coded)
    else:
        assert len(coded[3]) < len(coded[0]) // 20      # p = 0.001 compresses far below p = 0.5
    back, _ = A.streams_decode(ctx, coded, S, N, digit_bits, model)
    assert back == bits.tobytes()
    sample = [0, 1, 2, 3, S - 1]
    pth = os.path.join(tmpdir_path, "s.bin")
    open(pth, "wb").write(bits[sample].tobytes())
    out = subprocess.run([os.path.join(REF, "avr_oracle"), "streams", str(digit_bits), str(model), str(len(sample)), str(N), pth],
                         stdout=subprocess.PIPE, check=True).stdout
    want = [l[2:].strip().decode() for l in out.splitlines() if l.startswith(b"A")]
    assert [coded[s].hex() for s in sample] == want


@pytest.mark.gpu
def test_gpu_streams_ragged_batch(ctx, tmpdir_path):
    """a stream count that is no multiple of the block (128) or of the compaction scan's step (4096) and a bin count that is
    no multiple of 8: every stream round-trips and first / middle / last streams equal the oracle"""
    import avr_ctypes as A
    S, N = 5003, 257
    rng = np.random.RandomState(9)
    bits = (rng.random_sample((S, N)) < 0.2).astype(np.uint8)
    for digit_bits, model in ((16, 0), (8, 1)):
        coded, _ = A.streams_encode(ctx, bits.tobytes(), S, N, digit_bits, model)
        back, _ = A.streams_decode(ctx, coded, S, N, digit_bits, model)
        assert back == bits.tobytes()
        sample = [0, 127, 128, 4095, 4096, 4097, S - 1]
        pth = os.path.join(tmpdir_path, "s.bin")
        open(pth, "wb").write(bits[sample].tobytes())
        out = subprocess.run([os.path.join(REF, "avr_oracle"), "streams", str(digit_bits), str(model), str(len(sample)), str(N), pth],
                             stdout=subprocess.PIPE, check=True).stdout
        assert [coded[s].hex() for s in sample] == [l[2:].strip().decode() for l in out.splitlines() if l.startswith(b"A")]


@pytest.mark.gpu
def test_gpu_streams_bad_arguments(ctx):
    import avr_ctypes as A
    with pytest.raises(A.AvrError):
        A.streams_encode(ctx, b"\x00" * 8, 1, 8, digit_bits=12)
    with pytest.raises(A.AvrError):
        A.streams_encode(ctx, b"\x00" * 8, 1, 8, model=5)
    coded, _ = A.streams_encode(ctx, b"", 0, 100)
    assert coded == []
    # truncated input decodes (reads past the end are zeros, arithmetic_code.h:281-285) without leaving its buffers
    bits = bytes([1, 0] * 500)
    coded, _ = A.streams_encode(ctx, bits, 1, 1000, 8, 1)
    back, _ = A.streams_decode(ctx, [coded[0][: len(coded[0]) // 2]], 1, 1000, 8, 1)
    assert len(back) == 1000
