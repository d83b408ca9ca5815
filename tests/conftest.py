import os
import subprocess
import sys
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "avrecode-ms_b200")
sys.path.insert(0, ROOT)
sys.path.insert(0, PKG)
sys.path.insert(0, os.path.join(ROOT, "tests"))
GOLDEN = os.path.join(ROOT, "tests", "golden")
REF = os.path.join(ROOT, "oracle", "_ref")
GOLDEN_NAMES = ["ip_basic", "multislice_multiref", "dct8x8", "dense_full_blocks", "skippy_idr", "b_slices", "b_direct_multiref"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session", autouse=True)
def built():
    need = [os.path.join(PKG, "libavrecode_b200.so"), os.path.join(PKG, "bin", "avrgen"),
            os.path.join(ROOT, "build", "avr_emul"), os.path.join(ROOT, "build", "avr_emul_asan"), os.path.join(ROOT, "build", "avr_emul_spec"),
            os.path.join(REF, "avr_oracle")]
    if not all(os.path.exists(p) for p in need):
        import __graft_entry__ as g
        g.build()
    return True


def have_ref():
    return os.path.exists(os.path.join(REF, "recode")) and os.path.exists(os.path.join(REF, "ref_harness"))


def run(cmd, **kw):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, **kw)
    return r.returncode, r.stdout, r.stderr


def gen_clip(path, **params):
    args = [os.path.join(PKG, "bin", "avrgen"), path] + [f"{k}={v}" for k, v in params.items()]
    rc, _, err = run(args)
    assert rc == 0, err
    return open(path, "rb").read()


def ref_trace(clip_path, out_path):
    with open(out_path, "wb") as f:
        subprocess.run([os.path.join(REF, "ref_harness"), "trace", clip_path], stdout=f, stderr=subprocess.DEVNULL, check=True)


def oracle_encode(trace_path):
    rc, out, err = run([os.path.join(REF, "avr_oracle"), "encode", trace_path])
    assert rc == 0, err
    return [bytes.fromhex(l[2:].decode()) for l in out.splitlines() if l.startswith(b"P ")]


def hex_payloads(out):
    return [bytes.fromhex(l[2:].decode()) for l in out.splitlines() if l.startswith(b"P ")]


def golden_trace(name, td):
    import gzip
    p = os.path.join(td, name + ".trace")
    with open(p, "wb") as f:
        f.write(gzip.open(os.path.join(GOLDEN, name + ".trace.gz")).read())
    return p


@pytest.fixture
def tmpdir_path():
    with tempfile.TemporaryDirectory() as td:
        yield td


def damage_slice(data, which, seed=1, keep=12):
    """overwrite the payload (all but the first `keep` bytes of the NAL) of the which-th slice NAL with bytes 0x10..0xFF: start codes,
    NAL boundaries and the slice header stay, the CABAC data becomes noise"""
    import random
    import re
    d = bytearray(data)
    r = random.Random(seed)
    idx = [m.start() for m in re.finditer(b"\x00\x00\x01", data)]
    nals = []
    for k, i in enumerate(idx):
        s, e = i + 3, idx[k + 1] if k + 1 < len(idx) else len(d)
        while e > s and d[e - 1] == 0:
            e -= 1
        nals.append((s, e))
    s, e = [(s, e) for s, e in nals if d[s] & 31 in (1, 5)][which]
    for i in range(s + keep, e - 1):
        d[i] = r.randrange(0x10, 0x100)
    return bytes(d)


def as_cavlc(data):
    """the same stream with entropy_coding_mode_flag = 0 in every PPS (pps_id = sps_id = 0: third bit of the PPS payload): its slices
    are no longer CABAC slices for anyone who parses the headers"""
    import re
    d = bytearray(data)
    for m in re.finditer(b"\x00\x00\x01", data):
        h = m.start() + 3
        if d[h] & 31 == 8:
            assert d[h + 1] & 0xE0 == 0xE0, hex(d[h + 1])
            d[h + 1] ^= 0x20
    return bytes(d)


def drop_slice_nal(data, which_slice):
    """the stream without its which_slice-th slice NAL (start code included)"""
    import re
    idx = [m.start() for m in re.finditer(b"\x00\x00\x01", data)]
    k = [i for i in idx if data[i + 3] & 31 in (1, 5)][which_slice]
    nxt = [i for i in idx if i > k]
    end = nxt[0] if nxt else len(data)
    while end > k and data[end - 1] == 0 and end - 1 > k + 3:       # the zero_byte of a 4-byte start code belongs to the next NAL
        end -= 1
    return data[:k] + data[end:]
