"""Device logic on the CPU: tests/host_emul.cpp compiles the kernel bodies for the host and runs
them as an emulated grid (test scaffold, not a product path).  Checks them against the golden
reference containers and the oracle."""
import os

import pytest

from avrec_wire import parse_avrec
from conftest import GOLDEN, GOLDEN_NAMES, ROOT, as_cavlc, damage_slice, drop_slice_nal, gen_clip, have_ref, hex_payloads, oracle_encode, ref_trace, run

EMUL = os.path.join(ROOT, "build", "avr_emul")
EMUL_ASAN = os.path.join(ROOT, "build", "avr_emul_asan")
EMUL_SPEC = os.path.join(ROOT, "build", "avr_emul_spec")


def golden_payloads(name):
    return [b["cabac"] for b in parse_avrec(open(os.path.join(GOLDEN, name + ".avrec"), "rb").read()) if "cabac" in b]


@pytest.mark.parametrize("mode", ["chain", "pipeline"])
@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_compress_bodies_match_reference_golden(name, mode):
    rc, out, err = run([EMUL, mode, os.path.join(GOLDEN, name + ".h264")])
    assert rc == 0, err
    assert hex_payloads(out) == golden_payloads(name)


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_decompress_body_restores_original_cabac_bytes(name):
    rc, out, err = run([EMUL, "roundtrip", os.path.join(GOLDEN, name + ".h264")])
    assert rc == 0 and out.startswith(b"OK"), out + err


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_device_parse_body_emits_the_host_recorders_records(name):
    """avr_parse.h (compress-side CABAC parse, one slice per warp) == avr_record.cpp on bins, blocks, counts"""
    rc, out, err = run([EMUL, "parse", os.path.join(GOLDEN, name + ".h264")])
    assert rc == 0 and out.startswith(b"OK"), out + err


def test_pipeline_batch_of_files_keeps_models_separate():
    """the estimator state is per file (recode.cpp:1314): other files in the batch must not leak in"""
    clips = [os.path.join(GOLDEN, n + ".h264") for n in ("dct8x8", "skippy_idr", "ip_basic")]
    rc, out, err = run([EMUL, "pipeline"] + clips)
    assert rc == 0, err
    assert hex_payloads(out) == golden_payloads("dct8x8")


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("params", [
    dict(seed=31, mb_w=20, mb_h=12, frames=5, slices_per_frame=4, p_skip=200),     # lone 0x80 stop-bit bytes
    dict(seed=32, mb_w=16, mb_h=9, frames=12, transform_8x8=1, num_ref=4),          # > 95 symbols on many keys
    dict(seed=33, mb_w=11, mb_h=7, frames=6, idr_period=2, qp=12, cabac_init_idc=1),
    dict(seed=34, mb_w=7, mb_h=5, frames=3, p_sig=252, p_last=1, p_cbf=254, p_cbp=254, transform_8x8=1),  # full blocks
    dict(seed=35, mb_w=13, mb_h=8, frames=7, p_bslice=200, num_ref=4, slices_per_frame=3, transform_8x8=1),  # B slices, two lists, 8x8
    dict(seed=36, mb_w=9, mb_h=9, frames=6, p_bslice=230, p_direct=160, p_skip=120, num_ref=2, cabac_init_idc=2),  # direct-heavy B
])
def test_bodies_match_oracle_on_fresh_clips(params, tmpdir_path):
    clip = os.path.join(tmpdir_path, "c.h264")
    gen_clip(clip, **params)
    tr = os.path.join(tmpdir_path, "c.trace")
    ref_trace(clip, tr)
    want = oracle_encode(tr)
    for mode in ("chain", "pipeline"):
        rc, out, err = run([EMUL, mode, clip])
        assert rc == 0, err
        assert hex_payloads(out) == want, mode
    for cmd in ("roundtrip", "parse", "files"):
        rc, out, err = run([EMUL, cmd, clip])
        assert rc == 0 and out.startswith(b"OK"), out + err
    # the container written around the compress body == the live reference compressor's
    if have_ref():
        ours, theirs = os.path.join(tmpdir_path, "ours.avrec"), os.path.join(tmpdir_path, "ref.avrec")
        assert run([EMUL, "files", clip, ours])[0] == 0
        rc, _, err = run([os.path.join(ROOT, "oracle", "_ref", "recode"), "compress", clip, theirs])
        assert rc == 0, err
        assert open(ours, "rb").read() == open(theirs, "rb").read()


@pytest.mark.parametrize("cmd", ["fuzzdec", "fuzzparse"])
@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_corrupted_input_ends_with_a_status_inside_its_buffers(name, cmd):
    """bit flips in the range-coded payloads (decompress body) or in the CABAC slice bytes (parse body):
    every slice must end -- with a status or a (wrong) result -- without reading or writing outside its
    buffers.  Runs the device bodies under ASan + UBSan; any report fails the test."""
    for seed, flips in ((1, 1), (2, 8), (3, 64), (4, 8)):
        rc, out, err = run([EMUL_ASAN, cmd, os.path.join(GOLDEN, name + ".h264"), str(seed), str(flips)], timeout=120)
        assert rc == 0 and b"TERMINATED" in out and b"runtime error" not in err and b"ERROR" not in err, out + err


def test_header_fuzz_and_advisor_repro_stay_inside_buffers(tmpdir_path):
    """`int id = br.ue()` let a 2^32-2 Exp-Golomb id through as a negative index (heap write before the Sps
    table).  The 23-byte repro and bit flips over whole files (headers included) run through scan_annexb + record_annexb
    under ASan/UBSan."""
    repro = bytes.fromhex("0000016742001e0000030001" + "ff" * 10 + "80")
    p = os.path.join(tmpdir_path, "repro.h264")
    open(p, "wb").write(repro)
    rc, out, err = run([EMUL_ASAN, "scanraw", p], timeout=60)
    assert rc == 0 and b"TERMINATED" in out and b"ERROR" not in err and b"runtime error" not in err, out + err
    # oversized picture: mb_w = 70000 must not size anything (unsupported -> literal), ue() ids at the type limit
    for name in GOLDEN_NAMES:
        for seed in range(1, 9):
            rc, out, err = run([EMUL_ASAN, "scanraw", os.path.join(GOLDEN, name + ".h264"), str(seed), "24"], timeout=120)
            assert rc == 0 and b"TERMINATED" in out and b"ERROR" not in err and b"runtime error" not in err, out + err


def test_repeated_frame_num_extended_and_strict_domain(tmpdir_path):
    """avr_record.h frame_spec: by default a picture that repeats frame_num gets its own store (flag 2, the bodies round-trip);
    AVR_STRICT_DOMAIN=1 keeps to the reference's domain (flag 1: the file API stores such a file as a literal block)"""
    clip = os.path.join(tmpdir_path, "rf.h264")
    gen_clip(clip, seed=83, mb_w=9, mb_h=7, frames=8, repeat_frame_num=1, p_bslice=150, num_ref=2)
    tool = os.path.join(ROOT, "avrecode-ms_b200", "bin", "avrtool")
    rc, _, err = run([tool, "stats", clip])
    assert rc == 0 and b"ood=2" in err, err
    rc, _, err = run([tool, "stats", clip], env=dict(os.environ, AVR_STRICT_DOMAIN="1"))
    assert b"ood=1" in err, err
    for cmd in ("roundtrip", "parse", "files"):      # files: the decompress-side scan applies the same every-picture-flips rule
        rc, out, err = run([EMUL, cmd, clip])
        assert rc == 0 and out.startswith(b"OK"), out + err


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_speculative_chain_body_restores_original_cabac_bytes(name):
    """the experimental warp-speculative decompress chain (avr_decode_spec.h, AVR_DEC_SPEC=1 on the device), 32 lanes emulated as
    a loop: same roundtrip property as the product chain"""
    if not os.path.exists(EMUL_SPEC):
        pytest.skip("build/avr_emul_spec not built")
    rc, out, err = run([EMUL_SPEC, "roundtrip", os.path.join(GOLDEN, name + ".h264")])
    assert rc == 0 and out.startswith(b"OK"), out + err


@pytest.mark.parametrize("which,seed", [(1, 1), (3, 2), (5, 3), (11, 4)])
def test_damaged_slice_is_left_literal_and_the_rest_is_coded(which, seed, tmp_path):
    """per-slice fallback (avr_record.h slice_abort): a slice whose walk does not end on rbsp_stop_one_bit becomes a skip_coded span
    (reference container form recode.cpp:1296-1303), the other slices are recorded as before and the file round-trips through
    the compress / decompress bodies"""
    src = open(os.path.join(GOLDEN, "multislice_multiref.h264"), "rb").read()
    bad = tmp_path / "damaged.h264"
    bad.write_bytes(damage_slice(src, which, seed))
    tool = os.path.join(ROOT, "avrecode-ms_b200", "bin", "avrtool")
    rc, _, err = run([tool, "stats", str(bad)])
    assert rc == 0 and b"slices=11 spans=12" in err and b"skipped=1" in err, err
    rc, _, err = run([tool, "stats", os.path.join(GOLDEN, "multislice_multiref.h264")])
    assert rc == 0 and b"slices=12 spans=12" in err and b"skipped=0" in err, err
    rc, out, err = run([EMUL, "roundtrip", str(bad)])
    assert rc == 0 and out.startswith(b"OK slices=11"), out + err
    # the device parse body flags the slice (the file then goes through the host recorder, avr_files.cpp)
    rc, out, err = run([EMUL, "parse", str(bad)])
    assert rc != 0 and b"slice status -5" in out + err


def test_slices_outside_the_supported_syntax_are_passed_over_one_by_one(tmp_path):
    """a CABAC stream followed by a CAVLC one (new PPS): the first part is recorded, the slices of the second take no span at all
    (scan_annexb_slices), instead of the whole file being refused"""
    a = open(os.path.join(GOLDEN, "ip_basic.h264"), "rb").read()
    b = as_cavlc(open(os.path.join(GOLDEN, "skippy_idr.h264"), "rb").read())
    mixed = tmp_path / "mixed.h264"
    mixed.write_bytes(a + b)
    tool = os.path.join(ROOT, "avrecode-ms_b200", "bin", "avrtool")
    rc, _, err = run([tool, "stats", str(mixed)])
    assert rc == 0 and b"slices=3 spans=3" in err and b"skipped=10" in err, err
    only = tmp_path / "cavlc.h264"
    only.write_bytes(b)
    rc, _, err = run([tool, "stats", str(only)])
    assert rc == 0 and b"slices=0 spans=0" in err and b"skipped=10" in err, err


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_file_level_host_logic_writes_the_reference_container_and_inverts_it(name, tmp_path):
    """avr_container.h (what avr_compress_files / avr_decompress_files do around the device batches) on the CPU: the container
    written from the recorder's spans and the compress body's payloads is the reference compressor's .avrec byte for byte, and
    from the container ALONE (reader, surrogate scan, slice headers re-read from the literal bytes, picture flips, block matching,
    decompress body, last-byte patch) the input file comes back"""
    out = tmp_path / (name + ".avrec")
    rc, o, err = run([EMUL, "files", os.path.join(GOLDEN, name + ".h264"), str(out)])
    assert rc == 0 and o.startswith(b"OK"), o + err
    assert out.read_bytes() == open(os.path.join(GOLDEN, name + ".avrec"), "rb").read()


def test_file_level_host_logic_on_fallback_and_degenerate_inputs(tmp_path):
    """skip_coded blocks (damaged slices, also ones that open a picture: the store flip then belongs to the next coded slice), slices
    passed over (CAVLC part between two CABAC parts), files without any coded slice, junk, the empty file"""
    g = {n: open(os.path.join(GOLDEN, n + ".h264"), "rb").read() for n in GOLDEN_NAMES}
    m = g["multislice_multiref"]
    cases = {
        "first": (damage_slice(m, 0, 4), b"coded=11 skip_coded=1"),
        "opens_picture": (damage_slice(m, 3, 2), b"coded=11 skip_coded=1"),
        "two_in_a_row": (damage_slice(damage_slice(m, 6, 5), 7, 6), b"coded=10 skip_coded=2"),
        "mixed": (g["ip_basic"] + as_cavlc(g["skippy_idr"]) + g["b_slices"], b"coded=13 skip_coded=0"),
        "cavlc_only": (as_cavlc(g["dct8x8"]), b"blocks=1 coded=0"),
        "junk": (bytes(range(256)) * 3, b"blocks=1 coded=0"),
        "empty": (b"", b"coded=0"),
    }
    for name, (data, want) in cases.items():
        src = tmp_path / (name + ".h264")
        src.write_bytes(data)
        rc, o, err = run([EMUL_ASAN, "files", str(src)])
        assert rc == 0 and o.startswith(b"OK") and want in o, (name, o, err)


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built")
def test_random_clips_file_level_against_the_live_reference(tmpdir_path):
    """a short run of tools/cpu_sweep.py (100 clips by hand: 0 differences): random geometry, slice counts, I/P/B mix, 8x8, skip
    and coefficient densities; the container equals the reference compressor's and the file comes back from it"""
    import random
    rnd = random.Random(20261105)
    ref = os.path.join(ROOT, "oracle", "_ref", "recode")
    for i in range(10):
        p = dict(seed=rnd.randrange(1, 10 ** 6), mb_w=rnd.randrange(2, 20), mb_h=rnd.randrange(2, 12), frames=rnd.randrange(1, 6),
                 slices_per_frame=rnd.randrange(1, 5), transform_8x8=rnd.randrange(2), num_ref=rnd.randrange(1, 5),
                 p_skip=rnd.choice([0, 60, 200, 250]), qp=rnd.randrange(10, 45), cabac_init_idc=rnd.randrange(3))
        if rnd.random() < 0.5:
            p.update(p_bslice=rnd.choice([100, 200, 240]), p_direct=rnd.choice([0, 80, 200]))
        clip = os.path.join(tmpdir_path, f"r{i}.h264")
        gen_clip(clip, **p)
        ours, theirs = clip + ".ours", clip + ".ref"
        rc, out, err = run([EMUL, "files", clip, ours])
        assert rc == 0 and out.startswith(b"OK"), (p, out, err)
        rc, _, err = run([ref, "compress", clip, theirs])
        assert rc == 0, (p, err)
        assert open(ours, "rb").read() == open(theirs, "rb").read(), p


def test_stream_that_begins_in_the_middle_of_a_picture(tmp_path):
    """no slice with first_mb = 0 before the first coded slice (a cut stream, or a first slice whose header is refused): the first
    slice seen starts the picture on both directions (avr_record.h starts_picture, avr_container.h plan_decompress) instead of
    leaving coded slices without a picture store"""
    tool = os.path.join(ROOT, "avrecode-ms_b200", "bin", "avrtool")
    src = open(os.path.join(GOLDEN, "multislice_multiref.h264"), "rb").read()
    # the IDR's first slice missing: its second slice starts the picture; the first slice of the second picture missing: that
    # picture's second slice starts it (its first_mb does not lie beyond the previous slice's)
    for drop, frames in ((0, b"frames=4"), (3, b"frames=4")):
        clip = tmp_path / f"cut{drop}.h264"
        clip.write_bytes(drop_slice_nal(src, drop))
        rc, _, err = run([tool, "stats", str(clip)])
        assert rc == 0 and b"slices=11" in err and frames in err, err
        for cmd in ("roundtrip", "parse", "files"):
            rc, out, err = run([EMUL_ASAN, cmd, str(clip)])
            assert rc == 0 and out.startswith(b"OK"), (drop, cmd, out, err)
        # every macroblock position is coded once per picture store: the slice-parallel passes agree with the sequential chain
        assert run([EMUL, "chain", str(clip)])[1] == run([EMUL, "pipeline", str(clip)])[1]


def test_slice_qp_out_of_range_is_refused_like_libavcodec_does(tmpdir_path):
    """SliceQPY outside 0..51: libavcodec's slice header parse fails (no hook sees the slice, the reference writes no block for it);
    the generator's per-slice QP variation around qp=0 / qp=51 produces such headers"""
    tool = os.path.join(ROOT, "avrecode-ms_b200", "bin", "avrtool")
    refused = 0
    for qp in (0, 51):
        clip = os.path.join(tmpdir_path, f"q{qp}.h264")
        gen_clip(clip, seed=5, mb_w=8, mb_h=6, frames=3, slices_per_frame=2, num_ref=2, qp=qp)
        rc, _, err = run([tool, "stats", clip])
        assert rc == 0, err
        refused += b"slices=6" not in err
        for cmd in ("roundtrip", "parse", "files"):
            rc, out, err = run([EMUL, cmd, clip])
            assert rc == 0 and out.startswith(b"OK"), (qp, cmd, out, err)
    assert refused, "the generator no longer produces out-of-range slice QPs: pick other parameters"


def test_slice_that_ends_without_end_of_slice_flag_is_left_literal(tmp_path):
    """found by tools/cpu_bitflip_sweep.py: bit 2 of byte 2046 of ip_basic.h264 damages the last macroblock so that the walk ends on
    the right bit by running out of macroblocks, end_of_slice_flag read as 0 -- the decompress chain could never close that slice.
    Both parsers now require the flag (avr_record.cpp, avr_parse.h): the slice becomes a skip_coded block and the file round-trips"""
    d = bytearray(open(os.path.join(GOLDEN, "ip_basic.h264"), "rb").read())
    d[2046] ^= 1 << 2
    clip = tmp_path / "flip.h264"
    clip.write_bytes(bytes(d))
    rc, out, err = run([EMUL_ASAN, "files", str(clip)])
    assert rc == 0 and out.startswith(b"OK") and b"coded=2 skip_coded=1" in out, out + err
    rc, out, err = run([EMUL, "parse", str(clip)])
    assert rc != 0 and b"slice status -5" in out + err           # the device parse body flags it: the file goes to the host recorder
    assert run([EMUL, "chain", str(clip)])[1] == run([EMUL, "pipeline", str(clip)])[1]


def test_random_bit_flips_round_trip(tmp_path):
    """a short seeded run of tools/cpu_bitflip_sweep.py (3100 cases by hand: 0 failures): lossless on arbitrary input"""
    import random
    rnd = random.Random(99)
    for i in range(40):
        name = rnd.choice(GOLDEN_NAMES)
        d = bytearray(open(os.path.join(GOLDEN, name + ".h264"), "rb").read())
        for _ in range(1 + i % 3):
            d[rnd.randrange(len(d))] ^= 1 << rnd.randrange(8)
        clip = tmp_path / "f.h264"
        clip.write_bytes(bytes(d))
        rc, out, err = run([EMUL, "files", str(clip)])
        assert rc == 0 and out.startswith(b"OK"), (name, i, out, err)


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_corrupted_container_ends_with_a_status_inside_its_buffers(name):
    """bit flips in the .avrec itself (what avr_decompress_files receives from outside): container reader, surrogate scan, plan and
    the decompress body under ASan/UBSan must end with a status or a wrong file, never outside their buffers
    (420 cases by hand: none)"""
    for seed in range(1, 9):
        rc, out, err = run([EMUL_ASAN, "filesfuzz", os.path.join(GOLDEN, name + ".h264"), str(seed), str(1 + seed % 6)])
        assert rc == 0 and out.startswith(b"TERMINATED"), (seed, out, err[-400:])


def test_slices_of_one_picture_must_not_overlap(tmp_path):
    """found by tools/cpu_bitflip_sweep.py (1 of 2000 four-flip cases): two flips inside the first slice of a picture of skippy_idr.h264 leave
    it ending cleanly (right bit, end_of_slice_flag = 1) but running past the next slice's first macroblock; positions coded twice in
    one picture store break the slice-parallel compress passes (chain != pipeline).  The recorder refuses the file (stored literally),
    the file API's device path flags it from the parse results (avr_api.cu)"""
    d = bytearray(open(os.path.join(GOLDEN, "skippy_idr.h264"), "rb").read())
    d[3447] ^= 1 << 6
    d[3270] ^= 1 << 3
    clip = tmp_path / "overlap.h264"
    clip.write_bytes(bytes(d))
    tool = os.path.join(ROOT, "avrecode-ms_b200", "bin", "avrtool")
    rc, _, err = run([tool, "stats", str(clip)])
    assert rc != 0 and b"rc=-4" in err, err
    rc, out, err = run([EMUL_ASAN, "files", str(clip)])
    assert rc == 0 and out.startswith(b"OK") and b"whole file literal" in out, out + err
