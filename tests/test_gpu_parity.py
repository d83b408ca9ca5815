"""Parity tests proper: the CUDA path through the C ABI against the oracle, the golden reference
containers, and size-independent properties (roundtrip == input)."""
import os

import pytest

from avrec_wire import parse_avrec
from conftest import GOLDEN, GOLDEN_NAMES, REF, ROOT, gen_clip, have_ref, oracle_encode, ref_trace, run, golden_trace

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import avr_ctypes as A
    c = A.Context(0)
    yield c
    c.close()


def golden_payloads(name):
    return [b["cabac"] for b in parse_avrec(open(os.path.join(GOLDEN, name + ".avrec"), "rb").read()) if "cabac" in b]


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_payloads_match_reference_golden(ctx, name, mode):
    import avr_ctypes as A
    rec = A.Record(open(os.path.join(GOLDEN, name + ".h264"), "rb").read())
    assert rec.rc == 0
    b = A.CBatch(ctx, [rec])
    b.run(mode)
    b.fetch()
    assert b.payloads()[0] == golden_payloads(name)
    b.close()
    rec.close()


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_container_is_byte_identical_to_reference(ctx, name):
    """.avrec bytes == the unmodified reference's `recode compress` output (golden)."""
    import avr_ctypes as A
    data = open(os.path.join(GOLDEN, name + ".h264"), "rb").read()
    comp, _ = A.compress_files(ctx, [data])
    assert comp[0] == open(os.path.join(GOLDEN, name + ".avrec"), "rb").read()


def test_device_parse_batch_matches_reference_golden(ctx):
    """raw files in, CABAC parse on the GPU (avr_k_parse): payloads == the reference's, spans == host recorder's"""
    import avr_ctypes as A
    datas = [open(os.path.join(GOLDEN, n + ".h264"), "rb").read() for n in GOLDEN_NAMES]
    b = A.CBatch(ctx, files=datas)
    assert b.file_status == [0] * len(datas)
    b.run(A.MODE_PIPELINE)
    b.fetch()
    got = b.payloads()
    for n, g in zip(GOLDEN_NAMES, got):
        assert g == golden_payloads(n), n
    t = b.timing()
    assert t["parse_ms"] > 0 and t["launches"] > 2
    b.close()


def test_host_parse_path_gives_the_same_containers(ctx, tmpdir_path, monkeypatch):
    """the fallback (host recorder) and the device parse write identical containers"""
    import avr_ctypes as A
    datas = [gen_clip(os.path.join(tmpdir_path, f"h{i}.h264"), seed=70 + i, mb_w=9 + i, mb_h=5 + i, frames=4,
                      slices_per_frame=1 + i, transform_8x8=i & 1, p_skip=120 + 40 * i) for i in range(3)]
    dev, td = A.compress_files(ctx, datas)
    monkeypatch.setenv("AVR_HOST_PARSE", "1")
    host, th = A.compress_files(ctx, datas)
    monkeypatch.delenv("AVR_HOST_PARSE")
    assert dev == host
    assert td["parse_ms"] > 0 and th["parse_ms"] == 0
    assert td["coded_symbols"] == th["coded_symbols"]


def test_device_parse_overflow_falls_back_to_host_recorder(ctx, monkeypatch):
    """record capacity exceeded on the device -> that file goes through the host recorder; same container"""
    import avr_ctypes as A
    datas = [open(os.path.join(GOLDEN, n + ".h264"), "rb").read() for n in GOLDEN_NAMES]
    monkeypatch.setenv("AVR_TEST_PARSE_CAP", "1")
    b = A.CBatch(ctx, files=datas)
    assert any(st == -3 for st in b.file_status), b.file_status      # AVR_E_OVERFLOW
    b.close()
    comp, t = A.compress_files(ctx, datas)
    monkeypatch.delenv("AVR_TEST_PARSE_CAP")
    for n, c in zip(GOLDEN_NAMES, comp):
        assert c == open(os.path.join(GOLDEN, n + ".avrec"), "rb").read(), n


def test_experimental_speculative_chain_restores_the_same_files(ctx, tmpdir_path, monkeypatch):
    """AVR_DEC_SPEC=1: the warp-speculative decompress kernel (avr_kernels_spec.cu; measured slower, not the product) gives back the
    same bytes as the product chain, on the goldens (I/P/B, 8x8) and on generated clips with several slices per picture"""
    import avr_ctypes as A
    datas = [open(os.path.join(GOLDEN, n + ".h264"), "rb").read() for n in GOLDEN_NAMES]
    datas += [gen_clip(os.path.join(tmpdir_path, f"s{i}.h264"), seed=170 + i, mb_w=11 + 3 * i, mb_h=6 + i, frames=5,
                       slices_per_frame=1 + i, transform_8x8=i & 1, p_skip=60 + 60 * i) for i in range(3)]
    comp, _ = A.compress_files(ctx, datas)
    monkeypatch.setenv("AVR_DEC_SPEC", "1")
    back, t = A.decompress_files(ctx, comp)
    monkeypatch.delenv("AVR_DEC_SPEC")
    assert back == datas
    assert t["kernel_ms"] > 0


def test_per_slice_fallback_in_the_file_api(ctx):
    """damaged slices and slices outside the supported syntax stay literal ONE BY ONE (skip_coded spans / no span), the rest of
    the file is coded and every file comes back byte for byte; golden neighbours in the same batch are untouched"""
    import avr_ctypes as A
    from avrec_wire import parse_avrec
    from conftest import as_cavlc, damage_slice
    g = {n: open(os.path.join(GOLDEN, n + ".h264"), "rb").read() for n in GOLDEN_NAMES}
    # slices 0, 3, 6, 9 open a picture (first_mb = 0: the store flip happens at a slice that is then left literal)
    damaged = [damage_slice(g["multislice_multiref"], w, seed) for w, seed in ((1, 1), (3, 2), (11, 3))]
    mixed = g["ip_basic"] + as_cavlc(g["skippy_idr"]) + g["b_slices"]
    first = damage_slice(g["multislice_multiref"], 0, 4)          # the IDR's first slice
    two = damage_slice(damage_slice(g["multislice_multiref"], 6, 5), 7, 6)   # two in a row, the first of them opening a picture
    datas = [g["ip_basic"]] + damaged + [mixed, as_cavlc(g["dct8x8"]), g["multislice_multiref"], first, two]
    comp, _ = A.compress_files(ctx, datas)
    assert comp[0] == open(os.path.join(GOLDEN, "ip_basic.avrec"), "rb").read()
    assert comp[6] == open(os.path.join(GOLDEN, "multislice_multiref.avrec"), "rb").read()
    assert sum(1 for b in parse_avrec(comp[7]) if b.get("skip_coded")) == 1
    assert sum(1 for b in parse_avrec(comp[8]) if b.get("skip_coded")) == 2 and sum(1 for b in parse_avrec(comp[8]) if "cabac" in b) == 10
    for c in comp[1:4]:                                           # 11 coded slices + one size-only block
        blocks = parse_avrec(c)
        assert sum(1 for b in blocks if "cabac" in b) == 11 and sum(1 for b in blocks if b.get("skip_coded")) == 1
    blocks = parse_avrec(comp[4])
    assert sum(1 for b in blocks if "cabac" in b) == 3 + 10       # ip_basic + b_slices; the CAVLC part takes no block
    assert all("literal" in b for b in parse_avrec(comp[5]))      # nothing to code: literal only
    assert len(comp[4]) < len(mixed)
    back, _ = A.decompress_files(ctx, comp)
    assert back == datas
    # the device parse refuses exactly the damaged files (they go through the host recorder); everything else stays on the device
    b = A.CBatch(ctx, files=datas)
    assert [st != 0 for st in b.file_status] == [False, True, True, True, False, False, False, True, True], b.file_status
    b.close()


def test_streams_that_begin_in_the_middle_of_a_picture(ctx, tmpdir_path):
    """no first_mb = 0 slice before the first coded slice (a cut stream; a first slice with a refused header, here QP out of range):
    the first slice seen starts the picture on both directions -- no slice is left without a picture store of its size"""
    import avr_ctypes as A
    from conftest import drop_slice_nal
    src = open(os.path.join(GOLDEN, "multislice_multiref.h264"), "rb").read()
    datas = [drop_slice_nal(src, 0), drop_slice_nal(src, 3), drop_slice_nal(drop_slice_nal(src, 0), 0)]
    datas += [gen_clip(os.path.join(tmpdir_path, f"q{qp}.h264"), seed=5, mb_w=8, mb_h=6, frames=3, slices_per_frame=2, num_ref=2, qp=qp)
              for qp in (0, 51)]
    datas.append(open(os.path.join(GOLDEN, "ip_basic.h264"), "rb").read())
    comp, _ = A.compress_files(ctx, datas)
    assert comp[-1] == open(os.path.join(GOLDEN, "ip_basic.avrec"), "rb").read()
    assert all(len(c) < len(d) + 64 for c, d in zip(comp, datas))
    back, _ = A.decompress_files(ctx, comp)
    assert back == datas


def test_device_parse_accepts_every_slice_of_generated_clips(ctx, tmpdir_path):
    """the end-of-slice length check of the device parse (avr_parse.h) must not send a conforming file to the host recorder:
    clips of the bench family (several sizes, 1-4 slices per picture, I/P/B, 8x8 on and off)"""
    import avr_ctypes as A
    datas = [gen_clip(os.path.join(tmpdir_path, f"v{i}.h264"), seed=900 + i, mb_w=8 + 5 * i, mb_h=6 + 3 * i, frames=3 + (i & 3),
                      slices_per_frame=1 + i % 4, transform_8x8=(i >> 1) & 1, p_skip=30 * i, num_ref=1 + i % 3,
                      **({"p_bslice": 500} if i % 3 == 2 else {})) for i in range(9)]
    b = A.CBatch(ctx, files=datas)
    assert not any(b.file_status), b.file_status
    b.close()


def test_cli_recode_test_directory(ctx, tmpdir_path):
    """`recode test <dir>` (test.cpp): decompressed copies + metrics.csv with the reference's columns"""
    import shutil
    d = os.path.join(tmpdir_path, "corpus")
    os.makedirs(d)
    for n in GOLDEN_NAMES[:3]:
        shutil.copy(os.path.join(GOLDEN, n + ".h264"), d)
    rc, out, err = run([os.path.join(ROOT, "avrecode-ms_b200", "bin", "recode"), "test", d])
    assert rc == 0, out + err
    rows = open(os.path.join(d, "output", "metrics.csv")).read().splitlines()
    assert rows[0].startswith("File,Duration,Initial size (MB),Compressed size (MB),Compression rate (%),Space saving (%),Total time (ms)")
    assert len(rows) == 4
    for n in GOLDEN_NAMES[:3]:
        assert open(os.path.join(d, "output", n + ".h264"), "rb").read() == open(os.path.join(GOLDEN, n + ".h264"), "rb").read()


def test_cli_matches_reference_cli_shape(ctx, tmpdir_path):
    """compress / decompress / roundtrip: same arguments, container bytes and exit codes as the reference's CLI"""
    exe = os.path.join(ROOT, "avrecode-ms_b200", "bin", "recode")
    src = os.path.join(GOLDEN, "multislice.h264") if os.path.exists(os.path.join(GOLDEN, "multislice.h264")) else os.path.join(GOLDEN, GOLDEN_NAMES[0] + ".h264")
    name = os.path.basename(src)[:-5]
    avrec, back = os.path.join(tmpdir_path, "x.avrec"), os.path.join(tmpdir_path, "x.h264")
    assert run([exe, "compress", src, avrec])[0] == 0
    assert open(avrec, "rb").read() == open(os.path.join(GOLDEN, name + ".avrec"), "rb").read()
    assert run([exe, "decompress", avrec, back])[0] == 0
    assert open(back, "rb").read() == open(src, "rb").read()
    rc, out, _ = run([exe, "compress", src])                       # stdout when no output is named (recode.cpp:1659)
    assert rc == 0 and out == open(avrec, "rb").read()
    assert run([exe, "roundtrip", src])[0] == 0
    assert run([exe, "decompress", src, back])[0] == 1             # not a container -> exit 1
    assert run([exe, "frobnicate", src])[0] == 1
    assert run([exe, "compress", os.path.join(tmpdir_path, "missing.h264")])[0] == 1


def test_decompress_reference_containers(ctx):
    """our decompressor restores the input from containers written by the REFERENCE compressor"""
    import avr_ctypes as A
    blobs = [open(os.path.join(GOLDEN, n + ".avrec"), "rb").read() for n in GOLDEN_NAMES]
    out, t = A.decompress_files(ctx, blobs)
    for n, o in zip(GOLDEN_NAMES, out):
        assert o == open(os.path.join(GOLDEN, n + ".h264"), "rb").read(), n
    assert t["bins"] > 0


def test_batch_pipeline_equals_chain_and_oracle(ctx, tmpdir_path):
    import avr_ctypes as A
    params = [dict(seed=40 + i, mb_w=8 + 3 * i, mb_h=6 + i, frames=3 + i, slices_per_frame=1 + i % 3,
                   transform_8x8=i % 2, num_ref=1 + i % 3, p_bslice=(0, 0, 180, 220)[i % 4]) for i in range(6)]
    datas, wants = [], []
    for i, p in enumerate(params):
        clip = os.path.join(tmpdir_path, f"c{i}.h264")
        datas.append(gen_clip(clip, **p))
        if have_ref():
            tr = os.path.join(tmpdir_path, f"c{i}.trace")
            ref_trace(clip, tr)
            wants.append(oracle_encode(tr))
    recs = [A.Record(d) for d in datas]
    res = []
    for mode in (A.MODE_PIPELINE, A.MODE_CHAIN):
        b = A.CBatch(ctx, recs)
        b.run(mode)
        b.fetch()
        res.append(b.payloads())
        b.close()
    assert res[0] == res[1]
    if wants:
        assert res[0] == wants
    comp, _ = A.compress_files(ctx, datas)
    back, _ = A.decompress_files(ctx, comp)
    assert back == datas


def test_edge_inputs(ctx, tmpdir_path):
    """empty file, junk, a stream with no slices, slices too short to recode (< 8 bytes -> skip_coded)"""
    import avr_ctypes as A
    clip = gen_clip(os.path.join(tmpdir_path, "s.h264"), seed=50, mb_w=2, mb_h=1, frames=6, p_skip=250)
    sps_pps_only = clip[:clip.find(b"\x00\x00\x01\x65") if b"\x00\x00\x01\x65" in clip else 40]
    blobs = [b"", b"junk that is not h264" * 10, sps_pps_only, clip]
    comp, _ = A.compress_files(ctx, blobs)
    back, _ = A.decompress_files(ctx, comp)
    assert back == blobs
    kinds = [k for b in parse_avrec(comp[3]) for k in b if k in ("skip_coded", "cabac")]
    assert "skip_coded" in kinds or "cabac" in kinds


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built")
def test_container_identical_to_live_reference_with_skipped_slices(ctx, tmpdir_path):
    import avr_ctypes as A
    clip = os.path.join(tmpdir_path, "c.h264")
    data = gen_clip(clip, seed=51, mb_w=4, mb_h=3, frames=10, p_skip=235, slices_per_frame=2)
    avrec = os.path.join(tmpdir_path, "c.avrec")
    rc, _, err = run([os.path.join(REF, "recode"), "compress", clip, avrec])
    assert rc == 0, err
    comp, _ = A.compress_files(ctx, [data])
    assert comp[0] == open(avrec, "rb").read()


def test_large_roundtrip_property(ctx, tmpdir_path):
    """bench-sized property check: decompress(compress(x)) == x on a 1080p multi-slice clip, and
    pipeline == chain on the same records"""
    import avr_ctypes as A
    data = gen_clip(os.path.join(tmpdir_path, "big.h264"), seed=60, mb_w=120, mb_h=68, frames=6, slices_per_frame=8,
                    transform_8x8=1, num_ref=2)
    comp, tc = A.compress_files(ctx, [data])
    back, td = A.decompress_files(ctx, comp)
    assert back[0] == data
    assert tc["bins"] == td["bins"] > 1_000_000
    rec = A.Record(data)
    res = []
    for mode in (0, 1):
        b = A.CBatch(ctx, [rec]); b.run(mode); b.fetch(); res.append(b.payloads()); b.close()
    assert res[0] == res[1]


@pytest.mark.parametrize("shape", [
    dict(seed=61, mb_w=240, mb_h=135, frames=2, slices_per_frame=68, p_sig=150, p_cbf=230, p_cbp=240, transform_8x8=1),   # config 5: 2160p, 68 slices/frame, dense residuals
    dict(seed=62, mb_w=120, mb_h=68, frames=12, slices_per_frame=8, num_ref=3),                                          # config 3: 1080p, 8 slices/frame
    dict(seed=2, mb_w=120, mb_h=68, frames=5, slices_per_frame=1),                                                       # config 2: 1080p, 1 slice/frame
])
def test_baseline_config_shapes_roundtrip(ctx, tmpdir_path, shape):
    """BASELINE configs[1] / configs[2] / configs[4] picture and slice geometry (fewer frames): roundtrip == input, device parse ==
    host parse, pipeline == chain"""
    import avr_ctypes as A
    data = gen_clip(os.path.join(tmpdir_path, "cfg.h264"), **shape)
    comp, tc = A.compress_files(ctx, [data])
    back, td = A.decompress_files(ctx, comp)
    assert back[0] == data
    assert tc["parse_ms"] > 0 and tc["bins"] == td["bins"] > 1_000_000
    rec = A.Record(data)
    assert rec.rc == 0 and rec.slices >= shape["frames"] * (shape["slices_per_frame"] - 1)
    b = A.CBatch(ctx, [rec]); b.run(A.MODE_CHAIN); b.fetch()
    want = b.payloads()[0]
    b.close(); rec.close()
    got = [blk["cabac"] for blk in parse_avrec(comp[0]) if "cabac" in blk]
    assert got == want


IN_DOMAIN_SHAPES = [
    dict(seed=61, mb_w=240, mb_h=135, frames=2, slices_per_frame=68, p_sig=150, p_cbf=230, p_cbp=240),   # config 5 geometry, no 8x8
    dict(seed=62, mb_w=120, mb_h=68, frames=6, slices_per_frame=8, num_ref=3),                          # config 3 geometry
    dict(seed=2, mb_w=120, mb_h=68, frames=5, slices_per_frame=1),                                       # config 2 geometry
    dict(seed=63, mb_w=120, mb_h=68, frames=4, slices_per_frame=8, num_ref=3, p_bslice=200),             # config 3 geometry with B slices
]


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("shape", IN_DOMAIN_SHAPES)
def test_baseline_config_shapes_against_live_reference(ctx, tmpdir_path, shape):
    """parity pinned at the benchmarked picture sizes.  In-domain variants (no 8x8 transform) of the
    BASELINE config 2/3/5 shapes: our container == `oracle/_ref/recode compress` byte for byte (estimator halving after
    long runs, key widths, > 1 M sort keys per file), the REFERENCE decompressor restores the input from OUR container,
    and our decompressor restores it from the reference's."""
    import avr_ctypes as A
    clip = os.path.join(tmpdir_path, "cfg.h264")
    data = gen_clip(clip, **shape)
    ref_avrec = os.path.join(tmpdir_path, "ref.avrec")
    rc, _, err = run([os.path.join(REF, "recode"), "compress", clip, ref_avrec], timeout=600)
    assert rc == 0, err
    want = open(ref_avrec, "rb").read()
    comp, tc = A.compress_files(ctx, [data])
    assert tc["bins"] > 1_000_000
    assert comp[0] == want, "container differs from the reference compressor's"
    ours = os.path.join(tmpdir_path, "ours.avrec")
    open(ours, "wb").write(comp[0])
    back_ref = os.path.join(tmpdir_path, "back_ref.h264")
    rc, _, err = run([os.path.join(REF, "recode"), "decompress", ours, back_ref], timeout=600)
    assert rc == 0, err
    assert open(back_ref, "rb").read() == data, "reference decompressor does not restore the input from our container"
    back, td = A.decompress_files(ctx, [want])
    assert back[0] == data and td["bins"] == tc["bins"]


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_reference_decompressor_accepts_our_container(ctx, tmpdir_path, name):
    """the survey's third oracle: a repo-written container fed to the reference DEcompressor (in-domain clips only: the
    reference cannot invert 8x8 blocks)"""
    import avr_ctypes as A
    if name in ("dct8x8", "dense_full_blocks"):  # (the B-slice goldens are in the reference's domain: both run)    # both generated with transform_8x8=1 (tests/golden/make_golden.py)
        pytest.skip("reference decompressor does not invert 8x8-transform streams (recode.cpp:1215-1219 vs 1485-1487)")
    data = open(os.path.join(GOLDEN, name + ".h264"), "rb").read()
    comp, _ = A.compress_files(ctx, [data])
    p = os.path.join(tmpdir_path, "o.avrec")
    open(p, "wb").write(comp[0])
    back = os.path.join(tmpdir_path, "o.h264")
    rc, _, err = run([os.path.join(REF, "recode"), "decompress", p, back], timeout=120)
    assert rc == 0, err
    assert open(back, "rb").read() == data


@pytest.mark.parametrize("kind", ["empty", "junk", "sps_pps_only", "mp4ish"])
def test_single_file_batches_without_coded_slices(ctx, tmpdir_path, kind):
    """a batch in which NO file has a coded slice (ns == 0) crashed avr_cbatch_create_files; the reference writes
    a literal-only container for such input"""
    import avr_ctypes as A
    clip = gen_clip(os.path.join(tmpdir_path, "s.h264"), seed=50, mb_w=2, mb_h=1, frames=2)
    blob = {"empty": b"", "junk": b"junk that is not h264" * 10,
            "sps_pps_only": clip[:clip.find(b"\x00\x00\x01\x65")] if b"\x00\x00\x01\x65" in clip else clip[:40],
            "mp4ish": b"\x00\x00\x00\x18ftypmp42" + bytes(range(200))}[kind]
    comp, _ = A.compress_files(ctx, [blob])
    blocks = parse_avrec(comp[0])
    assert all("cabac" not in b for b in blocks)
    back, _ = A.decompress_files(ctx, comp)
    assert back == [blob]
    # the same through the CLI
    exe = os.path.join(ROOT, "avrecode-ms_b200", "bin", "recode")
    src, dst, rt = (os.path.join(tmpdir_path, n) for n in ("in.bin", "out.avrec", "rt.bin"))
    open(src, "wb").write(blob)
    assert run([exe, "compress", src, dst])[0] == 0
    assert open(dst, "rb").read() == comp[0]
    assert run([exe, "decompress", dst, rt])[0] == 0
    assert open(rt, "rb").read() == blob


def test_lanes_and_multi_context_give_the_same_containers(ctx, tmpdir_path, monkeypatch):
    """avr_*_files cut their input into sub-batches over several lanes (AVR_LANES) and avr_*_files_multi over several contexts
    (one per GPU): the containers and the restored files must not depend on either"""
    import avr_ctypes as A
    import torch
    datas = [gen_clip(os.path.join(tmpdir_path, f"m{i}.h264"), seed=300 + i, mb_w=6 + i % 5, mb_h=4 + i % 3, frames=2 + i % 3,
                      slices_per_frame=1 + i % 2, transform_8x8=i & 1) for i in range(160)]
    datas[7] = b"not a stream"                      # a literal-only file in the middle of a batch
    monkeypatch.setenv("AVR_LANES", "1")
    want, _ = A.compress_files(ctx, datas)
    for lanes, bpl in (("2", "2"), ("3", "1")):
        monkeypatch.setenv("AVR_LANES", lanes)
        monkeypatch.setenv("AVR_DLANES", lanes)
        monkeypatch.setenv("AVR_BATCHES_PER_LANE", bpl)
        monkeypatch.setenv("AVR_DBATCHES_PER_LANE", bpl)
        got, t = A.compress_files(ctx, datas)
        assert got == want
        back, _ = A.decompress_files(ctx, got)
        assert back == datas
    other = A.Context(1 if torch.cuda.device_count() > 1 else 0)
    try:
        got, _ = A.compress_files_multi([ctx, other], datas)
        assert got == want
        back, _ = A.decompress_files_multi([ctx, other], got)
        assert back == datas
    finally:
        other.close()
    # the CLI form
    exe = os.path.join(ROOT, "avrecode-ms_b200", "bin", "recode")
    paths = [os.path.join(tmpdir_path, f"m{i}.h264") for i in range(0, 160, 2) if i != 7]
    assert run([exe, "roundtrip-batch", "--gpus", "2"] + paths)[0] == 0


@pytest.mark.parametrize("params", [
    dict(seed=81, mb_w=12, mb_h=9, frames=9, repeat_frame_num=1, slices_per_frame=2),                       # non-reference P pictures
    dict(seed=82, mb_w=10, mb_h=8, frames=8, repeat_frame_num=1, p_bslice=180, num_ref=2, transform_8x8=1),   # non-reference B pictures
])
def test_repeated_frame_num_round_trips_where_the_reference_cannot(ctx, tmpdir_path, params):
    """consecutive pictures sharing frame_num (non-reference pictures): the reference skips the store reset (recode.cpp:832-834),
    counts stale flags and cannot decompress its own output.  Here every picture gets its own store in both
    directions: the stream is coded (not stored literally) and round-trips.  AVR_STRICT_DOMAIN=1 is covered by
    tests/test_host_logic.py (a fresh process: the switch is read once)."""
    import avr_ctypes as A
    clip = os.path.join(tmpdir_path, "rf.h264")
    data = gen_clip(clip, **params)
    if have_ref():
        assert run([os.path.join(REF, "recode"), "roundtrip", clip])[0] != 0, "the reference round-trips this clip after all?"
    comp, tc = A.compress_files(ctx, [data])
    assert any("cabac" in b for b in parse_avrec(comp[0])) and len(comp[0]) < len(data)
    back, _ = A.decompress_files(ctx, comp)
    assert back[0] == data
    rec = A.Record(data)
    assert rec.rc == 0 and rec.out_of_domain == 2
    rec.close()
    # the reference's domain only: a literal block, still a valid container
    exe = os.path.join(ROOT, "avrecode-ms_b200", "bin", "recode")
    lit, rt = os.path.join(tmpdir_path, "lit.avrec"), os.path.join(tmpdir_path, "rt.h264")
    assert run([exe, "compress", clip, lit], env=dict(os.environ, AVR_STRICT_DOMAIN="1"))[0] == 0
    assert all("cabac" not in b for b in parse_avrec(open(lit, "rb").read()))
    assert run([exe, "decompress", lit, rt])[0] == 0 and open(rt, "rb").read() == data
