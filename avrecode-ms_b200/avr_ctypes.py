"""ctypes binding of the C ABI (include/avrecode_b200.h) for tests/ and bench.py.

This is the same stub a maintainer of the reference would write for any FFI (INTEGRATION.md shows
the C++ one); it adds no logic.  Every compute call goes to libavrecode_b200.so -> CUDA; if the
library is missing or no sm_100 GPU is usable the calls raise, there is no fallback.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libavrecode_b200.so")

AVR_OK = 0
MODE_PIPELINE = 0
MODE_CHAIN = 1

EXPORTS = [
    "avr_version", "avr_last_error", "avr_context_create", "avr_context_destroy", "avr_context_stream",
    "avr_record_new", "avr_record_free", "avr_rec_frame_spec", "avr_rec_slice_begin", "avr_rec_slice_end",
    "avr_rec_mb_xy", "avr_rec_begin_sub_mb", "avr_rec_end_sub_mb", "avr_rec_begin_coding_type",
    "avr_rec_end_coding_type", "avr_rec_bin", "avr_record_annexb", "avr_record_counts", "avr_record_span",
    "avr_cbatch_create", "avr_cbatch_create_files", "avr_cbatch_span_count", "avr_cbatch_span", "avr_cbatch_run", "avr_cbatch_fetch", "avr_cbatch_payload", "avr_cbatch_timing",
    "avr_cbatch_destroy", "avr_dbatch_create", "avr_dbatch_run", "avr_dbatch_fetch", "avr_dbatch_slice",
    "avr_dbatch_timing", "avr_dbatch_destroy", "avr_compress_files", "avr_decompress_files", "avr_buffer_free", "avr_compress_files_multi", "avr_decompress_files_multi",
    "avr_streams_encode", "avr_streams_decode", "avr_streams_encode_static", "avr_measure_recurrence_floor",
]
STREAM_FIXED_HALF = 0
STREAM_ADAPTIVE = 1


class SliceParams(C.Structure):
    _fields_ = [("slice_type", C.c_int32), ("first_mb", C.c_int32), ("mb_w", C.c_int32), ("mb_h", C.c_int32),
                ("num_ref", C.c_int32 * 2), ("cabac_init_idc", C.c_int32), ("slice_qp", C.c_int32),
                ("transform_8x8_mode", C.c_int32), ("chroma_format_idc", C.c_int32),
                ("direct_8x8_inference", C.c_int32)]


class DSlice(C.Structure):
    _fields_ = [("payload", C.c_void_p), ("payload_len", C.c_uint32), ("size", C.c_uint32), ("flip", C.c_uint32),
                ("sp", SliceParams)]


class Timing(C.Structure):
    _fields_ = [("h2d_ms", C.c_double), ("kernel_ms", C.c_double), ("d2h_ms", C.c_double),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64), ("bins", C.c_uint64),
                ("coded_symbols", C.c_uint64), ("launches", C.c_uint32), ("coder_launch_ms_x1000", C.c_uint32),
                ("parse_ms", C.c_double), ("host_ms", C.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class Buffer(C.Structure):
    _fields_ = [("data", C.POINTER(C.c_uint8)), ("size", C.c_size_t)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'`")
        L = C.CDLL(LIB_PATH)
        L.avr_last_error.restype = C.c_char_p
        L.avr_context_stream.restype = C.c_void_p
        L.avr_record_new.restype = C.c_void_p
        for name in EXPORTS:
            getattr(L, name)  # AttributeError if the ABI lost a symbol
        L.avr_context_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
        L.avr_context_destroy.argtypes = [C.c_void_p]
        L.avr_context_stream.argtypes = [C.c_void_p]
        L.avr_record_free.argtypes = [C.c_void_p]
        L.avr_record_annexb.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
        L.avr_record_counts.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
        L.avr_record_span.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_uint64), C.POINTER(C.c_uint32),
                                      C.POINTER(C.c_int32), C.POINTER(C.c_uint32), C.POINTER(SliceParams)]
        L.avr_cbatch_create.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_int, C.POINTER(C.c_void_p)]
        L.avr_cbatch_create_files.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.c_int, C.c_int,
                                              C.POINTER(C.c_void_p), C.POINTER(C.c_int)]
        L.avr_cbatch_span_count.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint32)]
        L.avr_cbatch_span.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.POINTER(C.c_uint64), C.POINTER(C.c_uint32),
                                      C.POINTER(C.c_int32)]
        L.avr_cbatch_run.argtypes = [C.c_void_p, C.c_int]
        L.avr_cbatch_fetch.argtypes = [C.c_void_p]
        L.avr_cbatch_payload.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.POINTER(C.POINTER(C.c_uint8)),
                                         C.POINTER(C.c_uint32)]
        L.avr_cbatch_timing.argtypes = [C.c_void_p, C.POINTER(Timing)]
        L.avr_cbatch_destroy.argtypes = [C.c_void_p]
        L.avr_dbatch_create.argtypes = [C.c_void_p, C.POINTER(DSlice), C.POINTER(C.c_uint32), C.c_int,
                                        C.POINTER(C.c_void_p)]
        L.avr_dbatch_run.argtypes = [C.c_void_p]
        L.avr_dbatch_fetch.argtypes = [C.c_void_p]
        L.avr_dbatch_slice.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.POINTER(C.c_uint8)), C.POINTER(C.c_uint32),
                                       C.POINTER(C.c_uint32)]
        L.avr_dbatch_timing.argtypes = [C.c_void_p, C.POINTER(Timing)]
        L.avr_dbatch_destroy.argtypes = [C.c_void_p]
        L.avr_compress_files.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.c_int, C.c_int,
                                         C.POINTER(Buffer), C.POINTER(Timing)]
        L.avr_decompress_files.argtypes = L.avr_compress_files.argtypes
        L.avr_compress_files_multi.argtypes = [C.POINTER(C.c_void_p), C.c_int] + L.avr_compress_files.argtypes[1:]
        L.avr_decompress_files_multi.argtypes = L.avr_compress_files_multi.argtypes
        L.avr_buffer_free.argtypes = [C.POINTER(Buffer)]
        L.avr_streams_encode.argtypes = [C.c_void_p, C.c_char_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.POINTER(Buffer),
                                         C.POINTER(C.c_uint64), C.POINTER(Timing)]
        L.avr_streams_decode.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_uint64), C.c_uint32, C.c_uint32, C.c_int, C.c_int,
                                         C.c_char_p, C.POINTER(Timing)]
        L.avr_streams_encode_static.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_void_p,
                                                C.c_uint64, C.c_void_p]
        L.avr_rec_frame_spec.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.avr_rec_slice_begin.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_uint32, C.POINTER(SliceParams)]
        L.avr_rec_slice_end.argtypes = [C.c_void_p]
        L.avr_rec_mb_xy.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.avr_rec_begin_sub_mb.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        L.avr_rec_end_sub_mb.argtypes = [C.c_void_p]
        L.avr_rec_begin_coding_type.argtypes = [C.c_void_p, C.c_int]
        L.avr_rec_end_coding_type.argtypes = [C.c_void_p, C.c_int]
        L.avr_rec_bin.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.avr_measure_recurrence_floor.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_double)]
        _lib = L
    return _lib


class AvrError(RuntimeError):
    pass


def _check(rc):
    if rc != AVR_OK:
        raise AvrError(f"avr error {rc}: {lib().avr_last_error().decode()}")


class Context:
    def __init__(self, device=0):
        self.h = C.c_void_p()
        _check(lib().avr_context_create(device, C.byref(self.h)))

    def close(self):
        if self.h:
            lib().avr_context_destroy(self.h)
            self.h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class Record:
    """Host records of one file (native Annex-B front end)."""

    def __init__(self, data: bytes):
        self.h = C.c_void_p(lib().avr_record_new())
        self.rc = lib().avr_record_annexb(self.h, data, len(data))
        c = (C.c_uint64 * 6)()
        lib().avr_record_counts(self.h, c)
        self.bins, self.blocks, self.slices, self.nspans, self.pictures, self.out_of_domain = [int(x) for x in c]

    def spans(self):
        out = []
        for i in range(self.nspans):
            off, size, idx, flip, sp = C.c_uint64(), C.c_uint32(), C.c_int32(), C.c_uint32(), SliceParams()
            _check(lib().avr_record_span(self.h, i, C.byref(off), C.byref(size), C.byref(idx), C.byref(flip), C.byref(sp)))
            out.append(dict(payload_off=off.value, size=size.value, slice_index=idx.value, flip=flip.value, sp=sp))
        return out

    def close(self):
        if self.h:
            lib().avr_record_free(self.h)
            self.h = C.c_void_p()


class HookRecord(Record):
    """A record filled through the AVCodecHooks-shaped sinks (avr_rec_*), the way a libavcodec-hooks Driver fills it
    (oracle/gpu_driver.cpp is the compiled one).  `events` = lines of the hook trace `oracle/_ref/ref_harness trace` prints from the
    reference's own av_decoder<> (format in oracle/ref_harness.cpp)."""

    PIP_SIGNIFICANCE_MAP = 2            # coding_hooks.h:10-22

    def __init__(self, events):
        L = lib()
        self.h = C.c_void_p(L.avr_record_new())
        self.rc = 0
        fnum, last_frame, open_slice, w, h = 0, None, False, 0, 0
        for ln in events:
            t = ln.split()
            if not t:
                continue
            k = t[0]
            if k == "c":
                L.avr_rec_bin(self.h, int(t[1]), int(t[2]))
            elif k == "b":
                L.avr_rec_bin(self.h, 1024, int(t[1]))
            elif k == "t":
                L.avr_rec_bin(self.h, 1025, int(t[1]))
            elif k == "M":
                L.avr_rec_mb_xy(self.h, int(t[1]), int(t[2]))
            elif k == "B":
                L.avr_rec_begin_sub_mb(self.h, *[int(x) for x in t[1:6]])
            elif k == "[":
                L.avr_rec_begin_coding_type(self.h, self.PIP_SIGNIFICANCE_MAP)
            elif k == "]":
                L.avr_rec_end_coding_type(self.h, self.PIP_SIGNIFICANCE_MAP)
                L.avr_rec_end_sub_mb(self.h)
            elif k == "F":                      # the model's cur_frame after ModelHooks::frame_spec: a flip shows as a new value
                cur = (int(t[1]), int(t[2]), int(t[3]))
                if cur != last_frame:
                    last_frame, fnum = cur, fnum + 1
                    w, h = cur[1], cur[2]
                    L.avr_rec_frame_spec(self.h, fnum, w, h)
            elif k == "S":                      # CABACCodingHooks::init_decoder + the Driver's block-finder decision
                if open_slice:
                    L.avr_rec_slice_end(self.h)
                sp = SliceParams()
                sp.mb_w, sp.mb_h = w, h
                off = int(t[2])
                L.avr_rec_slice_begin(self.h, 1 if off >= 0 else 0, max(off, 0), int(t[1]), C.byref(sp))
                open_slice = True
        if open_slice:
            L.avr_rec_slice_end(self.h)
        c = (C.c_uint64 * 6)()
        L.avr_record_counts(self.h, c)
        self.bins, self.blocks, self.slices, self.nspans, self.pictures, self.out_of_domain = [int(x) for x in c]


class CBatch:
    def __init__(self, ctx: Context, records=None, files=None, threads=0):
        """records: list of Record (host parse)  |  files: list of bytes (device-side parse)"""
        self.h = C.c_void_p()
        if files is not None:
            n = len(files)
            self._files = files
            ins = (C.c_char_p * n)(*files)
            lens = (C.c_size_t * n)(*[len(b) for b in files])
            st = (C.c_int * n)()
            _check(lib().avr_cbatch_create_files(ctx.h, ins, lens, n, threads or (os.cpu_count() or 1), C.byref(self.h), st))
            self.file_status = list(st)
            self.records = None
            self.nslices = []
            for f in range(n):
                c = C.c_uint32()
                _check(lib().avr_cbatch_span_count(self.h, f, C.byref(c)))
                k = 0
                for i in range(c.value):
                    idx = C.c_int32()
                    lib().avr_cbatch_span(self.h, f, i, None, None, C.byref(idx))
                    k += idx.value >= 0
                self.nslices.append(k)
            return
        self.records = records
        self.nslices = [r.slices for r in records]
        arr = (C.c_void_p * len(records))(*[r.h for r in records])
        _check(lib().avr_cbatch_create(ctx.h, arr, len(records), C.byref(self.h)))

    def run(self, mode=MODE_PIPELINE):
        _check(lib().avr_cbatch_run(self.h, mode))

    def fetch(self):
        _check(lib().avr_cbatch_fetch(self.h))

    def payload(self, file, slice_):
        p, n = C.POINTER(C.c_uint8)(), C.c_uint32()
        _check(lib().avr_cbatch_payload(self.h, file, slice_, C.byref(p), C.byref(n)))
        return C.string_at(p, n.value)

    def payloads(self):
        return [[self.payload(f, s) for s in range(k)] for f, k in enumerate(self.nslices)]

    def timing(self):
        t = Timing()
        lib().avr_cbatch_timing(self.h, C.byref(t))
        return t.as_dict()

    def close(self):
        if self.h:
            lib().avr_cbatch_destroy(self.h)
            self.h = C.c_void_p()


class DBatch:
    """slices_per_file: list (per file) of dicts(payload=bytes, size, flip, sp=SliceParams)."""

    def __init__(self, ctx: Context, slices_per_file):
        flat = [s for f in slices_per_file for s in f]
        self._keep = [C.create_string_buffer(s["payload"], len(s["payload"])) for s in flat]
        arr = (DSlice * max(1, len(flat)))()
        for i, s in enumerate(flat):
            arr[i].payload = C.cast(self._keep[i], C.c_void_p)
            arr[i].payload_len = len(s["payload"])
            arr[i].size = s["size"]
            arr[i].flip = s["flip"]
            arr[i].sp = s["sp"]
        per = (C.c_uint32 * max(1, len(slices_per_file)))(*[len(f) for f in slices_per_file])
        self.n = len(flat)
        self.h = C.c_void_p()
        _check(lib().avr_dbatch_create(ctx.h, arr, per, len(slices_per_file), C.byref(self.h)))

    def run(self):
        _check(lib().avr_dbatch_run(self.h))

    def fetch(self):
        _check(lib().avr_dbatch_fetch(self.h))

    def slice(self, i):
        p, n, b = C.POINTER(C.c_uint8)(), C.c_uint32(), C.c_uint32()
        _check(lib().avr_dbatch_slice(self.h, i, C.byref(p), C.byref(n), C.byref(b)))
        return C.string_at(p, n.value), b.value

    def timing(self):
        t = Timing()
        lib().avr_dbatch_timing(self.h, C.byref(t))
        return t.as_dict()

    def close(self):
        if self.h:
            lib().avr_dbatch_destroy(self.h)
            self.h = C.c_void_p()


class FileResults:
    """Library-owned output buffers of avr_compress_files / avr_decompress_files, exposed without a copy (what a C or
    C++ caller gets): views[i] is a memoryview over file i; close() hands the buffers back (avr_buffer_free)."""

    def __init__(self, outs, n):
        self._outs, self._n = outs, n
        self.views = [memoryview((C.c_uint8 * outs[i].size).from_address(C.addressof(outs[i].data.contents))).cast("B")
                      if outs[i].size else memoryview(b"") for i in range(n)]

    def close(self):
        self.views = []
        for i in range(self._n):
            lib().avr_buffer_free(C.byref(self._outs[i]))
        self._n = 0

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class _Multi:
    """adapter: fn(ctxs, nctx, ...) behind the single-context call shape of _files_call"""

    def __init__(self, fn, ctxs):
        self.fn, self.arr, self.n = fn, (C.c_void_p * len(ctxs))(*[c.h for c in ctxs]), len(ctxs)
        self.h = None

    def __call__(self, _h, *rest):
        return self.fn(self.arr, self.n, *rest)


def compress_files_multi(ctxs, blobs, threads=0, copy=True):
    m = _Multi(lib().avr_compress_files_multi, ctxs)
    return _files_call(m, m, blobs, threads or (os.cpu_count() or 1), copy)


def decompress_files_multi(ctxs, blobs, threads=0, copy=True):
    m = _Multi(lib().avr_decompress_files_multi, ctxs)
    return _files_call(m, m, blobs, threads or (os.cpu_count() or 1), copy)


def _files_call(fn, ctx, blobs, threads, copy=True):
    n = len(blobs)
    ins = (C.c_void_p * n)()
    keep = []
    for i, b in enumerate(blobs):            # bytes or any contiguous buffer (e.g. the views of a previous call)
        if isinstance(b, bytes):
            ins[i] = C.cast(C.c_char_p(b), C.c_void_p)
        else:
            arr = (C.c_uint8 * len(b)).from_buffer(b) if not b.readonly else (C.c_uint8 * len(b)).from_buffer_copy(b)
            keep.append(arr)
            ins[i] = C.addressof(arr)
    lens = (C.c_size_t * n)(*[len(b) for b in blobs])
    outs = (Buffer * n)()
    t = Timing()
    _check(fn(ctx.h, C.cast(ins, C.POINTER(C.c_char_p)), lens, n, threads, outs, C.byref(t)))
    if not copy:
        return FileResults(outs, n), t.as_dict()
    res = [C.string_at(outs[i].data, outs[i].size) for i in range(n)]
    for i in range(n):
        lib().avr_buffer_free(C.byref(outs[i]))
    return res, t.as_dict()


def compress_files(ctx, blobs, threads=0, copy=True):
    """`recode compress` for a batch of Annex-B H.264 byte strings -> .avrec byte strings (copy=False: FileResults)."""
    return _files_call(lib().avr_compress_files, ctx, blobs, threads or (os.cpu_count() or 1), copy)


def decompress_files(ctx, blobs, threads=0, copy=True):
    """`recode decompress` for a batch of .avrec byte strings -> original byte strings (copy=False: FileResults)."""
    return _files_call(lib().avr_decompress_files, ctx, blobs, threads or (os.cpu_count() or 1), copy)


def streams_encode(ctx, bits, nstreams, nbins, digit_bits=16, model=STREAM_FIXED_HALF):
    """bits: bytes of nstreams * nbins 0/1 values (stream-major).  Returns ([coded bytes per stream], timing dict)."""
    assert len(bits) == nstreams * nbins
    buf, t = Buffer(), Timing()
    off = (C.c_uint64 * (nstreams + 1))()
    _check(lib().avr_streams_encode(ctx.h, bytes(bits), nstreams, nbins, digit_bits, model, C.byref(buf), off, C.byref(t)))
    blob = C.string_at(buf.data, buf.size) if buf.size else b""
    lib().avr_buffer_free(C.byref(buf))
    return [blob[off[s]:off[s + 1]] for s in range(nstreams)], t.as_dict()


def streams_decode(ctx, coded, nstreams, nbins, digit_bits=16, model=STREAM_FIXED_HALF):
    """coded: list of per-stream bytes.  Returns (bytes of nstreams * nbins 0/1 values, timing dict)."""
    assert len(coded) == nstreams
    off = (C.c_uint64 * (nstreams + 1))()
    for s, c in enumerate(coded):
        off[s + 1] = off[s] + len(c)
    out = C.create_string_buffer(max(nstreams * nbins, 1))
    t = Timing()
    _check(lib().avr_streams_decode(ctx.h, b"".join(coded), off, nstreams, nbins, digit_bits, model, out, C.byref(t)))
    return out.raw[:nstreams * nbins], t.as_dict()


STREAMS_STATIC_SCRATCH_BYTES = 8 << 20


def streams_coded_bound(nbins, model):
    """AVR_STREAMS_CODED_BOUND of include/avrecode_b200.h"""
    return (((nbins // 8 + 32) if model == STREAM_FIXED_HALF else (nbins + 32)) + 15) & ~15


def streams_encode_static(ctx, d_bits_ptr, nstreams, nbins, digit_bits, model, d_coded_ptr, coded_cap, d_coded_off_ptr):
    """avr_streams_encode_static: raw DEVICE pointers (ints) in and out; allocates nothing.  Returns the C return code."""
    return lib().avr_streams_encode_static(ctx.h, d_bits_ptr, nstreams, nbins, digit_bits, model, d_coded_ptr, coded_cap, d_coded_off_ptr)


def recurrence_floor(ctx, steps=1 << 20):
    """SM cycles per bin of the bare low/range recurrence on one thread (avr_measure_recurrence_floor)."""
    v = C.c_double()
    _check(lib().avr_measure_recurrence_floor(ctx.h, steps, C.byref(v)))
    return v.value
