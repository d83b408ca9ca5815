/* avrecode entropy hot path, B200 (sm_100a) -- C ABI.
 *
 * Drop-in boundary for ONE path of pbluc/avrecode-ms: per-bin context pick -> adaptive 64-bit
 * range coder (compress), and range decode -> CABAC re-encode (decompress).  Plain pointers and
 * sizes only; no exceptions cross this boundary (the reference throws C++ exceptions through
 * libavcodec's C frames, recode.cpp:169,1372,1556 -- here every call returns AVR_OK or a
 * negative AVR_E_* code and avr_last_error() holds the message).
 *
 * There is NO CPU implementation behind these calls: every compute entry point fails with
 * AVR_E_CUDA when no sm_100-class device is usable.
 *
 * What each group replaces in the reference (file:line into pbluc/avrecode-ms):
 *   recorder      the hook sinks compressor::cabac_decoder::get/get_bypass/get_terminate/
 *                 begin_coding_type/end_coding_type (recode.cpp:1182-1236) and
 *                 h264_model::update_frame_spec / mb_coord / sub_mb_* setters (recode.cpp:173-202,
 *                 831-850), i.e. the AVCodecHooks table of ffmpeg/libavcodec/coding_hooks.h:27-55
 *   compress      h264_symbol::execute + h264_model + recoded_code::encoder
 *                 (recode.cpp:1075-1103, 625-1066; arithmetic_code.h:106-190)
 *   decompress    decompressor::cabac_decoder::get* + h264_model + recoded_code::decoder +
 *                 cabac::encoder + libavcodec's slice_data() walk that drives them
 *                 (recode.cpp:1418-1527; arithmetic_code.h:220-288; cabac_code.h:30-67;
 *                 ffmpeg/libavcodec/h264_cabac.c:1926-2480)
 *   files         compressor::run / decompressor::run incl. the recode.proto container
 *                 (recode.cpp:1109-1139, 1282-1304, 1331-1416; recode.proto:1-19)
 */
#ifndef AVRECODE_B200_H
#define AVRECODE_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define AVR_OK 0
#define AVR_E_ARG -1
#define AVR_E_CUDA -2
#define AVR_E_OVERFLOW -3
#define AVR_E_UNSUPPORTED -4
#define AVR_E_CORRUPT -5
#define AVR_E_NOMEM -6

typedef struct avr_context avr_context;   /* one per GPU; not thread-safe, use one per host thread */
typedef struct avr_record avr_record;     /* flat per-file record arrays (host) */
typedef struct avr_cbatch avr_cbatch;     /* a staged compress batch (device resident) */
typedef struct avr_dbatch avr_dbatch;     /* a staged decompress batch (device resident) */

/* Slice header fields the device-side slice_data() walk needs (what libavcodec holds in
 * H264SliceContext / PPS / SPS when it calls the hooks). */
typedef struct avr_slice_params {
  int32_t slice_type;            /* 0 = P, 1 = B, 2 = I (slice_type % 5) */
  int32_t first_mb;
  int32_t mb_w, mb_h;
  int32_t num_ref[2];
  int32_t cabac_init_idc;
  int32_t slice_qp;
  int32_t transform_8x8_mode;
  int32_t chroma_format_idc;
  int32_t direct_8x8_inference;
} avr_slice_params;

/* One coded slice handed to decompress (reference: one Recoded::Block with `cabac`). */
typedef struct avr_dslice {
  const uint8_t* payload;        /* Block.cabac */
  uint32_t payload_len;
  uint32_t size;                 /* Block.size = CABAC bytes of the original slice */
  uint32_t flip;                 /* pictures started since the file's previous coded slice: 0, 1, or 2 (= 2 or more / size change) */
  avr_slice_params sp;
} avr_dslice;

typedef struct avr_timing {
  double h2d_ms, kernel_ms, d2h_ms;       /* CUDA-event times on the context's stream */
  uint64_t h2d_bytes, d2h_bytes;
  uint64_t bins;                          /* source bins (get/get_bypass/get_terminate calls) */
  uint64_t coded_symbols;                 /* symbols through the range coder (EOB bins excluded, nnz bits included) */
  uint32_t launches;                      /* kernels launched by the run */
  uint32_t coder_launch_ms_x1000;         /* duration of the dominant kernel (range-coder pass / chain kernel), microseconds */
  double parse_ms;                        /* device-side parse kernel (avr_cbatch_create_files), CUDA events */
  double host_ms;                         /* host wall time inside the call(s): scan, staging, framing */
} avr_timing;

/* ---- context ---------------------------------------------------------------------------- */
int avr_version(void);
const char* avr_last_error(void);                         /* thread-local */
int avr_context_create(int device, avr_context** out);    /* AVR_E_CUDA if the device is unusable */
void avr_context_destroy(avr_context* ctx);
void* avr_context_stream(avr_context* ctx);               /* cudaStream_t the library launches on */

/* ---- recorder: AVCodecHooks-shaped sinks (coding_hooks.h:27-55) ------------------------- */
avr_record* avr_record_new(void);
void avr_record_free(avr_record* r);
void avr_rec_frame_spec(avr_record* r, int frame_num, int mb_width, int mb_height);        /* ModelHooks::frame_spec */
/* CABACCodingHooks::init_decoder.  coded / payload_off / size are the Driver's block-finder decision (recode.cpp:1282-1304: where
 * the slice's CABAC bytes sit in the original file, or coded = 0 for a slice that stays literal / skip_coded); init_decoder's own
 * arguments give buf and size.  sp is stored for avr_record_span only: the compress path never reads it (decompress derives it from
 * the container's literal header bytes), so a hook Driver passes the picture size and zeros (oracle/gpu_driver.cpp does). */
void avr_rec_slice_begin(avr_record* r, int coded, uint64_t payload_off, uint32_t size,
                         const avr_slice_params* sp);
void avr_rec_slice_end(avr_record* r);
void avr_rec_mb_xy(avr_record* r, int x, int y);                                           /* ModelHooks::mb_xy */
void avr_rec_begin_sub_mb(avr_record* r, int cat, int scan8index, int max_coeff, int is_dc, int chroma422);
void avr_rec_end_sub_mb(avr_record* r);
void avr_rec_begin_coding_type(avr_record* r, int coding_type);                            /* only PIP_SIGNIFICANCE_MAP (2) acts */
void avr_rec_end_coding_type(avr_record* r, int coding_type);
void avr_rec_bin(avr_record* r, int ctx_code, int symbol);  /* get: ctxIdx 0..1023; get_bypass: 1024; get_terminate: 1025 */
/* native front end: Annex-B H.264 bytes -> records, no libavcodec */
int avr_record_annexb(avr_record* r, const uint8_t* data, size_t n);
/* counts[0..5] = source bins, residual blocks, coded slices, slice spans (coded + skipped), pictures, out_of_domain flags */
void avr_record_counts(const avr_record* r, uint64_t counts[6]);
int avr_record_span(const avr_record* r, uint32_t i, uint64_t* payload_off, uint32_t* size, int32_t* slice_index,
                    uint32_t* flip, avr_slice_params* sp);

/* ---- compress: records -> range-coded payloads ------------------------------------------- */
#define AVR_MODE_PIPELINE 0   /* slice-parallel pipeline (product path) */
#define AVR_MODE_CHAIN 1      /* one sequential chain per file (on-device cross-check) */
int avr_cbatch_create(avr_context* ctx, const avr_record* const* recs, int nfiles, avr_cbatch** out);  /* H2D */
/* Same, from raw Annex-B files: the host only scans NALs/headers; the CABAC decode + slice_data() walk that
 * produces the records runs on the GPU (one warp per slice).  file_status[i] != AVR_OK marks a file the device
 * parse does not cover (unsupported syntax, record capacity exceeded): it is left out of the batch and the
 * caller decides (avr_compress_files falls back to the host recorder, then to a literal block). */
int avr_cbatch_create_files(avr_context* ctx, const uint8_t* const* files, const size_t* lens, int nfiles,
                            int host_threads, avr_cbatch** out, int* file_status);
/* spans (coded + skipped slices, in stream order) of file `file` of a batch made by avr_cbatch_create_files */
int avr_cbatch_span_count(const avr_cbatch* b, int file, uint32_t* nspans);
int avr_cbatch_span(const avr_cbatch* b, int file, uint32_t i, uint64_t* payload_off, uint32_t* size, int32_t* slice_index);
int avr_cbatch_run(avr_cbatch* b, int mode);                                                           /* kernels only */
int avr_cbatch_fetch(avr_cbatch* b);                                                                   /* D2H */
int avr_cbatch_payload(const avr_cbatch* b, int file, uint32_t slice, const uint8_t** data, uint32_t* len);
void avr_cbatch_timing(const avr_cbatch* b, avr_timing* t);
void avr_cbatch_destroy(avr_cbatch* b);

/* ---- decompress: payloads -> original CABAC slice bytes ---------------------------------- */
int avr_dbatch_create(avr_context* ctx, const avr_dslice* slices, const uint32_t* slices_per_file, int nfiles,
                      avr_dbatch** out);                                                               /* H2D */
int avr_dbatch_run(avr_dbatch* b);
int avr_dbatch_fetch(avr_dbatch* b);
/* slice = index into the `slices` array given at creation; data = re-encoded CABAC bytes BEFORE the
 * length_parity/last_byte patch of recode.cpp:1352-1362 */
int avr_dbatch_slice(const avr_dbatch* b, uint32_t slice, const uint8_t** data, uint32_t* len, uint32_t* bins);
void avr_dbatch_timing(const avr_dbatch* b, avr_timing* t);
void avr_dbatch_destroy(avr_dbatch* b);

/* ---- files: `recode compress` / `recode decompress` on memory buffers --------------------- */
typedef struct avr_buffer { uint8_t* data; size_t size; } avr_buffer;   /* library-owned; avr_buffer_free */
/* in: n Annex-B H.264 files; out: n .avrec containers (recode.proto wire format) */
int avr_compress_files(avr_context* ctx, const uint8_t* const* in, const size_t* in_len, int n, int host_threads,
                       avr_buffer* out, avr_timing* t);
int avr_decompress_files(avr_context* ctx, const uint8_t* const* in, const size_t* in_len, int n, int host_threads,
                         avr_buffer* out, avr_timing* t);
/* `out` is output-only: zeroed on entry, all-null when the call fails.  Internally the files are cut into sub-batches that run on
 * lanes (stream + host thread each; compress AVR_LANES, default 2; decompress AVR_DLANES, default 1) so that one batch's copies,
 * host framing and kernel tails overlap another batch's kernels. */
void avr_buffer_free(avr_buffer* b);
/* The same over several contexts, normally one per GPU of the box: sub-batches are spread over the contexts by
 * greedy longest-first on bytes; files share nothing (recode.cpp:1314,1597), so nothing moves between GPUs.  The outputs do
 * not depend on the number of contexts.  `recode compress|decompress|roundtrip --gpus N` uses these. */
int avr_compress_files_multi(avr_context* const* ctxs, int nctx, const uint8_t* const* in, const size_t* in_len, int n,
                             int host_threads, avr_buffer* out, avr_timing* t);
int avr_decompress_files_multi(avr_context* const* ctxs, int nctx, const uint8_t* const* in, const size_t* in_len, int n,
                               int host_threads, avr_buffer* out, avr_timing* t);

/* ---- streams: the reference's coder test on S independent bin streams (test/arithmetic_code.cpp:93-110) ---- */
/* bits: nstreams * nbins bytes, stream-major, one byte per bin (zero / non-zero).  digit_bits selects
 * arithmetic_code<uint64_t,uint16_t> (16, the test's coder) or <uint64_t,uint8_t> (8, recode.cpp:322-323).
 * model 0: p1 = range/2 (test/arithmetic_code.cpp:97); model 1: one adaptive estimator per stream,
 * p1 = (range/(pos+neg))*pos with the counts of recode.cpp:1037-1054 (limit 0x60).
 * coded = the streams' bytes back to back, coded_off[nstreams + 1] = where each begins (caller-allocated). */
#define AVR_STREAM_FIXED_HALF 0
#define AVR_STREAM_ADAPTIVE 1
int avr_streams_encode(avr_context* ctx, const uint8_t* bits, uint32_t nstreams, uint32_t nbins, int digit_bits, int model,
                       avr_buffer* coded, uint64_t* coded_off, avr_timing* t);
int avr_streams_decode(avr_context* ctx, const uint8_t* coded, const uint64_t* coded_off, uint32_t nstreams, uint32_t nbins,
                       int digit_bits, int model, uint8_t* bits_out, avr_timing* t);

/* ---- measurement aid: SM cycles per bin of the range decoder's bare recurrence on one thread (arithmetic_code.h:232-248 with
 * a constant estimator, nothing else); the denominator bench.py reports the chain kernel's cycles per bin against. */
int avr_measure_recurrence_floor(avr_context* ctx, uint32_t steps, double* cycles_per_bin);

/* ---- streams, no-allocation variant ------------------------------------------------------------------------ */
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  avr_streams_encode with
 * DEVICE buffers in and out and every intermediate buffer (per-stream scratch, lengths, status, offsets) inside one
 * file-scope `static __device__` array of AVR_STREAMS_STATIC_SCRATCH_BYTES that each GPU loading the library holds until
 * the process exits.  The call makes no cudaMalloc / cudaFree / host allocation and launches on the context's stream,
 * which it synchronises before returning.
 *   d_bits       device, nstreams * nbins bytes, stream-major, one byte per bin (zero / non-zero)
 *   d_coded      device, coded_cap >= nstreams * AVR_STREAMS_CODED_BOUND(nbins, model) bytes; the streams back to back
 *   d_coded_off  device, nstreams + 1 entries: where each stream begins, [nstreams] = total bytes
 * Limit: (nstreams * AVR_STREAMS_CODED_BOUND + 20 * nstreams + ~2 KB) <= AVR_STREAMS_STATIC_SCRATCH_BYTES, else AVR_E_ARG
 * with the sizes in avr_last_error() (use avr_streams_encode for larger batches).  All calls share the array, which
 * keeps the previous call's contents (the call clears what must start at zero): calls that overlap on the GPU -- two
 * contexts of one device, two host threads -- race and must be serialised by the caller. */
#define AVR_STREAMS_STATIC_SCRATCH_BYTES (8u << 20)
#define AVR_STREAMS_CODED_BOUND(nbins, model) \
  (((((model) == AVR_STREAM_FIXED_HALF) ? (uint64_t)(nbins) / 8 + 32 : (uint64_t)(nbins) + 32) + 15) & ~(uint64_t)15)
int avr_streams_encode_static(avr_context* ctx, const uint8_t* d_bits, uint32_t nstreams, uint32_t nbins, int digit_bits,
                              int model, uint8_t* d_coded, uint64_t coded_cap, uint64_t* d_coded_off);

#ifdef __cplusplus
}
#endif
#endif
