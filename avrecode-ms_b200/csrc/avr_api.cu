// C-ABI implementation (include/avrecode_b200.h): context, recorder bindings, staged
// compress / decompress batches.  Owns device memory and the stream; all compute is in
// avr_kernels.cu.  No host implementation of the hot path exists here.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <mutex>
#include <string>
#include <thread>
#include <algorithm>
#include <vector>
#include "../../include/avrecode_b200.h"
#include "avr_api_internal.h"
#include "avr_host_tables.h"
#include "avr_kernels.h"
#include "avr_record.h"

static thread_local std::string g_err;
void avr_set_error(const char* fmt, ...) {
  char buf[512];
  va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
  g_err = buf;
}
extern "C" const char* avr_last_error(void) { return g_err.c_str(); }
extern "C" int avr_version(void) { return 100; }

#define CU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { avr_set_error("%s: %s", #x, cudaGetErrorString(e_)); return AVR_E_CUDA; } } while (0)

// Device blocks released by a batch stay with the context and are handed to the next batch: cudaMalloc /
// cudaFree of multi-GB buffers cost 50-200 ms per batch, more than some of the kernels.  Everything runs on
// the context's one stream and batches synchronise before they release, so reuse needs no extra ordering.
// The lanes of one context (avr_context_lane) share ONE arena: a block released by a batch of one lane serves the next batch
// of any lane (separate caches doubled the footprint and ran a 180 GB device out of memory), hence the lock.
struct AvrArena {
  std::vector<std::pair<size_t, void*>> free_blocks;
  size_t cached = 0;
  std::mutex mu;
  void* take(size_t n, size_t* got) {
    std::lock_guard<std::mutex> g(mu);
    size_t best = (size_t)-1, bi = 0;
    for (size_t i = 0; i < free_blocks.size(); i++) {
      const size_t sz = free_blocks[i].first;
      if (sz >= n && sz <= 2 * n + (1u << 20) && sz < best) { best = sz; bi = i; }
    }
    if (best == (size_t)-1) return nullptr;
    void* p = free_blocks[bi].second;
    cached -= best; *got = best;
    free_blocks.erase(free_blocks.begin() + bi);
    return p;
  }
  void give(void* p, size_t n) { std::lock_guard<std::mutex> g(mu); free_blocks.push_back({n, p}); cached += n; }
  void flush() { std::lock_guard<std::mutex> g(mu); for (auto& b : free_blocks) cudaFree(b.second); free_blocks.clear(); cached = 0; }
};
// Page-locked host staging, cached the same way: D2H into pageable memory goes through the driver's own bounce
// buffers at a third of the bandwidth, and a fresh std::vector of the payload size is zero-filled first.
struct AvrPinned {
  std::vector<std::pair<size_t, void*>> free_blocks;
  void* take(size_t n, size_t* got) {
    for (size_t i = 0; i < free_blocks.size(); i++)
      if (free_blocks[i].first >= n) {
        void* p = free_blocks[i].second; *got = free_blocks[i].first;
        free_blocks.erase(free_blocks.begin() + i);
        return p;
      }
    for (auto& b : free_blocks) cudaFreeHost(b.second);      // too small: one growing buffer per role is enough
    free_blocks.clear();
    void* p = nullptr;
    n = (n + (n >> 2) + (1u << 20)) & ~(size_t)4095;
    if (cudaHostAlloc(&p, n, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    *got = n;
    return p;
  }
  void give(void* p, size_t n) { if (p) free_blocks.push_back({n, p}); }
  void flush() { for (auto& b : free_blocks) cudaFreeHost(b.second); free_blocks.clear(); }
};
struct avr_context {
  int device;
  cudaStream_t stream;
  AvrTables* d_tab;
  cudaEvent_t ev[4];
  AvrArena arena_own;
  AvrArena* arena = &arena_own;           // twins point at their parent's
  AvrPinned pinned;
  std::vector<avr_context*> twins;        // further lanes on the same device (own stream, events, caches): avr_context_lane
};

extern "C" int avr_context_create(int device, avr_context** out) {
  if (!out) return AVR_E_ARG;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { avr_set_error("no CUDA device: this library has no CPU path"); return AVR_E_CUDA; }
  if (device < 0 || device >= n) { avr_set_error("device %d out of range (%d devices)", device, n); return AVR_E_ARG; }
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) { avr_set_error("device %d is sm_%d%d; kernels are built for sm_100a only", device, prop.major, prop.minor); return AVR_E_CUDA; }
  CU(cudaSetDevice(device));
  avr_context* c = new avr_context();
  c->device = device; c->stream = nullptr; c->d_tab = nullptr;
  for (auto& e : c->ev) e = nullptr;
  AvrTables t;
  avr_fill_tables(&t);
  cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
  for (auto& ev : c->ev) if (e == cudaSuccess) e = cudaEventCreate(&ev);
  if (e == cudaSuccess) e = cudaMalloc(&c->d_tab, sizeof t);
  if (e == cudaSuccess) e = cudaMemcpy(c->d_tab, &t, sizeof t, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {                                  // nothing half-built is left behind
    avr_set_error("context on device %d: %s", device, cudaGetErrorString(e));
    avr_context_destroy(c);
    return AVR_E_CUDA;
  }
  *out = c;
  return AVR_OK;
}
extern "C" void avr_context_destroy(avr_context* c) {
  if (!c) return;
  for (avr_context* t : c->twins) avr_context_destroy(t);
  c->twins.clear();
  cudaSetDevice(c->device);
  c->arena_own.flush(); c->pinned.flush();
  if (c->d_tab) cudaFree(c->d_tab);
  for (auto& e : c->ev) if (e) cudaEventDestroy(e);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}
extern "C" void* avr_context_stream(avr_context* c) { return c ? (void*)c->stream : nullptr; }
int avr_context_device(const avr_context* c) { return c ? c->device : -1; }
// Lane k of a context: k = 0 is the context itself, k >= 1 a twin on the same device with its own stream, events, device-block
// and pinned caches, created on first use and destroyed with the context.  The file-level calls run sub-batches on several
// lanes so that one batch's copies and host work overlap another batch's kernels (avr_files.cpp).  Call from one thread.
avr_context* avr_context_lane(avr_context* c, int k) {
  if (!c || k < 0) return nullptr;
  if (k == 0) return c;
  while ((int)c->twins.size() < k) {
    avr_context* t = nullptr;
    if (avr_context_create(c->device, &t) != AVR_OK) return nullptr;
    t->arena = c->arena;
    c->twins.push_back(t);
  }
  return c->twins[(size_t)k - 1];
}

// ---- recorder bindings ------------------------------------------------------------------------
struct avr_record {
  avr::FileRecord rec;
  avr::Recorder recorder;
  avr_record() : recorder(&rec) {}
};
const avr::FileRecord* avr_record_view(const avr_record* r) { return &r->rec; }
avr::FileRecord* avr_record_mut(avr_record* r) { return &r->rec; }

static avr::SliceParams to_sp(const avr_slice_params* p) {
  avr::SliceParams s;
  s.slice_type = p->slice_type; s.first_mb = p->first_mb; s.mb_w = p->mb_w; s.mb_h = p->mb_h;
  s.num_ref[0] = p->num_ref[0]; s.num_ref[1] = p->num_ref[1]; s.cabac_init_idc = p->cabac_init_idc;
  s.slice_qp = p->slice_qp; s.transform_8x8_mode = p->transform_8x8_mode; s.chroma_format_idc = p->chroma_format_idc;
  s.direct_8x8_inference = p->direct_8x8_inference;
  return s;
}
static void from_sp(const avr::SliceParams& s, avr_slice_params* p) {
  p->slice_type = s.slice_type; p->first_mb = s.first_mb; p->mb_w = s.mb_w; p->mb_h = s.mb_h;
  p->num_ref[0] = s.num_ref[0]; p->num_ref[1] = s.num_ref[1]; p->cabac_init_idc = s.cabac_init_idc;
  p->slice_qp = s.slice_qp; p->transform_8x8_mode = s.transform_8x8_mode; p->chroma_format_idc = s.chroma_format_idc;
  p->direct_8x8_inference = s.direct_8x8_inference;
}

extern "C" avr_record* avr_record_new(void) { return new avr_record(); }
extern "C" void avr_record_free(avr_record* r) { delete r; }
extern "C" void avr_rec_frame_spec(avr_record* r, int fn, int w, int h) { r->recorder.frame_spec(fn, w, h); }
extern "C" void avr_rec_slice_begin(avr_record* r, int coded, uint64_t off, uint32_t size, const avr_slice_params* sp) {
  r->recorder.slice_begin(coded != 0, off, size, to_sp(sp));
}
extern "C" void avr_rec_slice_end(avr_record* r) { r->recorder.slice_end(); }
extern "C" void avr_rec_mb_xy(avr_record* r, int x, int y) { r->recorder.mb_xy(x, y); }
extern "C" void avr_rec_begin_sub_mb(avr_record* r, int cat, int n, int mc, int dc, int c422) { r->recorder.begin_sub_mb(cat, n, mc, dc, c422); }
extern "C" void avr_rec_end_sub_mb(avr_record*) {}
extern "C" void avr_rec_begin_coding_type(avr_record* r, int ct) { if (ct == 2) r->recorder.begin_sigmap(); }
extern "C" void avr_rec_end_coding_type(avr_record* r, int ct) { if (ct == 2) r->recorder.end_sigmap(); }
extern "C" void avr_rec_bin(avr_record* r, int ctx, int sym) { r->recorder.bin(ctx, sym); }
extern "C" int avr_record_annexb(avr_record* r, const uint8_t* data, size_t n) {
  if (!r || (!data && n)) return AVR_E_ARG;
  return avr::record_annexb(data, n, &r->rec);
}
extern "C" void avr_record_counts(const avr_record* r, uint64_t c[6]) {
  c[0] = r->rec.source_bins; c[1] = r->rec.blocks.size(); c[2] = r->rec.slices.size(); c[3] = r->rec.spans.size();
  c[4] = r->rec.frames.size(); c[5] = (uint64_t)r->rec.out_of_domain;
}
uint32_t avr_span_flip(const avr::FileRecord& rec, uint32_t i) {
  // pictures started since the previous CODED slice of the file
  const avr::SliceSpan& sp = rec.spans[i];
  int64_t last = -1;
  for (uint32_t k = i; k-- > 0;) if (rec.spans[k].slice_index >= 0) { last = rec.spans[k].frame; break; }
  if (last == (int64_t)sp.frame) return 0;
  int64_t k = (int64_t)sp.frame - last;
  return (k == 1 && rec.frames[sp.frame].prev_valid) ? 1u : 2u;
}
extern "C" int avr_record_span(const avr_record* r, uint32_t i, uint64_t* off, uint32_t* size, int32_t* slice_index,
                               uint32_t* flip, avr_slice_params* sp) {
  if (!r || i >= r->rec.spans.size()) return AVR_E_ARG;
  const avr::SliceSpan& s = r->rec.spans[i];
  if (off) *off = s.payload_off;
  if (size) *size = s.size;
  if (slice_index) *slice_index = s.slice_index;
  if (flip) *flip = avr_span_flip(r->rec, i);
  if (sp) from_sp(s.sp, sp);
  return AVR_OK;
}

// ---- device memory helper ---------------------------------------------------------------------
struct DevMem {
  std::vector<std::pair<size_t, void*>> ptrs;
  size_t bytes = 0;
  AvrArena* arena = nullptr;             // the owning context's block cache (set by the batch constructors)
  template <class T> int alloc(T** p, size_t count) {
    *p = nullptr;
    size_t n = (count ? count : 1) * sizeof(T) + 64;
    n = n >= (1u << 20) ? (n + (2u << 20) - 1) & ~(size_t)((2u << 20) - 1) : (n + 511) & ~(size_t)511;
    size_t got = n;
    void* q = arena ? arena->take(n, &got) : nullptr;
    if (!q) {
      cudaError_t e = cudaMalloc(&q, n);
      if (e != cudaSuccess && arena && arena->cached) { cudaGetLastError(); arena->flush(); e = cudaMalloc(&q, n); }
      if (e != cudaSuccess) { cudaGetLastError(); avr_set_error("cudaMalloc(%zu): %s", n, cudaGetErrorString(e)); return AVR_E_NOMEM; }
    }
    ptrs.push_back({got, q}); bytes += got; *p = (T*)q;
    return AVR_OK;
  }
  void release() {
    for (auto& b : ptrs) { if (arena) arena->give(b.second, b.first); else cudaFree(b.second); }
    ptrs.clear(); bytes = 0;
  }
};
#define TRY(x) do { int r_ = (x); if (r_ != AVR_OK) return r_; } while (0)

static float ev_ms(cudaEvent_t a, cudaEvent_t b) { float ms = 0; cudaEventElapsedTime(&ms, a, b); return ms; }

// ---- compress batch ---------------------------------------------------------------------------
// Pass D is serial per slice.  A slice of up to this many bins shares a warp with 31 others of similar length (one slice per
// lane, in lockstep: ~15x fewer issue slots per symbol); a longer one keeps a warp to itself, because a lockstep step costs
// ~4x the latency of a lone lane's step and the kernel lasts as long as its slowest warp (profiles/r02_summary.md).
// Threshold: a packed warp should not outlast the longest lone slice of the batch.  Measured on the 1024-clip bench corpus
// (longest slice 0.9 Mbins; pass D ms at thresholds 0 / 150k / 400k / 500k / 800k / all packed: 87 / 76 / 68 / 85 / 134 / 167), i.e. the
// optimum sits at ~0.45 x the longest slice; AVR_CODE_LANE_MAX overrides it.
static uint32_t code_long_of(const std::vector<AvrBatchSlice>& bsl, const std::vector<uint32_t>& order) {
  if (order.empty()) return 0;
  static const long env = [] { const char* e = getenv("AVR_CODE_LANE_MAX"); return e && *e ? atol(e) : -1L; }();
  const uint32_t longest = bsl[order[0]].bin_end - bsl[order[0]].bin_start;
  const uint32_t lane_max = env >= 0 ? (uint32_t)env : (uint32_t)(0.45 * longest);
  uint32_t n = 0;
  while (n < order.size() && bsl[order[n]].bin_end - bsl[order[n]].bin_start > lane_max) n++;
  return n;
}
static std::vector<uint32_t> code_order_of(const std::vector<AvrBatchSlice>& bsl) {
  std::vector<uint32_t> o(bsl.size());
  for (uint32_t i = 0; i < o.size(); i++) o[i] = i;
  std::stable_sort(o.begin(), o.end(), [&](uint32_t a, uint32_t b) { return bsl[a].bin_end - bsl[a].bin_start > bsl[b].bin_end - bsl[b].bin_start; });
  return o;
}
struct avr_cbatch {
  avr_context* ctx;
  DevMem mem;
  AvrDevBatch d;
  std::vector<uint32_t> file_slice_off;    // per file: first batch-global slice
  std::vector<uint32_t> h_len; std::vector<int> h_status; std::vector<uint64_t> h_off;
  uint8_t* h_out = nullptr; size_t h_out_cap = 0;   // page-locked, from the context's cache
  uint64_t scratch_bytes = 0, pic_mbs = 0, chain_mbs = 0;
  bool have_pipe = false, have_chain = false;
  uint32_t create_launches = 0;           // kernels launched by avr_cbatch_create_files (parse, block gather)
  avr_timing t;
  bool ran = false, fetched = false;
  // device-side mutable pointers kept for fetch
  uint64_t* d_scratch_off = nullptr; uint32_t* d_scratch_cap = nullptr;
};

extern "C" int avr_cbatch_create(avr_context* ctx, const avr_record* const* recs, int nfiles, avr_cbatch** out) {
  if (!ctx || !recs || nfiles < 0 || !out) return AVR_E_ARG;
  if (nfiles > AVR_MAX_BATCH_FILES) { avr_set_error("at most %d files per batch (sort key has 13 file bits)", AVR_MAX_BATCH_FILES); return AVR_E_ARG; }
  CU(cudaSetDevice(ctx->device));
  avr_cbatch* b = new avr_cbatch();
  b->ctx = ctx; b->mem.arena = ctx->arena;
  memset(&b->d, 0, sizeof b->d);
  memset(&b->t, 0, sizeof b->t);
  std::vector<AvrFileDesc> fds((size_t)nfiles);
  std::vector<AvrBatchSlice> bsl;
  std::vector<AvrSlice> sl;
  std::vector<uint64_t> soff; std::vector<uint32_t> scap;
  uint64_t nbins = 0, nblocks = 0, nframes = 0, emitted = 0, chain_mbs = 0, pic_mbs = 0, scratch = 0;
  for (int f = 0; f < nfiles; f++) {
    const avr::FileRecord& r = recs[f]->rec;
    if (r.error || (r.out_of_domain & 1)) { avr_set_error("file %d: record not compressible (error %d, out_of_domain %d)", f, r.error, r.out_of_domain); delete b; return AVR_E_UNSUPPORTED; }
    AvrFileDesc& fd = fds[f];
    memset(&fd, 0, sizeof fd);
    fd.bin_off = nbins; fd.nbins = (uint32_t)r.bins.size();
    fd.block_off = (uint32_t)nblocks; fd.nblocks = (uint32_t)r.blocks.size();
    fd.slice_off = (uint32_t)bsl.size(); fd.nslices = (uint32_t)r.slices.size();
    fd.frame_off = (uint32_t)nframes; fd.nframes = (uint32_t)r.frames.size();
    fd.pic_mb_base = (uint32_t)pic_mbs;
    for (auto& fr : r.frames) { uint32_t m = (uint32_t)fr.mb_w * fr.mb_h; if (m > fd.max_mbs) fd.max_mbs = m; pic_mbs += m; }
    fd.mb_base = (uint32_t)chain_mbs;
    chain_mbs += 2ull * fd.max_mbs;
    b->file_slice_off.push_back((uint32_t)bsl.size());
    for (auto& s : r.slices) {
      // every coded slice must end with end_of_slice_flag = 1 (recode.cpp:1099-1102)
      if (s.bin_end <= s.bin_start || AVR_BIN_CTX(r.bins[s.bin_end - 1]) != AVR_CTX_TERMINATE || !AVR_BIN_SYM(r.bins[s.bin_end - 1])) {
        avr_set_error("file %d: slice without terminating bin", f); delete b; return AVR_E_CORRUPT;
      }
      AvrBatchSlice bs = {(uint32_t)f, s.bin_start, s.bin_end, s.block_start + fd.block_off, s.block_end + fd.block_off};
      bsl.push_back(bs); sl.push_back(s);
      uint32_t cap = (s.bin_end - s.bin_start) + 6 * (s.block_end - s.block_start) + 16;
      soff.push_back(scratch); scap.push_back(cap); scratch += (cap + 15) & ~15ull;
    }
    nbins += r.bins.size(); nblocks += r.blocks.size(); nframes += r.frames.size(); emitted += r.emitted;
    b->t.bins += r.source_bins;
  }
  if (emitted >= (1ull << 31) || nblocks + bsl.size() >= (1ull << 31) || pic_mbs >= (1ull << 32) || chain_mbs >= (1ull << 32)) {
    avr_set_error("batch too large for one launch (%llu coded symbols); split it", (unsigned long long)emitted); delete b; return AVR_E_ARG;
  }
  b->file_slice_off.push_back((uint32_t)bsl.size());
  b->t.coded_symbols = emitted;
  b->scratch_bytes = scratch;
  b->pic_mbs = pic_mbs; b->chain_mbs = chain_mbs;
  AvrDevBatch& d = b->d;
  d.nfiles = nfiles; d.nslices = (uint32_t)bsl.size(); d.nblocks = (uint32_t)nblocks; d.nbins = nbins;
  d.emit_cap = (uint32_t)emitted; d.tab = ctx->d_tab;
  int fb = 0; while ((1 << fb) < nfiles) fb++;
  d.sort_end_bit = AVR_KEY_BITS + 1 + fb;
  // allocate
  AvrFileDesc* d_fds; uint16_t* d_bins; AvrBlock* d_blocks; AvrSlice* d_sl; AvrBatchSlice* d_bsl; AvrFrame* d_frames;
  DevMem& m = b->mem;
  int rc = AVR_OK;
  auto fail = [&](int r) { m.release(); delete b; return r; };
  if ((rc = m.alloc(&d_fds, fds.size())) || (rc = m.alloc(&d_bins, nbins)) || (rc = m.alloc(&d_blocks, nblocks)) ||
      (rc = m.alloc(&d_sl, sl.size())) || (rc = m.alloc(&d_bsl, bsl.size())) || (rc = m.alloc(&d_frames, nframes)) ||
      (rc = m.alloc(&b->d_scratch_off, soff.size())) || (rc = m.alloc(&b->d_scratch_cap, scap.size())) ||
      (rc = m.alloc(&d.scratch, scratch)) || (rc = m.alloc(&d.out_len, bsl.size())) || (rc = m.alloc(&d.status, bsl.size())) ||
      (rc = m.alloc(&d.out_off, bsl.size())) || (rc = m.alloc(&d.out_total, 1)) || (rc = m.alloc(&d.out, scratch)))
    return fail(rc);
  d.fds = d_fds; d.bins = d_bins; d.blocks = d_blocks; d.slices = d_sl; d.bslices = d_bsl; d.frames = d_frames;
  uint32_t* d_order; if ((rc = m.alloc(&d_order, bsl.size()))) return fail(rc);
  d.code_order = d_order;
  const std::vector<uint32_t> order = code_order_of(bsl);
  d.code_long = code_long_of(bsl, order);
  d.scratch_off = b->d_scratch_off; d.scratch_cap = b->d_scratch_cap;
  // upload (timed: this is the H2D leg of the e2e number)
  cudaStream_t st = ctx->stream;
  cudaEventRecord(ctx->ev[0], st);
  uint64_t h2d = 0;
  auto up = [&](void* dst, const void* src, size_t n) { if (n) cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, st); h2d += n; };
  up(d_fds, fds.data(), fds.size() * sizeof(AvrFileDesc));
  for (int f = 0; f < nfiles; f++) {
    const avr::FileRecord& r = recs[f]->rec;
    up(d_bins + fds[f].bin_off, r.bins.data(), r.bins.size() * sizeof(uint16_t));
    up(d_blocks + fds[f].block_off, r.blocks.data(), r.blocks.size() * sizeof(AvrBlock));
    up(d_frames + fds[f].frame_off, r.frames.data(), r.frames.size() * sizeof(AvrFrame));
  }
  up(d_sl, sl.data(), sl.size() * sizeof(AvrSlice));
  up(d_bsl, bsl.data(), bsl.size() * sizeof(AvrBatchSlice));
  up(d_order, order.data(), order.size() * sizeof(uint32_t));
  up(b->d_scratch_off, soff.data(), soff.size() * sizeof(uint64_t));
  up(b->d_scratch_cap, scap.data(), scap.size() * sizeof(uint32_t));
  cudaEventRecord(ctx->ev[1], st);
  cudaError_t e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) { avr_set_error("upload: %s", cudaGetErrorString(e)); return fail(AVR_E_CUDA); }
  b->t.h2d_ms = ev_ms(ctx->ev[0], ctx->ev[1]);
  b->t.h2d_bytes = h2d;
  b->h_off = soff;   // reused as host copy of scratch offsets until fetch overwrites
  *out = b;
  return AVR_OK;
}

extern "C" int avr_cbatch_run(avr_cbatch* b, int mode) {
  if (!b) return AVR_E_ARG;
  avr_context* ctx = b->ctx;
  CU(cudaSetDevice(ctx->device));
  AvrDevBatch& d = b->d;
  DevMem& m = b->mem;
  cudaStream_t st = ctx->stream;
  if (mode == AVR_MODE_PIPELINE && !b->have_pipe) {
    const uint32_t n = d.emit_cap, nent = d.nblocks + d.nslices;
    d.pic_store_bytes = (size_t)b->pic_mbs * AVR_MB_STRIDE + 64;
    d.cub_tmp_bytes = avr_pipeline_tmp_bytes(n, nent, d.sort_end_bit);
    uint8_t* tmp;
    TRY(m.alloc(&d.pic_store, d.pic_store_bytes)); TRY(m.alloc(&d.seg, (size_t)nent + 1));
    TRY(m.alloc(&d.keys_a, n)); TRY(m.alloc(&d.keys_b, n)); TRY(m.alloc(&d.idx_a, n)); TRY(m.alloc(&d.idx_b, n));
    TRY(m.alloc(&d.head, n)); TRY(m.alloc(&d.psum, n)); TRY(m.alloc(&d.fn, n)); TRY(m.alloc(&d.prob, n));
    TRY(m.alloc(&tmp, d.cub_tmp_bytes)); d.cub_tmp = tmp;
    b->have_pipe = true;
  }
  if (mode == AVR_MODE_CHAIN && !d.slices) { avr_set_error("chain mode needs a batch made from host records"); return AVR_E_ARG; }
  if (mode == AVR_MODE_CHAIN && !b->have_chain) {
    TRY(m.alloc(&d.est_pool, (size_t)d.nfiles * AVR_KEY_COUNT));
    TRY(m.alloc(&d.store_pool, (size_t)b->chain_mbs * AVR_MB_STRIDE + 64));
    b->have_chain = true;
  }
  int launches = 0, rc;
  cudaEventRecord(ctx->ev[0], st);
  if (mode == AVR_MODE_CHAIN) {
    CU(cudaMemsetAsync(d.est_pool, 0x01, (size_t)d.nfiles * AVR_KEY_COUNT * sizeof(uint16_t), st));   // pos = neg = 1
    cudaEventRecord(ctx->ev[2], st);
    rc = avr_launch_compress_chain(d, st); launches = 1;
    cudaEventRecord(ctx->ev[3], st);
  } else if (mode == AVR_MODE_PIPELINE) {
    rc = avr_launch_compress_pipeline(d, st, &launches, ctx->ev[2], ctx->ev[3]);
  } else return AVR_E_ARG;
  if (rc) { avr_set_error("compress launch: %s", cudaGetErrorString((cudaError_t)rc)); return AVR_E_CUDA; }
  rc = avr_launch_compact(d.scratch, d.scratch_off, d.out_len, d.out_off, d.out_total, d.out, d.nslices, st); launches += 2;
  if (rc) { avr_set_error("compact launch: %s", cudaGetErrorString((cudaError_t)rc)); return AVR_E_CUDA; }
  cudaEventRecord(ctx->ev[1], st);
  cudaError_t e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) { avr_set_error("compress kernels: %s", cudaGetErrorString(e)); return AVR_E_CUDA; }
  b->t.kernel_ms = ev_ms(ctx->ev[0], ctx->ev[1]);
  b->t.coder_launch_ms_x1000 = (uint32_t)(ev_ms(ctx->ev[2], ctx->ev[3]) * 1000.0f);
  b->t.launches = (uint32_t)launches + b->create_launches;
  b->ran = true; b->fetched = false;
  return AVR_OK;
}

extern "C" int avr_cbatch_fetch(avr_cbatch* b) {
  if (!b || !b->ran) return AVR_E_ARG;
  avr_context* ctx = b->ctx;
  CU(cudaSetDevice(ctx->device));
  AvrDevBatch& d = b->d;
  cudaStream_t st = ctx->stream;
  const size_t ns = d.nslices;
  b->h_len.resize(ns); b->h_status.resize(ns); b->h_off.resize(ns);
  uint64_t total = 0;
  cudaEventRecord(ctx->ev[0], st);
  if (ns) {
    cudaMemcpyAsync(b->h_len.data(), d.out_len, ns * sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(b->h_status.data(), d.status, ns * sizeof(int), cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(b->h_off.data(), d.out_off, ns * sizeof(uint64_t), cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(&total, d.out_total, sizeof total, cudaMemcpyDeviceToHost, st);
    CU(cudaStreamSynchronize(st));
    if (total > b->scratch_bytes) { avr_set_error("payload total %llu exceeds scratch", (unsigned long long)total); return AVR_E_OVERFLOW; }
    if (total > b->h_out_cap) {
      ctx->pinned.give(b->h_out, b->h_out_cap);
      b->h_out = (uint8_t*)ctx->pinned.take(total, &b->h_out_cap);
      if (!b->h_out) { b->h_out_cap = 0; avr_set_error("cudaHostAlloc(%llu) failed", (unsigned long long)total); return AVR_E_NOMEM; }
    }
    if (total) cudaMemcpyAsync(b->h_out, d.out, total, cudaMemcpyDeviceToHost, st);
  }
  cudaEventRecord(ctx->ev[1], st);
  CU(cudaStreamSynchronize(st));
  b->t.d2h_ms = ev_ms(ctx->ev[0], ctx->ev[1]);
  b->t.d2h_bytes = total + ns * 16 + 8;
  for (size_t s = 0; s < ns; s++)
    if (b->h_status[s]) { avr_set_error("slice %zu: device status %d", s, b->h_status[s]); return b->h_status[s]; }
  b->fetched = true;
  return AVR_OK;
}

extern "C" int avr_cbatch_payload(const avr_cbatch* b, int file, uint32_t slice, const uint8_t** data, uint32_t* len) {
  if (!b || !b->fetched || file < 0 || file >= b->d.nfiles) return AVR_E_ARG;
  uint32_t s = b->file_slice_off[file] + slice;
  if (s >= b->file_slice_off[file + 1]) return AVR_E_ARG;
  *data = b->h_out + b->h_off[s];
  *len = b->h_len[s];
  return AVR_OK;
}
extern "C" void avr_cbatch_timing(const avr_cbatch* b, avr_timing* t) { if (b && t) *t = b->t; }
void avr_cbatch_files_release(avr_cbatch* b);
extern "C" void avr_cbatch_destroy(avr_cbatch* b) {
  if (!b) return;
  avr_cbatch_files_release(b);
  cudaSetDevice(b->ctx->device);
  b->ctx->pinned.give(b->h_out, b->h_out_cap);
  b->mem.release();
  delete b;
}

// ---- decompress batch -------------------------------------------------------------------------
struct avr_dbatch {
  avr_context* ctx;
  DevMem mem;
  AvrDevDBatch d;
  std::vector<uint64_t> out_off; std::vector<uint32_t> out_cap;
  std::vector<uint32_t> h_len, h_bins; std::vector<int> h_status;
  uint8_t* h_out = nullptr; size_t h_out_cap = 0;   // page-locked, from the context's cache
  uint64_t out_bytes = 0;
  avr_timing t;
  bool ran = false, fetched = false;
};

extern "C" int avr_dbatch_create(avr_context* ctx, const avr_dslice* slices, const uint32_t* per_file, int nfiles, avr_dbatch** out) {
  if (!ctx || (!slices && nfiles) || !per_file || nfiles < 0 || !out) return AVR_E_ARG;
  CU(cudaSetDevice(ctx->device));
  avr_dbatch* b = new avr_dbatch();
  b->ctx = ctx; b->mem.arena = ctx->arena;
  memset(&b->d, 0, sizeof b->d);
  memset(&b->t, 0, sizeof b->t);
  std::vector<AvrFileDesc> fds((size_t)nfiles);
  std::vector<AvrDecSlice> ds;
  uint64_t pay = 0, outb = 0, chain_mbs = 0;
  uint32_t s = 0;
  int has_b = 0;
  for (int f = 0; f < nfiles; f++) {
    AvrFileDesc& fd = fds[f];
    memset(&fd, 0, sizeof fd);
    fd.slice_off = s; fd.nslices = per_file[f];
    for (uint32_t k = 0; k < per_file[f]; k++, s++) {
      const avr_dslice& in = slices[s];
      if (in.sp.mb_w <= 0 || in.sp.mb_h <= 0 || in.sp.mb_w > 1024 || in.sp.mb_h > 1024 || (!in.payload && in.payload_len)) {
        avr_set_error("slice %u: bad parameters", s); delete b; return AVR_E_ARG;
      }
      if (k == 0 && in.flip == 0) { avr_set_error("slice %u: first slice of a file must start a picture", s); delete b; return AVR_E_ARG; }
      AvrDecSlice d;
      memset(&d, 0, sizeof d);
      d.payload_off = pay; d.payload_len = in.payload_len; pay += ((uint64_t)in.payload_len + 15) & ~15ull;
      d.out_off = outb; d.out_cap = in.size + 64; outb += ((uint64_t)d.out_cap + 15) & ~15ull;
      d.flip = in.flip; d.sp = to_sp(&in.sp);
      if (d.sp.slice_type == avr::SLICE_B) has_b = 1;
      uint32_t mbs = (uint32_t)in.sp.mb_w * (uint32_t)in.sp.mb_h;
      if (mbs > fd.max_mbs) fd.max_mbs = mbs;
      ds.push_back(d);
      b->out_off.push_back(d.out_off); b->out_cap.push_back(d.out_cap);
    }
    fd.mb_base = (uint32_t)chain_mbs;
    chain_mbs += 2ull * (fd.max_mbs > 1024 ? fd.max_mbs : 1024);   // also hosts the file's MbInfo row (mb_w <= 1024)
  }
  if (chain_mbs >= (1ull << 32)) { avr_set_error("batch too large"); delete b; return AVR_E_ARG; }
  b->out_bytes = outb;
  AvrDevDBatch& d = b->d;
  DevMem& m = b->mem;
  d.nfiles = nfiles; d.nslices = (uint32_t)ds.size(); d.tab = ctx->d_tab; d.has_b = has_b;
  avr_trace(nullptr);
  AvrFileDesc* d_fds; AvrDecSlice* d_ds; uint8_t* d_pay;
  int rc;
  auto fail = [&](int r) { m.release(); delete b; return r; };
  if ((rc = m.alloc(&d_fds, fds.size())) || (rc = m.alloc(&d_ds, ds.size())) || (rc = m.alloc(&d_pay, pay + 16)) ||
      (rc = m.alloc(&d.est_pool, (size_t)nfiles * AVR_KEY_COUNT)) || (rc = m.alloc(&d.store_pool, (size_t)chain_mbs * AVR_MB_STRIDE)) ||
      (rc = m.alloc(&d.row_pool, (size_t)chain_mbs)) || (rc = m.alloc(&d.out, outb)) || (rc = m.alloc(&d.out_len, ds.size())) ||
      (rc = m.alloc(&d.out_bins, ds.size())) || (rc = m.alloc(&d.status, ds.size())))
    return fail(rc);
  avr_trace("dc: alloc");
  d.fds = d_fds; d.slices = d_ds; d.payloads = d_pay;
  uint32_t* d_forder; if ((rc = m.alloc(&d_forder, (size_t)nfiles))) return fail(rc);
  d.file_order = d_forder;
  std::vector<uint32_t> forder((size_t)nfiles);
  {
    std::vector<uint64_t> fbytes((size_t)nfiles, 0);
    for (int f = 0; f < nfiles; f++) for (uint32_t k = 0; k < fds[f].nslices; k++) fbytes[f] += ds[fds[f].slice_off + k].payload_len;
    for (int f = 0; f < nfiles; f++) forder[f] = (uint32_t)f;
    std::stable_sort(forder.begin(), forder.end(), [&](uint32_t a, uint32_t b) { return fbytes[a] > fbytes[b]; });
  }
  cudaStream_t st = ctx->stream;
  cudaEventRecord(ctx->ev[0], st);
  uint64_t h2d = 0;
  auto up = [&](void* dst, const void* src, size_t n) { if (n) cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, st); h2d += n; };
  up(d_fds, fds.data(), fds.size() * sizeof(AvrFileDesc));
  up(d_ds, ds.data(), ds.size() * sizeof(AvrDecSlice));
  up(d_forder, forder.data(), forder.size() * sizeof(uint32_t));
  {
    // one page-locked staging buffer and one copy instead of a pageable copy per slice
    size_t cap = 0;
    uint8_t* stage = pay ? (uint8_t*)ctx->pinned.take(pay, &cap) : nullptr;
    if (stage) {
      const int nth = pay > (4u << 20) ? 8 : 1;
      std::vector<std::thread> th;
      for (int t = 0; t < nth; t++) th.emplace_back([&, t] {
        for (size_t i = (size_t)t; i < ds.size(); i += (size_t)nth) if (slices[i].payload_len) memcpy(stage + ds[i].payload_off, slices[i].payload, slices[i].payload_len);
      });
      for (auto& t : th) t.join();
      up(d_pay, stage, pay);
      cudaStreamSynchronize(st);
      ctx->pinned.give(stage, cap);
    } else {
      for (size_t i = 0; i < ds.size(); i++) up(d_pay + ds[i].payload_off, slices[i].payload, slices[i].payload_len);
    }
  }
  cudaEventRecord(ctx->ev[1], st);
  cudaError_t e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) { avr_set_error("upload: %s", cudaGetErrorString(e)); return fail(AVR_E_CUDA); }
  avr_trace("dc: upload");
  b->t.h2d_ms = ev_ms(ctx->ev[0], ctx->ev[1]);
  b->t.h2d_bytes = h2d;
  *out = b;
  return AVR_OK;
}

extern "C" int avr_dbatch_run(avr_dbatch* b) {
  if (!b) return AVR_E_ARG;
  avr_context* ctx = b->ctx;
  CU(cudaSetDevice(ctx->device));
  AvrDevDBatch& d = b->d;
  cudaStream_t st = ctx->stream;
  cudaEventRecord(ctx->ev[0], st);
  CU(cudaMemsetAsync(d.est_pool, 0x01, (size_t)d.nfiles * AVR_KEY_COUNT * sizeof(uint16_t), st));
  cudaEventRecord(ctx->ev[2], st);
  int rc = avr_launch_decompress_chain(d, st);
  cudaEventRecord(ctx->ev[3], st);
  if (rc) { avr_set_error("decompress launch: %s", cudaGetErrorString((cudaError_t)rc)); return AVR_E_CUDA; }
  cudaEventRecord(ctx->ev[1], st);
  cudaError_t e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) { avr_set_error("decompress kernel: %s", cudaGetErrorString(e)); return AVR_E_CUDA; }
  b->t.kernel_ms = ev_ms(ctx->ev[0], ctx->ev[1]);
  b->t.coder_launch_ms_x1000 = (uint32_t)(ev_ms(ctx->ev[2], ctx->ev[3]) * 1000.0f);
  b->t.launches = 1;
  b->ran = true; b->fetched = false;
  return AVR_OK;
}

extern "C" int avr_dbatch_fetch(avr_dbatch* b) {
  if (!b || !b->ran) return AVR_E_ARG;
  avr_context* ctx = b->ctx;
  CU(cudaSetDevice(ctx->device));
  AvrDevDBatch& d = b->d;
  cudaStream_t st = ctx->stream;
  const size_t ns = d.nslices;
  b->h_len.resize(ns); b->h_bins.resize(ns); b->h_status.resize(ns);
  if (b->out_bytes > b->h_out_cap) {
    ctx->pinned.give(b->h_out, b->h_out_cap);
    b->h_out = (uint8_t*)ctx->pinned.take(b->out_bytes, &b->h_out_cap);
    if (!b->h_out) { b->h_out_cap = 0; avr_set_error("cudaHostAlloc(%llu) failed", (unsigned long long)b->out_bytes); return AVR_E_NOMEM; }
  }
  cudaEventRecord(ctx->ev[0], st);
  if (ns) {
    cudaMemcpyAsync(b->h_len.data(), d.out_len, ns * sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(b->h_bins.data(), d.out_bins, ns * sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(b->h_status.data(), d.status, ns * sizeof(int), cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(b->h_out, d.out, b->out_bytes, cudaMemcpyDeviceToHost, st);
  }
  cudaEventRecord(ctx->ev[1], st);
  CU(cudaStreamSynchronize(st));
  b->t.d2h_ms = ev_ms(ctx->ev[0], ctx->ev[1]);
  b->t.d2h_bytes = b->out_bytes + ns * 12;
  b->t.bins = 0;
  for (size_t s = 0; s < ns; s++) {
    b->t.bins += b->h_bins[s];
    if (b->h_status[s]) { avr_set_error("slice %zu: device status %d", s, b->h_status[s]); return b->h_status[s]; }
  }
  b->fetched = true;
  return AVR_OK;
}

extern "C" int avr_dbatch_slice(const avr_dbatch* b, uint32_t slice, const uint8_t** data, uint32_t* len, uint32_t* bins) {
  if (!b || !b->fetched || slice >= b->d.nslices) return AVR_E_ARG;
  if (data) *data = b->h_out + b->out_off[slice];
  if (len) *len = b->h_len[slice];
  if (bins) *bins = b->h_bins[slice];
  return AVR_OK;
}
extern "C" void avr_dbatch_timing(const avr_dbatch* b, avr_timing* t) { if (b && t) *t = b->t; }
extern "C" void avr_dbatch_destroy(avr_dbatch* b) {
  if (!b) return;
  cudaSetDevice(b->ctx->device);
  b->ctx->pinned.give(b->h_out, b->h_out_cap);
  b->mem.release();
  delete b;
}

// ---- compress batch from raw files: device-side parse (avr_parse.h) ---------------------------
#include <atomic>
#include <mutex>
#include <chrono>
#include <thread>
#include "avr_parse.h"

namespace {
template <class F>
void par_for(int n, int threads, F&& f) {
  if (threads <= 1 || n <= 1) { for (int i = 0; i < n; i++) f(i); return; }
  std::atomic<int> next(0);
  std::vector<std::thread> th;
  for (int t = 0; t < (threads < n ? threads : n); t++) th.emplace_back([&] { for (int i; (i = next.fetch_add(1)) < n;) f(i); });
  for (auto& t : th) t.join();
}
double wall_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
}  // namespace

struct AvrFileScan {
  avr::FileRecord rec;                  // spans, frames (no bins)
  std::vector<avr::ScanSlice> coded;
  int rc = 0;
};
// kept by the batch for container building
struct avr_cbatch_files { std::vector<AvrFileScan> scans; };
static std::vector<std::pair<avr_cbatch*, avr_cbatch_files*>>& files_registry() { static std::vector<std::pair<avr_cbatch*, avr_cbatch_files*>> r; return r; }
static std::mutex& files_registry_mu() { static std::mutex m; return m; }   // contexts of different GPUs may live on different threads

extern "C" int avr_cbatch_create_files(avr_context* ctx, const uint8_t* const* files, const size_t* lens, int nfiles,
                                       int host_threads, avr_cbatch** out, int* file_status) {
  if (!ctx || !files || !lens || nfiles < 0 || !out || !file_status) return AVR_E_ARG;
  if (nfiles > AVR_MAX_BATCH_FILES) { avr_set_error("at most %d files per batch (sort key has 13 file bits)", AVR_MAX_BATCH_FILES); return AVR_E_ARG; }
  CU(cudaSetDevice(ctx->device));
  const double t_host0 = wall_ms();
  avr_trace(nullptr);
  avr_cbatch_files* bf = new avr_cbatch_files();
  bf->scans.resize((size_t)nfiles);
  // The input goes to the device as ONE copy from a page-locked staging buffer: a cudaMemcpyAsync per file from the caller's
  // (usually pageable) memory ran at 2.8 GB/s on the bench corpus.  Each scanning thread copies its file right after the scan.
  std::vector<uint64_t> raw_base((size_t)nfiles, 0);
  uint64_t raw_bytes = 0;
  for (int f = 0; f < nfiles; f++) { raw_base[f] = raw_bytes; raw_bytes += (lens[f] + 31) & ~15ull; }
  size_t stage_cap = 0;
  uint8_t* stage = raw_bytes ? (uint8_t*)ctx->pinned.take(raw_bytes + 64, &stage_cap) : nullptr;
  par_for(nfiles, host_threads, [&](int f) {
    AvrFileScan& sc = bf->scans[f];
    sc.rc = avr::scan_annexb(files[f], lens[f], &sc.rec, &sc.coded);
    if (!sc.rc && (sc.rec.out_of_domain & 1)) sc.rc = AVR_E_UNSUPPORTED;
    if (!sc.rc && stage && lens[f]) memcpy(stage + raw_base[f], files[f], lens[f]);
  });
  avr_trace("cf: host scan");
  // layout of the pools.  AVR_TEST_PARSE_CAP (tests only) lowers the record capacity so that the overflow -> host
  // recorder fallback can be exercised with ordinary clips.
  uint32_t bins_per_byte = 24u;
  if (const char* e = getenv("AVR_TEST_PARSE_CAP")) { const int v = atoi(e); if (v >= 1 && v <= 24) bins_per_byte = (uint32_t)v; }
  std::vector<AvrParseSlice> ps; std::vector<uint64_t> row_off; std::vector<int> ps_file;
  uint64_t bin_pool = 0, blk_pool = 0, rows = 0;
  for (int f = 0; f < nfiles; f++) {
    file_status[f] = bf->scans[f].rc;
    if (file_status[f]) continue;
    for (const avr::ScanSlice& c : bf->scans[f].coded) {
      AvrParseSlice p; memset(&p, 0, sizeof p);
      p.raw_off = raw_base[f] + c.raw_off; p.size = c.size; memcpy(p.tail, c.tail, 4);
      // records per CABAC byte: ~11 bins typical; skip runs cost ~2 bins per macroblock and almost no bytes
      const uint32_t mbs = (uint32_t)c.sp.mb_w * (uint32_t)c.sp.mb_h;
      p.bin_cap = bins_per_byte * c.size + 4u * (mbs > (uint32_t)c.sp.first_mb ? mbs - (uint32_t)c.sp.first_mb : 0u) + 2048u; p.blk_cap = p.bin_cap / 6;
      p.bin_off = bin_pool; bin_pool += p.bin_cap; p.blk_off = blk_pool; blk_pool += p.blk_cap;
      p.frame = c.frame; p.sp = c.sp;
      ps.push_back(p); ps_file.push_back(f); row_off.push_back(rows); rows += (uint64_t)c.sp.mb_w;
    }
  }
  avr_cbatch* b = new avr_cbatch();
  b->ctx = ctx; b->mem.arena = ctx->arena; memset(&b->d, 0, sizeof b->d); memset(&b->t, 0, sizeof b->t);
  DevMem& m = b->mem;
  DevMem tmp; tmp.arena = ctx->arena;   // returned to the context before this call ends
  auto fail = [&](int r) { if (stage) cudaStreamSynchronize(ctx->stream); ctx->pinned.give(stage, stage_cap); stage = nullptr; tmp.release(); m.release(); delete b; delete bf; return r; };
  const uint32_t ns = (uint32_t)ps.size();
  uint8_t* d_raw; uint16_t* d_bins; AvrBlock* d_blkpool; avr::MbInfo* d_rows; uint64_t* d_row_off; AvrParseSlice* d_ps; AvrParseResult* d_res;
  uint32_t* d_order;
  std::vector<uint32_t> order(ps.size());
  for (uint32_t i = 0; i < order.size(); i++) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return ps[a].size > ps[b].size; });
  int rc;
  if ((rc = tmp.alloc(&d_raw, raw_bytes + 64)) || (rc = m.alloc(&d_bins, bin_pool)) || (rc = tmp.alloc(&d_blkpool, blk_pool)) ||
      (rc = tmp.alloc(&d_rows, rows)) || (rc = tmp.alloc(&d_row_off, ns)) || (rc = tmp.alloc(&d_ps, ns)) || (rc = tmp.alloc(&d_res, ns)) ||
      (rc = tmp.alloc(&d_order, ns)))
    return fail(rc);
  avr_trace("cf: pool alloc");
  cudaStream_t st = ctx->stream;
  // file-local bin numbering = position in the file's region of the bins pool
  {
    uint64_t file_first = 0; int cur = -1;
    for (uint32_t s = 0; s < ns; s++) {
      if (ps_file[s] != cur) { cur = ps_file[s]; file_first = ps[s].bin_off; }
      ps[s].file_bin_base = (uint32_t)(ps[s].bin_off - file_first);
    }
  }
  uint64_t h2d = 0;
  cudaEventRecord(ctx->ev[0], st);
  if (stage) { cudaMemcpyAsync(d_raw, stage, raw_bytes, cudaMemcpyHostToDevice, st); h2d += raw_bytes; }
  else
    for (int f = 0; f < nfiles; f++)
      if (!file_status[f] && lens[f]) { cudaMemcpyAsync(d_raw + raw_base[f], files[f], lens[f], cudaMemcpyHostToDevice, st); h2d += lens[f]; }
  cudaMemcpyAsync(d_ps, ps.data(), ns * sizeof(AvrParseSlice), cudaMemcpyHostToDevice, st); h2d += ns * sizeof(AvrParseSlice);
  cudaMemcpyAsync(d_row_off, row_off.data(), ns * sizeof(uint64_t), cudaMemcpyHostToDevice, st); h2d += ns * 8;
  cudaMemcpyAsync(d_order, order.data(), ns * sizeof(uint32_t), cudaMemcpyHostToDevice, st); h2d += ns * 4;
  cudaEventRecord(ctx->ev[1], st);
  int has_b = 0;
  for (const AvrParseSlice& q : ps) if (q.sp.slice_type == avr::SLICE_B) has_b = 1;
  // slices longer than AVR_PARSE_LONG percent of the longest keep a warp to themselves (avr_k_parse)
  uint32_t nlong = 0;
  {
    int pct = 35;
    if (const char* e = getenv("AVR_PARSE_LONG")) { const int v = atoi(e); if (v >= 0 && v <= 100) pct = v; }
    const uint64_t lim = ns ? (uint64_t)ps[order[0]].size * (uint64_t)pct / 100u : 0;
    while (nlong < ns && ps[order[nlong]].size > lim) nlong++;
  }
  rc = avr_launch_parse(d_ps, ns, d_raw, d_bins, d_blkpool, d_rows, d_row_off, ctx->d_tab, d_res, d_order, nlong, has_b, st);
  cudaEventRecord(ctx->ev[2], st);
  if (rc) { avr_set_error("parse launch: %s", cudaGetErrorString((cudaError_t)rc)); return fail(AVR_E_CUDA); }
  std::vector<AvrParseResult> res(ns);
  cudaMemcpyAsync(res.data(), d_res, ns * sizeof(AvrParseResult), cudaMemcpyDeviceToHost, st);
  cudaError_t e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) { avr_set_error("parse kernel: %s", cudaGetErrorString(e)); return fail(AVR_E_CUDA); }
  ctx->pinned.give(stage, stage_cap); stage = nullptr;
  avr_trace("cf: upload+parse+results");
  b->t.h2d_ms = ev_ms(ctx->ev[0], ctx->ev[1]); b->t.parse_ms = ev_ms(ctx->ev[1], ctx->ev[2]); b->t.launches = 1;
  // files with a failed slice drop out of the batch
  for (uint32_t s = 0; s < ns; s++) if (res[s].status && !file_status[ps_file[s]]) file_status[ps_file[s]] = res[s].status;
  // Slices of one picture must not overlap (a damaged slice can run past the next one's first macroblock and still end cleanly):
  // the slice-parallel passes need every macroblock position coded at most once per picture store.  Such a file goes to the host
  // recorder, which stores it literally (avr_record.cpp record_annexb).
  for (uint32_t s = 1; s < ns; s++)
    if (ps_file[s] == ps_file[s - 1] && ps[s].frame == ps[s - 1].frame && !file_status[ps_file[s]] &&
        (uint64_t)ps[s - 1].sp.first_mb + res[s - 1].mbs > (uint64_t)ps[s].sp.first_mb)
      file_status[ps_file[s]] = AVR_E_CORRUPT;
  // dense layout of what remains
  std::vector<AvrFileDesc> fds((size_t)nfiles);
  std::vector<AvrBatchSlice> bsl; std::vector<AvrFrame> frames;
  std::vector<uint64_t> soff; std::vector<uint32_t> scap, blk_cnt(ns, 0), blk_dst(ns, 0);
  uint64_t nblocks = 0, emitted = 0, pic_mbs = 0, scratch = 0, nbins = 0;
  uint32_t s = 0;
  for (int f = 0; f < nfiles; f++) {
    AvrFileDesc& fd = fds[f]; memset(&fd, 0, sizeof fd);
    const AvrFileScan& sc = bf->scans[f];
    b->file_slice_off.push_back((uint32_t)bsl.size());
    if (bf->scans[f].rc) continue;                      // never entered the pools
    const uint32_t s0 = s; s += (uint32_t)sc.coded.size();
    if (file_status[f]) continue;
    fd.bin_off = s0 < ns ? ps[s0].bin_off : 0; fd.slice_off = (uint32_t)bsl.size(); fd.nslices = (uint32_t)sc.coded.size();
    fd.frame_off = (uint32_t)frames.size(); fd.nframes = (uint32_t)sc.rec.frames.size(); fd.block_off = (uint32_t)nblocks;
    fd.pic_mb_base = (uint32_t)pic_mbs;
    for (auto& fr : sc.rec.frames) { uint32_t mm = (uint32_t)fr.mb_w * fr.mb_h; if (mm > fd.max_mbs) fd.max_mbs = mm; pic_mbs += mm; }
    frames.insert(frames.end(), sc.rec.frames.begin(), sc.rec.frames.end());
    for (uint32_t k = s0; k < s; k++) {
      AvrBatchSlice bs = {(uint32_t)f, ps[k].file_bin_base, ps[k].file_bin_base + res[k].nbins, (uint32_t)nblocks, (uint32_t)nblocks + res[k].nblocks};
      blk_cnt[k] = res[k].nblocks; blk_dst[k] = (uint32_t)nblocks;
      nblocks += res[k].nblocks; emitted += res[k].emitted; nbins += res[k].nbins;
      bsl.push_back(bs);
      uint32_t cap = res[k].nbins + 6 * res[k].nblocks + 16;
      soff.push_back(scratch); scap.push_back(cap); scratch += (cap + 15) & ~15ull;
    }
    fd.nblocks = (uint32_t)nblocks - fd.block_off;
  }
  b->file_slice_off.push_back((uint32_t)bsl.size());
  if (emitted >= (1ull << 31) || nblocks + bsl.size() >= (1ull << 31) || pic_mbs >= (1ull << 32)) {
    avr_set_error("batch too large for one launch (%llu coded symbols); split it", (unsigned long long)emitted); return fail(AVR_E_OVERFLOW);
  }
  b->t.bins = nbins; b->t.coded_symbols = emitted; b->scratch_bytes = scratch; b->pic_mbs = pic_mbs;
  AvrDevBatch& d = b->d;
  d.nfiles = nfiles; d.nslices = (uint32_t)bsl.size(); d.nblocks = (uint32_t)nblocks; d.nbins = nbins; d.emit_cap = (uint32_t)emitted;
  d.tab = ctx->d_tab; d.bins = d_bins;
  int fb = 0; while ((1 << fb) < nfiles) fb++;
  d.sort_end_bit = AVR_KEY_BITS + 1 + fb;
  AvrFileDesc* d_fds; AvrBlock* d_blocks; AvrBatchSlice* d_bsl; AvrFrame* d_frames; uint32_t* d_cnt; uint32_t* d_dst;
  if ((rc = m.alloc(&d_fds, fds.size())) || (rc = m.alloc(&d_blocks, nblocks)) || (rc = m.alloc(&d_bsl, bsl.size())) ||
      (rc = m.alloc(&d_frames, frames.size())) || (rc = tmp.alloc(&d_cnt, ns)) || (rc = tmp.alloc(&d_dst, ns)) ||
      (rc = m.alloc(&b->d_scratch_off, soff.size())) || (rc = m.alloc(&b->d_scratch_cap, scap.size())) ||
      (rc = m.alloc(&d.scratch, scratch)) || (rc = m.alloc(&d.out_len, bsl.size())) || (rc = m.alloc(&d.status, bsl.size())) ||
      (rc = m.alloc(&d.out_off, bsl.size())) || (rc = m.alloc(&d.out_total, 1)) || (rc = m.alloc(&d.out, scratch)))
    return fail(rc);
  avr_trace("cf: layout+alloc2");
  d.fds = d_fds; d.blocks = d_blocks; d.bslices = d_bsl; d.frames = d_frames; d.slices = nullptr;
  uint32_t* d_corder; if ((rc = m.alloc(&d_corder, bsl.size()))) return fail(rc);
  d.code_order = d_corder;
  const std::vector<uint32_t> corder = code_order_of(bsl);
  d.code_long = code_long_of(bsl, corder);
  d.scratch_off = b->d_scratch_off; d.scratch_cap = b->d_scratch_cap;
  cudaEventRecord(ctx->ev[0], st);
  auto up = [&](void* dst, const void* src, size_t n) { if (n) cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, st); h2d += n; };
  up(d_fds, fds.data(), fds.size() * sizeof(AvrFileDesc)); up(d_bsl, bsl.data(), bsl.size() * sizeof(AvrBatchSlice));
  up(d_corder, corder.data(), corder.size() * sizeof(uint32_t));
  up(d_frames, frames.data(), frames.size() * sizeof(AvrFrame)); up(d_cnt, blk_cnt.data(), ns * 4); up(d_dst, blk_dst.data(), ns * 4);
  up(b->d_scratch_off, soff.data(), soff.size() * 8); up(b->d_scratch_cap, scap.data(), scap.size() * 4);
  rc = avr_launch_gather_blocks(d_blkpool, d_ps, d_cnt, d_dst, d_blocks, ns, st);
  cudaEventRecord(ctx->ev[1], st);
  e = cudaStreamSynchronize(st);
  if (rc || e != cudaSuccess) { avr_set_error("block gather: %s", cudaGetErrorString(rc ? (cudaError_t)rc : e)); return fail(AVR_E_CUDA); }
  b->t.h2d_ms += ev_ms(ctx->ev[0], ctx->ev[1]); b->t.h2d_bytes = h2d; b->t.launches = b->create_launches = 2;
  avr_trace("cf: gather");
  tmp.release();
  b->t.host_ms = wall_ms() - t_host0 - b->t.h2d_ms - b->t.parse_ms;
  avr_trace("cf: free tmp");
  { std::lock_guard<std::mutex> g(files_registry_mu()); files_registry().push_back({b, bf}); }
  *out = b;
  return AVR_OK;
}

static avr_cbatch_files* files_of(const avr_cbatch* b) {
  std::lock_guard<std::mutex> g(files_registry_mu());
  for (auto& p : files_registry()) if (p.first == b) return p.second;
  return nullptr;
}
extern "C" int avr_cbatch_span_count(const avr_cbatch* b, int file, uint32_t* n) {
  avr_cbatch_files* bf = files_of(b);
  if (!bf || file < 0 || file >= (int)bf->scans.size() || !n) return AVR_E_ARG;
  *n = (uint32_t)bf->scans[file].rec.spans.size();
  return AVR_OK;
}
extern "C" int avr_cbatch_span(const avr_cbatch* b, int file, uint32_t i, uint64_t* off, uint32_t* size, int32_t* idx) {
  avr_cbatch_files* bf = files_of(b);
  if (!bf || file < 0 || file >= (int)bf->scans.size() || i >= bf->scans[file].rec.spans.size()) return AVR_E_ARG;
  const avr::SliceSpan& s = bf->scans[file].rec.spans[i];
  if (off) *off = s.payload_off;
  if (size) *size = s.size;
  if (idx) *idx = s.slice_index;
  return AVR_OK;
}
void avr_cbatch_files_release(avr_cbatch* b) {
  std::lock_guard<std::mutex> g(files_registry_mu());
  auto& r = files_registry();
  for (size_t i = 0; i < r.size(); i++) if (r[i].first == b) { delete r[i].second; r.erase(r.begin() + i); return; }
}
