// File-level drop-in for `recode compress` / `recode decompress` on memory buffers: the
// recode.proto container (reference recode.proto:1-19; block sequence rules recode.cpp:1282-1304,
// 1129-1131; reader rules recode.cpp:1366-1416; x264 padding patch recode.cpp:1352-1362) around
// the device batches of avr_api.cu.  Host work here is framing only -- every bin is coded on
// the GPU.
#include <string.h>
#include <algorithm>
#include <atomic>
#include <mutex>
#include <string>
#include <thread>
#include <chrono>
#include <vector>
#include "../../include/avrecode_b200.h"
#include "avr_api_internal.h"

namespace {

// A container under construction: malloc-backed so that the finished bytes become the caller's avr_buffer without a copy.
struct Buf {
  uint8_t* p = nullptr; size_t n = 0, cap = 0; bool bad = false;
  Buf() = default;
  Buf(const Buf&) = delete; Buf& operator=(const Buf&) = delete;
  Buf(Buf&& o) noexcept : p(o.p), n(o.n), cap(o.cap), bad(o.bad) { o.p = nullptr; o.n = o.cap = 0; }
  ~Buf() { free(p); }
  void reserve(size_t c) {
    if (c <= cap) return;
    uint8_t* q = (uint8_t*)realloc(p, c);
    if (!q) { bad = true; return; }
    p = q; cap = c;
  }
  bool room(size_t k) { if (n + k > cap) reserve((n + k) + (n + k) / 2 + 64); return n + k <= cap; }
  void push_back(uint8_t v) { if (room(1)) p[n++] = v; }
  void append(const uint8_t* s, size_t k) { if (k && room(k)) { memcpy(p + n, s, k); n += k; } }
  void clear() { n = 0; }
  size_t size() const { return n; }
};

// ---- proto2 wire writer (fields in tag order, set-but-empty fields are emitted) ----------------
void put_varint(Buf& o, uint64_t v) {
  while (v >= 0x80) { o.push_back((uint8_t)(v | 0x80)); v >>= 7; }
  o.push_back((uint8_t)v);
}
size_t varint_len(uint64_t v) { size_t n = 1; while (v >= 0x80) { v >>= 7; n++; } return n; }
void put_bytes_field(Buf& o, uint8_t tag, const uint8_t* p, size_t n) {
  o.push_back(tag); put_varint(o, n); o.append(p, n);
}
void put_block_literal(Buf& o, const uint8_t* p, size_t n) {
  o.push_back(0x12); put_varint(o, 1 + varint_len(n) + n);
  put_bytes_field(o, 0x12, p, n);
}
void put_block_skip(Buf& o, uint64_t size) {
  o.push_back(0x12); put_varint(o, 1 + varint_len(size) + 2);
  o.push_back(0x08); put_varint(o, size);
  o.push_back(0x18); o.push_back(0x01);
}
void put_block_cabac(Buf& o, uint64_t size, const uint8_t* cabac, size_t n, int parity, const uint8_t* last_byte) {
  size_t body = 1 + varint_len(size) + 1 + varint_len(n) + n + 2 + (last_byte ? 3 : 0);
  o.push_back(0x12); put_varint(o, body);
  o.push_back(0x08); put_varint(o, size);
  put_bytes_field(o, 0x22, cabac, n);
  o.push_back(0x28); o.push_back((uint8_t)parity);
  if (last_byte) { o.push_back(0x32); o.push_back(0x01); o.push_back(*last_byte); }
}

// ---- reader -----------------------------------------------------------------------------------
struct Block {
  bool has_size = false, has_literal = false, has_skip = false, has_cabac = false, has_parity = false, has_last = false;
  uint64_t size = 0; bool skip = false; int parity = 0;
  const uint8_t* literal = nullptr; size_t literal_len = 0;
  const uint8_t* cabac = nullptr; size_t cabac_len = 0;
  const uint8_t* last = nullptr; size_t last_len = 0;
};
bool get_varint(const uint8_t* p, size_t n, size_t& i, uint64_t* v) {
  uint64_t r = 0; int sh = 0;
  while (i < n && sh < 64) {
    uint8_t b = p[i++];
    r |= (uint64_t)(b & 0x7F) << sh; sh += 7;
    if (!(b & 0x80)) { *v = r; return true; }
  }
  return false;
}
bool skip_field(const uint8_t* p, size_t n, size_t& i, int wt) {
  uint64_t v;
  switch (wt) {
    case 0: return get_varint(p, n, i, &v);
    case 1: if (i + 8 > n) return false; i += 8; return true;
    case 2: if (!get_varint(p, n, i, &v) || v > n - i) return false; i += (size_t)v; return true;
    case 5: if (i + 4 > n) return false; i += 4; return true;
    default: return false;
  }
}
bool parse_container(const uint8_t* p, size_t n, std::vector<Block>* blocks) {
  size_t i = 0;
  while (i < n) {
    uint64_t tag, len;
    if (!get_varint(p, n, i, &tag)) return false;
    if ((tag >> 3) != 2 || (tag & 7) != 2) { if (!skip_field(p, n, i, (int)(tag & 7))) return false; continue; }
    if (!get_varint(p, n, i, &len) || len > n - i) return false;
    const uint8_t* b = p + i; size_t bn = (size_t)len, j = 0;
    i += bn;
    Block blk;
    while (j < bn) {
      uint64_t t, v;
      if (!get_varint(b, bn, j, &t)) return false;
      int f = (int)(t >> 3), wt = (int)(t & 7);
      if (wt == 0 && (f == 1 || f == 3 || f == 5)) {
        if (!get_varint(b, bn, j, &v)) return false;
        if (f == 1) { blk.has_size = true; blk.size = v; }
        else if (f == 3) { blk.has_skip = true; blk.skip = v != 0; }
        else { blk.has_parity = true; blk.parity = v != 0; }
      } else if (wt == 2 && (f == 2 || f == 4 || f == 6)) {
        if (!get_varint(b, bn, j, &v) || v > bn - j) return false;
        if (f == 2) { blk.has_literal = true; blk.literal = b + j; blk.literal_len = (size_t)v; }
        else if (f == 4) { blk.has_cabac = true; blk.cabac = b + j; blk.cabac_len = (size_t)v; }
        else { blk.has_last = true; blk.last = b + j; blk.last_len = (size_t)v; }
        j += (size_t)v;
      } else if (!skip_field(b, bn, j, wt)) return false;
    }
    blocks->push_back(blk);
  }
  return true;
}

template <class F>
void parallel_for(int n, int threads, F&& f) {
  if (threads <= 1 || n <= 1) { for (int i = 0; i < n; i++) f(i); return; }
  std::atomic<int> next(0);
  std::vector<std::thread> th;
  for (int t = 0; t < std::min(threads, n); t++)
    th.emplace_back([&] { for (int i; (i = next.fetch_add(1)) < n;) f(i); });
  for (auto& t : th) t.join();
}

void add_timing(avr_timing* acc, const avr_timing& t) {
  acc->h2d_ms += t.h2d_ms; acc->kernel_ms += t.kernel_ms; acc->d2h_ms += t.d2h_ms;
  acc->h2d_bytes += t.h2d_bytes; acc->d2h_bytes += t.d2h_bytes; acc->bins += t.bins;
  acc->coded_symbols += t.coded_symbols; acc->launches += t.launches; acc->coder_launch_ms_x1000 += t.coder_launch_ms_x1000;
  acc->parse_ms += t.parse_ms; acc->host_ms += t.host_ms;
}
}  // namespace
void avr_trace(const char* what) {
  static const bool on = [] { const char* e = getenv("AVR_TRACE"); return e && e[0] == '1'; }();
  static thread_local double last = 0;
  if (!on) return;
  const double t = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
  if (what) fprintf(stderr, "[avr] %-28s %9.2f ms\n", what, last ? t - last : 0.0);
  last = t;
}
namespace {
double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

bool to_buffer(Buf& v, avr_buffer* b) {          // hands the bytes over: no copy
  if (v.bad) return false;
  if (!v.p) { v.reserve(1); if (v.bad) return false; }
  b->size = v.n; b->data = v.p;
  v.p = nullptr; v.n = v.cap = 0;
  return true;
}

}  // namespace

extern "C" void avr_buffer_free(avr_buffer* b) { if (b) { free(b->data); b->data = nullptr; b->size = 0; } }

// ---- lanes -------------------------------------------------------------------------------------------------------
// The file-level calls cut their input into sub-batches and run them on several LANES: a lane is a context (own stream,
// events, device-block and pinned caches) driven by its own host thread.  One batch's NAL scan, staging copies, D2H and
// container framing then overlap another batch's kernels, and the tail of one batch's serial kernels (device parse, pass D,
// chains: each lasts as long as its longest slice / file) is filled by the other lane's work.  With several contexts
// (avr_*_files_multi) the lanes also span GPUs: files share nothing (recode.cpp:1314,1597), so no data moves between them.
static int env_int(const char* name, int dflt, int lo, int hi) {
  const char* e = getenv(name);
  if (!e || !*e) return dflt;
  const int v = atoi(e);
  return v < lo ? lo : v > hi ? hi : v;
}
// Measured on B200, 2048 clips (tools/e2e_sweep.py, profiles/r02_e2e_lanes.md): compress gains from two concurrent batches
// (566 against 666 ms; more batches or lanes only add tails), the chain kernel does not (two concurrent launches of 1024
// chains take 1000 ms where one launch of 2048 takes 680 ms), so decompress defaults to one lane and one batch per context.
static int make_lanes(avr_context* const* ctxs, int nctx, int per_ctx, std::vector<avr_context*>* lanes) {
  for (int k = 0; k < per_ctx; k++)            // lane k of every context before lane k + 1: batches are handed out in this order
    for (int c = 0; c < nctx; c++) {
      avr_context* l = avr_context_lane(ctxs[c], k);
      if (!l) return AVR_E_CUDA;
      lanes->push_back(l);
    }
  return AVR_OK;
}
// files -> nb sub-batches of similar weight: heaviest first onto the lightest batch (greedy LPT)
static std::vector<std::vector<int>> lpt_batches(const std::vector<int>& ids, const std::vector<uint64_t>& weight, int nb) {
  std::vector<std::vector<int>> out((size_t)nb);
  std::vector<uint64_t> load((size_t)nb, 0);
  std::vector<int> order(ids);
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return weight[a] > weight[b]; });
  for (int f : order) {
    size_t best = 0;
    for (size_t k = 1; k < out.size(); k++) if (load[k] < load[best]) best = k;
    out[best].push_back(f); load[best] += weight[f] + 1;
  }
  for (auto& b : out) std::sort(b.begin(), b.end());
  out.erase(std::remove_if(out.begin(), out.end(), [](const std::vector<int>& b) { return b.empty(); }), out.end());
  return out;
}
// runs job(batch index, lane context, host threads of the lane) for every batch; batches are pulled in order by one host
// thread per lane.  Returns the first error (its message is re-raised on the calling thread: avr_last_error is thread-local).
template <class Job>
static int run_on_lanes(const std::vector<avr_context*>& lanes, int nbatches, int host_threads, Job&& job) {
  const int nl = std::max(1, std::min((int)lanes.size(), nbatches));
  const int lane_threads = std::max(1, host_threads / nl);
  if (nl == 1) { for (int b = 0; b < nbatches; b++) { const int rc = job(b, lanes[0], host_threads); if (rc) return rc; } return AVR_OK; }
  std::atomic<int> next(0), first_rc(AVR_OK);
  std::string msg; std::mutex mu;
  std::vector<std::thread> th;
  for (int l = 0; l < nl; l++)
    th.emplace_back([&, l] {
      for (int b; first_rc.load() == AVR_OK && (b = next.fetch_add(1)) < nbatches;) {
        const int rc = job(b, lanes[(size_t)l], lane_threads);
        if (rc) { std::lock_guard<std::mutex> g(mu); if (first_rc.load() == AVR_OK) { first_rc = rc; msg = avr_last_error(); } }
      }
    });
  for (auto& t : th) t.join();
  if (first_rc.load() != AVR_OK) avr_set_error("%s", msg.c_str());
  return first_rc.load();
}

// compressor::run (recode.cpp:1122-1132) for n files at once.
// Phase 1 of avr_compress_files: batches whose CABAC parse runs on the GPU (avr_cbatch_create_files).
// Marks done[f] for every file it produced a container for; the rest fall to the host recorder.
static int compress_device_parse(avr_context* ctx, const uint8_t* const* in, const size_t* in_len, const std::vector<int>& ids,
                                 int host_threads, std::vector<Buf>& containers, std::vector<char>& done,
                                 avr_timing* tt) {
  if (ids.empty()) return AVR_OK;
  std::vector<const uint8_t*> fin; std::vector<size_t> flen;
  for (int f : ids) { fin.push_back(in[f]); flen.push_back(in_len[f]); }
  std::vector<int> status(ids.size(), 0);
  avr_cbatch* b = nullptr;
  int rc = avr_cbatch_create_files(ctx, fin.data(), flen.data(), (int)ids.size(), host_threads, &b, status.data());
  if (rc == AVR_E_OVERFLOW) {                     // too many coded symbols for one launch: halve
    if (ids.size() == 1) return AVR_OK;           // left to the host path, which reports the limit
    std::vector<int> lo(ids.begin(), ids.begin() + ids.size() / 2), hi(ids.begin() + ids.size() / 2, ids.end());
    rc = compress_device_parse(ctx, in, in_len, lo, host_threads, containers, done, tt);
    return rc ? rc : compress_device_parse(ctx, in, in_len, hi, host_threads, containers, done, tt);
  }
  avr_trace(nullptr);
  if (rc == AVR_OK) rc = avr_cbatch_run(b, AVR_MODE_PIPELINE);
  avr_trace("c: run");
  if (rc == AVR_OK) rc = avr_cbatch_fetch(b);
  avr_trace("c: fetch");
  if (rc == AVR_OK) {
    avr_timing t; avr_cbatch_timing(b, &t); add_timing(tt, t);
    const double t0 = now_ms();
    parallel_for((int)ids.size(), host_threads, [&](int k) {
      if (status[k]) return;
      const int f = ids[k];
      Buf& o = containers[f];
      o.reserve(in_len[f] + in_len[f] / 16 + 64);
      uint32_t nspans = 0; avr_cbatch_span_count(b, k, &nspans);
      size_t prev_end = 0;
      for (uint32_t i = 0; i < nspans; i++) {
        uint64_t off; uint32_t size; int32_t idx;
        avr_cbatch_span(b, k, i, &off, &size, &idx);
        if (idx < 0) { put_block_skip(o, size); continue; }
        put_block_literal(o, in[f] + prev_end, (size_t)off - prev_end);
        prev_end = (size_t)off + size;
        const uint8_t* p; uint32_t len;
        avr_cbatch_payload(b, k, (uint32_t)idx, &p, &len);
        put_block_cabac(o, size, p, len, (int)(size & 1), size > 1 ? in[f] + prev_end - 1 : nullptr);
      }
      put_block_literal(o, in[f] + prev_end, in_len[f] - prev_end);
      done[f] = 1;
    });
    tt->host_ms += now_ms() - t0;
    avr_trace("c: containers");
  }
  avr_cbatch_destroy(b);
  avr_trace("c: destroy");
  return rc;
}

static int compress_core(const std::vector<avr_context*>& lanes, const uint8_t* const* in, const size_t* in_len, int n, int host_threads,
                         avr_buffer* out, avr_timing* tacc) {
  avr_context* ctx = lanes[0];
  avr_timing tt; memset(&tt, 0, sizeof tt);
  const double t_call0 = now_ms();
  std::vector<Buf> containers((size_t)n);
  std::vector<char> done((size_t)n, 0);
  int rc = AVR_OK;
  // AVR_HOST_PARSE=1 keeps the CABAC parse on the host cores (A/B measurement, bench.py)
  const char* hp = getenv("AVR_HOST_PARSE");
  if (!(hp && hp[0] == '1') && n > 0) {
    // sub-batches: at most AVR_MAX_BATCH_FILES files and 160 MB of input each (~8.4 coded symbols per input byte on High-profile
    // streams, the launch limit is 2^31); a batch that still exceeds the limit is halved (AVR_E_OVERFLOW).  One batch per lane
    // unless the limits ask for more: every further batch adds the tail of its longest slice.
    std::vector<int> ids((size_t)n); std::vector<uint64_t> w((size_t)n);
    uint64_t bytes = 0;
    for (int f = 0; f < n; f++) { ids[f] = f; w[f] = in_len[f]; bytes += in_len[f]; }
    int nb = (int)lanes.size() * env_int("AVR_BATCHES_PER_LANE", 1, 1, 16);
    nb = std::min(nb, std::max(1, n / 64));
    nb = std::max(nb, (int)((bytes + (160u << 20) - 1) / (160u << 20)));
    nb = std::max(nb, (n + AVR_MAX_BATCH_FILES - 1) / AVR_MAX_BATCH_FILES);
    const std::vector<std::vector<int>> batches = lpt_batches(ids, w, nb);
    std::mutex mu;
    rc = run_on_lanes(lanes, (int)batches.size(), host_threads, [&](int b, avr_context* lane, int threads) {
      avr_timing t; memset(&t, 0, sizeof t);
      const int r = compress_device_parse(lane, in, in_len, batches[(size_t)b], threads, containers, done, &t);
      std::lock_guard<std::mutex> g(mu);
      add_timing(&tt, t);
      return r;
    });
    if (rc != AVR_OK) return rc;
  }
  // host recorder for what the device parse left (record capacity exceeded, syntax it rejects)
  std::vector<int> rest;
  for (int f = 0; f < n; f++) if (!done[f]) rest.push_back(f);
  const double t_host0 = now_ms();
  std::vector<avr_record*> recs((size_t)n, nullptr);
  std::vector<int> rec_rc((size_t)n, 0);
  parallel_for((int)rest.size(), host_threads, [&](int k) {
    const int i = rest[k];
    recs[i] = avr_record_new();
    rec_rc[i] = avr_record_annexb(recs[i], in[i], in_len[i]);
    if (avr_record_view(recs[i])->out_of_domain & 1) rec_rc[i] = AVR_E_UNSUPPORTED;
  });
  tt.host_ms += now_ms() - t_host0;
  auto cleanup = [&]() { for (auto r : recs) if (r) avr_record_free(r); };
  size_t r0 = 0;
  while (r0 < rest.size() && rc == AVR_OK) {
    std::vector<int> files;
    uint64_t sym = 0;
    size_t i = r0;
    for (; i < rest.size() && files.size() < AVR_MAX_BATCH_FILES; i++) {
      const int f = rest[i];
      if (rec_rc[f]) continue;
      uint64_t e = avr_record_view(recs[f])->emitted;
      if (!files.empty() && sym + e > 1900000000ull) break;
      files.push_back(f); sym += e;
    }
    r0 = i;
    if (files.empty()) continue;
    std::vector<const avr_record*> br;
    for (int f : files) br.push_back(recs[f]);
    avr_cbatch* b = nullptr;
    rc = avr_cbatch_create(ctx, br.data(), (int)br.size(), &b);
    if (rc == AVR_OK) rc = avr_cbatch_run(b, AVR_MODE_PIPELINE);
    if (rc == AVR_OK) rc = avr_cbatch_fetch(b);
    if (rc == AVR_OK) {
      avr_timing t; avr_cbatch_timing(b, &t); add_timing(&tt, t);
      // containers of this batch (recode.cpp:1282-1304 block sequence)
      parallel_for((int)files.size(), host_threads, [&](int k) {
        const int f = files[k];
        const avr::FileRecord& r = *avr_record_view(recs[f]);
        Buf& o = containers[f];
        o.reserve(in_len[f] + in_len[f] / 16 + 64);
        size_t prev_end = 0;
        for (const avr::SliceSpan& sp : r.spans) {
          if (sp.slice_index < 0) { put_block_skip(o, sp.size); continue; }
          put_block_literal(o, in[f] + prev_end, (size_t)sp.payload_off - prev_end);
          prev_end = (size_t)sp.payload_off + sp.size;
          const uint8_t* p; uint32_t len;
          avr_cbatch_payload(b, k, (uint32_t)sp.slice_index, &p, &len);
          put_block_cabac(o, sp.size, p, len, (int)(sp.size & 1), sp.size > 1 ? in[f] + prev_end - 1 : nullptr);
        }
        put_block_literal(o, in[f] + prev_end, in_len[f] - prev_end);
      });
    }
    avr_cbatch_destroy(b);
  }
  if (rc != AVR_OK) { cleanup(); return rc; }
  avr_trace(nullptr);
  std::atomic<int> nomem(0);
  parallel_for(n, host_threads, [&](int f) {
    // what neither path could record (a launch limit, AVR_STRICT_DOMAIN with repeated frame_num) is stored whole as one literal
    // block: a valid container that restores the input.  Slices outside the covered syntax (CAVLC, field/MBAFF, 4:2:2+, I_PCM,
    // damaged data) are left literal one by one instead (avr_record.h slice_abort, scan_annexb_slices).
    if (!done[f] && rec_rc[f]) { containers[f].clear(); put_block_literal(containers[f], in[f], in_len[f]); }
    if (!to_buffer(containers[f], &out[f])) nomem = 1;
  });
  cleanup();
  avr_trace("c: to_buffer");
  if (nomem) { for (int f = 0; f < n; f++) { free(out[f].data); out[f].data = nullptr; out[f].size = 0; } return AVR_E_NOMEM; }
  // with several lanes the device times of concurrent batches add up beyond the wall time: host_ms is then what is left, not below 0
  tt.host_ms = std::max(0.0, now_ms() - t_call0 - (tt.h2d_ms + tt.parse_ms + tt.kernel_ms + tt.d2h_ms));
  if (tacc) *tacc = tt;
  return AVR_OK;
}
extern "C" int avr_compress_files_multi(avr_context* const* ctxs, int nctx, const uint8_t* const* in, const size_t* in_len, int n,
                                        int host_threads, avr_buffer* out, avr_timing* tacc) {
  if (!ctxs || nctx <= 0 || !in || !in_len || n < 0 || !out) return AVR_E_ARG;
  for (int c = 0; c < nctx; c++) if (!ctxs[c]) return AVR_E_ARG;
  if (n) memset(out, 0, (size_t)n * sizeof *out);          // output-only: all-null on every failure path
  std::vector<avr_context*> lanes;
  if (make_lanes(ctxs, nctx, env_int("AVR_LANES", 2, 1, 8), &lanes)) return AVR_E_CUDA;
  return compress_core(lanes, in, in_len, n, host_threads, out, tacc);
}
extern "C" int avr_compress_files(avr_context* ctx, const uint8_t* const* in, const size_t* in_len, int n, int host_threads,
                                  avr_buffer* out, avr_timing* tacc) {
  return avr_compress_files_multi(&ctx, 1, in, in_len, n, host_threads, out, tacc);
}

// decompressor::run (recode.cpp:1345-1364) for n containers at once.
namespace {
struct DFile {
  std::vector<Block> blocks;
  std::vector<uint8_t> surrogate;             // literals + 'X' placeholders (recode.cpp:1534-1551 uses marker + 'X')
  std::vector<avr_dslice> slices;             // coded slices in decode order
  std::vector<int> slice_block;               // block index of each
  int rc = AVR_OK;
};
}  // namespace

static int decompress_core(const std::vector<avr_context*>& lanes, const uint8_t* const* in, const size_t* in_len, int n, int host_threads,
                           avr_buffer* out, avr_timing* tacc) {
  avr_timing tt; memset(&tt, 0, sizeof tt);
  const double t_call0 = now_ms();
  avr_trace(nullptr);
  std::vector<DFile> files((size_t)n);
  parallel_for(n, host_threads, [&](int f) {
    DFile& d = files[f];
    if (!parse_container(in[f], in_len[f], &d.blocks)) { d.rc = AVR_E_CORRUPT; return; }
    std::vector<size_t> coded_pos; std::vector<int> coded_blk;
    size_t total = 0;
    for (auto& b : d.blocks) total += b.has_literal ? b.literal_len : b.has_cabac ? (size_t)b.size : 0;
    d.surrogate.reserve(total);
    for (size_t i = 0; i < d.blocks.size(); i++) {
      const Block& b = d.blocks[i];
      if ((int)b.has_literal + (int)b.has_cabac + (int)b.has_skip != 1) { d.rc = AVR_E_CORRUPT; return; }   // recode.cpp:1371-1373
      if (b.has_literal) d.surrogate.insert(d.surrogate.end(), b.literal, b.literal + b.literal_len);
      else if (b.has_cabac) {
        if (!b.has_size || b.size > (1u << 30)) { d.rc = AVR_E_CORRUPT; return; }                          // recode.cpp:1384-1386
        coded_pos.push_back(d.surrogate.size()); coded_blk.push_back((int)i);
        d.surrogate.insert(d.surrogate.end(), (size_t)b.size, (uint8_t)'X');
      } else if (b.skip) { coded_pos.push_back((size_t)-1); coded_blk.push_back((int)i); }
      else { d.rc = AVR_E_CORRUPT; return; }
    }
    bool any_cabac = false;
    for (auto& b : d.blocks) any_cabac |= b.has_cabac;
    if (!any_cabac) return;
    size_t next = 0;
    bool have = false; int w = 0, h = 0, fnum = 0;
    int nflips = 0; bool last_prev_valid = false;
    int unsupported = 0;
    int rc = avr::scan_annexb_slices(d.surrogate.data(), d.surrogate.size(), [&](const avr::SliceLoc& loc) -> int {
      const avr::SliceHeaderInfo& sh = loc.sh;
      if (sh.sp.first_mb == 0 && !(avr::strict_domain() && have && sh.sp.mb_w == w && sh.sp.mb_h == h && sh.frame_num == fnum)) {   // update_frame_spec; every picture flips unless AVR_STRICT_DOMAIN (avr_record.h)
        last_prev_valid = have && sh.sp.mb_w == w && sh.sp.mb_h == h;
        nflips++;
        have = true; w = sh.sp.mb_w; h = sh.sp.mb_h; fnum = sh.frame_num;
      }
      if (next >= coded_blk.size()) return AVR_E_CORRUPT;
      const Block& b = d.blocks[coded_blk[next]];
      const size_t pos = coded_pos[next];
      next++;
      if (b.has_skip) return b.size == loc.size ? 0 : AVR_E_CORRUPT;              // recode.cpp:1573-1575
      if (pos != loc.raw_off || b.size != loc.size) return AVR_E_CORRUPT;          // recode.cpp:1564-1566
      avr_dslice s;
      memset(&s, 0, sizeof s);
      s.payload = b.cabac; s.payload_len = (uint32_t)b.cabac_len; s.size = (uint32_t)b.size;
      s.flip = nflips == 0 ? 0u : (nflips == 1 && last_prev_valid) ? 1u : 2u;
      nflips = 0;
      s.sp.slice_type = sh.sp.slice_type; s.sp.first_mb = sh.sp.first_mb; s.sp.mb_w = sh.sp.mb_w; s.sp.mb_h = sh.sp.mb_h;
      s.sp.num_ref[0] = sh.sp.num_ref[0]; s.sp.num_ref[1] = sh.sp.num_ref[1]; s.sp.cabac_init_idc = sh.sp.cabac_init_idc;
      s.sp.slice_qp = sh.sp.slice_qp; s.sp.transform_8x8_mode = sh.sp.transform_8x8_mode;
      s.sp.chroma_format_idc = sh.sp.chroma_format_idc; s.sp.direct_8x8_inference = sh.sp.direct_8x8_inference;
      d.slices.push_back(s); d.slice_block.push_back(coded_blk[next - 1]);
      return 0;
    }, &unsupported);
    // a container whose coded blocks belong to slices this library passes over (field/MBAFF, 4:2:2+ ... written by the reference's
    // compressor) cannot line up with the scan: say unsupported rather than corrupt
    if (rc) { d.rc = unsupported ? AVR_E_UNSUPPORTED : rc; return; }
    for (; next < coded_blk.size(); next++)
      if (d.blocks[coded_blk[next]].has_cabac) { d.rc = unsupported ? AVR_E_UNSUPPORTED : AVR_E_CORRUPT; return; }   // "Not all blocks were decoded."
  });
  for (int f = 0; f < n; f++)
    if (files[f].rc) { avr_set_error("container %d: invalid (%d)", f, files[f].rc); return files[f].rc; }

  avr_trace("d: containers+scan");
  // Output files are assembled straight from the batch's device-fetched slice bytes (one copy per byte): literal
  // blocks from the container, coded blocks from the batch with the x264 padding patch applied in place
  // (recode.cpp:1352-1362, 1387-1391: the last byte is replaced, or appended when the length parity differs).
  std::atomic<int> nomem(0);
  auto patched_len = [](const Block& blk, uint32_t len) -> size_t {
    if (blk.has_parity && blk.has_last && blk.last_len > 0 && blk.parity != (int)(len & 1)) return (size_t)len + 1;
    return len;
  };
  auto assemble = [&](int f, const avr_dbatch* b, uint32_t first_slice) {
    DFile& d = files[f];
    std::vector<int> blk_slice(d.blocks.size(), -1);
    for (size_t k = 0; k < d.slice_block.size(); k++) blk_slice[d.slice_block[k]] = (int)k;
    size_t total = 0;
    for (size_t i = 0; i < d.blocks.size(); i++) {
      const Block& blk = d.blocks[i];
      if (blk.has_literal) total += blk.literal_len;
      else if (blk.has_cabac) {
        const uint8_t* p; uint32_t len;
        avr_dbatch_slice(b, first_slice + (uint32_t)blk_slice[i], &p, &len, nullptr);
        total += patched_len(blk, len);
      }
    }
    out[f].size = total;
    out[f].data = (uint8_t*)malloc(total ? total : 1);
    if (!out[f].data) { nomem = 1; return; }
    uint8_t* o = out[f].data;
    for (size_t i = 0; i < d.blocks.size(); i++) {
      const Block& blk = d.blocks[i];
      if (blk.has_literal) { if (blk.literal_len) memcpy(o, blk.literal, blk.literal_len); o += blk.literal_len; }
      else if (blk.has_cabac) {
        const uint8_t* p; uint32_t len;
        avr_dbatch_slice(b, first_slice + (uint32_t)blk_slice[i], &p, &len, nullptr);
        if (len) memcpy(o, p, len);
        if (blk.has_parity && blk.has_last && blk.last_len > 0) {
          if (blk.parity != (int)(len & 1)) o[len] = blk.last[0];
          else if (len) o[len - 1] = blk.last[0];
        }
        o += patched_len(blk, len);
      }
    }
  };
  std::vector<char> assembled((size_t)n, 0);
  int rc = AVR_OK;
  {
    // sub-batches over the lanes: every chain kernel lasts as long as its longest file, and a chain runs the faster the fewer
    // chains share its SM, so the files are spread by payload bytes (greedy LPT) and at most AVR_MAX_BATCH_FILES go into one launch
    std::vector<int> ids; std::vector<uint64_t> w((size_t)n, 0);
    uint64_t bytes = 0;
    for (int f = 0; f < n; f++) {
      if (files[f].slices.empty()) continue;
      for (auto& sl : files[f].slices) w[f] += sl.payload_len + sl.size;
      ids.push_back(f); bytes += w[f];
    }
    int nb = (int)lanes.size() * env_int("AVR_DBATCHES_PER_LANE", 1, 1, 16);
    nb = std::min(nb, std::max(1, (int)ids.size() / 64));
    nb = std::max(nb, (int)((bytes + (8ull << 30) - 1) / (8ull << 30)));
    nb = std::max(nb, ((int)ids.size() + AVR_MAX_BATCH_FILES - 1) / AVR_MAX_BATCH_FILES);
    const std::vector<std::vector<int>> batches = ids.empty() ? std::vector<std::vector<int>>() : lpt_batches(ids, w, nb);
    {
      int inflight = 0;                                         // chains resident at once: the first batch of every lane
      for (size_t b = 0; b < batches.size() && b < lanes.size(); b++) inflight += (int)batches[b].size();
      avr_set_chain_concurrency_hint(inflight);
    }
    std::mutex mu;
    rc = run_on_lanes(lanes, (int)batches.size(), host_threads, [&](int bi, avr_context* lane, int threads) {
      const std::vector<int>& bids = batches[(size_t)bi];
      std::vector<avr_dslice> sl; std::vector<uint32_t> per;
      for (int f : bids) { per.push_back((uint32_t)files[f].slices.size()); sl.insert(sl.end(), files[f].slices.begin(), files[f].slices.end()); }
      avr_dbatch* b = nullptr;
      int r = avr_dbatch_create(lane, sl.data(), per.data(), (int)bids.size(), &b);
      if (r == AVR_OK) r = avr_dbatch_run(b);
      if (r == AVR_OK) r = avr_dbatch_fetch(b);
      if (r == AVR_OK) {
        avr_timing t; avr_dbatch_timing(b, &t);
        { std::lock_guard<std::mutex> g(mu); add_timing(&tt, t); }
        std::vector<uint32_t> first(bids.size(), 0);
        for (size_t k = 1; k < bids.size(); k++) first[k] = first[k - 1] + per[k - 1];
        parallel_for((int)bids.size(), threads, [&](int k) { assemble(bids[(size_t)k], b, first[(size_t)k]); assembled[bids[(size_t)k]] = 1; });
      }
      avr_dbatch_destroy(b);
      return r;
    });
    avr_set_chain_concurrency_hint(0);
  }
  if (rc != AVR_OK) { for (int f = 0; f < n; f++) { free(out[f].data); out[f].data = nullptr; out[f].size = 0; } return rc; }
  parallel_for(n, host_threads, [&](int f) { if (!assembled[f]) assemble(f, nullptr, 0); });   // files without coded blocks: literals only
  if (nomem) { for (int f = 0; f < n; f++) { free(out[f].data); out[f].data = nullptr; out[f].size = 0; } return AVR_E_NOMEM; }
  avr_trace("d: assemble");
  tt.host_ms = std::max(0.0, now_ms() - t_call0 - (tt.h2d_ms + tt.kernel_ms + tt.d2h_ms));
  if (tacc) *tacc = tt;
  return AVR_OK;
}
extern "C" int avr_decompress_files_multi(avr_context* const* ctxs, int nctx, const uint8_t* const* in, const size_t* in_len, int n,
                                          int host_threads, avr_buffer* out, avr_timing* tacc) {
  if (!ctxs || nctx <= 0 || !in || !in_len || n < 0 || !out) return AVR_E_ARG;
  for (int c = 0; c < nctx; c++) if (!ctxs[c]) return AVR_E_ARG;
  if (n) memset(out, 0, (size_t)n * sizeof *out);          // output-only: all-null on every failure path
  std::vector<avr_context*> lanes;
  if (make_lanes(ctxs, nctx, env_int("AVR_DLANES", 1, 1, 8), &lanes)) return AVR_E_CUDA;
  return decompress_core(lanes, in, in_len, n, host_threads, out, tacc);
}
extern "C" int avr_decompress_files(avr_context* ctx, const uint8_t* const* in, const size_t* in_len, int n, int host_threads,
                                    avr_buffer* out, avr_timing* tacc) {
  return avr_decompress_files_multi(&ctx, 1, in, in_len, n, host_threads, out, tacc);
}
