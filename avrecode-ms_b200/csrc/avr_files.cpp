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
#include "avr_container.h"

namespace {
using avrc::Buf; using avrc::Block; using avrc::DFile; using avrc::put_block_literal;

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
      uint32_t nspans = 0; avr_cbatch_span_count(b, k, &nspans);
      avrc::write_container(in[f], in_len[f], nspans,
                            [&](uint32_t i, uint64_t* off, uint32_t* size, int32_t* idx) { avr_cbatch_span(b, k, i, off, size, idx); },
                            [&](uint32_t idx, const uint8_t** p, uint32_t* len) { avr_cbatch_payload(b, k, idx, p, len); }, containers[f]);
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
        avrc::write_container(in[f], in_len[f], (uint32_t)r.spans.size(),
                              [&](uint32_t i, uint64_t* off, uint32_t* size, int32_t* idx) { *off = r.spans[i].payload_off; *size = r.spans[i].size; *idx = r.spans[i].slice_index; },
                              [&](uint32_t idx, const uint8_t** p, uint32_t* len) { avr_cbatch_payload(b, k, idx, p, len); }, containers[f]);
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

static int decompress_core(const std::vector<avr_context*>& lanes, const uint8_t* const* in, const size_t* in_len, int n, int host_threads,
                           avr_buffer* out, avr_timing* tacc) {
  avr_timing tt; memset(&tt, 0, sizeof tt);
  const double t_call0 = now_ms();
  avr_trace(nullptr);
  std::vector<DFile> files((size_t)n);
  parallel_for(n, host_threads, [&](int f) { avrc::plan_decompress(in[f], in_len[f], &files[f]); });
  for (int f = 0; f < n; f++)
    if (files[f].rc) { avr_set_error("container %d: invalid (%d)", f, files[f].rc); return files[f].rc; }

  avr_trace("d: containers+scan");
  // Output files are assembled straight from the batch's device-fetched slice bytes (one copy per byte): literal
  // blocks from the container, coded blocks from the batch with the x264 padding patch applied in place
  // (recode.cpp:1352-1362, 1387-1391: the last byte is replaced, or appended when the length parity differs).
  std::atomic<int> nomem(0);
  auto assemble = [&](int f, const avr_dbatch* b, uint32_t first_slice) {
    const bool ok = avrc::assemble(files[f], [&](uint32_t k, const uint8_t** p, uint32_t* len) { avr_dbatch_slice(b, first_slice + k, p, len, nullptr); },
                                   &out[f].data, &out[f].size);
    if (!ok) nomem = 1;
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
