// avrbench: native timing / profiling harness over the C ABI (no Python in the process, so it can
// run under ncu and compute-sanitizer).
//   avrbench <reps> <mode: 0 pipeline | 1 chain | 2 decompress only | 3 file API (device parse)> <clip.h264>...
//   avrbench <reps> 4 <streams> <bins> <digit_bits 8|16> <model 0|1>     coder test on independent bin streams (config 1)
// Stages the clips' records once, runs compress `reps` times, stages the payloads, runs
// decompress `reps` times, verifies the re-encoded CABAC bytes against the input files and
// prints one line of JSON with the CUDA-event times.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "../../include/avrecode_b200.h"

static bool read_file(const char* fn, std::vector<uint8_t>* out) {
  FILE* f = fopen(fn, "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
  out->resize((size_t)n);
  size_t got = n ? fread(out->data(), 1, (size_t)n, f) : 0;
  fclose(f);
  return got == (size_t)n;
}
#define CHECK(x) do { int rc_ = (x); if (rc_ != AVR_OK) { fprintf(stderr, "%s -> %d: %s\n", #x, rc_, avr_last_error()); return 1; } } while (0)

int main(int argc, char** argv) {
  if (argc < 4) { fprintf(stderr, "usage: %s <reps> <mode> <clip>...\n", argv[0]); return 2; }
  const int reps = atoi(argv[1]), mode = atoi(argv[2]);
  if (mode == 4) {
    if (argc < 7) { fprintf(stderr, "usage: %s <reps> 4 <streams> <bins> <digit_bits> <model>\n", argv[0]); return 2; }
    const uint32_t S = (uint32_t)atol(argv[3]), N = (uint32_t)atol(argv[4]);
    const int digit_bits = atoi(argv[5]), model = atoi(argv[6]);
    std::vector<uint8_t> bits((size_t)S * N), back((size_t)S * N);
    uint64_t x = 88172645463325252ull;                       // xorshift64; p(1) cycles through 1/2, 1/8, 1/64, 1/1024 by stream
    for (uint32_t s = 0; s < S; s++) {
      static const int k[4] = {1, 3, 6, 10};
      for (uint32_t i = 0; i < N; i++) { x ^= x << 13; x ^= x >> 7; x ^= x << 17; bits[(size_t)s * N + i] = (x >> 20 & ((1u << k[s & 3]) - 1)) == 0; }
    }
    avr_context* ctx;
    CHECK(avr_context_create(0, &ctx));
    std::vector<uint64_t> off((size_t)S + 1);
    avr_timing te, td; double e_ms = 0, d_ms = 0; uint64_t coded_bytes = 0; int bad = 0;
    for (int r = -1; r < reps; r++) {                         // r = -1: untimed (the first launch of a kernel loads its module)
      avr_buffer coded;
      CHECK(avr_streams_encode(ctx, bits.data(), S, N, digit_bits, model, &coded, off.data(), &te));
      CHECK(avr_streams_decode(ctx, coded.data, off.data(), S, N, digit_bits, model, back.data(), &td));
      if (r >= 0) { e_ms += te.kernel_ms; d_ms += td.kernel_ms; }
      coded_bytes = coded.size;
      bad += memcmp(back.data(), bits.data(), bits.size()) != 0;
      avr_buffer_free(&coded);
    }
    printf("{\"streams\": %u, \"bins_per_stream\": %u, \"digit_bits\": %d, \"model\": %d, \"coded_bytes\": %llu, \"encode_ms\": %.3f, "
           "\"decode_ms\": %.3f, \"encode_gbins_s\": %.2f, \"decode_gbins_s\": %.2f, \"bad\": %d}\n", S, N, digit_bits, model,
           (unsigned long long)coded_bytes, e_ms / reps, d_ms / reps, (double)S * N / 1e6 / (e_ms / reps), (double)S * N / 1e6 / (d_ms / reps), bad);
    avr_context_destroy(ctx);
    return bad ? 1 : 0;
  }
  const int n = argc - 3;
  std::vector<std::vector<uint8_t>> data((size_t)n);
  if (mode == 5) {
    // avr_compress_files alone (device parse + pipeline), wall time of the call next to its device times: <reps> 5 <files...>
    uint64_t bytes = 0;
    for (int i = 0; i < n; i++) { if (!read_file(argv[3 + i], &data[i])) { fprintf(stderr, "cannot read %s\n", argv[3 + i]); return 2; } bytes += data[i].size(); }
    avr_context* ctx;
    CHECK(avr_context_create(0, &ctx));
    std::vector<const uint8_t*> in((size_t)n); std::vector<size_t> len((size_t)n);
    for (int i = 0; i < n; i++) { in[i] = data[i].data(); len[i] = data[i].size(); }
    std::vector<avr_buffer> comp((size_t)n);
    avr_timing tc; double best = 1e30;
    for (int r = -1; r < reps; r++) {
      const auto t0 = std::chrono::steady_clock::now();
      CHECK(avr_compress_files(ctx, in.data(), len.data(), n, 8, comp.data(), &tc));
      const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
      if (r >= 0 && ms < best) best = ms;
      for (int i = 0; i < n; i++) avr_buffer_free(&comp[i]);
    }
    printf("{\"files\": %d, \"bins\": %llu, \"bytes\": %llu, \"parse_ms\": %.3f, \"compress_kernel_ms\": %.3f, \"h2d_ms\": %.3f, \"compress_wall_ms\": %.3f}\n",
           n, (unsigned long long)tc.bins, (unsigned long long)bytes, tc.parse_ms, tc.kernel_ms, tc.h2d_ms, best);
    avr_context_destroy(ctx);
    return 0;
  }
  std::vector<avr_record*> recs((size_t)n);
  uint64_t bins = 0, bytes = 0;
  for (int i = 0; i < n; i++) {
    if (!read_file(argv[3 + i], &data[i])) { fprintf(stderr, "cannot read %s\n", argv[3 + i]); return 2; }
    recs[i] = avr_record_new();
    CHECK(avr_record_annexb(recs[i], data[i].data(), data[i].size()));
    uint64_t c[6]; avr_record_counts(recs[i], c);
    bins += c[0]; bytes += data[i].size();
  }
  avr_context* ctx;
  CHECK(avr_context_create(0, &ctx));
  if (mode == 3) {
    // the public file-level calls: raw files in, containers out, and back; every kernel of the product path launches
    std::vector<const uint8_t*> in((size_t)n), cin((size_t)n); std::vector<size_t> len((size_t)n), clen((size_t)n);
    for (int i = 0; i < n; i++) { in[i] = data[i].data(); len[i] = data[i].size(); }
    std::vector<avr_buffer> comp((size_t)n), back((size_t)n);
    avr_timing tc, td; int bad = 0;
    for (int r = 0; r < reps; r++) {
      CHECK(avr_compress_files(ctx, in.data(), len.data(), n, 8, comp.data(), &tc));
      for (int i = 0; i < n; i++) { cin[i] = comp[i].data; clen[i] = comp[i].size; }
      CHECK(avr_decompress_files(ctx, cin.data(), clen.data(), n, 8, back.data(), &td));
      for (int i = 0; i < n; i++) {
        bad += back[i].size != len[i] || memcmp(back[i].data, in[i], len[i]) != 0;
        avr_buffer_free(&comp[i]); avr_buffer_free(&back[i]);
      }
    }
    printf("{\"files\": %d, \"bins\": %llu, \"bytes\": %llu, \"parse_ms\": %.3f, \"compress_kernel_ms\": %.3f, \"coder_ms\": %.3f, "
           "\"decompress_kernel_ms\": %.3f, \"launches\": %u, \"bad_files\": %d}\n", n, (unsigned long long)tc.bins,
           (unsigned long long)bytes, tc.parse_ms, tc.kernel_ms, tc.coder_launch_ms_x1000 / 1000.0, td.kernel_ms, tc.launches + td.launches, bad);
    avr_context_destroy(ctx);
    return bad ? 1 : 0;
  }
  avr_cbatch* cb;
  CHECK(avr_cbatch_create(ctx, recs.data(), n, &cb));
  avr_timing tc; memset(&tc, 0, sizeof tc);
  double c_ms = 0, coder_ms = 0;
  const int cmode = mode == 1 ? AVR_MODE_CHAIN : AVR_MODE_PIPELINE;
  const int creps = mode == 2 ? 1 : reps;
  for (int r = 0; r < creps; r++) {
    CHECK(avr_cbatch_run(cb, cmode));
    avr_cbatch_timing(cb, &tc);
    c_ms += tc.kernel_ms; coder_ms += tc.coder_launch_ms_x1000 / 1000.0;
  }
  CHECK(avr_cbatch_fetch(cb));
  // decompress batch from the payloads
  std::vector<avr_dslice> sl; std::vector<uint32_t> per((size_t)n, 0);
  std::vector<std::pair<int, uint64_t>> where;   // (file, payload_off) per coded slice
  for (int i = 0; i < n; i++) {
    uint64_t c[6]; avr_record_counts(recs[i], c);
    for (uint32_t k = 0; k < c[3]; k++) {
      uint64_t off; uint32_t size, flip; int32_t idx; avr_slice_params sp;
      avr_record_span(recs[i], k, &off, &size, &idx, &flip, &sp);
      if (idx < 0) continue;
      avr_dslice d; memset(&d, 0, sizeof d);
      avr_cbatch_payload(cb, i, (uint32_t)idx, &d.payload, &d.payload_len);
      d.size = size; d.flip = flip; d.sp = sp;
      sl.push_back(d); per[i]++; where.push_back({i, off});
    }
  }
  avr_dbatch* db;
  CHECK(avr_dbatch_create(ctx, sl.data(), per.data(), n, &db));
  avr_timing td; memset(&td, 0, sizeof td);
  double d_ms = 0;
  const int dreps = mode == 1 ? 1 : reps;
  for (int r = 0; r < dreps; r++) {
    CHECK(avr_dbatch_run(db));
    avr_dbatch_timing(db, &td);
    d_ms += td.kernel_ms;
  }
  CHECK(avr_dbatch_fetch(db));
  int bad = 0;
  for (size_t s = 0; s < sl.size(); s++) {
    const uint8_t* p; uint32_t len, b;
    avr_dbatch_slice(db, (uint32_t)s, &p, &len, &b);
    const uint8_t* orig = data[where[s].first].data() + where[s].second;
    // all bytes but the last (recode.cpp:1352-1362 patches it) must match
    uint32_t cmp = len < sl[s].size ? len : sl[s].size;
    if (len + 1 < sl[s].size || len > sl[s].size || (cmp > 1 && memcmp(p, orig, cmp - 1))) bad++;
  }
  printf("{\"files\": %d, \"bins\": %llu, \"bytes\": %llu, \"slices\": %zu, \"compress_ms\": %.3f, \"coder_ms\": %.3f, "
         "\"decompress_ms\": %.3f, \"compress_mbins_s\": %.1f, \"decompress_mbins_s\": %.1f, \"bad_slices\": %d}\n",
         n, (unsigned long long)bins, (unsigned long long)bytes, sl.size(), c_ms / creps, coder_ms / creps, d_ms / dreps,
         bins / 1e3 / (c_ms / creps), bins / 1e3 / (d_ms / dreps), bad);
  avr_dbatch_destroy(db); avr_cbatch_destroy(cb); avr_context_destroy(ctx);
  return bad ? 1 : 0;
}
