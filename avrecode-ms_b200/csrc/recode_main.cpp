// `recode` CLI, drop-in for the reference's (recode.cpp:1642-1677):
//   recode compress|decompress|roundtrip <input> [output]      exit 0 / 1
//   recode test <directory>                                     roundtrip of every regular file, <dir>/output/metrics.csv
// plus the batch form the GPU path exists for:
//   recode compress-batch|roundtrip-batch <file>...            outputs <file>.avrec / report only
//   --gpus N (anywhere on the line, or RECODE_GPUS=N)          files are spread over the first N GPUs of the box, greedy
//                                                              longest-first (avr_*_files_multi); same bytes out for any N
// All entropy coding runs on the GPU through include/avrecode_b200.h; there is no CPU path.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <filesystem>
#include <string>
#include <thread>
#include <vector>
#include "../../include/avrecode_b200.h"

static bool read_file(const std::string& fn, std::vector<uint8_t>* out) {
  FILE* f = fopen(fn.c_str(), "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
  out->resize((size_t)n);
  size_t got = n ? fread(out->data(), 1, (size_t)n, f) : 0;
  fclose(f);
  return got == (size_t)n;
}
static bool write_out(const char* fn, const uint8_t* p, size_t n) {
  FILE* f = fn ? fopen(fn, "wb") : stdout;
  if (!f) return false;
  bool ok = fwrite(p, 1, n, f) == n;
  if (fn) fclose(f); else fflush(f);
  return ok;
}
static int fail(const std::string& what) {
  fprintf(stderr, "Exception (avr_error): %s\n", what.c_str());   // shape of recode.cpp:1672-1675
  return 1;
}
// one context per GPU asked for (--gpus N); fewer when the box has fewer
static bool open_contexts(int want, std::vector<avr_context*>* ctxs) {
  for (int d = 0; d < want; d++) {
    avr_context* c = nullptr;
    if (avr_context_create(d, &c) != AVR_OK) { if (d == 0) return false; break; }
    ctxs->push_back(c);
  }
  return true;
}
static void close_contexts(std::vector<avr_context*>& ctxs) { for (auto c : ctxs) avr_context_destroy(c); ctxs.clear(); }
static double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// `recode test <dir>` (test.cpp:20-148): roundtrip of every regular file of the directory, decompressed copies in
// <dir>/output/, results in <dir>/output/metrics.csv with the reference's columns (test.cpp:32) plus Bins and
// Gbins/s.  The files go through the GPU as ONE batch, so the per-file times are the batch times apportioned by file
// size; Duration / Video stream / Frames per second come from libavformat's dump in the reference and are "n/a" here.
static int run_test(const std::string& dir, int threads, int gpus) {
  namespace fs = std::filesystem;
  std::vector<std::string> names;
  std::error_code ec;
  for (const auto& e : fs::directory_iterator(dir, ec)) if (e.is_regular_file()) names.push_back(e.path().string());
  if (ec) return fail("Failed to open directory: " + dir);
  std::sort(names.begin(), names.end());
  fs::create_directory(dir + "/output", ec);
  const int n = (int)names.size();
  std::vector<std::vector<uint8_t>> data((size_t)n);
  for (int i = 0; i < n; i++) if (!read_file(names[i], &data[i])) return fail("Failed to open file: " + names[i]);
  std::vector<avr_context*> ctxs;
  if (!open_contexts(gpus, &ctxs)) return fail(avr_last_error());
  std::vector<const uint8_t*> in, cin; std::vector<size_t> len, clen;
  for (auto& d : data) { in.push_back(d.data()); len.push_back(d.size()); }
  std::vector<avr_buffer> comp((size_t)n), dec((size_t)n);
  avr_timing tc, td;
  const double t0 = now_ms();
  if (avr_compress_files_multi(ctxs.data(), (int)ctxs.size(), in.data(), len.data(), n, threads, comp.data(), &tc) != AVR_OK) return fail(avr_last_error());
  const double t1 = now_ms();
  for (auto& c : comp) { cin.push_back(c.data); clen.push_back(c.size); }
  if (avr_decompress_files_multi(ctxs.data(), (int)ctxs.size(), cin.data(), clen.data(), n, threads, dec.data(), &td) != AVR_OK) return fail(avr_last_error());
  const double t2 = now_ms();
  double total = 0;
  for (auto& d : data) total += (double)d.size();
  FILE* csv = fopen((dir + "/output/metrics.csv").c_str(), "w");
  if (!csv) return fail("Failed to open file: " + dir + "/output/metrics.csv");
  fprintf(csv, "File,Duration,Initial size (MB),Compressed size (MB),Compression rate (%%),Space saving (%%),Total time (ms),Compression time (ms),"
               "Compression speed (MB/s),Decompression time (ms),Decompression speed (MB/s),Video stream,Frames per second\n");
  int failed = 0;
  for (int i = 0; i < n; i++) {
    const bool same = dec[i].size == data[i].size() && !memcmp(dec[i].data, data[i].data(), data[i].size());
    if (!same) { failed++; continue; }                       // the reference leaves failed files out of the table (test.cpp:36-39)
    const std::string base = names[i].substr(names[i].find_last_of('/') + 1);
    write_out((dir + "/output/" + base).c_str(), dec[i].data, dec[i].size);
    const double mb = data[i].size() / 1e6, share = total > 0 ? data[i].size() / total : 0.0;
    const double cms = (t1 - t0) * share, dms = (t2 - t1) * share, rate = data[i].empty() ? 0.0 : 100.0 * comp[i].size / data[i].size();
    fprintf(csv, "\"%s\",n/a,%g,%g,%g,%g,%g,%g,%g,%g,%g,n/a,n/a\n", names[i].c_str(), mb, mb * rate / 100.0, rate, 100.0 - rate, cms + dms, cms,
            cms > 0 ? mb / (cms / 1e3) : 0.0, dms, dms > 0 ? mb / (dms / 1e3) : 0.0);
  }
  fclose(csv);
  FILE* log = fopen((dir + "/output/log.txt").c_str(), "w");
  if (log) {
    fprintf(log, "files %d  bytes %.0f  bins %llu\ncompress %.1f ms (parse %.1f, kernels %.1f)  decompress %.1f ms (kernel %.1f)\n", n, total,
            (unsigned long long)tc.bins, t1 - t0, tc.parse_ms, tc.kernel_ms, t2 - t1, td.kernel_ms);
    fclose(log);
  }
  printf("%d files, %.1f MB: compress %.1f MB/s, decompress %.1f MB/s, %.2f Gbins/s roundtrip\n", n, total / 1e6, total / 1e3 / (t1 - t0),
         total / 1e3 / (t2 - t1), tc.bins / 1e6 / (t2 - t0));
  if (failed) printf("Compress-decompress roundtrip failed on %d / %d files\n", failed, n);
  close_contexts(ctxs);
  return failed ? 1 : 0;
}

int main(int argc, char** argv) {
  int gpus = 1;
  if (const char* e = getenv("RECODE_GPUS")) gpus = atoi(e);
  {                                                  // --gpus N is ours, not the reference's: taken off the line before its parse
    int w = 1;
    for (int i = 1; i < argc; i++) {
      if (!strcmp(argv[i], "--gpus") && i + 1 < argc) { gpus = atoi(argv[++i]); continue; }
      argv[w++] = argv[i];
    }
    argc = w;
  }
  if (gpus < 1) gpus = 1;
  if (argc < 3) {
    fprintf(stderr, "Usage: %s [compress|decompress|roundtrip|test] <input> [output]\n", argv[0]);
    return 1;
  }
  std::string cmd = argv[1];
  int threads = (int)std::thread::hardware_concurrency();
  if (threads < 1) threads = 1;
  if (cmd == "test") return run_test(argv[2], threads, gpus);
  bool batch = cmd == "compress-batch" || cmd == "roundtrip-batch";
  if (!batch && cmd != "compress" && cmd != "decompress" && cmd != "roundtrip") {
    fprintf(stderr, "Unknown command: %s\n", cmd.c_str());
    return 1;
  }
  std::vector<std::string> names;
  if (batch) for (int i = 2; i < argc; i++) names.push_back(argv[i]); else names.push_back(argv[2]);
  const char* outname = (!batch && argc > 3) ? argv[3] : nullptr;
  std::vector<std::vector<uint8_t>> data(names.size());
  for (size_t i = 0; i < names.size(); i++)
    if (!read_file(names[i], &data[i])) return fail("Failed to open file: " + names[i]);
  std::vector<avr_context*> ctxs;
  if (!open_contexts(gpus, &ctxs)) return fail(avr_last_error());
  const int nctx = (int)ctxs.size();
  std::vector<const uint8_t*> in; std::vector<size_t> len;
  for (auto& d : data) { in.push_back(d.data()); len.push_back(d.size()); }
  const int n = (int)names.size();
  std::vector<avr_buffer> comp((size_t)n), dec((size_t)n);
  int rc = 0;
  if (cmd == "decompress") {
    if (avr_decompress_files_multi(ctxs.data(), nctx, in.data(), len.data(), 1, threads, dec.data(), nullptr) != AVR_OK) return fail(avr_last_error());
    if (!write_out(outname, dec[0].data, dec[0].size)) return fail("write failed");
    return 0;
  }
  avr_timing tc, td;
  double t0 = now_ms();
  if (avr_compress_files_multi(ctxs.data(), nctx, in.data(), len.data(), n, threads, comp.data(), &tc) != AVR_OK) return fail(avr_last_error());
  double t1 = now_ms();
  if (cmd == "compress") { if (!write_out(outname, comp[0].data, comp[0].size)) return fail("write failed"); return 0; }
  if (cmd == "compress-batch") {
    for (int i = 0; i < n; i++) if (!write_out((names[i] + ".avrec").c_str(), comp[i].data, comp[i].size)) return fail("write failed");
    return 0;
  }
  std::vector<const uint8_t*> cin; std::vector<size_t> clen;
  for (auto& c : comp) { cin.push_back(c.data); clen.push_back(c.size); }
  if (avr_decompress_files_multi(ctxs.data(), nctx, cin.data(), clen.data(), n, threads, dec.data(), &td) != AVR_OK) return fail(avr_last_error());
  double t2 = now_ms();
  size_t orig = 0, csz = 0;
  for (int i = 0; i < n; i++) {
    orig += data[i].size(); csz += comp[i].size;
    bool same = dec[i].size == data[i].size() && !memcmp(dec[i].data, data[i].data(), data[i].size());
    if (!same) { fprintf(stderr, "%s: roundtrip MISMATCH\n", names[i].c_str()); rc = 1; }
  }
  // report in the reference's terms (recode.cpp:1618-1635)
  printf("%s\n", rc ? "Compress-decompress roundtrip failed." : "Compress-decompress roundtrip succeeded:");
  printf(" compression ratio: %.2f%%\n", orig ? 100.0 * csz / orig : 0.0);
  printf(" compress: %.1f ms (%.1f MB/s)  decompress: %.1f ms (%.1f MB/s)  bins: %llu\n", t1 - t0, orig / 1e3 / (t1 - t0),
         t2 - t1, orig / 1e3 / (t2 - t1), (unsigned long long)tc.bins);
  close_contexts(ctxs);
  return rc;
}
