// avrtool: host-side debugging / test utility for the native front end.
//   avrtool trace <in.h264>      hook-event trace in the format of oracle/ref_harness.cpp
//   avrtool stats <in.h264>      record counts
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "avr_record.h"

static bool read_file(const char* fn, std::vector<uint8_t>* out) {
  FILE* f = fopen(fn, "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
  out->resize((size_t)n);
  size_t got = fread(out->data(), 1, (size_t)n, f);
  fclose(f);
  return got == (size_t)n;
}

int main(int argc, char** argv) {
  if (argc < 3) { fprintf(stderr, "usage: %s trace|stats <file>\n", argv[0]); return 1; }
  std::string cmd = argv[1];
  std::vector<uint8_t> data;
  if (!read_file(argv[2], &data)) { fprintf(stderr, "cannot read %s\n", argv[2]); return 1; }
  avr::FileRecord rec;
  int rc = avr::record_annexb(data.data(), data.size(), &rec, cmd == "trace" ? stdout : nullptr);
  if (cmd == "stats" || rc != 0)
    fprintf(stderr, "rc=%d bins=%zu blocks=%zu slices=%zu spans=%zu frames=%zu ood=%d skipped=%d\n", rc, rec.bins.size(),
            rec.blocks.size(), rec.slices.size(), rec.spans.size(), rec.frames.size(), rec.out_of_domain, rec.skipped_slices);
  return rc ? 1 : 0;
}
