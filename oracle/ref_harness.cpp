// TEST INFRASTRUCTURE -- drives the UNMODIFIED reference classes; never linked into the product.
//
// Includes the reference's recode.cpp as-is (main renamed) and adds a third `Driver` for its
// av_decoder<> template (reference recode.cpp:80-237) next to `compressor`/`decompressor`:
//   ref_harness trace <file.h264>      : dump every hook event libavcodec-hooks delivers
//                                        (the ground truth for avrecode-ms_b200/csrc/h264_syntax.h)
// Output, one event per line:
//   F cur_frame w h            (after a frame flip is observed)
//   S size off                 (init_decoder: new slice, `size` bytes of CABAC payload; off = file
//                               offset found by the compressor's memmem rule, recode.cpp:1282-1304,
//                               or -1 when the slice stays literal / skip_coded: its bins are not traced)
//   I <2048 hex>               (the 1024 CABAC context states at the slice's first bin)
//   M x y                      (macroblock change)
//   B cat n max_coeff is_dc c422   (begin_sub_mb observed)
//   [ / ]                      (begin/end_coding_type SIGNIFICANCE_MAP)
//   c <ctxIdx> <sym> | b <sym> | t <sym>
//   ref_harness arith <digit_bits 8|16> <model 0|1> <nstreams> <nbins> <bits file>
//                                      : the reference's arithmetic_code<uint64_t,Digit> exactly as
//                                        test/arithmetic_code.cpp:93-110 drives it (make_encoder over a
//                                        vector<uint8_t>, put(bit, lambda), finish, make_decoder, get), one
//                                        stream after the other; model 0 = the test's `range/2`, model 1 = one
//                                        adaptive estimator per stream with recode.cpp's counts (:823-827,
//                                        :1037-1054).  Prints "A <hex>" per stream, exit 1 if its own decoder
//                                        does not return the bits.
#define main recode_main
#include "recode.cpp"
#undef main

class tracer {
 public:
  tracer(const std::string& fn) : input_filename(fn) {
    if (av_file_map(fn.c_str(), &bytes, &size, 0, NULL) < 0) throw std::invalid_argument("open " + fn);
  }
  void run() {
    av_decoder<tracer> d(this, input_filename);
    d.decode_video();
  }
  int read_packet(uint8_t* out, int n) {
    n = std::min(n, int(size - (size_t)off));
    memcpy(out, bytes + off, n);
    off += n;
    return n;
  }
  h264_model* get_model() { return &model; }

  class cabac_decoder {
   public:
    cabac_decoder(tracer* t, CABACContext* ctx_in, const uint8_t* buf, int size) : t(t) {
      h264_model* m = &t->model;
      printf("F %d %u %u\n", m->cur_frame, m->frames[m->cur_frame].width(), m->frames[m->cur_frame].height());
      // same decision as compressor::find_next_coded_block_and_emit_literal (recode.cpp:1282-1304)
      uint8_t* found = static_cast<uint8_t*>(memmem(t->bytes + t->prev_end, t->off - t->prev_end, buf, size));
      if (!(found && size >= SURROGATE_MARKER_BYTES)) {
        printf("S %d -1\n", size);
        ctx_in->coding_hooks = nullptr;
        ctx_in->coding_hooks_opaque = nullptr;
        ::ff_reset_cabac_decoder(ctx_in, buf, size);
        skipped = true;
        return;
      }
      printf("S %d %ld\n", size, (long)(found - t->bytes));
      t->prev_end = (found - t->bytes) + size;
      ctx = *ctx_in;
      ctx.coding_hooks = nullptr;
      ctx.coding_hooks_opaque = nullptr;
      ::ff_reset_cabac_decoder(&ctx, buf, size);
      base = (const uint8_t*)ctx_in + sizeof(CABACContext);  // H264SliceContext::cabac_state follows ::cabac (h264.h:510-511)
      last_x = last_y = -1;
    }
    void sync() {
      h264_model* m = &t->model;
      if (!states_dumped) {
        states_dumped = true;
        printf("I ");
        for (int i = 0; i < 1024; i++) printf("%02x", base[i]);
        printf("\n");
      }
      if (m->mb_coord.mb_x != last_x || m->mb_coord.mb_y != last_y) {
        last_x = m->mb_coord.mb_x; last_y = m->mb_coord.mb_y;
        printf("M %d %d\n", last_x, last_y);
      }
    }
    int get(uint8_t* state) {
      sync();
      int s = ::ff_get_cabac(&ctx, state);
      printf("c %d %d\n", int(state - base), s);
      return s;
    }
    int get_bypass() { sync(); int s = ::ff_get_cabac_bypass(&ctx); printf("b %d\n", s); return s; }
    int get_terminate() { sync(); int s = ::ff_get_cabac_terminate(&ctx) != 0; printf("t %d\n", s); return s; }
    void begin_coding_type(CodingType ct, int, int, int) {
      if (skipped) return;
      sync();
      h264_model* m = &t->model;
      printf("B %d %d %d %d %d\n", m->sub_mb_cat, m->mb_coord.scan8_index, m->sub_mb_size, m->sub_mb_is_dc, m->sub_mb_chroma422);
      printf("[\n");
    }
    void end_coding_type(CodingType ct) { if (!skipped) printf("]\n"); }
   private:
    tracer* t;
    CABACContext ctx;
    const uint8_t* base;
    int last_x, last_y;
    bool skipped = false, states_dumped = false;
  };

 private:
  std::string input_filename;
  uint8_t* bytes = nullptr;
  size_t size = 0;
  long off = 0, prev_end = 0;
  h264_model model;
};

struct arith_counts {                      // estimator of recode.cpp:1064 with the update of :1037-1054 (limit 0x60)
  int pos = 1, neg = 1;
  void update(int symbol) {
    if (symbol) pos++; else neg++;
    if (pos + neg > 0x60) { pos = (pos + 1) / 2; neg = (neg + 1) / 2; }
  }
};
template <typename Digit>
static int arith_streams(int model, long nstreams, long nbins, const std::vector<uint8_t>& bits) {
  typedef arithmetic_code<uint64_t, Digit> code;
  for (long s = 0; s < nstreams; s++) {
    const uint8_t* b = bits.data() + s * nbins;
    std::vector<uint8_t> out;
    {
      auto encoder = make_encoder<code>(&out);
      arith_counts c;
      for (long i = 0; i < nbins; i++) {
        int bit = b[i] != 0;
        if (model == 0) encoder.put(bit, [](uint64_t range) { return range / 2; });
        else { encoder.put(bit, [&c](uint64_t range) { return (range / (c.pos + c.neg)) * c.pos; }); c.update(bit); }
      }
      encoder.finish();
    }
    printf("A ");
    for (uint8_t x : out) printf("%02x", x);
    printf("\n");
    auto decoder = make_decoder<code>(out);
    arith_counts c;
    for (long i = 0; i < nbins; i++) {
      int bit;
      if (model == 0) bit = decoder.get([](uint64_t range) { return range / 2; });
      else { bit = decoder.get([&c](uint64_t range) { return (range / (c.pos + c.neg)) * c.pos; }); c.update(bit); }
      if (bit != (b[i] != 0)) { fprintf(stderr, "stream %ld: mismatch at bit %ld\n", s, i); return 1; }
    }
  }
  return 0;
}

int main(int argc, char** argv) {
  if (argc >= 7 && std::string(argv[1]) == "arith") {
    const int digit_bits = atoi(argv[2]), model = atoi(argv[3]);
    const long nstreams = atol(argv[4]), nbins = atol(argv[5]);
    std::vector<uint8_t> bits((size_t)(nstreams * nbins));
    FILE* f = fopen(argv[6], "rb");
    if (!f || fread(bits.data(), 1, bits.size(), f) != bits.size()) { fprintf(stderr, "cannot read %s\n", argv[6]); return 2; }
    fclose(f);
    return digit_bits == 16 ? arith_streams<uint16_t>(model, nstreams, nbins, bits) : arith_streams<uint8_t>(model, nstreams, nbins, bits);
  }
  av_register_all();
  av_log_set_level(AV_LOG_ERROR);
  if (argc >= 3 && std::string(argv[1]) == "trace") {
    try {
      tracer t(argv[2]);
      t.run();
    } catch (const std::exception& e) {
      fprintf(stderr, "Exception: %s\n", e.what());
      return 1;
    }
    return 0;
  }
  fprintf(stderr, "usage: %s trace <file> | arith <digit_bits> <model> <nstreams> <nbins> <bits file>\n", argv[0]);
  return 1;
}
