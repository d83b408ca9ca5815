// TEST INFRASTRUCTURE -- the lower boundary driven for real; never linked into the product.
//
// INTEGRATION.md's `gpu_compressor`, compiled: a third Driver for the UNMODIFIED reference's av_decoder<>
// (reference recode.cpp:80-237) next to its `compressor` / `decompressor`.  The reference's libavformat demux and
// libavcodec-hooks parse run as they are; every hook event goes to the C ABI's recorder sinks
// (avr_rec_frame_spec / slice_begin / mb_xy / begin_sub_mb / begin_coding_type / bin / ..., include/avrecode_b200.h);
// the bins are range-coded on the GPU (avr_cbatch_*); the payloads go back into the reference's own container
// (Recoded protobuf through oracle/shim/recode.pb.h, block sequence of recode.cpp:1282-1304, 1129-1131).
//
//   gpu_driver compress <input> <output.avrec>      any container libavformat demuxes (Annex-B, MP4, MKV, TS)
//   gpu_driver records <input>                      hooks -> recorder only (no GPU): prints the record counts
//
// What it settles (VERDICT r1, row b): the sink signatures ARE fillable from what init_decoder and the model hooks
// deliver.  `coded`, `payload_off` and `size` come from the block finder the Driver runs itself, exactly like
// compressor::cabac_decoder (recode.cpp:1143-1164); avr_slice_params is not needed on the compress side at all (the
// device never reads it there; decompress re-derives it from the literal header bytes of the container) and is passed
// with the picture size only.  av_decoder<> calls the model's setters directly (recode.cpp:173-202), so the Driver
// observes the model (frame flip, mb_coord, sub_mb_*) at the next bin, the way oracle/ref_harness.cpp does.
#define main recode_main
#include "recode.cpp"
#undef main
extern "C" {
#include "../include/avrecode_b200.h"
}

class gpu_compressor {
 public:
  gpu_compressor(const std::string& fn) : input_filename(fn) {
    if (av_file_map(fn.c_str(), &original_bytes, &original_size, 0, NULL) < 0) throw std::invalid_argument("Failed to open file: " + fn);
    rec = avr_record_new();
  }
  ~gpu_compressor() { avr_record_free(rec); av_file_unmap(original_bytes, original_size); }

  int read_packet(uint8_t* buffer_out, int size) {                       // recode.cpp:1134-1139
    size = std::min(size, int(original_size - read_offset));
    memcpy(buffer_out, &original_bytes[read_offset], size);
    read_offset += size;
    return size;
  }
  h264_model* get_model() { return &model; }

  // recode.cpp:1282-1304, with the offset kept for the recorder
  Recoded::Block* find_next_coded_block_and_emit_literal(const uint8_t* buf, int size, long* found_off) {
    uint8_t* found = static_cast<uint8_t*>(memmem(&original_bytes[prev_coded_block_end], read_offset - prev_coded_block_end, buf, size));
    if (found && size >= SURROGATE_MARKER_BYTES) {
      size_t gap = found - &original_bytes[prev_coded_block_end];
      out.add_block()->set_literal(&original_bytes[prev_coded_block_end], gap);
      prev_coded_block_end += gap + size;
      *found_off = found - original_bytes;
      Recoded::Block* b = out.add_block();
      b->set_length_parity(size & 1);
      if (size > 1) b->set_last_byte(&(buf[size - 1]), 1);
      return b;
    }
    Recoded::Block* b = out.add_block();
    b->set_skip_coded(true);
    b->set_size(size);
    *found_off = -1;
    return nullptr;
  }

  class cabac_decoder {
   public:
    cabac_decoder(gpu_compressor* c, CABACContext* ctx_in, const uint8_t* buf, int size) : c(c) {
      c->close_slice();
      h264_model* m = &c->model;
      // ModelHooks::frame_spec reached the model before this slice (h264_slice.c:570-573): a flip shows as a new cur_frame
      if (m->cur_frame != c->seen_frame || (int)m->frames[m->cur_frame].width() != c->seen_w || (int)m->frames[m->cur_frame].height() != c->seen_h) {
        c->seen_frame = m->cur_frame; c->seen_w = m->frames[m->cur_frame].width(); c->seen_h = m->frames[m->cur_frame].height();
        avr_rec_frame_spec(c->rec, ++c->frame_counter, c->seen_w, c->seen_h);
      }
      long off = -1;
      block = c->find_next_coded_block_and_emit_literal(buf, size, &off);
      avr_slice_params sp;
      memset(&sp, 0, sizeof sp);
      sp.mb_w = c->seen_w; sp.mb_h = c->seen_h;
      avr_rec_slice_begin(c->rec, block != nullptr, off >= 0 ? (uint64_t)off : 0, (uint32_t)size, &sp);
      c->slice_open = true;
      if (!block) {                                                      // recode.cpp:1146-1151
        ctx_in->coding_hooks = nullptr;
        ctx_in->coding_hooks_opaque = nullptr;
        ::ff_reset_cabac_decoder(ctx_in, buf, size);
        return;
      }
      block->set_size(size);
      c->coded_blocks.push_back(block);
      ctx = *ctx_in;
      ctx.coding_hooks = nullptr;
      ctx.coding_hooks_opaque = nullptr;
      ::ff_reset_cabac_decoder(&ctx, buf, size);
      base = (const uint8_t*)ctx_in + sizeof(CABACContext);              // H264SliceContext::cabac_state follows ::cabac (h264.h:510-511)
      last_x = last_y = -1;
      m->reset();
    }
    void sync() {
      h264_model* m = &c->model;
      if (m->mb_coord.mb_x != last_x || m->mb_coord.mb_y != last_y) {
        last_x = m->mb_coord.mb_x; last_y = m->mb_coord.mb_y;
        avr_rec_mb_xy(c->rec, last_x, last_y);                           // ModelHooks::mb_xy
      }
    }
    int get(uint8_t* state) { sync(); int s = ::ff_get_cabac(&ctx, state); avr_rec_bin(c->rec, int(state - base), s); return s; }
    int get_bypass() { sync(); int s = ::ff_get_cabac_bypass(&ctx); avr_rec_bin(c->rec, 1024, s); return s; }
    int get_terminate() { sync(); int s = ::ff_get_cabac_terminate(&ctx) != 0; avr_rec_bin(c->rec, 1025, s); return s; }
    void begin_coding_type(CodingType ct, int, int, int) {
      if (!block) return;
      sync();
      h264_model* m = &c->model;                                         // ModelHooks::begin_sub_mb set these (recode.cpp:182-189)
      avr_rec_begin_sub_mb(c->rec, m->sub_mb_cat, m->mb_coord.scan8_index, m->sub_mb_size, m->sub_mb_is_dc, m->sub_mb_chroma422);
      avr_rec_begin_coding_type(c->rec, (int)ct);
    }
    void end_coding_type(CodingType ct) {
      if (!block) return;
      avr_rec_end_coding_type(c->rec, (int)ct);
      avr_rec_end_sub_mb(c->rec);
    }
   private:
    gpu_compressor* c;
    Recoded::Block* block = nullptr;
    CABACContext ctx;
    const uint8_t* base = nullptr;
    int last_x = -1, last_y = -1;
  };

  void close_slice() { if (slice_open) { avr_rec_slice_end(rec); slice_open = false; } }

  void record() {
    {
      av_decoder<gpu_compressor> d(this, input_filename);
      d.decode_video();
    }
    close_slice();
  }
  void print_counts() {
    uint64_t c[6];
    avr_record_counts(rec, c);
    printf("bins %llu blocks %llu slices %llu spans %llu pictures %llu out_of_domain %llu\n", (unsigned long long)c[0], (unsigned long long)c[1],
           (unsigned long long)c[2], (unsigned long long)c[3], (unsigned long long)c[4], (unsigned long long)c[5]);
  }
  int run(std::ostream& out_stream) {
    record();
    // the hot path: records -> range-coded payloads on the GPU
    avr_context* ctx = nullptr;
    if (avr_context_create(0, &ctx) != AVR_OK) { fprintf(stderr, "gpu_driver: %s\n", avr_last_error()); return 1; }
    avr_cbatch* b = nullptr;
    const avr_record* recs[1] = {rec};
    int rc = avr_cbatch_create(ctx, recs, 1, &b);
    if (rc == AVR_OK) rc = avr_cbatch_run(b, AVR_MODE_PIPELINE);
    if (rc == AVR_OK) rc = avr_cbatch_fetch(b);
    if (rc != AVR_OK) { fprintf(stderr, "gpu_driver: %s\n", avr_last_error()); return 1; }
    for (size_t i = 0; i < coded_blocks.size(); i++) {
      const uint8_t* p; uint32_t n;
      if (avr_cbatch_payload(b, 0, (uint32_t)i, &p, &n) != AVR_OK) { fprintf(stderr, "gpu_driver: payload %zu missing\n", i); return 1; }
      coded_blocks[i]->set_cabac(p, n);
    }
    avr_cbatch_destroy(b);
    avr_context_destroy(ctx);
    out.add_block()->set_literal(&original_bytes[prev_coded_block_end], original_size - prev_coded_block_end);   // recode.cpp:1129-1131
    out_stream << out.SerializeAsString();
    return 0;
  }

 private:
  std::string input_filename;
  uint8_t* original_bytes = nullptr;
  size_t original_size = 0;
  int read_offset = 0;
  int prev_coded_block_end = 0;
  h264_model model;
  Recoded out;
  avr_record* rec = nullptr;
  std::vector<Recoded::Block*> coded_blocks;
  bool slice_open = false;
  int seen_frame = -1, seen_w = -1, seen_h = -1, frame_counter = 0;
};

int main(int argc, char** argv) {
  av_register_all();
  if (argc >= 3 && std::string(argv[1]) == "records") {
    try { gpu_compressor c(argv[2]); c.record(); c.print_counts(); return 0; }
    catch (const std::exception& e) { fprintf(stderr, "Exception (%s): %s\n", typeid(e).name(), e.what()); return 1; }
  }
  if (argc < 4 || std::string(argv[1]) != "compress") { fprintf(stderr, "usage: %s compress <input> <output.avrec> | records <input>\n", argv[0]); return 2; }
  try {
    gpu_compressor c(argv[2]);
    std::ofstream out(argv[3], std::ios::binary);
    return c.run(out);
  } catch (const std::exception& e) {
    fprintf(stderr, "Exception (%s): %s\n", typeid(e).name(), e.what());
    return 1;
  }
}
