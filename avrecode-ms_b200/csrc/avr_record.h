// Host recorder: turns the hook event stream of one file into the flat record arrays of
// avr_common.h.  Its methods are 1:1 with the reference's AVCodecHooks callbacks
// (ffmpeg/libavcodec/coding_hooks.h:27-55, trampolines at recode.cpp:150-215), so it can be
// driven either by libavcodec-hooks itself or by the native parser
// (avr_record.cpp).
#pragma once
#include <stdlib.h>
#include <stdint.h>
#include <stdio.h>
#include <vector>
#include "avr_common.h"
#include "h264_syntax.h"

namespace avr {

// One init_decoder call, in decode order (reference recode.cpp:1143-1164, 1282-1304).
struct SliceSpan {
  uint64_t payload_off;   // file offset of the first CABAC byte
  uint32_t size;          // CABAC payload size libavcodec reports (h264_slice.c:2361-2364)
  int32_t slice_index;    // index into FileRecord::slices, or -1 if skip_coded
  uint32_t frame;         // frame-flip index current at this slice
  SliceParams sp;         // header fields the device-side parser needs on decompress
};

struct FileRecord {
  std::vector<uint16_t> bins;
  std::vector<AvrBlock> blocks;
  std::vector<AvrSlice> slices;
  std::vector<AvrFrame> frames;
  std::vector<SliceSpan> spans;
  uint64_t source_bins = 0;   // get/get_bypass/get_terminate calls on coded slices (SURVEY 8d "bin")
  uint64_t emitted = 0;       // symbols that go through the range coder: bins minus EOB bins plus nnz side-channel bits
  int out_of_domain = 0;      // bit0: consecutive pictures share frame_num and AVR_STRICT_DOMAIN=1: not compressible (literal)
                              // bit1: they do, and the picture was given its own store (extended domain, see frame_spec)
  int error = 0;              // AVR_E_* from the parse
  int skipped_slices = 0;     // slices left uncoded one by one: walk failed (slice_abort) or header outside the supported syntax
};

inline bool strict_domain() {
  static const bool on = [] { const char* e = getenv("AVR_STRICT_DOMAIN"); return e && e[0] == '1'; }();
  return on;
}

class Recorder {
 public:
  explicit Recorder(FileRecord* r, FILE* trace = nullptr) : r_(r), trace_(trace) {}

  // ModelHooks::frame_spec -> h264_model::update_frame_spec (recode.cpp:831-850)
  void frame_spec(int frame_num, int mb_w, int mb_h) {
    if (have_ && mb_w == w_ && mb_h == h_ && frame_num == frame_num_) {
      // The reference does NOT flip here (recode.cpp:832-834), keeps the previous picture's significance flags and
      // counts them (:942-954): its compressor's output for such streams (non-reference B pictures, all-IDR streams)
      // cannot be decompressed by anyone, its own decompressor included (as shipped it aborts on the assert at
      // :1040).  There is no reference result to be identical to, so by default every picture gets its own zeroed store
      // like any other -- both directions of this library agree on that rule and the stream round-trips.
      // AVR_STRICT_DOMAIN=1 keeps to the reference's domain instead: the file is stored as a literal block.
      if (strict_domain()) { r_->out_of_domain |= 1; return; }
      r_->out_of_domain |= 2;
    }
    AvrFrame f;
    f.mb_w = (uint16_t)mb_w; f.mb_h = (uint16_t)mb_h;
    f.prev_valid = (have_ && mb_w == w_ && mb_h == h_) ? 1u : 0u;
    f.mb_base = r_->frames.empty() ? 0u : r_->frames.back().mb_base + (uint32_t)r_->frames.back().mb_w * r_->frames.back().mb_h;
    r_->frames.push_back(f);
    have_ = true; w_ = mb_w; h_ = mb_h; frame_num_ = frame_num;
  }
  uint32_t cur_frame() const { return r_->frames.empty() ? 0u : (uint32_t)r_->frames.size() - 1; }
  // A slice starts a picture when it is the picture's first (first_mb == 0), and also when no picture has been started yet, when the
  // picture size has changed under it, or when its first_mb does not lie beyond the previous slice's (CABAC profiles have no
  // arbitrary slice order, so that slice belongs to a NEW picture whose first slice is missing or was refused).  Streams cut in
  // the middle of a picture or with lost slices then never leave a slice without a picture store of its size, and never code a
  // macroblock position twice into one store -- the slice-parallel compress passes rely on every position being visited once.
  // avr_container.h plan_decompress applies the same rule from the same header fields.
  bool starts_picture(int first_mb, int mb_w, int mb_h) {
    const bool yes = first_mb == 0 || !have_ || mb_w != w_ || mb_h != h_ || first_mb <= last_first_mb_;
    last_first_mb_ = first_mb;
    return yes;
  }

  // CABACCodingHooks::init_decoder. coded=false is the skip_coded path (hooks disabled).
  void slice_begin(bool coded, uint64_t payload_off, uint32_t size, const SliceParams& sp) {
    SliceSpan s;
    s.payload_off = payload_off; s.size = size; s.sp = sp; s.frame = cur_frame();
    s.slice_index = coded ? (int32_t)r_->slices.size() : -1;
    r_->spans.push_back(s);
    coded_ = coded;
    if (trace_) {
      fprintf(trace_, "F %d %d %d\n", (int)(r_->frames.size() & 1), w_, h_);
      fprintf(trace_, "S %u\n", size);
    }
    if (coded) {
      AvrSlice sl;
      sl.bin_start = sl.bin_end = (uint32_t)r_->bins.size();
      sl.block_start = sl.block_end = (uint32_t)r_->blocks.size();
      r_->slices.push_back(sl);
      mark_emitted_ = r_->emitted; mark_source_ = r_->source_bins;
    }
    in_sigmap_ = false;
  }
  // Per-slice fallback: the walk of the slice that slice_begin(coded = true) opened failed (I_PCM, a value out of range in a
  // damaged slice).  Its records are dropped and its span becomes a skip_coded one (recode.cpp:1296-1303: the bytes stay in the
  // literal, a size-only block keeps the block sequence aligned), so the model never sees the slice -- on decompress it is not
  // walked either -- and the rest of the file is still coded.
  void slice_abort() {
    if (!coded_) return;
    const AvrSlice sl = r_->slices.back();
    r_->slices.pop_back();
    r_->bins.resize(sl.bin_start);
    r_->blocks.resize(sl.block_start);
    r_->emitted = mark_emitted_; r_->source_bins = mark_source_;
    r_->spans.back().slice_index = -1;
    r_->skipped_slices++;
    coded_ = false; in_sigmap_ = false;
  }
  void slice_end() {
    if (!coded_) return;
    r_->slices.back().bin_end = (uint32_t)r_->bins.size();
    r_->slices.back().block_end = (uint32_t)r_->blocks.size();
  }
  void mb_xy(int x, int y) {
    mb_addr_ = (uint32_t)(x + y * w_);
    if (trace_) fprintf(trace_, "M %d %d\n", x, y);
  }
  void begin_sub_mb(int cat, int n, int max_coeff, int is_dc, int c422) {
    cat_ = cat; n_ = n; max_coeff_ = max_coeff; is_dc_ = is_dc; c422_ = c422;
  }
  void begin_sigmap() {
    if (trace_) fprintf(trace_, "B %d %d %d %d %d\n[\n", cat_, n_, max_coeff_, is_dc_, c422_);
    if (!coded_) return;
    AvrBlock b;
    b.bin_start = (uint32_t)r_->bins.size();
    b.mb_addr = mb_addr_;
    b.frame = cur_frame();
    b.n = (uint8_t)n_;
    b.cat_flags = (uint8_t)(cat_ | (is_dc_ << 4) | (c422_ << 5));
    b.max_coeff = (uint8_t)max_coeff_;
    b.nbins = 0;
    r_->blocks.push_back(b);
    in_sigmap_ = true; expect_eob_ = false;
    r_->emitted += (uint64_t)avr_nz_bits(max_coeff_);
  }
  void end_sigmap() {
    if (trace_) fprintf(trace_, "]\n");
    in_sigmap_ = false;
  }
  // get / get_bypass / get_terminate
  void bin(int ctx, int sym) {
    if (trace_) {
      if (ctx == AVR_CTX_BYPASS) fprintf(trace_, "b %d\n", sym);
      else if (ctx == AVR_CTX_TERMINATE) fprintf(trace_, "t %d\n", sym);
      else fprintf(trace_, "c %d %d\n", ctx, sym);
    }
    if (!coded_) return;
    int kind = AVR_KIND_DEFAULT;
    if (in_sigmap_) {
      kind = expect_eob_ ? AVR_KIND_SIG_EOB : AVR_KIND_SIG_MAP;
      expect_eob_ = !expect_eob_ && sym;   // MAP bin = 1 -> next is its EOB bin (recode.cpp:998-1000)
      r_->blocks.back().nbins++;
    }
    r_->bins.push_back(AVR_BIN(sym, ctx, kind));
    if (kind != AVR_KIND_SIG_EOB) r_->emitted++;
    r_->source_bins++;
  }

 private:
  FileRecord* r_;
  FILE* trace_;
  bool have_ = false; int w_ = 0, h_ = 0, frame_num_ = 0, last_first_mb_ = -1;
  bool coded_ = false, in_sigmap_ = false, expect_eob_ = false;
  uint64_t mark_emitted_ = 0, mark_source_ = 0;
  uint32_t mb_addr_ = 0;
  int cat_ = 0, n_ = 0, max_coeff_ = 0, is_dc_ = 0, c422_ = 0;
};

}  // namespace avr
#include <functional>
#include "h264_headers.h"
namespace avr {

// One CABAC slice located in an Annex-B stream: what libavcodec knows when it calls
// CABACCodingHooks::init_decoder (h264_slice.c:2357-2364).
struct SliceLoc {
  SliceHeaderInfo sh;
  size_t raw_off = 0;          // file offset of the first CABAC byte (if the payload has no escapes before it)
  size_t size = 0;             // CABAC payload bytes libavcodec reports
  const uint8_t* payload = nullptr;   // unescaped payload (valid during the callback)
  size_t payload_avail = 0;    // unescaped bytes available from `payload` to the end of the NAL
};
// Walks the NALs, tracks SPS/PPS, calls on_slice for every CABAC slice with a non-empty payload.
// Returns AVR_OK or the callback's non-zero result.
// Slices whose header is outside the supported syntax (CAVLC, SP/SI, field/MBAFF, FMO, 4:2:2+, high bit depth) are passed over --
// no callback, *unsupported counts them: they stay literal bytes and take no block of the container, on both directions alike.
int scan_annexb_slices(const uint8_t* data, size_t n, const std::function<int(const SliceLoc&)>& on_slice, int* unsupported = nullptr);

// One coded slice found by scan_annexb: what the device-side parse (avr_parse.h) needs.
struct ScanSlice {
  uint64_t raw_off;        // file offset of the first CABAC byte (the block-finder's match)
  uint32_t size;
  uint8_t tail[4];         // unescaped bytes after the payload that libavcodec's CABAC reader may peek at
  uint32_t frame;          // picture (flip) index
  int32_t slice_index;     // index among the file's coded slices
  SliceParams sp;
};
int scan_annexb(const uint8_t* data, size_t n, FileRecord* out, std::vector<ScanSlice>* coded);

// Native front end: parse an Annex-B H.264 byte stream without libavcodec.
// Returns AVR_OK, AVR_E_UNSUPPORTED (stream uses syntax outside h264_syntax.h) or AVR_E_CORRUPT.
int record_annexb(const uint8_t* data, size_t n, FileRecord* out, FILE* trace = nullptr);

}  // namespace avr
