// Native compress-side front end: Annex-B H.264 -> record arrays, no libavcodec.
// Does what libavcodec-hooks + compressor::cabac_decoder do together in the reference
// (recode.cpp:1141-1199: a private standard CABAC decode of the ORIGINAL bytes whose bins are
// handed to the model), plus compressor::find_next_coded_block_and_emit_literal's decision
// (recode.cpp:1282-1304) of which slices are recoded and which stay literal.
#include "avr_record.h"
#include <string.h>
#include <functional>
#include "cabac_std.h"
#include "h264_headers.h"

namespace avr {
namespace {

struct RecordIO {
  static constexpr bool kFastResidual = false;
  int residual_block(int, int, int, int, int, int, int) { return 0; }
  CabacDecoder dec;
  uint8_t st[1024];
  Recorder& rec;
  explicit RecordIO(Recorder& r) : rec(r) {}
  int get(int ctx) { int s = dec.decode(&st[ctx]); rec.bin(ctx, s); return s; }
  int bypass(int) { int s = dec.bypass(); rec.bin(AVR_CTX_BYPASS, s); return s; }
  int terminate(int) { int s = dec.terminate(); rec.bin(AVR_CTX_TERMINATE, s); return s; }
  void mb_begin(int x, int y) { rec.mb_xy(x, y); }
  void sub_mb_begin(int cat, int n, int mc, int dc, int c422) { rec.begin_sub_mb(cat, n, mc, dc, c422); }
  void sub_mb_end() {}
  void sigmap_begin() { rec.begin_sigmap(); }
  void sigmap_end() { rec.end_sigmap(); }
  void hint(int, int) {}
};

}  // namespace

// The block finder's memmem (recode.cpp:1283-1286): FIRST occurrence of the unescaped payload in the original bytes
// not yet emitted, here bounded by the end of the slice's own NAL.  A payload without emulation-prevention bytes sits
// at its raw offset, so that place is compared first (memcmp) and the general search (glibc's two-way memmem, slow
// on needles of tens of KB) runs only when that fails or when the payload's first 8 bytes also occur earlier in the
// window, i.e. only when an earlier occurrence is possible at all.
static bool find_payload(const uint8_t* data, size_t n, size_t prev_coded_end, const SliceLoc& loc, size_t* found_at) {
  const size_t hay_end = loc.raw_off + loc.size <= n ? loc.raw_off + loc.size : n;
  if (loc.raw_off + loc.size <= n && memcmp(data + loc.raw_off, loc.payload, loc.size) == 0) {
    const size_t before = loc.raw_off - prev_coded_end;        // bytes in front of the raw offset
    const size_t span = before + 7 < hay_end - prev_coded_end ? before + 7 : hay_end - prev_coded_end;
    const void* m8 = before ? memmem(data + prev_coded_end, span, loc.payload, 8) : nullptr;   // a match must start before raw_off
    if (!m8) { *found_at = loc.raw_off; return true; }
  }
  const void* m = memmem(data + prev_coded_end, hay_end - prev_coded_end, loc.payload, loc.size);
  if (!m) return false;
  *found_at = (size_t)((const uint8_t*)m - data);
  return true;
}

int scan_annexb_slices(const uint8_t* data, size_t n, const std::function<int(const SliceLoc&)>& on_slice, int* unsupported) {
  std::vector<Sps> sps(32);
  std::vector<Pps> pps(256);
  std::vector<uint8_t> rbsp;
  std::vector<uint32_t> map;
  size_t pos = 0;
  NalSpan nal;
  while (next_nal(data, n, pos, &nal)) {
    pos = nal.end;
    const uint8_t hdr = data[nal.start];
    const int type = hdr & 31, ref_idc = hdr >> 5;
    if (type != 1 && type != 5 && type != 7 && type != 8) continue;
    unescape_nal(data + nal.start + 1, nal.end - nal.start - 1, &rbsp, &map);
    size_t len = rbsp.size();
    while (len > 0 && rbsp[len - 1] == 0) len--;   // decode_nal_units trailing-zero strip
    if (len == 0) continue;
    if (type == 7) { Sps s; int id; if (parse_sps(rbsp.data(), len, &s, &id)) sps[id] = s; continue; }
    if (type == 8) { Pps p; int id; if (parse_pps(rbsp.data(), len, sps.data(), &p, &id)) pps[id] = p; continue; }
    // slice
    int r = 0; { uint8_t v = rbsp[len - 1]; while (!(v & 1)) { v >>= 1; r++; } r++; }  // decode_rbsp_trailing
    const size_t bit_length = 8 * len - (size_t)r;
    SliceLoc loc;
    if (!parse_slice_header(rbsp.data(), bit_length, type, ref_idc, sps.data(), pps.data(), &loc.sh)) continue;
    if (!loc.sh.cabac || !loc.sh.supported) { if (unsupported) ++*unsupported; continue; }   // CAVLC: nothing to recode; syntax not covered
    const size_t start = (loc.sh.header_bits + 7) / 8;
    if (bit_length <= 8 * start) continue;
    loc.size = (bit_length - 8 * start + 7) / 8;
    loc.payload = rbsp.data() + start;
    // libavcodec's CABAC reader may look one or two bytes past `size` (a lone 0x80 stop-bit byte
    // is excluded from size but still in its buffer, cabac_functions.h refill): give ours the same view
    loc.payload_avail = rbsp.size() - start;
    loc.raw_off = nal.start + 1 + rbsp_to_raw(map, start);
    int rc = on_slice(loc);
    if (rc) return rc;
  }
  return AVR_OK;
}

// Host half of the device-side parse (avr_parse.h): everything record_annexb does EXCEPT the CABAC
// decode -- NAL scan, headers, update_frame_spec bookkeeping, the block-finder decision.
int scan_annexb(const uint8_t* data, size_t n, FileRecord* out, std::vector<ScanSlice>* coded) {
  Recorder rec(out, nullptr);
  size_t prev_coded_end = 0;
  int rc = scan_annexb_slices(data, n, [&](const SliceLoc& loc) -> int {
    const SliceHeaderInfo& sh = loc.sh;
    if (rec.starts_picture(sh.sp.first_mb, sh.sp.mb_w, sh.sp.mb_h)) rec.frame_spec(sh.frame_num, sh.sp.mb_w, sh.sp.mb_h);
    bool found = false;
    size_t found_at = 0;
    if (loc.size >= 8 && loc.raw_off >= prev_coded_end) found = find_payload(data, n, prev_coded_end, loc, &found_at);
    rec.slice_begin(found, found ? found_at : loc.raw_off, (uint32_t)loc.size, sh.sp);
    if (!found) return 0;
    prev_coded_end = found_at + loc.size;
    ScanSlice s;
    s.raw_off = found_at; s.size = (uint32_t)loc.size; s.frame = rec.cur_frame(); s.sp = sh.sp;
    s.slice_index = (int32_t)out->slices.size() - 1;
    for (int k = 0; k < 4; k++) s.tail[k] = loc.size + k < loc.payload_avail ? loc.payload[loc.size + k] : 0;
    coded->push_back(s);
    rec.slice_end();
    return 0;
  }, &out->skipped_slices);
  out->error = rc;
  return rc;
}

int record_annexb(const uint8_t* data, size_t n, FileRecord* out, FILE* trace) {
  Recorder rec(out, trace);
  std::vector<MbInfo> row;
  uint32_t uid = 0, cover_frame = ~0u, cover_end = 0;
  size_t prev_coded_end = 0;
  int rc = scan_annexb_slices(data, n, [&](const SliceLoc& loc) -> int {
    const SliceHeaderInfo& sh = loc.sh;
    if (rec.starts_picture(sh.sp.first_mb, sh.sp.mb_w, sh.sp.mb_h)) rec.frame_spec(sh.frame_num, sh.sp.mb_w, sh.sp.mb_h);  // h264_frame_start
    // recode.cpp:1283-1286: memmem of the unescaped payload in the not-yet-emitted original bytes
    bool found = false;
    size_t found_at = 0;
    if (loc.size >= 8 && loc.raw_off >= prev_coded_end) found = find_payload(data, n, prev_coded_end, loc, &found_at);
    rec.slice_begin(found, found ? found_at : loc.raw_off, (uint32_t)loc.size, sh.sp);
    if (!found) return 0;
    const size_t prev_coded_end_before = prev_coded_end;
    prev_coded_end = found_at + loc.size;
    if ((int)row.size() < sh.sp.mb_w) { row.resize((size_t)sh.sp.mb_w); for (auto& m : row) m.uid = 0; }
    RecordIO io(rec);
    io.dec.init(loc.payload, loc.payload_avail);
    cabac_init_states(io.st, sh.sp.slice_type == SLICE_I, sh.sp.cabac_init_idc, sh.sp.slice_qp);
    SliceParser<RecordIO> parser(io, sh.sp, row.data(), ++uid);
    int prc = parser.run();
    // A walk that does not end on the last bit of slice_data() is not one the decompressor can reproduce (damaged data, syntax
    // the template does not cover): for a conforming slice the last bit read is rbsp_stop_one_bit (cabac_std.h), which `size` of
    // libavcodec's init_decoder call does not count.
    if (prc >= 0 && (io.dec.bits_consumed() + 6) / 8 != loc.size) prc = SYN_E_LENGTH;
    // ... and through end_of_slice_flag = 1, not by running out of macroblocks (a damaged last macroblock can end on the right
    // bit with the flag read as 0): the decompress chain could not close such a slice
    if (prc >= 0 && (out->bins.empty() || out->bins.back() != AVR_BIN(1, AVR_CTX_TERMINATE, AVR_KIND_DEFAULT))) prc = SYN_E_LENGTH;
    if (prc < 0) {                         // also I_PCM or a value out of range: this slice alone stays literal
      rec.slice_abort();
      prev_coded_end = prev_coded_end_before;
      return 0;
    }
    rec.slice_end();
    // Slices of one picture must not overlap: a damaged slice can run past the next one's first macroblock and still end cleanly, and
    // the slice-parallel compress passes need every macroblock position coded at most once per picture store.  The damaged one is
    // already recorded by then: the whole file is stored literally (the caller's fallback for a non-zero result).
    if (rec.cur_frame() == cover_frame && (uint32_t)sh.sp.first_mb < cover_end) return AVR_E_UNSUPPORTED;
    cover_frame = rec.cur_frame(); cover_end = (uint32_t)sh.sp.first_mb + (uint32_t)prc;
    return 0;
  }, &out->skipped_slices);
  out->error = rc;
  return rc;
}

}  // namespace avr
