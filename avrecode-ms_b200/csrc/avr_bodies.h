// Kernel bodies of the sequential ("chain") paths, one call per file.  __global__ wrappers
// live in avr_kernels.cu; tests/ compiles this header for the host and runs it as an
// emulated grid.
#pragma once
#include "avr_chain.h"
#include "h264_syntax.h"
#include "h264_std_tables.h"

// Constant tables every kernel needs, uploaded once per context (avr_api.cu).
struct AvrHotTabs {            // the per-bin tables: copied to shared memory by the chain kernels
  AvrDivTab div;
  uint8_t range_lps[256];     // ITU-T H.264 Table 9-44, [pStateIdx*4 + qCodIRangeIdx]
  uint8_t trans_lps[64];      // Table 9-45
  uint8_t pad[6];
};
struct AvrTables {
  AvrHotTabs hot;
  int8_t ctx_m[4][1024];      // Tables 9-12..9-33: plane 0 = I, 1..3 = cabac_init_idc 0..2
  int8_t ctx_n[4][1024];
};

// Batch layout: all files' record arrays are concatenated; a descriptor per file.
struct AvrFileDesc {
  uint64_t bin_off;                  // into bins[]
  uint32_t nbins;
  uint32_t block_off, nblocks;       // into blocks[]
  uint32_t slice_off, nslices;       // into slices[] (compress) / dec_slices[] (decompress)
  uint32_t frame_off, nframes;       // into frames[]
  uint32_t max_mbs;                  // largest picture of the file, in macroblocks
  uint32_t mb_base;                  // chain kernels: first macroblock slot of this file (a file owns 2*max_mbs slots)
  uint32_t pic_mb_base;              // pipeline: first macroblock slot of this file's per-picture stores
};

// Compress-side slice in batch coordinates (built by the host when a batch is staged).
struct AvrBatchSlice {
  uint32_t file;                     // index into fds[]
  uint32_t bin_start, bin_end;       // file-local, as in AvrSlice
  uint32_t block_start, block_end;   // batch-global block indices
};

// One coded slice on the decompress side (reference decompressor::cabac_decoder, recode.cpp:1418-1527).
struct AvrDecSlice {
  uint64_t payload_off;    // range-coded payload (Block.cabac) in the payload pool
  uint64_t out_off;        // where the re-encoded CABAC bytes go in the output pool
  uint32_t payload_len;
  uint32_t out_cap;
  uint32_t flip;           // pictures started since the previous coded slice of the file:
                           // 0 none, 1 one flip (previous picture kept), 2 flip with both stores zeroed
  avr::SliceParams sp;
};

// ---- single-thread compress of one file (cross-check path) ----------------------------------
// Emission order per residual block (reference recode.cpp:1211-1262):
// nnz side-channel bits, then the MAP bins in scan order; EOB bins are not coded.
static AVR_HDN void avr_compress_chain_body(const AvrFileDesc& fd, const uint16_t* bins_all, const AvrBlock* blocks_all,
                                            const AvrSlice* slices_all, const AvrFrame* frames_all,
                                            uint16_t* est, uint8_t* store0, uint8_t* store1, const AvrTables* tab,
                                            uint8_t* scratch, const uint64_t* scratch_off, const uint32_t* scratch_cap,
                                            uint32_t* out_len, int* status) {
  const uint16_t* bins = bins_all + fd.bin_off;
  const AvrBlock* blocks = blocks_all + fd.block_off;
  const AvrFrame* frames = frames_all + fd.frame_off;
  AvrChain ch;
  ch.init(est, store0, store1, fd.max_mbs, &tab->hot.div);
  int curf = -1;
  for (uint32_t s = 0; s < fd.nslices; s++) {
    const AvrSlice sl = slices_all[fd.slice_off + s];
    AvrRangeEnc enc;
    enc.init(scratch + scratch_off[fd.slice_off + s], scratch_cap[fd.slice_off + s]);
    uint32_t b = sl.block_start;
    uint32_t i = sl.bin_start;
    int finished = 0;
    while (i < sl.bin_end && !finished) {
      if (b < sl.block_end && blocks[b].bin_start == i) {
        const AvrBlock blk = blocks[b++];
        while (curf < (int)blk.frame) {
          curf++;
          ch.flip(frames[curf].mb_w, frames[curf].mb_h, (int)frames[curf].prev_valid);
        }
        const int cat = blk.cat_flags & 15, is_dc = (blk.cat_flags >> 4) & 1, c422 = (blk.cat_flags >> 5) & 1;
        const int size = blk.max_coeff;
        // count (end_coding_type, recode.cpp:942-954): MAP ones + the inferred last coefficient
        int ones = 0;
        uint16_t lastbin = 0;
        for (uint32_t k = 0; k < blk.nbins; k++) {
          lastbin = bins[i + k];
          if (AVR_BIN_KIND(lastbin) == AVR_KIND_SIG_MAP) ones += AVR_BIN_SYM(lastbin);
        }
        int count = ones + !(AVR_BIN_KIND(lastbin) == AVR_KIND_SIG_EOB && AVR_BIN_SYM(lastbin));
        uint8_t* mb = ch.mb(blk.mb_addr);
        if (size > 32) mb[AVR_MB_IS8X8] = 1;
        const int stored = count & 63;                       // six bits kept, recode.cpp:928-931
        mb[blk.n] = (uint8_t)stored;
        AvrNzCtx nc;
        avr_nz_ctx(&nc, ch.store[ch.cur], ch.store[ch.cur ^ 1], blk.mb_addr, ch.mb_w, blk.n, size, cat, is_dc, c422);
        const int nbits = avr_nz_bits(size);
        int so_far = 0;
        for (int k = 0; k < nbits; k++) {
          int bit = (stored >> k) & 1;
          ch.encode(enc, avr_nz_bit_key(nc, k, so_far), bit, 0x60);
          so_far |= bit << k;
        }
        int zz = 0, observed = 0;
        for (uint32_t k = 0; k < blk.nbins; k++) {
          uint16_t rec = bins[i + k];
          int sym = AVR_BIN_SYM(rec);
          if (AVR_BIN_KIND(rec) == AVR_KIND_SIG_MAP) {
            ch.encode(enc, avr_key_sig(cat, stored, observed, avr_zz_off(zz, size, is_dc, c422)), sym, 0x50);
            observed += sym;
            if (!sym) zz++;
          } else if (!sym) {
            zz++;
          }
        }
        i += blk.nbins;
        continue;
      }
      uint16_t rec = bins[i++];
      int ctx = AVR_BIN_CTX(rec), sym = AVR_BIN_SYM(rec);
      ch.encode(enc, (uint32_t)ctx, sym, 0x60);
      if (ctx == AVR_CTX_TERMINATE && sym) { enc.finish(); finished = 1; }   // recode.cpp:1099-1102
    }
    out_len[fd.slice_off + s] = enc.n;
    status[fd.slice_off + s] = enc.err ? enc.err : (finished ? AVR_OK : AVR_E_CORRUPT);
  }
}

// ---- decompress: one chain per file --------------------------------------------------------
// Bin I/O policy for avr::SliceParser (h264_syntax.h): every bin the slice_data() walk asks
// for is range-decoded with the model's estimator (or inferred, for EOB bins), re-encoded
// into real CABAC bytes and fed back to the walk, which picks the next context from it --
// the lock-step loop of reference recode.cpp:1442-1499.
struct AvrDecodeIO {
  AvrChain* ch;
  AvrRangeDec dec;
  AvrCabacEnc cab;
  uint8_t* st;                  // [1024] CABAC context states of this slice
  const AvrHotTabs* hot;
  uint32_t mb_addr;
  int cat, n, size, is_dc, c422;
  int in_sigmap, expect_eob, zz, observed, stored, ones, ended_by_eob;
  uint32_t bins;
  int done;

  AVR_HD int get(int ctx) {
    bins++;
    int sym;
    if (in_sigmap) {
      if (expect_eob) {
        sym = (stored == observed);                         // recode.cpp:1444-1445, 809-816
        expect_eob = 0;
        if (sym) ended_by_eob = 1; else zz++;
      } else {
        sym = ch->decode(dec, avr_key_sig(cat, stored, observed, avr_zz_off(zz, size, is_dc, c422)), 0x50);
        observed += sym; ones += sym;
        if (sym) expect_eob = 1; else zz++;
      }
    } else {
      sym = ch->decode(dec, (uint32_t)ctx, 0x60);
    }
    cab.put(sym, &st[ctx], hot->range_lps, hot->trans_lps);
    return sym;
  }
  AVR_HD int bypass(int) {
    bins++;
    int sym = ch->decode(dec, AVR_CTX_BYPASS, 0x60);
    cab.put_bypass(sym);
    return sym;
  }
  AVR_HD int terminate(int) {
    bins++;
    int sym = ch->decode(dec, AVR_CTX_TERMINATE, 0x60);
    if (cab.put_terminate(sym)) done = 1;
    return sym;
  }
  AVR_HD void mb_begin(int x, int y) { mb_addr = (uint32_t)(x + y * ch->mb_w); }
  AVR_HD void sub_mb_begin(int cat_, int n_, int mc, int dc, int c422_) { cat = cat_; n = n_; size = mc; is_dc = dc; c422 = c422_; }
  AVR_HD void sub_mb_end() {}
  // begin_coding_type on decompress: the nnz bits come first (recode.cpp:1483-1499).  The count
  // is recovered completely (a block that codes as 0 in 2/4 bits is a full one) so that the
  // store holds what the COMPRESSOR held (exact inverse, not the reference's
  // decompressor quirks).
  AVR_HDN void sigmap_begin() {
    AvrNzCtx nc;
    avr_nz_ctx(&nc, ch->store[ch->cur], ch->store[ch->cur ^ 1], mb_addr, ch->mb_w, n, size, cat, is_dc, c422);
    const int nbits = avr_nz_bits(size);
    int v = 0;
    for (int k = 0; k < nbits; k++) v |= ch->decode(dec, avr_nz_bit_key(nc, k, v), 0x60) << k;
    if (v == 0 && size <= 16) v = size;
    stored = v;
    in_sigmap = 1; expect_eob = 0; zz = 0; observed = 0; ones = 0; ended_by_eob = 0;
  }
  AVR_HD void sigmap_end() {
    uint8_t* mb = ch->mb(mb_addr);
    if (size > 32) mb[AVR_MB_IS8X8] = 1;
    mb[n] = (uint8_t)((ones + !ended_by_eob) & 63);
    in_sigmap = 0;
  }
  AVR_HD void hint(int, int) {}
};

// CABAC context initialisation (ITU-T H.264 9.3.1.1; reference ffmpeg h264_cabac.c:1267-1286).
static AVR_HDN void avr_cabac_init_states(uint8_t* st, const AvrTables* tab, int is_intra, int init_idc, int qp,
                                          int lane, int nlanes) {
  const int pl = is_intra ? 0 : 1 + init_idc;
  qp = qp < 0 ? 0 : qp > 51 ? 51 : qp;
  for (int i = lane; i < 1024; i += nlanes) {
    int pre = ((tab->ctx_m[pl][i] * qp) >> 4) + tab->ctx_n[pl][i];
    pre = pre < 1 ? 1 : pre > 126 ? 126 : pre;
    st[i] = pre <= 63 ? (uint8_t)((63 - pre) << 1) : (uint8_t)(((pre - 64) << 1) | 1);
  }
}

// Per-file decompress state that survives from slice to slice (lives in lane 0's registers /
// the block's shared memory in the kernel).
struct AvrDecState {
  AvrChain ch;
  uint32_t uid;
  int failed;
};

// One coded slice, executed by ONE thread (the chain).  st[] must already hold the slice's initial
// CABAC states and the picture stores must already be flipped (both done cooperatively by the
// caller: avr_k_decompress_chain / the host emulation).
static AVR_HDN void avr_decompress_slice(AvrDecState& S, const AvrDecSlice& ds, const uint8_t* payloads, avr::MbInfo* row,
                                         uint8_t* st, const AvrHotTabs* hot, uint8_t* out, uint32_t* out_len,
                                         uint32_t* out_bins, int* status) {
  if (S.failed) { *out_len = 0; *out_bins = 0; *status = AVR_E_CORRUPT; return; }   // model state is meaningless after a failed slice
  if (S.uid == 0) for (int x = 0; x < ds.sp.mb_w; x++) row[x].uid = 0;
  AvrDecodeIO io;
  io.ch = &S.ch; io.st = st; io.hot = hot;
  io.dec.init(payloads + ds.payload_off, ds.payload_len);
  io.cab.init(out + ds.out_off, ds.out_cap);
  io.in_sigmap = 0; io.expect_eob = 0; io.bins = 0; io.done = 0; io.mb_addr = 0;
  avr::SliceParser<AvrDecodeIO> parser(io, ds.sp, row, ++S.uid);
  int rc = parser.run();
  int err = rc < 0 ? AVR_E_CORRUPT : io.cab.err ? io.cab.err : !io.done ? AVR_E_CORRUPT
            : io.dec.pos > ds.payload_len + 16 ? AVR_E_CORRUPT : AVR_OK;
  *out_len = io.cab.n; *out_bins = io.bins; *status = err;
  if (err) S.failed = 1;
}

// Whole file on one thread (host emulation; the kernel interleaves the cooperative steps itself).
static AVR_HDN void avr_decompress_chain_body(const AvrFileDesc& fd, const AvrDecSlice* slices_all, const uint8_t* payloads,
                                              uint16_t* est, uint8_t* store0, uint8_t* store1, avr::MbInfo* row,
                                              uint8_t* st, const AvrTables* tab, uint8_t* out, uint32_t* out_len,
                                              uint32_t* out_bins, int* status) {
  AvrDecState S;
  S.ch.init(est, store0, store1, fd.max_mbs, &tab->hot.div);
  S.uid = 0; S.failed = 0;
  for (uint32_t s = 0; s < fd.nslices; s++) {
    const AvrDecSlice& ds = slices_all[fd.slice_off + s];
    if (ds.flip) S.ch.flip(ds.sp.mb_w, ds.sp.mb_h, ds.flip == 1);
    avr_cabac_init_states(st, tab, ds.sp.slice_type == avr::SLICE_I, ds.sp.cabac_init_idc, ds.sp.slice_qp, 0, 1);
    avr_decompress_slice(S, ds, payloads, row, st, &tab->hot, out, &out_len[fd.slice_off + s], &out_bins[fd.slice_off + s],
                         &status[fd.slice_off + s]);
  }
}
