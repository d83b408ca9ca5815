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

