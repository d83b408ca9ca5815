// Slice-parallel compress pipeline (bit-exact decomposition of the reference's sequential
// per-file model).  On compress every bin value is known up front, so:
//
//  A  count   (thread / residual block)  nnz of each block -> picture stores; size of the
//             block's segment of the emitted stream            recode.cpp:942-954, 865
//  -  scan    exclusive scan of segment sizes -> position of every coded symbol
//  B  keys    (thread / segment) model key of every coded symbol in emission order: nnz
//             side-channel bits, MAP bins, default bins          recode.cpp:683-822, 851-937
//  -  sort    stable radix sort of (file, key) carrying the emitted index: each key's symbols
//             become one contiguous run in file order
//  C  estimate the adaptive counts of a key depend only on the symbols of that key
//             (recode.cpp:1037-1054).  Halvings happen at FIXED ranks of the run (the total
//             goes 2 -> limit+1 once, then (limit+2)/2 -> limit+1 forever), so with c_j = ones
//             in period j the count at a period start is pos' = ceil((pos + c_j) / 2); these
//             maps compose as pos -> ceil((pos + a) / 2^k), an associative monoid, and the
//             whole trajectory is three scans instead of a serial walk.
//  D  code    (thread / slice) the only inherently serial part: low/range of the 64-bit range
//             coder over the slice's (pos,total,symbol) stream    arithmetic_code.h:106-180
//  -  compact exclusive scan of payload sizes + gather (avr_kernels.cu)
//
// Bodies are host+device (tests/host_emul.cpp runs them as an emulated grid); kernels and the
// CUB plumbing (device radix sort / scans) are in avr_kernels.cu.
#pragma once
#include "avr_bodies.h"

// Emitted-stream segments: entry e = block_index + slice_index + 1 for a residual block (its nnz
// bits, MAP bins and the default bins that follow it up to the next block / end of slice), and
// e = block_start(slice) + slice_index for the default bins that open a slice.  All indices are
// batch-global; an entry finds its slice by binary search in the (small) slice table.
struct AvrPipe {
  const AvrFileDesc* fds;
  const AvrBatchSlice* bslices; uint32_t nslices;
  const AvrBlock* blocks; uint32_t nblocks;
  const uint16_t* bins;
  const AvrFrame* frames;
  uint8_t* store;          // per-picture nnz stores of the whole batch, zeroed before pass A
  uint32_t* seg;           // [nblocks + nslices + 1] segment sizes, then (scanned) emitted positions
  uint32_t* keys;          // [nemit] (file << (AVR_KEY_BITS + 1)) | (key << 1) | symbol, emission order
  uint32_t nemit;
  uint16_t* prob;          // [nemit] pass C output
  const AvrTables* tab;
  uint8_t* scratch; const uint64_t* scratch_off; const uint32_t* scratch_cap;
  uint32_t* out_len; int* status;
};

static AVR_HD uint32_t avr_entry_slice(const AvrPipe& a, uint32_t e) {
  uint32_t lo = 0, hi = a.nslices;          // largest s with block_start[s] + s <= e
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (a.bslices[mid].block_start + mid <= e) lo = mid; else hi = mid;
  }
  return lo;
}

// ---- pass A: per residual block -----------------------------------------------------------
// returns the number of coded symbols of the block's segment (nnz bits + MAP bins + the
// default bins that follow it up to the next block / end of slice)
static AVR_HD uint32_t avr_pass_count(const AvrBlock& blk, uint32_t seg_end_bin, const uint16_t* bins /*file base*/,
                                      uint8_t* store /*file base*/, const AvrFrame* frames /*file base*/,
                                      uint8_t* nnz_out) {
  int ones = 0;
  uint16_t lastbin = 0;
  for (uint32_t k = 0; k < blk.nbins; k++) {
    lastbin = bins[blk.bin_start + k];
    if (AVR_BIN_KIND(lastbin) == AVR_KIND_SIG_MAP) ones += AVR_BIN_SYM(lastbin);
  }
  int count = ones + !(AVR_BIN_KIND(lastbin) == AVR_KIND_SIG_EOB && AVR_BIN_SYM(lastbin));
  uint8_t* mb = store + ((size_t)frames[blk.frame].mb_base + blk.mb_addr) * AVR_MB_STRIDE;
  if (blk.max_coeff > 32) mb[AVR_MB_IS8X8] = 1;
  mb[blk.n] = (uint8_t)(count & 63);
  *nnz_out = (uint8_t)(count & 63);
  // every MAP one is followed by exactly one EOB bin, which is not coded
  return (uint32_t)avr_nz_bits(blk.max_coeff) + (blk.nbins - (uint32_t)ones) + (seg_end_bin - (blk.bin_start + blk.nbins));
}

// ---- pass B: keys of one segment, written at the segment's emitted position ----------------
// out[i] = (file << (AVR_KEY_BITS + 1)) | (key << 1) | symbol
static AVR_HD void avr_pass_keys_default(const uint16_t* bins, uint32_t b0, uint32_t b1, uint32_t filebits, uint32_t* out) {
  for (uint32_t i = b0; i < b1; i++) {
    uint16_t rec = bins[i];
    *out++ = filebits | ((uint32_t)AVR_BIN_CTX(rec) << 1) | AVR_BIN_SYM(rec);
  }
}
static AVR_HD void avr_pass_keys_block(const AvrBlock& blk, int stored, uint32_t seg_end_bin, const uint16_t* bins,
                                       const uint8_t* store, const AvrFrame* frames, uint32_t filebits, uint32_t* out) {
  const int cat = blk.cat_flags & 15, is_dc = (blk.cat_flags >> 4) & 1, c422 = (blk.cat_flags >> 5) & 1;
  const int size = blk.max_coeff;
  const AvrFrame fr = frames[blk.frame];
  const uint8_t* cur = store + (size_t)fr.mb_base * AVR_MB_STRIDE;
  const uint8_t* prev = fr.prev_valid ? store + (size_t)frames[blk.frame - 1].mb_base * AVR_MB_STRIDE : nullptr;
  AvrNzCtx nc;
  avr_nz_ctx(&nc, cur, prev, blk.mb_addr, fr.mb_w, blk.n, size, cat, is_dc, c422);
  const int nbits = avr_nz_bits(size);
  int so_far = 0;
  for (int k = 0; k < nbits; k++) {
    int bit = (stored >> k) & 1;
    *out++ = filebits | (avr_nz_bit_key(nc, k, so_far) << 1) | (uint32_t)bit;
    so_far |= bit << k;
  }
  int zz = 0, observed = 0;
  for (uint32_t k = 0; k < blk.nbins; k++) {
    uint16_t rec = bins[blk.bin_start + k];
    int sym = AVR_BIN_SYM(rec);
    if (AVR_BIN_KIND(rec) == AVR_KIND_SIG_MAP) {
      *out++ = filebits | (avr_key_sig(cat, stored, observed, avr_zz_off(zz, size, is_dc, c422)) << 1) | (uint32_t)sym;
      observed += sym;
      if (!sym) zz++;
    } else if (!sym) {
      zz++;
    }
  }
  avr_pass_keys_default(bins, blk.bin_start + blk.nbins, seg_end_bin, filebits, out);
}

// ---- pass C: estimator trajectory from ranks ------------------------------------------------
// Period structure of a key's run (reference recode.cpp:1048-1052): limit 0x50 for sig-map
// keys, 0x60 otherwise; first halving after limit-1 symbols, then every limit+1-(limit+2)/2.
struct AvrPeriod { uint32_t L0, L, t0; };
static AVR_HD AvrPeriod avr_period_of(uint32_t key) {
  AvrPeriod p;
  uint32_t limit = avr_key_is_sig(key) ? 0x50u : 0x60u;
  p.L0 = limit - 1; p.t0 = (limit + 2) / 2; p.L = limit + 1 - p.t0;
  return p;
}
static AVR_HD int avr_is_period_head(uint32_t r, const AvrPeriod& p) { return r == 0 || (r >= p.L0 && (r - p.L0) % p.L == 0); }
static AVR_HD int avr_is_period_tail(uint32_t r, const AvrPeriod& p) { return r + 1 >= p.L0 && (r + 1 - p.L0) % p.L == 0; }
static AVR_HD uint32_t avr_total_at(uint32_t r, const AvrPeriod& p) { return r < p.L0 ? 2 + r : p.t0 + (r - p.L0) % p.L; }

// pos -> ceil((pos + a) / 2^k) on the domain pos in [1,96]; kept normalised to k <= 7
struct AvrHalf { uint32_t a, k; };
static AVR_HD AvrHalf avr_half_compose(AvrHalf f, AvrHalf g /* applied after f */) {
  AvrHalf h;
  h.a = f.a + (g.a << f.k);
  h.k = f.k + g.k;
  if (h.k > 7) {
    uint32_t m = 1u << h.k;
    uint32_t v = (1 + h.a + m - 1) >> h.k;       // value at pos = 1
    uint32_t th = v * m - h.a;                   // largest pos that still maps to v
    if (th > 96) th = 96;
    h.a = 128 * v - th; h.k = 7;
  }
  return h;
}
static AVR_HD uint32_t avr_half_apply(AvrHalf f, uint32_t pos) { return (pos + f.a + (1u << f.k) - 1) >> f.k; }

// prob record for pass D: bit0 symbol, bits 1..7 pos, bits 8..14 total
static AVR_HD uint16_t avr_prob_pack(uint32_t sym, uint32_t pos, uint32_t total) { return (uint16_t)(sym | (pos << 1) | (total << 8)); }

// ---- pass D: one slice ------------------------------------------------------------------------
// prob[] is read four records (one 8-byte word) at a time and the next word is requested before
// the current one is coded, so the load latency stays off the low/range dependency chain.
static AVR_HDN void avr_pass_code(const uint16_t* prob, uint32_t e0, uint32_t e1, const AvrDivTab* dt, uint8_t* scratch,
                                  uint32_t cap, uint32_t* out_len, int* status) {
  AvrRangeEnc enc;
  enc.init(scratch, cap);
  uint32_t e = e0;
#define AVR_CODE_STEP(r_) do { const uint32_t r__ = (uint32_t)(r_); \
    enc.put((int)(r__ & 1), avr_range_of_1(enc.range, (r__ >> 1) & 0x7F, (r__ >> 8) & 0x7F, dt)); } while (0)
  while (e < e1 && (e & 3)) { AVR_CODE_STEP(prob[e]); e++; }
  const uint64_t* p = reinterpret_cast<const uint64_t*>(prob + e);
  uint64_t nxt = e + 4 <= e1 ? p[0] : 0;
  while (e + 4 <= e1) {
    const uint64_t w = nxt;
    p++; e += 4;
    if (e + 4 <= e1) nxt = p[0];
    // the divisor constants of all four records first: they depend on the records only, so their shared-memory
    // latency overlaps the first symbol instead of opening every symbol's dependency chain
    const uint32_t t0 = (uint32_t)(w >> 8) & 0x7F, t1 = (uint32_t)(w >> 24) & 0x7F, t2 = (uint32_t)(w >> 40) & 0x7F, t3 = (uint32_t)(w >> 56) & 0x7F;
    const uint64_t m0 = dt->m[t0], m1 = dt->m[t1], m2 = dt->m[t2], m3 = dt->m[t3];
    const uint32_t s0 = dt->s[t0], s1 = dt->s[t1], s2 = dt->s[t2], s3 = dt->s[t3];
    enc.put((int)(w & 1), (AVR_MULHI64(enc.range, m0) >> s0) * (uint64_t)((w >> 1) & 0x7F));
    enc.put((int)((w >> 16) & 1), (AVR_MULHI64(enc.range, m1) >> s1) * (uint64_t)((w >> 17) & 0x7F));
    enc.put((int)((w >> 32) & 1), (AVR_MULHI64(enc.range, m2) >> s2) * (uint64_t)((w >> 33) & 0x7F));
    enc.put((int)((w >> 48) & 1), (AVR_MULHI64(enc.range, m3) >> s3) * (uint64_t)((w >> 49) & 0x7F));
  }
  while (e < e1) { AVR_CODE_STEP(prob[e]); e++; }
#undef AVR_CODE_STEP
  enc.finish();
  *out_len = enc.n;
  *status = enc.err;
}

// ---- entry-level bodies (one emulated / CUDA thread each) -----------------------------------
static AVR_HD void avr_body_count(const AvrPipe& a, uint32_t e) {
  const uint32_t s = avr_entry_slice(a, e);
  const AvrBatchSlice sl = a.bslices[s];
  const AvrFileDesc& fd = a.fds[sl.file];
  const AvrBlock* fblocks = a.blocks;            // batch-global indexing
  if (e == sl.block_start + s) {                 // slice head: default bins before the first block
    uint32_t end = sl.block_start < sl.block_end ? fblocks[sl.block_start].bin_start : sl.bin_end;
    a.seg[e] = end - sl.bin_start;
    return;
  }
  const uint32_t b = e - s - 1;
  const AvrBlock blk = fblocks[b];
  const uint32_t seg_end = b + 1 < sl.block_end ? fblocks[b + 1].bin_start : sl.bin_end;
  uint8_t nnz;
  a.seg[e] = avr_pass_count(blk, seg_end, a.bins + fd.bin_off, a.store + (size_t)fd.pic_mb_base * AVR_MB_STRIDE,
                            a.frames + fd.frame_off, &nnz);
}

static AVR_HD void avr_body_keys(const AvrPipe& a, uint32_t e) {
  const uint32_t s = avr_entry_slice(a, e);
  const AvrBatchSlice sl = a.bslices[s];
  const AvrFileDesc& fd = a.fds[sl.file];
  const uint16_t* bins = a.bins + fd.bin_off;
  uint32_t* out = a.keys + a.seg[e];
  const uint32_t filebits = sl.file << (AVR_KEY_BITS + 1);
  if (e == sl.block_start + s) {
    uint32_t end = sl.block_start < sl.block_end ? a.blocks[sl.block_start].bin_start : sl.bin_end;
    avr_pass_keys_default(bins, sl.bin_start, end, filebits, out);
    return;
  }
  const uint32_t b = e - s - 1;
  const AvrBlock blk = a.blocks[b];
  const uint32_t seg_end = b + 1 < sl.block_end ? a.blocks[b + 1].bin_start : sl.bin_end;
  const AvrFrame* frames = a.frames + fd.frame_off;
  const uint8_t* store = a.store + (size_t)fd.pic_mb_base * AVR_MB_STRIDE;
  const int stored = store[((size_t)frames[blk.frame].mb_base + blk.mb_addr) * AVR_MB_STRIDE + blk.n];
  avr_pass_keys_block(blk, stored, seg_end, bins, store, frames, filebits, out);
}

// After the sort: sorted[i] = (file,key,symbol), idx[i] = emitted index.  Three plain scans over
// the sorted array give everything the trajectory needs:
//   P[i]    inclusive prefix sum of the symbol bit           (ones of ANY range = difference)
//   head[i] index of the first element of i's key run        (max-scan of head positions)
//   fn[i]   composition of the halving maps of the run's periods completed up to and
//           including i                                       (segmented monoid scan)
static AVR_HD uint32_t avr_ones_between(const uint32_t* P, uint32_t first, uint32_t last_incl) {
  return P[last_incl] - (first ? P[first - 1] : 0u);
}
// element of the fn scan: bit 31 of kflag = starts a key run
struct AvrHalfSeg { uint32_t a, kflag; };
static AVR_HD AvrHalfSeg avr_fn_element(uint32_t i, const uint32_t* sorted, const uint32_t* head, const uint32_t* P) {
  const uint32_t key = (sorted[i] >> 1) & AVR_KEY_MASK;
  const uint32_t h = head[i], r = i - h;
  const AvrPeriod p = avr_period_of(key);
  AvrHalfSeg e;
  e.a = 0; e.kflag = r == 0 ? 0x80000000u : 0u;
  if (avr_is_period_tail(r, p)) {
    const uint32_t first = r + 1 == p.L0 ? h : i + 1 - p.L;     // first element of the period that ends at i
    e.a = avr_ones_between(P, first, i);
    e.kflag |= 1;
  }
  return e;
}
struct AvrHalfSegOp {
  AVR_HD AvrHalfSeg operator()(const AvrHalfSeg& l, const AvrHalfSeg& r) const {
    if (r.kflag & 0x80000000u) return r;
    AvrHalf f = {l.a, l.kflag & 0x7FFFFFFFu}, g = {r.a, r.kflag};
    AvrHalf h = avr_half_compose(f, g);
    AvrHalfSeg o = {h.a, h.k | (l.kflag & 0x80000000u)};
    return o;
  }
};
struct AvrMaxOp { AVR_HD uint32_t operator()(uint32_t a, uint32_t b) const { return a > b ? a : b; } };

static AVR_HD void avr_body_estimate(uint32_t i, const uint32_t* sorted, const uint32_t* idx, const uint32_t* head,
                                     const uint32_t* P, const AvrHalfSeg* fn, uint16_t* prob) {
  const uint32_t w = sorted[i], sym = w & 1, key = (w >> 1) & AVR_KEY_MASK;
  const uint32_t h = head[i], r = i - h;
  const AvrPeriod p = avr_period_of(key);
  uint32_t ob = 0, pos0 = 1;
  if (r > 0) {
    const uint32_t pstart = r < p.L0 ? h : i - (r - p.L0) % p.L;   // first element of i's period
    if (pstart < i) ob = avr_ones_between(P, pstart, i - 1);
    AvrHalf f = {fn[i - 1].a, fn[i - 1].kflag & 0x7FFFFFFFu};
    pos0 = avr_half_apply(f, 1);
  }
  prob[idx[i]] = avr_prob_pack(sym, pos0 + ob, avr_total_at(r, p));
}

static AVR_HDN void avr_body_code(const AvrPipe& a, uint32_t s, const AvrDivTab* dt) {
  const AvrBatchSlice sl = a.bslices[s];
  const uint32_t e0 = a.seg[sl.block_start + s];
  const uint32_t e1 = s + 1 < a.nslices ? a.seg[a.bslices[s + 1].block_start + s + 1] : a.nemit;
  avr_pass_code(a.prob, e0, e1, dt, a.scratch + a.scratch_off[s], a.scratch_cap[s], &a.out_len[s], &a.status[s]);
}
