// sm_100a kernels of the entropy hot path.  Launchers are declared in avr_kernels.h and take
// raw device pointers + a stream; no torch types, no host fallback.
//
// Sequential ("chain") kernels: one warp per file, lane 0 walks the chain (the walk is a
// strict per-bin dependency), the other lanes help with bulk work (zeroing the
// picture store, CABAC state init).  Slice-parallel compress pipeline: see avr_pipeline.h.
#include <cuda_runtime.h>
#include <stdlib.h>
#include "avr_bodies.h"
#include "avr_kernels.h"
#include "avr_pipeline.h"
#include "avr_decode.h"
#include "avr_parse.h"

// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) avr_k_compress_chain(const AvrFileDesc* fds, int nfiles, const uint16_t* bins,
                                                           const AvrBlock* blocks, const AvrSlice* slices,
                                                           const AvrFrame* frames, uint16_t* est_pool, uint8_t* store_pool,
                                                           const AvrTables* tab, uint8_t* scratch, const uint64_t* scratch_off,
                                                           const uint32_t* scratch_cap, uint32_t* out_len, int* status) {
  int f = blockIdx.x;
  if (f >= nfiles || threadIdx.x != 0) return;
  const AvrFileDesc fd = fds[f];
  uint8_t* s0 = store_pool + (size_t)fd.mb_base * AVR_MB_STRIDE;
  uint8_t* s1 = s0 + (size_t)fd.max_mbs * AVR_MB_STRIDE;
  avr_compress_chain_body(fd, bins, blocks, slices, frames, est_pool + (size_t)f * AVR_KEY_COUNT, s0, s1, tab, scratch,
                          scratch_off, scratch_cap, out_len, status);
}

// One block (two warps) per file.  Thread 0 is the chain (strict per-bin dependency), thread 32 is
// the CABAC encoder fed through a shared-memory ring (avr_decode.h); all 64 threads do the bulk
// steps between slices: zeroing the picture stores on a flip, CABAC state initialisation, staging
// the per-bin tables and the default-key estimators in shared memory.
#define AVR_DEC_THREADS 64
// Which of the block's two warps is the chain.  A block's warps take consecutive hardware warp slots and a slot's scheduler
// is slot % 4, so with the chain always on warp 0 every chain of an SM sits on two of its four schedulers and the sleeping
// encoders on the other two (ncu, 2048 chains: 720 M instructions on the busiest scheduler against 354 M on the idlest).
// That is the FASTER arrangement: -DAVR_MIX_WARPS (roles swapped in every other block by a hash of the block index, chains on
// all four schedulers, each next to polling encoders) runs the 2048-clip corpus in 770 ms against 588 ms
// (profiles/r02_ab_warp_roles.txt).
#if defined(AVR_ROLE4)
// experiment: four warps per block, the chain on one of warps 0..2 (hash of the block index), the encoder on warp 3, the others idle:
// chains on three schedulers, encoders on the fourth
#undef AVR_DEC_THREADS
#define AVR_DEC_THREADS 128
#define AVR_CHAIN_TID (32 * (int)(((blockIdx.x * 2654435761u) >> 15) % 3u))
#define AVR_ENC_TID 96
#elif defined(AVR_SWAP_WARPS)
#define AVR_CHAIN_TID 32
#define AVR_ENC_TID 0
#elif defined(AVR_MIX_WARPS)
#define AVR_CHAIN_TID (avr_block_swapped() ? 32 : 0)
#define AVR_ENC_TID (avr_block_swapped() ? 0 : 32)
static __device__ __forceinline__ bool avr_block_swapped() { return ((blockIdx.x * 2654435761u) >> 15) & 1u; }
#else
#define AVR_CHAIN_TID 0
#define AVR_ENC_TID 32
#endif
template <bool HAS_B>
__global__ void __launch_bounds__(AVR_DEC_THREADS) avr_k_decompress_chain(
    const AvrFileDesc* fds, int nfiles, const AvrDecSlice* slices, const uint8_t* payloads, uint16_t* est_pool,
    uint8_t* store_pool, avr::MbInfo* row_pool, const AvrTables* tab, uint8_t* out, uint32_t* out_len, uint32_t* out_bins,
    int* status, const uint32_t* file_order) {
  __shared__ int s_failed;
  const int tid = threadIdx.x;
  if ((int)blockIdx.x >= nfiles) return;
  // longest file first: consecutive blocks land on different SMs, so the long chains are spread instead of sharing
  // issue slots with each other (the run lasts as long as the slowest chain)
  const int f = file_order ? (int)file_order[blockIdx.x] : (int)blockIdx.x;
  const AvrFileDesc fd = fds[f];
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(&tab->hot);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&g_dec.hot);
    for (uint32_t i = tid; i < sizeof(AvrHotTabs) / 4; i += AVR_DEC_THREADS) dst[i] = src[i];
    for (uint32_t i = tid; i < AVR_NUM_CTX_KEYS; i += AVR_DEC_THREADS) g_dec.est_def[i] = AVR_EST_INIT;
    for (uint32_t i = tid; i < AVR_RING; i += AVR_DEC_THREADS) g_dec.ring[i] = AVR_RING_EMPTY;
    for (uint32_t i = tid; i < 98; i += AVR_DEC_THREADS) { g_dec.ms[i].m = tab->hot.div.m[i]; g_dec.ms[i].s = tab->hot.div.s[i]; }
  }
  if (tid == 0) {
    uint8_t* s0 = store_pool + (size_t)fd.mb_base * AVR_MB_STRIDE;
    avr_dec_file_begin(est_pool + (size_t)f * AVR_KEY_COUNT, s0, s0 + (size_t)fd.max_mbs * AVR_MB_STRIDE);
    s_failed = 0;
#if defined(AVR_ENC_MBAR)
    asm volatile("mbarrier.init.shared.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&g_dec.mbar)) : "memory");
    g_dec.mbar_phase = 0;
#endif
  }
  __syncthreads();
  uint32_t uid = 0;
  for (uint32_t s = 0; s < fd.nslices; s++) {
    const uint32_t g = fd.slice_off + s;
    const AvrDecSlice& ds = slices[g];
    if (s_failed) { if (tid == 0) { out_len[g] = 0; out_bins[g] = 0; status[g] = AVR_E_CORRUPT; } continue; }
    if (ds.flip) {
      if (tid == 0) avr_dec_flip_toggle(ds);
      __syncthreads();
      avr_dec_flip_zero(ds, tid, AVR_DEC_THREADS);
    }
    avr_cabac_init_states(g_dec.st, tab, ds.sp.slice_type == avr::SLICE_I, ds.sp.cabac_init_idc, ds.sp.slice_qp, tid, AVR_DEC_THREADS);
    __syncthreads();
    int rc = 0;
    if (tid == AVR_CHAIN_TID) rc = avr_dec_slice<HAS_B>(ds, payloads, row_pool + fd.mb_base, ++uid, &out_bins[g]);
    else if (tid == AVR_ENC_TID) avr_cab_encoder_thread(out + ds.out_off, ds.out_cap);
    __syncthreads();
    if (tid == AVR_CHAIN_TID) {
      if (!rc) rc = g_dec.cab_err;
      out_len[g] = g_dec.cab_n; status[g] = rc;
      if (rc) s_failed = 1;
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(128) avr_k_patch_tails(const AvrParseSlice* ps, uint32_t n, uint8_t* raw) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s < n) avr_body_patch_tail(ps[s], raw);
}
// Compress-side parse on the device (avr_parse.h).  The launch order is longest payload first.  The first `nlong` slices get a warp
// each (lane 0 walks the slice: a long slice sets the kernel's duration and must not wait for neighbours); the others are packed
// `pack` to a warp, neighbours in the order and therefore of similar length, every lane walking its own slice on its own
// AvrParShared: lanes that are at the same place of the syntax (same macroblock type, same block, same loop) issue together.
template <bool HAS_B>
__global__ void __launch_bounds__(32) avr_k_parse(const AvrParseSlice* ps, uint32_t n, const uint8_t* raw, uint16_t* bins_pool,
                                                  AvrBlock* blk_pool, avr::MbInfo* row_pool, const uint64_t* row_off,
                                                  const AvrTables* tab, AvrParseResult* res, const uint32_t* order, uint32_t nlong,
                                                  uint32_t pack) {
  const uint32_t lane = threadIdx.x;
  uint32_t i;
  if (blockIdx.x < nlong) { if (lane != 0) return; i = blockIdx.x; }
  else { if (lane >= pack) return; i = nlong + (blockIdx.x - nlong) * pack + lane; }
  if (i >= n) return;
  const uint32_t s = order[i];
  const AvrParseSlice& p = ps[s];
  avr_par_tables(tab, 0, 1);
  avr_cabac_init_states(g_par.st, tab, p.sp.slice_type == avr::SLICE_I, p.sp.cabac_init_idc, p.sp.slice_qp, 0, 1);
  avr_parse_slice<HAS_B>(p, raw, bins_pool, blk_pool, row_pool + row_off[s], &res[s]);
}
// capacity block pool -> dense block array
__global__ void __launch_bounds__(128) avr_k_gather_blocks(const AvrBlock* pool, const AvrParseSlice* ps, const uint32_t* count,
                                                           const uint32_t* dst_off, AvrBlock* dst, uint32_t n) {
  for (uint32_t s = blockIdx.x; s < n; s += gridDim.x) {
    const uint4* src = reinterpret_cast<const uint4*>(pool + ps[s].blk_off);
    uint4* d = reinterpret_cast<uint4*>(dst + dst_off[s]);
    for (uint32_t i = threadIdx.x; i < count[s]; i += blockDim.x) d[i] = src[i];
  }
}

// ---------------------------------------------------------------------------------------------
// Compaction of per-slice scratch into one contiguous payload buffer (north star: "per-slice
// scratch that a device-side exclusive scan compacts into the container").
__global__ void __launch_bounds__(1024) avr_k_exclusive_scan_u32(const uint32_t* len, uint64_t* off, uint32_t n, uint64_t* total) {
  __shared__ uint64_t warp_sum[32];
  __shared__ uint64_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (uint32_t base = 0; base < n; base += 4096) {            // four elements per thread and step
    const uint32_t i = base + threadIdx.x * 4;
    uint64_t e[4];
#pragma unroll
    for (int k = 0; k < 4; k++) e[k] = i + k < n ? len[i + k] : 0;
    const uint64_t v = e[0] + e[1] + e[2] + e[3];
    uint64_t x = v;
    for (int d = 1; d < 32; d <<= 1) { uint64_t y = __shfl_up_sync(0xFFFFFFFFu, x, d); if (lane >= d) x += y; }
    if (lane == 31) warp_sum[warp] = x;
    __syncthreads();
    if (warp == 0) {
      uint64_t w = warp_sum[lane], ws = w;
      for (int d = 1; d < 32; d <<= 1) { uint64_t y = __shfl_up_sync(0xFFFFFFFFu, ws, d); if (lane >= d) ws += y; }
      warp_sum[lane] = ws - w;
    }
    __syncthreads();
    uint64_t excl = carry + warp_sum[warp] + x - v;
#pragma unroll
    for (int k = 0; k < 4; k++) { if (i + k < n) off[i + k] = excl; excl += e[k]; }
    __syncthreads();
    if (threadIdx.x == 1023) carry = excl;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry;
}

__global__ void __launch_bounds__(256) avr_k_gather(const uint8_t* scratch, const uint64_t* scratch_off, const uint32_t* len,
                                                    const uint64_t* out_off, uint8_t* out, uint32_t n) {
  for (uint32_t s = blockIdx.x; s < n; s += gridDim.x) {
    const uint8_t* src = scratch + scratch_off[s];
    uint8_t* dst = out + out_off[s];
    for (uint32_t i = threadIdx.x; i < len[s]; i += blockDim.x) dst[i] = src[i];
  }
}

// ---------------------------------------------------------------------------------------------
#define AVR_CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return (int)e_; } while (0)

int avr_launch_compress_chain(const AvrDevBatch& b, cudaStream_t st) {
  if (b.nfiles == 0) return 0;
  avr_k_compress_chain<<<b.nfiles, 32, 0, st>>>(b.fds, b.nfiles, b.bins, b.blocks, b.slices, b.frames, b.est_pool, b.store_pool,
                                               b.tab, b.scratch, b.scratch_off, b.scratch_cap, b.out_len, b.status);
  return (int)cudaGetLastError();
}

#include <atomic>
static std::atomic<int> g_chain_hint(0);
void avr_set_chain_concurrency_hint(int files) { g_chain_hint = files; }

int avr_launch_decompress_chain(const AvrDevDBatch& b, cudaStream_t st) {
  if (b.nfiles == 0) return 0;
  if (const char* e = getenv("AVR_DEC_SPEC")) if (e[0] == '1') return avr_launch_decompress_chain_spec(b, st);   // experimental kernel (avr_kernels_spec.cu)
  {
    // L1 and shared memory are one 256 KB array per SM.  Left alone the driver carves out shared memory for as many
    // blocks as registers allow (14 x 10 KB), which leaves the chains' estimator cells ~90 KB of L1; a batch that puts
    // fewer chains on an SM asks only for what those need, the rest stays L1 (profiles/r01_contention.md).
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaFuncAttributes fa;
    int pct = -1;
    if (const char* e = getenv("AVR_DEC_CARVEOUT")) pct = atoi(e);
    else if (cudaFuncGetAttributes(&fa, b.has_b ? avr_k_decompress_chain<true> : avr_k_decompress_chain<false>) == cudaSuccess) {
      const int hint = g_chain_hint.load(), chains = hint > b.nfiles ? hint : b.nfiles;   // launches of other lanes share the SMs
      const int per_sm = (chains + sms - 1) / sms;
      const size_t need = (size_t)per_sm * (fa.sharedSizeBytes + 1024);      // + the per-block reservation
      pct = (int)((need * 100 + 233471) / 233472);
      // more than half of the array as shared memory costs more in L1 misses of the estimator cells than the extra resident
      // chains bring: 2048-clip corpus, 606 ms at 70 % (all chains resident), 594 at 60 %, 588 at 50 % (profiles/r02_summary.md)
      if (pct > 50) pct = 50;
    }
    if (pct >= 0) cudaFuncSetAttribute(b.has_b ? avr_k_decompress_chain<true> : avr_k_decompress_chain<false>, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
  }
  // the kernel with the B-slice parser instance only when the batch holds a B slice (avr_decode.h avr_dec_slice)
  if (b.has_b) avr_k_decompress_chain<true><<<b.nfiles, AVR_DEC_THREADS, 0, st>>>(b.fds, b.nfiles, b.slices, b.payloads, b.est_pool, b.store_pool, b.row_pool,
                                                 b.tab, b.out, b.out_len, b.out_bins, b.status, b.file_order);
  else avr_k_decompress_chain<false><<<b.nfiles, AVR_DEC_THREADS, 0, st>>>(b.fds, b.nfiles, b.slices, b.payloads, b.est_pool, b.store_pool, b.row_pool,
                                                 b.tab, b.out, b.out_len, b.out_bins, b.status, b.file_order);
  return (int)cudaGetLastError();
}

int avr_launch_compact(const uint8_t* scratch, const uint64_t* scratch_off, const uint32_t* len, uint64_t* out_off,
                       uint64_t* total, uint8_t* out, uint32_t n, cudaStream_t st) {
  if (n == 0) return 0;
  avr_k_exclusive_scan_u32<<<1, 1024, 0, st>>>(len, out_off, n, total);
  uint32_t grid = n < 148u * 8u ? n : 148u * 8u;
  avr_k_gather<<<grid, 256, 0, st>>>(scratch, scratch_off, len, out_off, out, n);
  return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Slice-parallel compress pipeline (avr_pipeline.h).  CUB's device radix sort and scans are the
// plumbing between the hand-written passes.
#include <cub/cub.cuh>

static AvrPipe make_pipe(const AvrDevBatch& b) {
  AvrPipe a;
  a.fds = b.fds; a.bslices = b.bslices; a.nslices = b.nslices; a.blocks = b.blocks; a.nblocks = b.nblocks;
  a.bins = b.bins; a.frames = b.frames; a.store = b.pic_store; a.seg = b.seg; a.keys = b.keys_a; a.nemit = b.emit_cap;
  a.prob = b.prob; a.tab = b.tab; a.scratch = b.scratch; a.scratch_off = b.scratch_off; a.scratch_cap = b.scratch_cap;
  a.out_len = b.out_len; a.status = b.status;
  return a;
}

__global__ void __launch_bounds__(256) avr_k_count(AvrPipe a, uint32_t nent) {
  uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < nent) avr_body_count(a, e);
  else if (e == nent) a.seg[e] = 0;
}
__global__ void __launch_bounds__(256) avr_k_keys(AvrPipe a, uint32_t nent, uint32_t* idx) {
  uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < nent) avr_body_keys(a, e);
}
__global__ void __launch_bounds__(256) avr_k_iota(uint32_t* idx, uint32_t n) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) idx[i] = i;
}
struct AvrSymBit {
  const uint32_t* k;
  __host__ __device__ uint32_t operator()(uint32_t i) const { return k[i] & 1u; }
};
struct AvrHeadPos {
  const uint32_t* k;
  __host__ __device__ uint32_t operator()(uint32_t i) const { return (i > 0 && (k[i] >> 1) != (k[i - 1] >> 1)) ? i : 0u; }
};
struct AvrFnElem {
  const uint32_t* k; const uint32_t* head; const uint32_t* P;
  __host__ __device__ AvrHalfSeg operator()(uint32_t i) const { return avr_fn_element(i, k, head, P); }
};
__global__ void __launch_bounds__(256) avr_k_estimate(const uint32_t* sorted, const uint32_t* idx, const uint32_t* head,
                                                      const uint32_t* P, const AvrHalfSeg* fn, uint16_t* prob, uint32_t n) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) avr_body_estimate(i, sorted, idx, head, P, fn, prob);
}
// Pass D, the one inherently serial pass.  Round 1 coded one slice per warp on lane 0 (one slice per LANE made every step of the
// warp pay for the renormalisation of whichever lane needed it, and the kernel lasted as long as the warp that held the longest
// slices).  Round 2 is a hybrid over the launch order (longest slice first): the first `nlong` slices keep a warp each, the others
// are packed 32 to a warp, neighbours in the order and therefore of similar length.  The exact-division table is staged in shared
// memory by the warp.
__global__ void __launch_bounds__(32) avr_k_code(AvrPipe a, const uint32_t* order, uint32_t nlong) {
  __shared__ AvrDivTab dt;
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(&a.tab->hot.div);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&dt);
    for (uint32_t i = threadIdx.x; i < sizeof(AvrDivTab) / 4; i += 32) dst[i] = src[i];
  }
  __syncwarp();
  if (blockIdx.x < nlong) {
    if (threadIdx.x == 0) avr_body_code(a, order ? order[blockIdx.x] : blockIdx.x, &dt);
    return;
  }
  const uint32_t i = nlong + (blockIdx.x - nlong) * 32u + threadIdx.x;
  if (i < a.nslices) avr_body_code(a, order ? order[i] : i, &dt);
}

size_t avr_pipeline_tmp_bytes(uint32_t nemit, uint32_t nentries, int sort_end_bit) {
  size_t need = 0, t = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, t, (uint32_t*)nullptr, (uint32_t*)nullptr, (int)(nentries + 1));
  need = t > need ? t : need;
  cub::DeviceRadixSort::SortPairs(nullptr, t, (uint32_t*)nullptr, (uint32_t*)nullptr, (uint32_t*)nullptr, (uint32_t*)nullptr,
                                  (int)nemit, 1, sort_end_bit);
  need = t > need ? t : need;
  cub::CountingInputIterator<uint32_t> cnt(0);
  cub::TransformInputIterator<uint32_t, AvrSymBit, cub::CountingInputIterator<uint32_t>> sym(cnt, AvrSymBit{nullptr});
  cub::DeviceScan::InclusiveSum(nullptr, t, sym, (uint32_t*)nullptr, (int)nemit);
  need = t > need ? t : need;
  cub::TransformInputIterator<uint32_t, AvrHeadPos, cub::CountingInputIterator<uint32_t>> hp(cnt, AvrHeadPos{nullptr});
  cub::DeviceScan::InclusiveScan(nullptr, t, hp, (uint32_t*)nullptr, AvrMaxOp(), (int)nemit);
  need = t > need ? t : need;
  cub::TransformInputIterator<AvrHalfSeg, AvrFnElem, cub::CountingInputIterator<uint32_t>> fe(cnt, AvrFnElem{nullptr, nullptr, nullptr});
  cub::DeviceScan::InclusiveScan(nullptr, t, fe, (AvrHalfSeg*)nullptr, AvrHalfSegOp(), (int)nemit);
  need = t > need ? t : need;
  return need + 256;
}

int avr_launch_compress_pipeline(const AvrDevBatch& b, cudaStream_t st, int* launches, cudaEvent_t coder_begin, cudaEvent_t coder_end) {
  if (b.nslices == 0) return 0;
  AvrPipe a = make_pipe(b);
  const uint32_t nent = b.nblocks + b.nslices, n = b.emit_cap;
  size_t tmp = b.cub_tmp_bytes;
  int nl = 0;
  AVR_CK(cudaMemsetAsync(b.pic_store, 0, b.pic_store_bytes, st));
  avr_k_count<<<(nent + 1 + 255) / 256, 256, 0, st>>>(a, nent); nl++;
  AVR_CK(cub::DeviceScan::ExclusiveSum(b.cub_tmp, tmp, b.seg, b.seg, (int)(nent + 1), st)); nl += 2;
  avr_k_keys<<<(nent + 255) / 256, 256, 0, st>>>(a, nent, b.idx_a); nl++;
  avr_k_iota<<<(n + 255) / 256, 256, 0, st>>>(b.idx_a, n); nl++;
  tmp = b.cub_tmp_bytes;
  AVR_CK(cub::DeviceRadixSort::SortPairs(b.cub_tmp, tmp, b.keys_a, b.keys_b, b.idx_a, b.idx_b, (int)n, 1, b.sort_end_bit, st));
  nl += 2 * ((b.sort_end_bit - 1 + 7) / 8) + 1;
  cub::CountingInputIterator<uint32_t> cnt(0);
  cub::TransformInputIterator<uint32_t, AvrSymBit, cub::CountingInputIterator<uint32_t>> sym(cnt, AvrSymBit{b.keys_b});
  tmp = b.cub_tmp_bytes;
  AVR_CK(cub::DeviceScan::InclusiveSum(b.cub_tmp, tmp, sym, b.psum, (int)n, st)); nl += 2;
  cub::TransformInputIterator<uint32_t, AvrHeadPos, cub::CountingInputIterator<uint32_t>> hp(cnt, AvrHeadPos{b.keys_b});
  tmp = b.cub_tmp_bytes;
  AVR_CK(cub::DeviceScan::InclusiveScan(b.cub_tmp, tmp, hp, b.head, AvrMaxOp(), (int)n, st)); nl += 2;
  cub::TransformInputIterator<AvrHalfSeg, AvrFnElem, cub::CountingInputIterator<uint32_t>> fe(cnt, AvrFnElem{b.keys_b, b.head, b.psum});
  tmp = b.cub_tmp_bytes;
  AVR_CK(cub::DeviceScan::InclusiveScan(b.cub_tmp, tmp, fe, b.fn, AvrHalfSegOp(), (int)n, st)); nl += 2;
  avr_k_estimate<<<(n + 255) / 256, 256, 0, st>>>(b.keys_b, b.idx_b, b.head, b.psum, b.fn, b.prob, n); nl++;
  if (coder_begin) cudaEventRecord(coder_begin, st);
  {
    const uint32_t nlong = b.code_order ? (b.code_long < b.nslices ? b.code_long : b.nslices) : b.nslices;
    avr_k_code<<<nlong + (b.nslices - nlong + 31) / 32, 32, 0, st>>>(a, b.code_order, nlong); nl++;
  }
  if (coder_end) cudaEventRecord(coder_end, st);
  if (launches) *launches = nl;
  return (int)cudaGetLastError();
}

int avr_launch_parse(const AvrParseSlice* ps, uint32_t n, uint8_t* raw, uint16_t* bins_pool, AvrBlock* blk_pool,
                     avr::MbInfo* row_pool, const uint64_t* row_off, const AvrTables* tab, AvrParseResult* res, const uint32_t* order,
                     uint32_t nlong, int has_b, cudaStream_t st) {
  if (n == 0) return 0;
  avr_k_patch_tails<<<(n + 127) / 128, 128, 0, st>>>(ps, n, raw);
  // AVR_PARSE_PACK = slices per warp behind the first `nlong` (1 = every slice its own warp, round 1's layout).  Sweep on the
  // 2048-clip corpus (profiles/r02_parse_pack_sweep.txt, parse ms summed over the two concurrent launches): 505 unpacked, 392-428
  // with 4 per warp, 405-452 with 6, 409-445 with 8 (511 when slices up to 45 % of the longest are packed), 620-690 with 16,
  // 1400 with 32: lanes only issue together while they are at the same place of the syntax, and every packed lane costs 3 KB
  // of shared memory, i.e. residency.
  uint32_t pack = 4;
  if (const char* e = getenv("AVR_PARSE_PACK")) { const int v = atoi(e); if (v >= 1 && v <= 32) pack = (uint32_t)v; }
  if (nlong > n || pack == 1) nlong = n;
  // Shared memory per block = one AvrParShared per packed lane.  With one-lane blocks only, registers would allow 28 blocks per
  // SM, but with that many walkers at different places of the syntax code the kernel waits for instruction fetch (round 1:
  // stall no_instruction 34 %); 16 KB per block caps the residency at 12 (148 -> 138 ms on 9 537 slices).
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  size_t dyn = (size_t)pack * AVR_PAR_STRIDE;
  if (dyn < 16000 && n > (uint32_t)sms * 12u) dyn = 16000;
  const uint32_t grid = nlong + (n - nlong + pack - 1) / pack;
  if (dyn > 48 * 1024) {                       // per device: set on every such launch (multi-GPU callers share the process)
    if (has_b) cudaFuncSetAttribute(avr_k_parse<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * (int)AVR_PAR_STRIDE);
    else cudaFuncSetAttribute(avr_k_parse<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * (int)AVR_PAR_STRIDE);
  }
  if (has_b) avr_k_parse<true><<<grid, 32, dyn, st>>>(ps, n, raw, bins_pool, blk_pool, row_pool, row_off, tab, res, order, nlong, pack);
  else avr_k_parse<false><<<grid, 32, dyn, st>>>(ps, n, raw, bins_pool, blk_pool, row_pool, row_off, tab, res, order, nlong, pack);
  return (int)cudaGetLastError();
}
int avr_launch_gather_blocks(const AvrBlock* pool, const AvrParseSlice* ps, const uint32_t* count, const uint32_t* dst_off,
                             AvrBlock* dst, uint32_t n, cudaStream_t st) {
  if (n == 0) return 0;
  avr_k_gather_blocks<<<n < 148u * 16u ? n : 148u * 16u, 128, 0, st>>>(pool, ps, count, dst_off, dst, n);
  return (int)cudaGetLastError();
}
