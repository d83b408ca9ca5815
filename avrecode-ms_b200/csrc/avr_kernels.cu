// sm_100a kernels of the entropy hot path.  Launchers are declared in avr_kernels.h and take
// raw device pointers + a stream; no torch types, no host fallback.
//
// Sequential ("chain") kernels: one warp per file, lane 0 walks the chain (the walk is a
// strict per-bin dependency), the other lanes help with bulk work (zeroing the
// picture store, CABAC state init).  Slice-parallel compress pipeline: see avr_pipeline.h.
#include <cuda_runtime.h>
#include "avr_bodies.h"
#include "avr_kernels.h"
#include "avr_pipeline.h"

// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) avr_k_compress_chain(const AvrFileDesc* fds, int nfiles, const uint16_t* bins,
                                                           const AvrBlock* blocks, const AvrSlice* slices,
                                                           const AvrFrame* frames, uint16_t* est_pool, uint8_t* store_pool,
                                                           const AvrTables* tab, uint8_t* scratch, const uint64_t* scratch_off,
                                                           const uint32_t* scratch_cap, uint32_t* out_len, int* status) {
  int f = blockIdx.x;
  if (f >= nfiles || threadIdx.x != 0) return;
  const AvrFileDesc fd = fds[f];
  uint8_t* s0 = store_pool + (size_t)fd.mb_base * 2 * AVR_MB_STRIDE;
  uint8_t* s1 = s0 + (size_t)fd.max_mbs * AVR_MB_STRIDE;
  avr_compress_chain_body(fd, bins, blocks, slices, frames, est_pool + (size_t)f * AVR_KEY_COUNT, s0, s1, tab, scratch,
                          scratch_off, scratch_cap, out_len, status);
}

__global__ void __launch_bounds__(32) avr_k_decompress_chain(const AvrFileDesc* fds, int nfiles, const AvrDecSlice* slices,
                                                             const uint8_t* payloads, uint16_t* est_pool, uint8_t* store_pool,
                                                             avr::MbInfo* row_pool, const AvrTables* tab, uint8_t* out,
                                                             uint32_t* out_len, uint32_t* out_bins, int* status) {
  __shared__ uint8_t st[1024];
  int f = blockIdx.x;
  if (f >= nfiles || threadIdx.x != 0) return;
  const AvrFileDesc fd = fds[f];
  uint8_t* s0 = store_pool + (size_t)fd.mb_base * 2 * AVR_MB_STRIDE;
  uint8_t* s1 = s0 + (size_t)fd.max_mbs * AVR_MB_STRIDE;
  avr_decompress_chain_body(fd, slices, payloads, est_pool + (size_t)f * AVR_KEY_COUNT, s0, s1, row_pool + fd.mb_base, st, tab,
                            out, out_len, out_bins, status);
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
  for (uint32_t base = 0; base < n; base += 1024) {
    uint32_t i = base + threadIdx.x;
    uint64_t v = i < n ? len[i] : 0, x = v;
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
    if (i < n) off[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry = excl + v;
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

int avr_launch_decompress_chain(const AvrDevDBatch& b, cudaStream_t st) {
  if (b.nfiles == 0) return 0;
  avr_k_decompress_chain<<<b.nfiles, 32, 0, st>>>(b.fds, b.nfiles, b.slices, b.payloads, b.est_pool, b.store_pool, b.row_pool,
                                                 b.tab, b.out, b.out_len, b.out_bins, b.status);
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
