// EXPERIMENTAL decompress kernel: one WARP per chain, lanes as speculative symbol paths (avr_decode_spec.h).
// Bit-exact (tests/test_host_logic.py with the lanes emulated; tests/test_gpu_parity.py runs it with AVR_DEC_SPEC=1) but measured
// slower than the scalar chain on B200 (profiles/r02_summary.md): kept as the measured answer to "use the idle
// lanes for speculative multi-bin decode".  Own translation unit: it shares names with avr_decode.h.
#include <cuda_runtime.h>
#include "avr_bodies.h"
#include "avr_kernels.h"
#include "avr_decode_spec.h"

// One block (two warps) per file.  Warp 0 is the chain: all 32 lanes walk slice_data() together and act as the 32
// speculative paths inside residual blocks (avr_decode.h); thread 32 is the CABAC encoder fed through a shared-memory
// ring; all 64 threads do the bulk steps between slices: zeroing the picture stores on a flip, CABAC context entries,
// staging the per-bin tables and the default-key estimators in shared memory.
#define AVR_DEC_THREADS 64
__global__ void __launch_bounds__(AVR_DEC_THREADS) avr_k_decompress_chain_spec(
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
  avr_ms_fill(tab, tid, AVR_DEC_THREADS);
  avr_cab_fill_trans(tab, tid, AVR_DEC_THREADS);
  for (uint32_t i = tid; i < AVR_NUM_CTX_KEYS; i += AVR_DEC_THREADS) g_dec.est_def[i] = AVR_EST_INIT;
  if (tid == 0) {
    uint8_t* s0 = store_pool + (size_t)fd.mb_base * AVR_MB_STRIDE;
    avr_dec_file_begin(est_pool + (size_t)f * AVR_KEY_COUNT, s0, s0 + (size_t)fd.max_mbs * AVR_MB_STRIDE);
    s_failed = 0;
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
    avr_cab_init_entries(tab, ds.sp.slice_type == avr::SLICE_I, ds.sp.cabac_init_idc, ds.sp.slice_qp, tid, AVR_DEC_THREADS);
    __syncthreads();
    int rc = 0;
    if (tid < 32) rc = avr_dec_slice(ds, payloads, row_pool + fd.mb_base, ++uid, &out_bins[g]);
    else if (tid == 32) avr_cab_encoder_thread(out + ds.out_off, ds.out_cap);
    __syncthreads();
    if (tid == 0) {
      if (!rc) rc = g_dec.cab_err;
      out_len[g] = g_dec.cab_n; status[g] = rc;
      if (rc) s_failed = 1;
    }
    __syncthreads();
  }
}


int avr_launch_decompress_chain_spec(const AvrDevDBatch& b, cudaStream_t st) {
  if (b.nfiles == 0) return 0;
  avr_k_decompress_chain_spec<<<b.nfiles, AVR_DEC_THREADS, 0, st>>>(b.fds, b.nfiles, b.slices, b.payloads, b.est_pool, b.store_pool, b.row_pool,
                                                                    b.tab, b.out, b.out_len, b.out_bins, b.status, b.file_order);
  return (int)cudaGetLastError();
}
