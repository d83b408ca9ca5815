// avr_streams_encode / avr_streams_decode: S independent bin streams through the generic range coder on the GPU
// (BASELINE configs[0]; bodies and reference citations in avr_streams.h).  One stream per thread, 128 streams per
// block; per-stream scratch is compacted by the same device-side exclusive scan + gather the slice payloads use.
#include <cuda_runtime.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "../../include/avrecode_b200.h"
#include "avr_api_internal.h"
#include "avr_kernels.h"
#include "avr_streams.h"

#define AVR_STREAM_BLOCK 128

template <int DIGIT_BITS, int MODEL>
__global__ void __launch_bounds__(AVR_STREAM_BLOCK) avr_k_stream_encode(const uint8_t* bits, uint32_t nstreams, uint32_t nbins,
                                                                        const AvrDivTab* div, uint8_t* scratch, uint32_t cap,
                                                                        uint32_t* len, int* status) {
  __shared__ AvrDivTab dt;
  if (MODEL == AVR_STREAM_ADAPTIVE) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(div);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&dt);
    for (uint32_t i = threadIdx.x; i < sizeof(AvrDivTab) / 4; i += AVR_STREAM_BLOCK) dst[i] = src[i];
    __syncthreads();
  }
  const uint32_t s = blockIdx.x * AVR_STREAM_BLOCK + threadIdx.x;
  if (s < nstreams) avr_body_stream_encode<DIGIT_BITS, MODEL>(s, bits, nbins, &dt, scratch, cap, len, status);
}

template <int DIGIT_BITS, int MODEL>
__global__ void __launch_bounds__(AVR_STREAM_BLOCK) avr_k_stream_decode(const uint8_t* coded, const uint64_t* coded_off, uint32_t nstreams,
                                                                        uint32_t nbins, const AvrDivTab* div, uint8_t* bits_out) {
  __shared__ AvrDivTab dt;
  if (MODEL == AVR_STREAM_ADAPTIVE) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(div);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&dt);
    for (uint32_t i = threadIdx.x; i < sizeof(AvrDivTab) / 4; i += AVR_STREAM_BLOCK) dst[i] = src[i];
    __syncthreads();
  }
  const uint32_t s = blockIdx.x * AVR_STREAM_BLOCK + threadIdx.x;
  if (s < nstreams) avr_body_stream_decode<DIGIT_BITS, MODEL>(s, coded, coded_off, nbins, &dt, bits_out);
}

__global__ void __launch_bounds__(256) avr_k_stream_scratch_off(uint64_t* off, uint32_t n, uint32_t cap) {
  const uint32_t s = blockIdx.x * 256 + threadIdx.x;
  if (s < n) off[s] = (uint64_t)s * cap;
}

#define CU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { avr_set_error("%s: %s", #x, cudaGetErrorString(e_)); rc = AVR_E_CUDA; goto done; } } while (0)

static int check_args(avr_context* ctx, uint32_t nstreams, uint32_t nbins, int digit_bits, int model) {
  if (!ctx) { avr_set_error("streams: no context (this library has no CPU path)"); return AVR_E_ARG; }
  if (digit_bits != 8 && digit_bits != 16) { avr_set_error("streams: digit_bits must be 8 or 16"); return AVR_E_ARG; }
  if (model != AVR_STREAM_FIXED_HALF && model != AVR_STREAM_ADAPTIVE) { avr_set_error("streams: unknown model %d", model); return AVR_E_ARG; }
  if ((uint64_t)nbins + 64 > 0xFFFFFFFFull / 2) { avr_set_error("streams: nbins too large"); return AVR_E_ARG; }
  (void)nstreams;
  return AVR_OK;
}
static float ev_ms(cudaEvent_t a, cudaEvent_t b) { float ms = 0; cudaEventElapsedTime(&ms, a, b); return ms; }

extern "C" int avr_streams_encode(avr_context* ctx, const uint8_t* bits, uint32_t nstreams, uint32_t nbins, int digit_bits, int model,
                                  avr_buffer* coded, uint64_t* coded_off, avr_timing* t) {
  int rc = check_args(ctx, nstreams, nbins, digit_bits, model);
  if (rc) return rc;
  if (!coded || !coded_off || (!bits && nstreams && nbins)) { avr_set_error("streams: null argument"); return AVR_E_ARG; }
  coded->data = nullptr; coded->size = 0;
  if (t) memset(t, 0, sizeof *t);
  if (nstreams == 0) { coded_off[0] = 0; return AVR_OK; }
  if (cudaSetDevice(avr_context_device(ctx)) != cudaSuccess) { avr_set_error("streams: cudaSetDevice failed"); return AVR_E_CUDA; }
  cudaStream_t st = (cudaStream_t)avr_context_stream(ctx);
  const uint32_t cap = avr_stream_cap(nbins, model);
  const size_t nbytes = (size_t)nstreams * nbins;
  uint8_t *d_bits = nullptr, *d_scratch = nullptr, *d_out = nullptr;
  uint32_t* d_len = nullptr; int* d_status = nullptr; uint64_t *d_off = nullptr, *d_soff = nullptr; AvrDivTab* d_div = nullptr;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  std::vector<int> status(nstreams);
  std::vector<uint32_t> len(nstreams);
  uint64_t total = 0;
  const AvrDivTab h_div = AVR_DIV_TAB_INIT;
  const uint32_t grid = (nstreams + AVR_STREAM_BLOCK - 1) / AVR_STREAM_BLOCK;
  for (auto& e : ev) CU(cudaEventCreate(&e));
  CU(cudaMalloc(&d_bits, nbytes + 8));
  CU(cudaMalloc(&d_scratch, (size_t)nstreams * cap));
  CU(cudaMalloc(&d_out, (size_t)nstreams * cap));
  CU(cudaMalloc(&d_len, sizeof(uint32_t) * nstreams));
  CU(cudaMalloc(&d_status, sizeof(int) * nstreams));
  CU(cudaMalloc(&d_off, sizeof(uint64_t) * (nstreams + 1)));
  CU(cudaMalloc(&d_soff, sizeof(uint64_t) * nstreams));
  CU(cudaMalloc(&d_div, sizeof(AvrDivTab)));
  CU(cudaEventRecord(ev[0], st));
  CU(cudaMemcpyAsync(d_div, &h_div, sizeof h_div, cudaMemcpyHostToDevice, st));
  if (nbytes) CU(cudaMemcpyAsync(d_bits, bits, nbytes, cudaMemcpyHostToDevice, st));
  CU(cudaEventRecord(ev[1], st));
#define LAUNCH_ENC(D, M) avr_k_stream_encode<D, M><<<grid, AVR_STREAM_BLOCK, 0, st>>>(d_bits, nstreams, nbins, d_div, d_scratch, cap, d_len, d_status)
  if (digit_bits == 8) { if (model == AVR_STREAM_ADAPTIVE) LAUNCH_ENC(8, AVR_STREAM_ADAPTIVE); else LAUNCH_ENC(8, AVR_STREAM_FIXED_HALF); }
  else { if (model == AVR_STREAM_ADAPTIVE) LAUNCH_ENC(16, AVR_STREAM_ADAPTIVE); else LAUNCH_ENC(16, AVR_STREAM_FIXED_HALF); }
#undef LAUNCH_ENC
  CU(cudaGetLastError());
  {
    // per-stream scratch -> one contiguous buffer (device-side exclusive scan of the lengths, then a gather)
    avr_k_stream_scratch_off<<<(nstreams + 255) / 256, 256, 0, st>>>(d_soff, nstreams, cap);
    const int lrc = avr_launch_compact(d_scratch, d_soff, d_len, d_off, d_off + nstreams, d_out, nstreams, st);
    if (lrc) { avr_set_error("streams: compaction failed (%d)", lrc); rc = AVR_E_CUDA; goto done; }
  }
  CU(cudaEventRecord(ev[2], st));
  CU(cudaMemcpyAsync(coded_off, d_off, sizeof(uint64_t) * (nstreams + 1), cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(status.data(), d_status, sizeof(int) * nstreams, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(len.data(), d_len, sizeof(uint32_t) * nstreams, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  for (uint32_t s = 0; s < nstreams; s++)
    if (status[s]) { avr_set_error("streams: stream %u failed with %d (scratch %u bytes, wrote %u)", s, status[s], cap, len[s]); rc = status[s]; goto done; }
  total = coded_off[nstreams];
  coded->data = (uint8_t*)malloc(total ? total : 1);
  if (!coded->data) { avr_set_error("streams: out of host memory"); rc = AVR_E_NOMEM; goto done; }
  coded->size = total;
  if (total) CU(cudaMemcpyAsync(coded->data, d_out, total, cudaMemcpyDeviceToHost, st));
  CU(cudaEventRecord(ev[3], st));
  CU(cudaStreamSynchronize(st));
  if (t) {
    t->h2d_ms = ev_ms(ev[0], ev[1]); t->kernel_ms = ev_ms(ev[1], ev[2]); t->d2h_ms = ev_ms(ev[2], ev[3]);
    t->h2d_bytes = nbytes; t->d2h_bytes = total + (uint64_t)nstreams * 16 + 8;
    t->bins = (uint64_t)nstreams * nbins; t->coded_symbols = t->bins; t->launches = 4;
    t->coder_launch_ms_x1000 = (uint32_t)(t->kernel_ms * 1000.0);
  }
done:
  if (rc && coded->data) { free(coded->data); coded->data = nullptr; coded->size = 0; }
  cudaFree(d_bits); cudaFree(d_scratch); cudaFree(d_out); cudaFree(d_len); cudaFree(d_status); cudaFree(d_off); cudaFree(d_soff); cudaFree(d_div);
  for (auto& e : ev) if (e) cudaEventDestroy(e);
  return rc;
}

extern "C" int avr_streams_decode(avr_context* ctx, const uint8_t* coded, const uint64_t* coded_off, uint32_t nstreams, uint32_t nbins,
                                  int digit_bits, int model, uint8_t* bits_out, avr_timing* t) {
  int rc = check_args(ctx, nstreams, nbins, digit_bits, model);
  if (rc) return rc;
  if (!coded_off || (nstreams && nbins && !bits_out)) { avr_set_error("streams: null argument"); return AVR_E_ARG; }
  if (t) memset(t, 0, sizeof *t);
  if (nstreams == 0 || nbins == 0) return AVR_OK;
  for (uint32_t s = 0; s < nstreams; s++)
    if (coded_off[s + 1] < coded_off[s] || coded_off[s + 1] - coded_off[s] > 0xFFFFFFFFull) { avr_set_error("streams: offsets not ascending at %u", s); return AVR_E_ARG; }
  const uint64_t total = coded_off[nstreams];
  if (total && !coded) { avr_set_error("streams: null argument"); return AVR_E_ARG; }
  if (cudaSetDevice(avr_context_device(ctx)) != cudaSuccess) { avr_set_error("streams: cudaSetDevice failed"); return AVR_E_CUDA; }
  cudaStream_t st = (cudaStream_t)avr_context_stream(ctx);
  const size_t nbytes = (size_t)nstreams * nbins;
  uint8_t *d_coded = nullptr, *d_bits = nullptr; uint64_t* d_off = nullptr; AvrDivTab* d_div = nullptr;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  const AvrDivTab h_div = AVR_DIV_TAB_INIT;
  const uint32_t grid = (nstreams + AVR_STREAM_BLOCK - 1) / AVR_STREAM_BLOCK;
  for (auto& e : ev) CU(cudaEventCreate(&e));
  CU(cudaMalloc(&d_coded, total + 8));
  CU(cudaMalloc(&d_bits, nbytes + 8));
  CU(cudaMalloc(&d_off, sizeof(uint64_t) * (nstreams + 1)));
  CU(cudaMalloc(&d_div, sizeof(AvrDivTab)));
  CU(cudaEventRecord(ev[0], st));
  CU(cudaMemcpyAsync(d_div, &h_div, sizeof h_div, cudaMemcpyHostToDevice, st));
  if (total) CU(cudaMemcpyAsync(d_coded, coded, total, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_off, coded_off, sizeof(uint64_t) * (nstreams + 1), cudaMemcpyHostToDevice, st));
  CU(cudaEventRecord(ev[1], st));
#define LAUNCH_DEC(D, M) avr_k_stream_decode<D, M><<<grid, AVR_STREAM_BLOCK, 0, st>>>(d_coded, d_off, nstreams, nbins, d_div, d_bits)
  if (digit_bits == 8) { if (model == AVR_STREAM_ADAPTIVE) LAUNCH_DEC(8, AVR_STREAM_ADAPTIVE); else LAUNCH_DEC(8, AVR_STREAM_FIXED_HALF); }
  else { if (model == AVR_STREAM_ADAPTIVE) LAUNCH_DEC(16, AVR_STREAM_ADAPTIVE); else LAUNCH_DEC(16, AVR_STREAM_FIXED_HALF); }
#undef LAUNCH_DEC
  CU(cudaGetLastError());
  CU(cudaEventRecord(ev[2], st));
  CU(cudaMemcpyAsync(bits_out, d_bits, nbytes, cudaMemcpyDeviceToHost, st));
  CU(cudaEventRecord(ev[3], st));
  CU(cudaStreamSynchronize(st));
  if (t) {
    t->h2d_ms = ev_ms(ev[0], ev[1]); t->kernel_ms = ev_ms(ev[1], ev[2]); t->d2h_ms = ev_ms(ev[2], ev[3]);
    t->h2d_bytes = total + (uint64_t)(nstreams + 1) * 8; t->d2h_bytes = nbytes;
    t->bins = (uint64_t)nstreams * nbins; t->coded_symbols = t->bins; t->launches = 1;
    t->coder_launch_ms_x1000 = (uint32_t)(t->kernel_ms * 1000.0);
  }
done:
  cudaFree(d_coded); cudaFree(d_bits); cudaFree(d_off); cudaFree(d_div);
  for (auto& e : ev) if (e) cudaEventDestroy(e);
  return rc;
}

// ---- avr_streams_encode_static: the encode above for callers that may not allocate ------------------------------
// No counterpart in the reference (include/avrecode_b200.h).
// Same kernels as avr_streams_encode (avr_k_stream_encode, scratch-offset fill, exclusive scan + gather); what differs
// is where the intermediate buffers live: the per-stream scratch, lengths, status words and offsets are carved out of
// ONE file-scope device array instead of cudaMalloc'ed blocks.  Every GPU that loads this library holds the array
// until the process exits.  All calls share it (it keeps the previous call's contents): calls that overlap on the GPU
// race, and the words that must start at zero are cleared by every call.
static __device__ __align__(256) unsigned char g_streams_static_scratch[AVR_STREAMS_STATIC_SCRATCH_BYTES];

// In host code `g_streams_static_scratch` is a host-side stand-in, not a device address: the device address is asked
// for once per device, with that device current, and cached.
#include <mutex>
static std::mutex g_static_mu;
static unsigned char* g_static_ptr[64];
static int static_scratch_ptr(int device, unsigned char** out) {
  if (device < 0 || device >= 64) { avr_set_error("streams_static: device %d out of range", device); return AVR_E_ARG; }
  std::lock_guard<std::mutex> g(g_static_mu);
  if (!g_static_ptr[device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_streams_static_scratch);      // the caller made `device` current
    if (e != cudaSuccess || !p) { avr_set_error("streams_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e)); return AVR_E_CUDA; }
    g_static_ptr[device] = (unsigned char*)p;
  }
  *out = g_static_ptr[device];
  return AVR_OK;
}
// first non-zero status of the streams -> *first (cleared by the caller)
__global__ void __launch_bounds__(256) avr_k_stream_first_status(const int* status, uint32_t n, int* first) {
  const uint32_t s = blockIdx.x * 256 + threadIdx.x;
  if (s < n && status[s]) atomicCAS(first, 0, status[s]);
}

extern "C" int avr_streams_encode_static(avr_context* ctx, const uint8_t* d_bits, uint32_t nstreams, uint32_t nbins, int digit_bits,
                                         int model, uint8_t* d_coded, uint64_t coded_cap, uint64_t* d_coded_off) {
  int rc = check_args(ctx, nstreams, nbins, digit_bits, model);
  if (rc) return rc;
  if (!d_coded_off || (nstreams && (!d_coded || (!d_bits && nbins)))) { avr_set_error("streams_static: null argument"); return AVR_E_ARG; }
  const uint32_t cap = avr_stream_cap(nbins, model);
  // layout inside the array (every part 256-byte aligned)
  auto up = [](uint64_t x) { return (x + 255) & ~255ull; };
  const uint64_t o_div = 0, o_first = up(sizeof(AvrDivTab)), o_len = o_first + 256, o_status = o_len + up(4ull * nstreams),
                 o_soff = o_status + up(4ull * nstreams), o_scratch = o_soff + up(8ull * nstreams);
  const uint64_t need = o_scratch + (uint64_t)nstreams * cap;
  if (need > AVR_STREAMS_STATIC_SCRATCH_BYTES) {
    avr_set_error("streams_static: %u streams x %u bins need %llu bytes of scratch, the static array holds %u (use avr_streams_encode)",
                  nstreams, nbins, (unsigned long long)need, (unsigned)AVR_STREAMS_STATIC_SCRATCH_BYTES);
    return AVR_E_ARG;
  }
  if (coded_cap < (uint64_t)nstreams * cap) {
    avr_set_error("streams_static: coded_cap %llu < nstreams * AVR_STREAMS_CODED_BOUND = %llu", (unsigned long long)coded_cap,
                  (unsigned long long)nstreams * cap);
    return AVR_E_ARG;
  }
  if (cudaSetDevice(avr_context_device(ctx)) != cudaSuccess) { avr_set_error("streams_static: cudaSetDevice failed"); return AVR_E_CUDA; }
  unsigned char* base = nullptr;
  if ((rc = static_scratch_ptr(avr_context_device(ctx), &base))) return rc;
  cudaStream_t st = (cudaStream_t)avr_context_stream(ctx);
  AvrDivTab* d_div = (AvrDivTab*)(base + o_div);
  int* d_first = (int*)(base + o_first);
  uint32_t* d_len = (uint32_t*)(base + o_len);
  int* d_status = (int*)(base + o_status);
  uint64_t* d_soff = (uint64_t*)(base + o_soff);
  uint8_t* d_scratch = base + o_scratch;
  static const AvrDivTab h_div = AVR_DIV_TAB_INIT;
  int first = 0;
  CU(cudaMemsetAsync(d_first, 0, 256, st));                                  // must start at zero; the rest is overwritten before it is read
  if (nstreams == 0) { CU(cudaMemsetAsync(d_coded_off, 0, sizeof(uint64_t), st)); CU(cudaStreamSynchronize(st)); return AVR_OK; }
  CU(cudaMemcpyAsync(d_div, &h_div, sizeof h_div, cudaMemcpyHostToDevice, st));
  {
    const uint32_t grid = (nstreams + AVR_STREAM_BLOCK - 1) / AVR_STREAM_BLOCK;
#define LAUNCH_ENC(D, M) avr_k_stream_encode<D, M><<<grid, AVR_STREAM_BLOCK, 0, st>>>(d_bits, nstreams, nbins, d_div, d_scratch, cap, d_len, d_status)
    if (digit_bits == 8) { if (model == AVR_STREAM_ADAPTIVE) LAUNCH_ENC(8, AVR_STREAM_ADAPTIVE); else LAUNCH_ENC(8, AVR_STREAM_FIXED_HALF); }
    else { if (model == AVR_STREAM_ADAPTIVE) LAUNCH_ENC(16, AVR_STREAM_ADAPTIVE); else LAUNCH_ENC(16, AVR_STREAM_FIXED_HALF); }
#undef LAUNCH_ENC
    CU(cudaGetLastError());
    avr_k_stream_scratch_off<<<(nstreams + 255) / 256, 256, 0, st>>>(d_soff, nstreams, cap);
    const int lrc = avr_launch_compact(d_scratch, d_soff, d_len, d_coded_off, d_coded_off + nstreams, d_coded, nstreams, st);
    if (lrc) { avr_set_error("streams_static: compaction failed (%d)", lrc); return AVR_E_CUDA; }
    avr_k_stream_first_status<<<(nstreams + 255) / 256, 256, 0, st>>>(d_status, nstreams, d_first);
    CU(cudaGetLastError());
  }
  CU(cudaMemcpyAsync(&first, d_first, sizeof first, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (first) { avr_set_error("streams_static: a stream failed with %d", first); rc = first; }
done:
  return rc;
}

// ---- avr_measure_recurrence_floor: what ONE thread can do per bin when nothing but the decoder's recurrence is left ----
// low/range -> r1 = (range / total) * pos -> symbol = low >= range - r1 -> low/range (arithmetic_code.h:232-248 with the
// exact multiply-high division of avr_div_tab.h), a CONSTANT estimator (pos 40 of total 81), digits from a register
// generator: no estimator load/update, no context selection, no CABAC re-encode, no ring.  The chain kernel's
// cycles-per-bin are reported against this number (bench.py roofline.serial_chain); it replaces a planning constant.
__global__ void __launch_bounds__(32) avr_k_recurrence_floor(uint32_t steps, uint64_t m, uint32_t sh, uint32_t pos, uint64_t seed,
                                                             unsigned long long* out) {
  if (threadIdx.x != 0) return;
  uint64_t low = seed >> 1, range = 1ull << 63, x = seed;
  uint32_t ones = 0;
  const long long t0 = clock64();
  for (uint32_t i = 0; i < steps; i++) {
    const uint64_t r1 = (__umul64hi(range, m) >> sh) * (uint64_t)pos;
    const uint64_t r0 = range - r1;
    const uint32_t sym = low >= r0;
    low -= sym ? r0 : 0ull;
    range = sym ? r1 : r0;
    ones += sym;
    if (range < (1ull << 51)) {
      while (range < (1ull << 55)) {
        x = x * 6364136223846793005ull + 1442695040888963407ull;          // the next digit (off the recurrence: does not depend on low/range)
        low = (low << 8) + (x >> 56);
        range <<= 8;
      }
    }
  }
  const long long t1 = clock64();
  out[0] = (unsigned long long)(t1 - t0); out[1] = ones; out[2] = low ^ range;
}
extern "C" int avr_measure_recurrence_floor(avr_context* ctx, uint32_t steps, double* cycles_per_bin) {
  if (!ctx || !cycles_per_bin || steps == 0) { avr_set_error("recurrence_floor: bad argument"); return AVR_E_ARG; }
  int rc = AVR_OK;
  if (cudaSetDevice(avr_context_device(ctx)) != cudaSuccess) { avr_set_error("recurrence_floor: cudaSetDevice failed"); return AVR_E_CUDA; }
  cudaStream_t st = (cudaStream_t)avr_context_stream(ctx);
  static const AvrDivTab dt = AVR_DIV_TAB_INIT;
  unsigned long long* d_out = nullptr; unsigned long long h[3] = {0, 0, 0};
  CU(cudaMalloc(&d_out, sizeof h));
  for (int rep = 0; rep < 2; rep++) {                                        // the first launch loads the module
    avr_k_recurrence_floor<<<1, 32, 0, st>>>(steps, dt.m[81], dt.s[81], 40, 0x9E3779B97F4A7C15ull, d_out);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(h, d_out, sizeof h, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
  }
  *cycles_per_bin = (double)h[0] / steps;
done:
  cudaFree(d_out);
  return rc;
}
