// Launchers of the sm_100a kernels (avr_kernels.cu).  Raw device pointers and a stream; the
// host API (avr_api.cu) owns the memory.
#pragma once
#include <cuda_runtime.h>
#include "avr_pipeline.h"
struct AvrParseSlice;
struct AvrParseResult;

// Device view of a staged compress batch.
struct AvrDevBatch {
  int nfiles;
  uint32_t nslices, nblocks;
  uint64_t nbins;
  const AvrFileDesc* fds;
  const uint16_t* bins;
  const AvrBlock* blocks;
  const AvrSlice* slices;          // file-local slice table (chain kernel)
  const AvrBatchSlice* bslices;    // batch-global slice table (pipeline)
  const uint32_t* code_order;      // slices by descending bin count (pass D launch order), or null
  uint32_t code_long;              // the first code_long slices of that order are coded one per WARP, the rest one per LANE
  const AvrFrame* frames;
  const AvrTables* tab;
  // chain kernel state
  uint16_t* est_pool;              // [nfiles][AVR_KEY_COUNT]
  uint8_t* store_pool;             // [sum 2*max_mbs][64]
  // pipeline state
  uint8_t* pic_store; size_t pic_store_bytes;
  uint32_t* seg;                   // [nblocks + nslices + 1]
  uint32_t* keys_a; uint32_t* keys_b; uint32_t* idx_a; uint32_t* idx_b;   // [emit_cap]
  uint32_t* head; uint32_t* psum; AvrHalfSeg* fn;                         // [emit_cap]
  uint16_t* prob;                  // [emit_cap]
  uint32_t emit_cap;               // exact number of coded symbols (host computes it while staging)
  void* cub_tmp; size_t cub_tmp_bytes;
  int sort_end_bit;                // AVR_KEY_BITS + 1 + bits(nfiles)
  // outputs
  uint8_t* scratch; const uint64_t* scratch_off; const uint32_t* scratch_cap;
  uint32_t* out_len; int* status;
  uint64_t* out_off; uint64_t* out_total; uint8_t* out;
};

struct AvrDevDBatch {
  int nfiles;
  int has_b;                       // a slice of the batch is a B slice: launch the kernel that holds the B parser instance
  uint32_t nslices;
  const AvrFileDesc* fds;
  const AvrDecSlice* slices;
  const uint32_t* file_order;      // files by descending payload bytes (block -> file), or null
  const uint8_t* payloads;
  const AvrTables* tab;
  uint16_t* est_pool;
  uint8_t* store_pool;
  avr::MbInfo* row_pool;
  uint8_t* out; uint32_t* out_len; uint32_t* out_bins; int* status;
};

int avr_launch_compress_chain(const AvrDevBatch& b, cudaStream_t st);
int avr_launch_compress_pipeline(const AvrDevBatch& b, cudaStream_t st, int* launches, cudaEvent_t coder_begin, cudaEvent_t coder_end);
size_t avr_pipeline_tmp_bytes(uint32_t nemit, uint32_t nentries, int sort_end_bit);
int avr_launch_decompress_chain(const AvrDevDBatch& b, cudaStream_t st);
int avr_launch_decompress_chain_spec(const AvrDevDBatch& b, cudaStream_t st);   // avr_kernels_spec.cu (experimental, AVR_DEC_SPEC=1)
int avr_launch_compact(const uint8_t* scratch, const uint64_t* scratch_off, const uint32_t* len, uint64_t* out_off,
                       uint64_t* total, uint8_t* out, uint32_t n, cudaStream_t st);
int avr_launch_parse(const AvrParseSlice* ps, uint32_t n, uint8_t* raw, uint16_t* bins_pool, AvrBlock* blk_pool,
                     avr::MbInfo* row_pool, const uint64_t* row_off, const AvrTables* tab, AvrParseResult* res, const uint32_t* order, uint32_t nlong, int has_b, cudaStream_t st);
int avr_launch_gather_blocks(const AvrBlock* pool, const AvrParseSlice* ps, const uint32_t* count, const uint32_t* dst_off,
                             AvrBlock* dst, uint32_t n, cudaStream_t st);
