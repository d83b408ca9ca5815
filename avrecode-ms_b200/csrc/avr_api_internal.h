// Internal C++ glue shared by avr_api.cu and avr_files.cpp (not part of the C ABI).
#pragma once
#include <stdint.h>
#include "avr_record.h"
struct avr_record;
void avr_set_error(const char* fmt, ...);
const avr::FileRecord* avr_record_view(const avr_record* r);
avr::FileRecord* avr_record_mut(avr_record* r);
uint32_t avr_span_flip(const avr::FileRecord& rec, uint32_t span_index);
// AVR_TRACE=1: stage wall times on stderr (diagnostics for the e2e path; off by default)
void avr_trace(const char* what);
struct avr_context;
int avr_context_device(const avr_context* ctx);   // CUDA device ordinal of the context
avr_context* avr_context_lane(avr_context* ctx, int k);   // lane k of a context (0 = itself), created on first use
// chains that will be resident at the same time across concurrent decompress launches (carve-out sizing); 0 = this launch only
void avr_set_chain_concurrency_hint(int files);
