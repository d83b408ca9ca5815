// Host-side fill of the constant tables uploaded to the device once per context.
#pragma once
#include <string.h>
#include "avr_bodies.h"

static inline void avr_fill_tables(AvrTables* t) {
  static const AvrDivTab div = AVR_DIV_TAB_INIT;
  t->hot.div = div;
  memcpy(t->hot.range_lps, avr_range_tab_lps, 256);
  memcpy(t->hot.trans_lps, avr_trans_idx_lps, 64);
  const int8_t* const M[4] = {avr_ctx_init_m0, avr_ctx_init_m1, avr_ctx_init_m2, avr_ctx_init_m3};
  const int8_t* const N[4] = {avr_ctx_init_n0, avr_ctx_init_n1, avr_ctx_init_n2, avr_ctx_init_n3};
  for (int p = 0; p < 4; p++) { memcpy(t->ctx_m[p], M[p], 1024); memcpy(t->ctx_n[p], N[p], 1024); }
}
