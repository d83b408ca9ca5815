// Address-space explicit accessors for the serial per-bin loops (avr_decode.h, avr_parse.h).
// On the device: ld/st.shared with a 32-bit shared-window address and ld/st.global.  With plain C++
// accesses ptxas re-derives the shared window base from SR_CgaCtaId inside the loops (an S2R per bin)
// and uses generic LD/ST with a per-iteration descriptor set-up for global pools (seen with
// tools/sass_model.py).  On the host (tests/host_emul.cpp): plain memory.
#pragma once
#include <stdint.h>
#if defined(__CUDA_ARCH__)
#define AVR_MEM_INL static __device__ __forceinline__
typedef uint32_t avr_sptr;
AVR_MEM_INL avr_sptr avr_sp(const void* p) { return (avr_sptr)__cvta_generic_to_shared(p); }
AVR_MEM_INL uint32_t avr_s_ld8(avr_sptr a) { uint32_t v; asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a) : "memory"); return v; }
AVR_MEM_INL uint32_t avr_s_ld16(avr_sptr a) { uint16_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a) : "memory"); return v; }
AVR_MEM_INL void avr_s_st8(avr_sptr a, uint32_t v) { asm volatile("st.shared.u8 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
AVR_MEM_INL void avr_s_st16(avr_sptr a, uint32_t v) { asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"((uint16_t)v) : "memory"); }
AVR_MEM_INL void avr_s_st32(avr_sptr a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
// read-only tables: not volatile, free to schedule
AVR_MEM_INL uint64_t avr_s_ld64_ro(avr_sptr a) { uint64_t v; asm("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(a)); return v; }
AVR_MEM_INL void avr_s_ld_2x64_ro(avr_sptr a, uint64_t& x, uint64_t& y) { asm("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(x), "=l"(y) : "r"(a)); }
AVR_MEM_INL uint32_t avr_g_ld16(const uint16_t* p) { uint16_t v; asm volatile("ld.global.u16 %0, [%1];" : "=h"(v) : "l"(p) : "memory"); return v; }
AVR_MEM_INL void avr_g_st16(uint16_t* p, uint32_t v) { asm volatile("st.global.u16 [%0], %1;" ::"l"(p), "h"((uint16_t)v) : "memory"); }
AVR_MEM_INL uint32_t avr_g_ld32(const void* p) { uint32_t v; asm volatile("ld.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
AVR_MEM_INL uint64_t avr_g_ld64(const void* p) { uint64_t v; asm volatile("ld.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
AVR_MEM_INL void avr_g_ld_2x64(const void* p, uint64_t& x, uint64_t& y) { asm volatile("ld.global.v2.u64 {%0, %1}, [%2];" : "=l"(x), "=l"(y) : "l"(p) : "memory"); }
#else
#if defined(__CUDACC__)
#define AVR_MEM_INL static __device__ __forceinline__
#else
#define AVR_MEM_INL static inline
#endif
typedef uintptr_t avr_sptr;
AVR_MEM_INL avr_sptr avr_sp(const void* p) { return (avr_sptr)p; }
AVR_MEM_INL uint32_t avr_s_ld8(avr_sptr a) { return *reinterpret_cast<const uint8_t*>(a); }
AVR_MEM_INL uint32_t avr_s_ld16(avr_sptr a) { return *reinterpret_cast<const uint16_t*>(a); }
AVR_MEM_INL void avr_s_st8(avr_sptr a, uint32_t v) { *reinterpret_cast<uint8_t*>(a) = (uint8_t)v; }
AVR_MEM_INL void avr_s_st16(avr_sptr a, uint32_t v) { *reinterpret_cast<uint16_t*>(a) = (uint16_t)v; }
AVR_MEM_INL void avr_s_st32(avr_sptr a, uint32_t v) { *reinterpret_cast<uint32_t*>(a) = v; }
AVR_MEM_INL uint64_t avr_s_ld64_ro(avr_sptr a) { return *reinterpret_cast<const uint64_t*>(a); }
AVR_MEM_INL void avr_s_ld_2x64_ro(avr_sptr a, uint64_t& x, uint64_t& y) { const uint64_t* e = reinterpret_cast<const uint64_t*>(a); x = e[0]; y = e[1]; }
AVR_MEM_INL uint32_t avr_g_ld16(const uint16_t* p) { return *p; }
AVR_MEM_INL void avr_g_st16(uint16_t* p, uint32_t v) { *p = (uint16_t)v; }
AVR_MEM_INL uint32_t avr_g_ld32(const void* p) { const uint16_t* q = reinterpret_cast<const uint16_t*>(p); return q[0] | ((uint32_t)q[1] << 16); }
AVR_MEM_INL uint64_t avr_g_ld64(const void* p) { return avr_g_ld32(p) | ((uint64_t)avr_g_ld32(reinterpret_cast<const uint16_t*>(p) + 2) << 32); }
AVR_MEM_INL void avr_g_ld_2x64(const void* p, uint64_t& x, uint64_t& y) { x = avr_g_ld64(p); y = avr_g_ld64(reinterpret_cast<const uint16_t*>(p) + 4); }
#endif
