// 32-bit multiply-add / add primitives with explicit carry flag.
//
// Device: one PTX instruction each (mad.lo.cc.u32, madc.hi.cc.u32, ...).  ptxas pairs an adjacent
// mad.lo/madc.hi on the same operands into a single IMAD.WIDE.U32(.X), which is what makes the
// even/odd accumulator layout in field.cuh cost one integer-pipe instruction per 32x32 product.
// Every instruction is `asm volatile` so the NVVM front end keeps carry chains in program order
// (it does not know about the CC register); ptxas tracks CC itself and is free to schedule.
//
// Host: the same functions emulate CC with a thread_local flag so the multi-limb algorithms in
// field.cuh / ec.cuh can be unit-tested on the CPU box (tests/host_check.cpp) instruction for
// instruction before they ever run on a GPU.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define KZG_HD __host__ __device__ __forceinline__
#define KZG_D __device__ __forceinline__
// group-law level routines are real calls: forcing ~10-40 inlined Montgomery products into every
// call site makes NVVM's compile time explode and buys nothing (a call costs a few cycles per ~4000
// instructions of work).  The hot accumulate loop opts back in explicitly where it matters.
#define KZG_HDN __host__ __device__ __noinline__
#else
#define KZG_HD inline
#define KZG_D inline
#define KZG_HDN __attribute__((noinline))
#endif

namespace kzgb200 {
namespace ptx {

#if defined(__CUDA_ARCH__)

KZG_D uint32_t mul_lo(uint32_t a, uint32_t b) { return a * b; }
KZG_D uint32_t mul_hi(uint32_t a, uint32_t b) { return __umulhi(a, b); }

// d = lo(a*b) + c ; CC = carry
KZG_D uint32_t mad_lo_cc(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d; asm volatile("mad.lo.cc.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d;
}
// d = lo(a*b) + c + CC ; CC = carry
KZG_D uint32_t madc_lo_cc(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d; asm volatile("madc.lo.cc.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d;
}
// d = hi(a*b) + c ; CC = carry
KZG_D uint32_t mad_hi_cc(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d; asm volatile("mad.hi.cc.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d;
}
// d = hi(a*b) + c + CC ; CC = carry
KZG_D uint32_t madc_hi_cc(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d; asm volatile("madc.hi.cc.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d;
}
// d = hi(a*b) + c + CC ; CC unchanged/unused afterwards
KZG_D uint32_t madc_hi(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d; asm volatile("madc.hi.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d;
}
KZG_D uint32_t add_cc(uint32_t a, uint32_t b) {
  uint32_t d; asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d;
}
KZG_D uint32_t addc_cc(uint32_t a, uint32_t b) {
  uint32_t d; asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d;
}
KZG_D uint32_t addc(uint32_t a, uint32_t b) {
  uint32_t d; asm volatile("addc.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d;
}
KZG_D uint32_t sub_cc(uint32_t a, uint32_t b) {
  uint32_t d; asm volatile("sub.cc.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d;
}
KZG_D uint32_t subc_cc(uint32_t a, uint32_t b) {
  uint32_t d; asm volatile("subc.cc.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d;
}
// d = a - b - borrow  (for sub: CC.CF = 1 means borrow in PTX semantics is handled by the ISA)
KZG_D uint32_t subc(uint32_t a, uint32_t b) {
  uint32_t d; asm volatile("subc.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d;
}

#else  // ---------------------------------------------------------------- host emulation

inline uint32_t& cc_flag() { static thread_local uint32_t cc = 0; return cc; }

inline uint32_t mul_lo(uint32_t a, uint32_t b) { return a * b; }
inline uint32_t mul_hi(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }

inline uint32_t emu_add3(uint32_t x, uint32_t y, uint32_t cin, bool set) {
  uint64_t s = (uint64_t)x + y + cin;
  if (set) cc_flag() = (uint32_t)(s >> 32);
  return (uint32_t)s;
}
inline uint32_t mad_lo_cc(uint32_t a, uint32_t b, uint32_t c) { return emu_add3(mul_lo(a, b), c, 0, true); }
inline uint32_t madc_lo_cc(uint32_t a, uint32_t b, uint32_t c) { return emu_add3(mul_lo(a, b), c, cc_flag(), true); }
inline uint32_t mad_hi_cc(uint32_t a, uint32_t b, uint32_t c) { return emu_add3(mul_hi(a, b), c, 0, true); }
inline uint32_t madc_hi_cc(uint32_t a, uint32_t b, uint32_t c) { return emu_add3(mul_hi(a, b), c, cc_flag(), true); }
inline uint32_t madc_hi(uint32_t a, uint32_t b, uint32_t c) { return emu_add3(mul_hi(a, b), c, cc_flag(), false); }
inline uint32_t add_cc(uint32_t a, uint32_t b) { return emu_add3(a, b, 0, true); }
inline uint32_t addc_cc(uint32_t a, uint32_t b) { return emu_add3(a, b, cc_flag(), true); }
inline uint32_t addc(uint32_t a, uint32_t b) { return emu_add3(a, b, cc_flag(), false); }
// borrow is kept in the same flag: 1 = borrow pending
inline uint32_t emu_sub3(uint32_t x, uint32_t y, uint32_t bin, bool set) {
  uint64_t d = (uint64_t)x - y - bin;
  if (set) cc_flag() = (uint32_t)((d >> 32) & 1);
  return (uint32_t)d;
}
inline uint32_t sub_cc(uint32_t a, uint32_t b) { return emu_sub3(a, b, 0, true); }
inline uint32_t subc_cc(uint32_t a, uint32_t b) { return emu_sub3(a, b, cc_flag(), true); }
inline uint32_t subc(uint32_t a, uint32_t b) { return emu_sub3(a, b, cc_flag(), false); }

#endif

}  // namespace ptx
}  // namespace kzgb200
