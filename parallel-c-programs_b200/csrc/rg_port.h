/*
 * rg_port.h -- the few primitives the engine's kernel bodies are written against.
 *
 * The kernel bodies in rg_core.h are compiled twice:
 *   - by nvcc for sm_100a (the product: RG_FN = __device__, CUDA intrinsics and atomics);
 *   - by g++ for tests/emu (a sequential, one-thread-at-a-time walk of the SAME bodies
 *     used only by the `not gpu` logic tests; plain loads/stores stand in for atomics).
 * There is no CPU execution path in the product library.
 */
#pragma once
#include <stdint.h>
#include <stddef.h>

#if defined(__CUDACC__)
#define RG_FN __device__ __forceinline__
#define RG_MFN __device__ __forceinline__
#define RG_HDFN __host__ __device__ __forceinline__
#define RG_DEVICE_BUILD 1
#else
#define RG_FN static inline
#define RG_MFN inline
#define RG_HDFN static inline
#define RG_DEVICE_BUILD 0
#include <string.h>
#include <stdio.h>
#include <stdlib.h>
#endif

namespace rg {

struct u32x4 { uint32_t x, y, z, w; };

#if RG_DEVICE_BUILD
RG_FN int popc(uint32_t x) { return __popc(x); }
RG_FN int clz(uint32_t x) { return __clz((int)x); }
RG_FN int ffs1(uint32_t x) { return __ffs((int)x); }          /* 1-based, 0 if x == 0 */
RG_FN uint32_t vabsdiff4(uint32_t a, uint32_t b) { return __vabsdiffu4(a, b); }                 /* one VABSDIFF4 */
RG_FN uint32_t funnel_l(uint32_t lo, uint32_t hi, int n) { return __funnelshift_l(lo, hi, n); }   /* (hi:lo << n) >> 32 */
RG_FN uint32_t absdiff(uint32_t a, uint32_t b) { return __sad((int)a, (int)b, 0u); }
RG_FN uint32_t atomic_or(uint32_t *p, uint32_t v) { return atomicOr(p, v); }
RG_FN uint32_t atomic_add(uint32_t *p, uint32_t v) { return atomicAdd(p, v); }
RG_FN uint32_t atomic_max(uint32_t *p, uint32_t v) { return atomicMax(p, v); }
RG_FN uint32_t cas32(uint32_t *p, uint32_t cmp, uint32_t val) { return atomicCAS(p, cmp, val); }
RG_FN uint16_t cas16(uint16_t *p, uint16_t cmp, uint16_t val)
{
    return atomicCAS((unsigned short *)p, (unsigned short)cmp, (unsigned short)val);
}
RG_FN uint16_t ldv16(const uint16_t *p) { return *(const volatile uint16_t *)p; }
RG_FN uint32_t ldv32(const uint32_t *p) { return *(const volatile uint32_t *)p; }
RG_FN void stv16(uint16_t *p, uint16_t v) { *(volatile uint16_t *)p = v; }
RG_FN void stv32(uint32_t *p, uint32_t v) { *(volatile uint32_t *)p = v; }
RG_FN u32x4 ld128(const void *p) { uint4 v = *(const uint4 *)p; return u32x4{v.x, v.y, v.z, v.w}; }
RG_FN void st128(void *p, u32x4 v) { *(uint4 *)p = make_uint4(v.x, v.y, v.z, v.w); }
/* streaming global accesses: the image is read once per pass and the mask written once */
RG_FN u32x4 ld128_stream(const void *p)
{
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return u32x4{v.x, v.y, v.z, v.w};
}
RG_FN void st128_stream(void *p, u32x4 v)
{
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
/* asynchronous global->shared copies (LDGSTS): many chunks in flight per thread, no register staging.
   `ok` == false zero-fills the destination (src-size 0) */
RG_FN void cp_async16(void *smem_dst, const void *gsrc, bool ok)
{
    uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(d), "l"(gsrc), "r"(ok ? 16 : 0) : "memory");
}
RG_FN void cp_async4(void *smem_dst, const void *gsrc, bool ok)
{
    uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" :: "r"(d), "l"(gsrc), "r"(ok ? 4 : 0) : "memory");
}
RG_FN void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
#else
RG_FN void cp_async16(void *smem_dst, const void *gsrc, bool ok) { if (ok) memcpy(smem_dst, gsrc, 16); else memset(smem_dst, 0, 16); }
RG_FN void cp_async4(void *smem_dst, const void *gsrc, bool ok) { if (ok) memcpy(smem_dst, gsrc, 4); else memset(smem_dst, 0, 4); }
RG_FN void cp_async_wait_all() {}
RG_FN int popc(uint32_t x) { return __builtin_popcount(x); }
RG_FN int clz(uint32_t x) { return x ? __builtin_clz(x) : 32; }
RG_FN int ffs1(uint32_t x) { return __builtin_ffs((int)x); }
RG_FN uint32_t vabsdiff4(uint32_t a, uint32_t b)
{
    uint32_t r = 0;
    for (int j = 0; j < 4; j++) {
        uint32_t x = (a >> (8 * j)) & 0xFF, y = (b >> (8 * j)) & 0xFF;
        r |= (x > y ? x - y : y - x) << (8 * j);
    }
    return r;
}
RG_FN uint32_t funnel_l(uint32_t lo, uint32_t hi, int n) { return (uint32_t)(((((uint64_t)hi << 32) | lo) << n) >> 32); }
RG_FN uint32_t absdiff(uint32_t a, uint32_t b) { return a > b ? a - b : b - a; }
RG_FN uint32_t atomic_or(uint32_t *p, uint32_t v) { uint32_t o = *p; *p = o | v; return o; }
RG_FN uint32_t atomic_add(uint32_t *p, uint32_t v) { uint32_t o = *p; *p = o + v; return o; }
RG_FN uint32_t atomic_max(uint32_t *p, uint32_t v) { uint32_t o = *p; if (v > o) *p = v; return o; }
RG_FN uint32_t cas32(uint32_t *p, uint32_t cmp, uint32_t val) { uint32_t o = *p; if (o == cmp) *p = val; return o; }
RG_FN uint16_t cas16(uint16_t *p, uint16_t cmp, uint16_t val) { uint16_t o = *p; if (o == cmp) *p = val; return o; }
RG_FN uint16_t ldv16(const uint16_t *p) { return *p; }
RG_FN uint32_t ldv32(const uint32_t *p) { return *p; }
RG_FN void stv16(uint16_t *p, uint16_t v) { *p = v; }
RG_FN void stv32(uint32_t *p, uint32_t v) { *p = v; }
/* the device versions are single 16-byte accesses and fault on a misaligned address: the CPU walk checks the same */
RG_FN void check16(const void *p) { if ((uintptr_t)p & 15u) { fprintf(stderr, "rg emulator: misaligned 16-byte access at %p\n", p); abort(); } }
RG_FN u32x4 ld128(const void *p) { check16(p); u32x4 v; memcpy(&v, p, 16); return v; }
RG_FN void st128(void *p, u32x4 v) { check16(p); memcpy(p, &v, 16); }
RG_FN u32x4 ld128_stream(const void *p) { return ld128(p); }
RG_FN void st128_stream(void *p, u32x4 v) { st128(p, v); }
#endif

/* Per-byte |a-b| < t for four voxels at once, as a 4-bit nibble (bit j <-> byte j).
 *
 * The SIMD-in-a-register compare of the toolkit (__vcmpltu4) is emulated with ~8 integer instructions on
 * sm_100a; this form needs 4 after the VABSDIFF4.  With d = |a-b| per byte, H = 0x80 in every byte and
 * t in 1..255:   u = (d | H) - c,  c = t (t <= 128) or t - 128 (t > 128) in every byte.
 * Bit 7 of every byte of d | H is set and c <= 128, so no byte borrows from its neighbour, and bit 7 of
 * u says "the low 7 bits of d are >= c".  For t <= 128:  d >= t  <=>  bit 7 of (u | d)  (a d with its own
 * bit 7 set is >= 128 >= t);  for t > 128:  d >= t  <=>  bit 7 of (u & d).  `sel` = all ones / zero picks
 * the two with one three-input logic operation.  The four flags (bits 7, 15, 23, 31 of `lt`) are gathered
 * into bits 28..31 by one multiplication: 2^21 + 2^14 + 2^7 + 1 moves them there and no two partial
 * products share a bit, so nothing carries. */
struct ByteCmp { uint32_t c, sel; };
RG_FN ByteCmp byte_cmp_for(uint32_t t)                         /* 1 <= t <= 255 */
{
    ByteCmp k; k.c = (t <= 128u ? t : t - 128u) * 0x01010101u; k.sel = t <= 128u ? 0xFFFFFFFFu : 0u;
    return k;
}
RG_FN uint32_t lt4_top_nibble(uint32_t a, uint32_t b, ByteCmp k)   /* result in bits 28..31, garbage below */
{
    const uint32_t H = 0x80808080u;
    const uint32_t d = vabsdiff4(a, b);
    const uint32_t u = (d | H) - k.c;
    const uint32_t ge = (u & d) | ((u | d) & k.sel);
    return (~ge & H) * 0x00204081u;
}

/* 4 mask bits -> 4 bytes of 0/1 */
RG_FN uint32_t expand4(uint32_t nib) { return (nib * 0x00204081u) & 0x01010101u; }

} // namespace rg
