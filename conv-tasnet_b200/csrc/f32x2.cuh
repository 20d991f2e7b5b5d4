// Packed fp32x2 arithmetic of sm_100 (fma.rn.f32x2, mul.rn.f32x2, add.rn.f32x2) for the
// HBM-bound marching kernels, whose scalar instruction streams were issue-bound.
#pragma once
#include "common.cuh"

namespace ctn {

#ifndef CTN_DM_SEG
#define CTN_DM_SEG 32
#endif
constexpr int kDmSeg = CTN_DM_SEG;    // frames per (phase, segment) unit of the marching kernels
#ifndef CTN_DM_WARPS
#define CTN_DM_WARPS 8
#endif
constexpr int kDmWarps = CTN_DM_WARPS;   // warps per CTA of the marching kernels

// ---- packed fp32x2 helpers (a 64-bit register pair holds two floats) ----
typedef unsigned long long f2_t;
__device__ __forceinline__ f2_t f2_pack(float lo, float hi) {
    f2_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void f2_unpack(f2_t v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f2_t f2_fma(f2_t a, f2_t b, f2_t c) {
    f2_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f2_t f2_mul(f2_t a, f2_t b) {
    f2_t d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f2_t f2_add(f2_t a, f2_t b) {
    f2_t d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
// (max(v, 0), min(v, 0)) of a pair: mx with two scalar FMNMX, mn = v - mx with one packed FMA
__device__ __forceinline__ void f2_split_sign(f2_t v, f2_t neg1, f2_t& mx, f2_t& mn) {
    float a, b;
    f2_unpack(v, a, b);
    mx = f2_pack(fmaxf(a, 0.f), fmaxf(b, 0.f));
    mn = f2_fma(mx, neg1, v);
}
// (hi, lo) bf16 split of a pair: hi = rn(v), lo = rn(v - hi); v.lo lands in the low half-word
__device__ __forceinline__ void f2_split_bf16(f2_t v, f2_t neg1, uint32_t& hi, uint32_t& lo) {
    float a, b;
    f2_unpack(v, a, b);
#if CTN_OPERAND_FP16
    (void)neg1;
    split_pack2(a, b, hi, lo);
    return;
#endif
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    hi = *reinterpret_cast<const uint32_t*>(&h);
    const f2_t hf = f2_pack(__uint_as_float(hi << 16), __uint_as_float(hi & 0xffff0000u));
    float c, d;
    f2_unpack(f2_fma(hf, neg1, v), c, d);
    const __nv_bfloat162 l = __floats2bfloat162_rn(c, d);
    lo = *reinterpret_cast<const uint32_t*>(&l);
}

}  // namespace ctn
