// Shared host/device helpers for libctn_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/ctn_b200.h"

namespace ctn {

constexpr float kEps = 1e-8f;  // conv_tasnet.py:10, pit_criterion.py:9

// ---------------------------------------------------------------- host side
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);  // records the message, returns CTN_ERR_CUDA

#define CTN_REQUIRE(cond, ...)            \
    do {                                  \
        if (!(cond)) {                    \
            ::ctn::set_error(__VA_ARGS__); \
            return CTN_ERR_INVALID;       \
        }                                 \
    } while (0)

#define CTN_LAUNCH_CHECK(what)                                     \
    do {                                                           \
        cudaError_t e__ = cudaGetLastError();                      \
        if (e__ != cudaSuccess) return ::ctn::cuda_fail(e__, what); \
    } while (0)

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
int pdl_enabled();  // 1 unless the environment says CTN_PDL=0 (read once)
// PDL is used for plain stream launches of SMALL problems only (rows = M * K frames): it hides the
// launch latency and prologue of latency-bound kernels (one 4 s utterance, eager: 2.08 -> 1.81 ms), but
// measured 3-7 % slower from 8 utterances up, and inside a CUDA-graph capture kernel-to-kernel edges
// have no launch latency to hide (programmatic edges: 2-4 % slower).  DESIGN.md 3.10.
constexpr long long kPdlMaxRows = 8192;
inline int pdl_allowed(cudaStream_t stream, long long rows) {
    if (!pdl_enabled() || rows > kPdlMaxRows) return 0;
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream, &st) != cudaSuccess) return 0;
    return st == cudaStreamCaptureStatusNone ? 1 : 0;
}

// Launch with the programmatic-dependent-launch attribute: the kernel may be scheduled while its
// predecessor in the stream drains (it must call ptx::grid_dep_wait() before touching anything the
// predecessor wrote).  Also valid under stream capture (becomes a programmatic graph edge).
template <typename... P, typename... A>
inline cudaError_t launch_pdl(long long rows, void (*kernel)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              A... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid, cfg.blockDim = block, cfg.dynamicSmemBytes = smem, cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl_allowed(stream, rows);
    cfg.attrs = attr, cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<P>(args)...);
}
#define CTN_LAUNCH_PDL(what, ...)                                   \
    do {                                                            \
        cudaError_t e__ = ::ctn::launch_pdl(__VA_ARGS__);           \
        if (e__ != cudaSuccess) return ::ctn::cuda_fail(e__, what); \
    } while (0)
int sm_count();  // multiprocessors of the current device (cached per device)
// Function attributes (dynamic shared-memory limit) are per DEVICE: a process that drives several GPUs
// (torch.nn.DataParallel, train.py:121) must set them on each.  Slot of the current device in a
// per-kernel `static bool done[kMaxDevices]` cache.
constexpr int kMaxDevices = 64;
inline int device_slot() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return 0;
    return dev;
}

// -------------------------------------------------------------- device side
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float prelu(float v, float slope) { return v >= 0.f ? v : slope * v; }

// Products per GEMM term on the tensor cores: 3 = a_hi w_hi + a_lo w_hi + a_hi w_lo (the product; fp32-grade
// results), 1 = a_hi w_hi only.  1 exists ONLY as a speed probe of single-MMA operand formats for the forward
// GEMMs (tools/precision_study.py: fp16 operands would keep 6e-4 rel-L2; with the bf16 operands of this
// build the result is 5e-3 and fails parity) - build with NVCC_FLAGS=-DCTN_MMA_TERMS=1, never shipped.
// CTN_OPERAND_FP16=1 (experiment build, implies one term): the 16-bit containers hold IEEE fp16 instead of bf16 -
// 11 significant bits per operand, 22 for a (hi, lo) pair such as the residual stream; tools/precision_study.py
// predicts 5.6e-4 rel-L2 for the forward.  Every site that forms or decodes a pair is switched below / in place;
// the default build is textually unchanged (#else branches).
#ifndef CTN_OPERAND_FP16
#define CTN_OPERAND_FP16 0
#endif
#ifndef CTN_MMA_TERMS
#if CTN_OPERAND_FP16
#define CTN_MMA_TERMS 1
#else
#define CTN_MMA_TERMS 3
#endif
#endif
#if CTN_OPERAND_FP16
#include <cuda_fp16.h>
static_assert(CTN_MMA_TERMS == 1, "fp16 operands are used with a single MMA per product");
#endif

// fp32 -> (hi, lo) bf16 pair, hi = rn(v), lo = rn(v - hi)
#if CTN_OPERAND_FP16
// value of one 16-bit container / of the two containers packed in a word (element 0 in the low half)
__device__ __forceinline__ float op16_value(__nv_bfloat16 v) {
    return __half2float(__ushort_as_half(__bfloat16_as_ushort(v)));
}
__device__ __forceinline__ void op16_value2(uint32_t w, float& v0, float& v1) {
    const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w));
    v0 = f.x, v1 = f.y;
}
__device__ __forceinline__ void split_bf16(float v, __nv_bfloat16& hi, __nv_bfloat16& lo) {
    const __half h = __float2half_rn(v);
    const __half l = __float2half_rn(v - __half2float(h));
    hi = __ushort_as_bfloat16(__half_as_ushort(h));
    lo = __ushort_as_bfloat16(__half_as_ushort(l));
}
#else
__device__ __forceinline__ void split_bf16(float v, __nv_bfloat16& hi, __nv_bfloat16& lo) {
    hi = __float2bfloat16_rn(v);
    lo = __float2bfloat16_rn(v - __bfloat162float(hi));
}
#endif
__device__ __forceinline__ uint32_t pack_bf16(__nv_bfloat16 a, __nv_bfloat16 b) {
    return static_cast<uint32_t>(__bfloat16_as_ushort(a)) | (static_cast<uint32_t>(__bfloat16_as_ushort(b)) << 16);
}
// split two floats, return packed hi pair and lo pair (v0 in the low half); 6 instructions:
// one packed cvt.rn.bf16x2 per pair, the hi halves are recovered with a shift / mask.
#if CTN_OPERAND_FP16
__device__ __forceinline__ void split_pack2(float v0, float v1, uint32_t& hi, uint32_t& lo) {
    const __half2 h = __floats2half2_rn(v0, v1);
    hi = *reinterpret_cast<const uint32_t*>(&h);
    const float2 hf = __half22float2(h);
    const __half2 l = __floats2half2_rn(v0 - hf.x, v1 - hf.y);
    lo = *reinterpret_cast<const uint32_t*>(&l);
}
#else
__device__ __forceinline__ void split_pack2(float v0, float v1, uint32_t& hi, uint32_t& lo) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(v0, v1);
    hi = *reinterpret_cast<const uint32_t*>(&h);
    const float f0 = __uint_as_float(hi << 16);
    const float f1 = __uint_as_float(hi & 0xffff0000u);
    const __nv_bfloat162 l = __floats2bfloat162_rn(v0 - f0, v1 - f1);
    lo = *reinterpret_cast<const uint32_t*>(&l);
}
#endif

// Whole-warp, fixed-order reduction of P (sum, sumsq) partials into (mean, rstd).
// Every lane returns the same value.  count = number of elements the partials cover.
__device__ __forceinline__ float2 gln_finalize_warp(const float2* __restrict__ part, int P, double count) {
    const int lane = threadIdx.x & 31;
    double s = 0.0, ss = 0.0;
    for (int i = lane; i < P; i += 32) {
        float2 v = __ldg(part + i);
        s += static_cast<double>(v.x);
        ss += static_cast<double>(v.y);
    }
    s = warp_sum(s);
    ss = warp_sum(ss);
    const double mean = s / count;
    double var = ss / count - mean * mean;
    var = var > 0.0 ? var : 0.0;
    return make_float2(static_cast<float>(mean), static_cast<float>(1.0 / sqrt(var + 1e-8)));
}

// Block-wide variant: every thread of the CTA loads its share of the partials at once (one
// round of memory latency instead of P/32), fixed reduction order.  red: >= 2 * (blockDim.x / 32)
// doubles of shared memory.  Returns (mean, rstd) to every thread; contains __syncthreads().
__device__ __forceinline__ float2 gln_finalize_block(const float2* __restrict__ part, int P, double count,
                                                     double* red) {
    double s = 0.0, ss = 0.0;
    for (int i = threadIdx.x; i < P; i += blockDim.x) {
        const float2 v = __ldg(part + i);
        s += static_cast<double>(v.x);
        ss += static_cast<double>(v.y);
    }
    s = warp_sum(s);
    ss = warp_sum(ss);
    const int nw = blockDim.x >> 5, w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) red[w] = s, red[nw + w] = ss;
    __syncthreads();
    s = 0.0, ss = 0.0;
    for (int i = 0; i < nw; ++i) s += red[i], ss += red[nw + i];
    const double mean = s / count;
    double var = ss / count - mean * mean;
    var = var > 0.0 ? var : 0.0;
    return make_float2(static_cast<float>(mean), static_cast<float>(1.0 / sqrt(var + 1e-8)));
}

}  // namespace ctn
