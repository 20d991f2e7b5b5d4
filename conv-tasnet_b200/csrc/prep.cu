// Weight preparation for the 1x1-convolution GEMMs: fold the affine of the norm
// that precedes the convolution into the weights and split them into bf16 (hi, lo).
//   W.(gamma*(y - mu)*rstd + beta) = rstd*(W diag(gamma)).y - rstd*mu*rowsum(W diag(gamma)) + W.beta
// (norm: conv_tasnet.py:326-328 / :352-354; convs: :169, :256).  Runs once per
// weight update; the weights total < 35 MB so its cost is irrelevant.
#include "common.cuh"

namespace ctn {

__global__ void weight_prep_kernel(const float* __restrict__ W, int rows, int cols, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, int transpose, __nv_bfloat16* __restrict__ w_hi,
                                   __nv_bfloat16* __restrict__ w_lo, float* __restrict__ col_g,
                                   float* __restrict__ col_bias) {
    const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (r >= rows) return;
    float sg = 0.f, sb = 0.f;
    for (int c = lane; c < cols; c += 32) {
        const float w = W[static_cast<long long>(r) * cols + c];
        const float wg = gamma ? w * gamma[c] : w;
        __nv_bfloat16 h, l;
        split_bf16(wg, h, l);
        const long long o = transpose ? static_cast<long long>(c) * rows + r : static_cast<long long>(r) * cols + c;
        w_hi[o] = h;
        w_lo[o] = l;
#if CTN_OPERAND_FP16
        sg += op16_value(h);   // the GEMM multiplies by w_hi only: the mean correction must use the same row sums
#elif CTN_MMA_TERMS == 1
        sg += __bfloat162float(h);
#else
        sg += __bfloat162float(h) + __bfloat162float(l);
#endif
        if (beta) sb = fmaf(w, beta[c], sb);
    }
    sg = warp_sum(sg);
    sb = warp_sum(sb);
    if (lane == 0) {
        if (col_g) col_g[r] = sg;
        if (col_bias) col_bias[r] = sb;
    }
}

// One launch for a whole table of weights (blockIdx.y = table entry): a training step refreshes
// 132 prepared weights, which as separate launches cost 1 ms of a 12 ms step.
// A warp owns a row and takes it in chunks of 256 columns, eight loads per lane in flight (the one-load-at-a-time
// loop this replaces was latency-bound: 2 x 95 us of a 6 ms training step); lane l still sums columns l, l + 32, ...
// in that order, so the row sums are bit-identical to weight_prep_kernel's.  Transposed copies leave through shared
// memory: the CTA's eight rows of one column are one 16-byte store instead of eight scattered 2-byte ones.
constexpr int kPrepWarps = 8, kPrepChunk = 256;
__global__ void __launch_bounds__(kPrepWarps * 32) weight_prep_batch_kernel(const ctn_prep_item* __restrict__ items) {
    __shared__ __align__(16) __nv_bfloat16 t_hi[kPrepChunk][kPrepWarps], t_lo[kPrepChunk][kPrepWarps];
    const ctn_prep_item it = items[blockIdx.y];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r0 = blockIdx.x * kPrepWarps, r = r0 + warp;
    if (r0 >= it.rows) return;
    __nv_bfloat16* w_hi = static_cast<__nv_bfloat16*>(it.w_hi);
    __nv_bfloat16* w_lo = static_cast<__nv_bfloat16*>(it.w_lo);
    // staged transposes need the CTA's eight rows to exist and 16-byte aligned destinations
    const bool staged = it.transpose && r0 + kPrepWarps <= it.rows && it.rows % 8 == 0 &&
                        ((reinterpret_cast<uintptr_t>(w_hi) | reinterpret_cast<uintptr_t>(w_lo)) & 15) == 0;
    const bool live = r < it.rows;
    float sg = 0.f, sb = 0.f;
    for (int c0 = 0; c0 < it.cols; c0 += kPrepChunk) {
        float w[kPrepChunk / 32], gm[kPrepChunk / 32], bt[kPrepChunk / 32];
#pragma unroll
        for (int j = 0; j < kPrepChunk / 32; ++j) {
            const int c = c0 + j * 32 + lane;
            const bool ok = live && c < it.cols;
            w[j] = ok ? it.W[static_cast<long long>(r) * it.cols + c] : 0.f;
            gm[j] = ok && it.gamma ? it.gamma[c] : 1.f;
            bt[j] = ok && it.beta ? it.beta[c] : 0.f;
        }
#pragma unroll
        for (int j = 0; j < kPrepChunk / 32; ++j) {
            const int c = c0 + j * 32 + lane;
            if (!(live && c < it.cols)) continue;
            const float wg = it.gamma ? w[j] * gm[j] : w[j];
            __nv_bfloat16 h, l;
            split_bf16(wg, h, l);
            if (staged) {
                t_hi[j * 32 + lane][warp] = h;
                t_lo[j * 32 + lane][warp] = l;
            } else {
                const long long o = it.transpose ? static_cast<long long>(c) * it.rows + r
                                                 : static_cast<long long>(r) * it.cols + c;
                w_hi[o] = h;
                w_lo[o] = l;
            }
#if CTN_OPERAND_FP16
            sg += op16_value(h);   // the GEMM multiplies by w_hi only: the mean correction must use the same row sums
#elif CTN_MMA_TERMS == 1
            sg += __bfloat162float(h);
#else
            sg += __bfloat162float(h) + __bfloat162float(l);
#endif
            if (it.beta) sb = fmaf(w[j], bt[j], sb);
        }
        if (staged) {       // uniform over the CTA
            __syncthreads();
            const int c = c0 + threadIdx.x;
            if (c < it.cols) {
                const long long o = static_cast<long long>(c) * it.rows + r0;
                *reinterpret_cast<uint4*>(w_hi + o) = *reinterpret_cast<const uint4*>(&t_hi[threadIdx.x][0]);
                *reinterpret_cast<uint4*>(w_lo + o) = *reinterpret_cast<const uint4*>(&t_lo[threadIdx.x][0]);
            }
            __syncthreads();
        }
    }
    if (!live) return;
    sg = warp_sum(sg);
    sb = warp_sum(sb);
    if (lane == 0) {
        if (it.col_g) it.col_g[r] = sg;
        if (it.col_bias) it.col_bias[r] = sb;
    }
}

}  // namespace ctn

extern "C" int ctn_weight_prep_batch(const ctn_prep_item* items_dev, int n_items, int max_rows, void* stream) {
    using namespace ctn;
    CTN_REQUIRE(items_dev != nullptr && n_items > 0 && n_items <= 65535 && max_rows > 0,
                "ctn_weight_prep_batch: bad arguments");
    dim3 grid(ceil_div(max_rows, kPrepWarps), n_items);
    weight_prep_batch_kernel<<<grid, kPrepWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(items_dev);
    CTN_LAUNCH_CHECK("weight_prep_batch_kernel");
    return CTN_OK;
}

extern "C" int ctn_weight_prep(const float* W, int rows, int cols, const float* gamma, const float* beta,
                               int transpose, void* w_hi, void* w_lo, float* col_g, float* col_bias, void* stream) {
    using namespace ctn;
    CTN_REQUIRE(W && w_hi && w_lo, "ctn_weight_prep: NULL pointer");
    CTN_REQUIRE(rows > 0 && cols > 0, "ctn_weight_prep: non-positive shape %d x %d", rows, cols);
    const int warps = 8;
    weight_prep_kernel<<<ceil_div(rows, warps), warps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
        W, rows, cols, gamma, beta, transpose, static_cast<__nv_bfloat16*>(w_hi), static_cast<__nv_bfloat16*>(w_lo),
        col_g, col_bias);
    CTN_LAUNCH_CHECK("weight_prep_kernel");
    return CTN_OK;
}

// nn.BatchNorm1d in inference mode (conv_tasnet.py:300-303 -> torch's batch_norm with running
// statistics): y = (x - running_mean) / sqrt(running_var + eps) * weight + bias, a per-channel affine
//   scale[c] = weight[c] / sqrt(running_var[c] + eps),  shift[c] = bias[c] - running_mean[c] * scale[c].
// The engine passes (scale, shift) wherever the gLN kernels take (gamma, beta) together with unit
// statistics (mean 0, rstd 1), so BN blocks run through the same fused kernels.
namespace ctn {
__global__ void bn_fold_kernel(const float* __restrict__ weight, const float* __restrict__ bias,
                               const float* __restrict__ mean, const float* __restrict__ var, float eps, int C,
                               float* __restrict__ scale, float* __restrict__ shift) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float s = (weight ? weight[c] : 1.f) / sqrtf(var[c] + eps);
    scale[c] = s;
    shift[c] = (bias ? bias[c] : 0.f) - mean[c] * s;
}
}  // namespace ctn

extern "C" int ctn_bn_fold(const float* weight, const float* bias, const float* running_mean, const float* running_var,
                           float eps, int C, float* scale, float* shift, void* stream) {
    using namespace ctn;
    CTN_REQUIRE(running_mean && running_var && scale && shift, "ctn_bn_fold: NULL pointer");
    CTN_REQUIRE(C > 0 && eps >= 0.f, "ctn_bn_fold: bad arguments");
    bn_fold_kernel<<<ceil_div(C, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(weight, bias, running_mean,
                                                                                    running_var, eps, C, scale, shift);
    CTN_LAUNCH_CHECK("bn_fold_kernel");
    return CTN_OK;
}

// Table-driven device-to-device copies in ONE launch: the state shift of the streaming separator
// (ctn_b200/_stream.py) moves ~70 small row ranges per chunk, which as individual memcpy nodes cost
// more than the data movement.  Source and destination ranges of one table must not overlap each other.
namespace ctn {
__global__ void __launch_bounds__(256) copy_table_kernel(const ctn_copy_item* __restrict__ items) {
    const ctn_copy_item it = items[blockIdx.y];
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    const uintptr_t both = reinterpret_cast<uintptr_t>(it.dst) | reinterpret_cast<uintptr_t>(it.src) |
                           static_cast<uintptr_t>(it.bytes);
    if ((both & 15) == 0) {
        const uint4* s = static_cast<const uint4*>(it.src);
        uint4* d = static_cast<uint4*>(it.dst);
        for (long long i = t; i < it.bytes / 16; i += stride) d[i] = s[i];
    } else {   // 4-byte granularity (float buffers)
        const uint32_t* s = static_cast<const uint32_t*>(it.src);
        uint32_t* d = static_cast<uint32_t*>(it.dst);
        for (long long i = t; i < it.bytes / 4; i += stride) d[i] = s[i];
    }
}
}  // namespace ctn

extern "C" int ctn_copy_table(const ctn_copy_item* items_dev, int n_items, long long max_bytes, void* stream) {
    using namespace ctn;
    CTN_REQUIRE(items_dev != nullptr && n_items > 0 && n_items <= 65535 && max_bytes > 0, "ctn_copy_table: bad arguments");
    int bx = static_cast<int>((max_bytes / 16 + 255) / 256);
    bx = bx < 1 ? 1 : (bx > 64 ? 64 : bx);
    copy_table_kernel<<<dim3(bx, n_items), 256, 0, static_cast<cudaStream_t>(stream)>>>(items_dev);
    CTN_LAUNCH_CHECK("copy_table_kernel");
    return CTN_OK;
}
