// nn.BatchNorm1d in TRAINING mode (conv_tasnet.py:300-303 builds it for every norm_type but gLN / cLN):
// per-channel statistics over every frame of every utterance of the batch, running-statistics update,
// and the per-channel terms of its backward.  The block kernels (depthwise, norm_split, 1x1 GEMMs, marching
// backward) are the gLN ones: they take the per-channel affine
//     scale[c] = weight[c] / sqrt(var[c] + eps),   shift[c] = bias[c] - mean[c] * scale[c]
// in place of (gamma, beta) with unit statistics, exactly as in inference (prep.cu: bn_fold_kernel); what
// training adds is (i) where (mean, var) come from and (ii) the two batch-mean terms of the backward,
//     dx = scale * (dy - mean_rows(dy) - xh * mean_rows(dy * xh)),   xh = (x - mean) * rstd,
// which bn_bwd_center_kernel subtracts from dy in place BEFORE the gLN kernels see it (they then apply
// `scale` with unit statistics and zero correction sums).  d weight = sum dy * xh and d bias = sum dy are the
// same two column sums.  Sums travel as per-CTA float partials reduced in fp64 in a fixed order
// (deterministic); across GPUs the [2, C] fp64 sums are all-reduced between the reduce and the fold /
// centre kernels (ctn_b200/_autograd.py), which makes a sharded batch equal to the single-process one.
#include "common.cuh"

namespace ctn {

constexpr int kBnRows = 64;      // rows per CTA of the column-sum kernel
constexpr int kBnThreads = 256;

// part[tile][0][c] = sum_r f0, part[tile][1][c] = sum_r f1 over the tile's rows:
//   MODE 0 (forward):  p = prelu(pre);                 f0 = p,  f1 = p * p
//   MODE 1 (backward): xh = (prelu(pre) - mean) * rstd; f0 = dy, f1 = dy * xh
template <int MODE>
__global__ void __launch_bounds__(kBnThreads) bn_colsums_kernel(const float* __restrict__ dy, const float* __restrict__ pre,
                                                                const float* __restrict__ slope_p,
                                                                const float* __restrict__ mean,
                                                                const float* __restrict__ rstd, long long R, int C,
                                                                float* __restrict__ part) {
    const long long r0 = static_cast<long long>(blockIdx.x) * kBnRows;
    const int rows = static_cast<int>(min(static_cast<long long>(kBnRows), R - r0));
    const float slope = slope_p ? __ldg(slope_p) : 1.f;
    for (int c = threadIdx.x; c < C; c += kBnThreads) {
        float s0 = 0.f, s1 = 0.f, mu = 0.f, rs = 1.f;
        if (MODE == 1) mu = __ldg(mean + c), rs = __ldg(rstd + c);
        const float* p = pre + r0 * C + c;
        const float* g = MODE == 1 ? dy + r0 * C + c : nullptr;
#pragma unroll 8
        for (int r = 0; r < rows; ++r) {
            const float v = prelu(__ldg(p + static_cast<long long>(r) * C), slope);
            if (MODE == 0) {
                s0 += v;
                s1 = fmaf(v, v, s1);
            } else {
                const float d = __ldg(g + static_cast<long long>(r) * C);
                s0 += d;
                s1 = fmaf(d, (v - mu) * rs, s1);
            }
        }
        float* o = part + static_cast<long long>(blockIdx.x) * 2 * C;
        o[c] = s0;
        o[C + c] = s1;
    }
}

// sums[j][c] = sum over tiles of part[tile][j][c], fp64, fixed order.  One CTA per 32 columns of the [2C] row:
// 8 warps walk interleaved tiles, then the 8 partials are added in warp order.
__global__ void __launch_bounds__(256) bn_reduce_kernel(const float* __restrict__ part, int n_tiles, int C2, double scale,
                                                        double* __restrict__ sums) {
    __shared__ double red[8][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int col = blockIdx.x * 32 + lane;
    double s = 0.0;
    if (col < C2)
        for (int t = w; t < n_tiles; t += 8) s += static_cast<double>(__ldg(part + static_cast<long long>(t) * C2 + col));
    red[w][lane] = s;
    __syncthreads();
    if (w == 0 && col < C2) {
        double tot = 0.0;
#pragma unroll
        for (int i = 0; i < 8; ++i) tot += red[i][lane];
        sums[col] = tot * scale;
    }
}

// (mean, biased var) from the batch sums (or the running statistics when use_running), the folded affine, and the
// buffer update of nn.BatchNorm1d(momentum): running_mean <- (1 - m) running_mean + m mean,
// running_var <- (1 - m) running_var + m var * n / (n - 1), num_batches_tracked += 1.
__global__ void bn_train_fold_kernel(const double* __restrict__ sums, double count, const float* __restrict__ weight,
                                     const float* __restrict__ bias, float eps, float momentum, int use_running,
                                     float* __restrict__ running_mean, float* __restrict__ running_var,
                                     long long* __restrict__ num_batches_tracked, int C, float* __restrict__ scale,
                                     float* __restrict__ shift, float* __restrict__ mean_out,
                                     float* __restrict__ rstd_out) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c == 0 && !use_running && num_batches_tracked) *num_batches_tracked += 1;
    if (c >= C) return;
    double mean, var;
    if (use_running) {
        mean = running_mean[c];
        var = running_var[c];
    } else {
        mean = sums[c] / count;
        var = sums[C + c] / count - mean * mean;
        var = var > 0.0 ? var : 0.0;
        const double unbiased = var * (count / (count > 1.0 ? count - 1.0 : 1.0));
        running_mean[c] = static_cast<float>((1.0 - momentum) * running_mean[c] + momentum * mean);
        running_var[c] = static_cast<float>((1.0 - momentum) * running_var[c] + momentum * unbiased);
    }
    const double rstd = 1.0 / sqrt(var + static_cast<double>(eps));
    const double s = (weight ? weight[c] : 1.f) * rstd;
    scale[c] = static_cast<float>(s);
    shift[c] = static_cast<float>((bias ? bias[c] : 0.f) - mean * s);
    mean_out[c] = static_cast<float>(mean);
    rstd_out[c] = static_cast<float>(rstd);
}

// dy[r, c] -= sums[0][c] / count + xh[r, c] * sums[1][c] / count (center != 0), in place;
// d weight[c] = sums[1][c], d bias[c] = sums[0][c] (written by the first CTA).
__global__ void __launch_bounds__(kBnThreads) bn_bwd_center_kernel(float* __restrict__ dy, const float* __restrict__ pre,
                                                                   const float* __restrict__ slope_p,
                                                                   const float* __restrict__ mean,
                                                                   const float* __restrict__ rstd,
                                                                   const double* __restrict__ sums, double count,
                                                                   int center, long long R, int C,
                                                                   float* __restrict__ dweight, float* __restrict__ dbias) {
    const long long r0 = static_cast<long long>(blockIdx.x) * kBnRows;
    const int rows = static_cast<int>(min(static_cast<long long>(kBnRows), R - r0));
    const float slope = slope_p ? __ldg(slope_p) : 1.f;
    for (int c = threadIdx.x; c < C; c += kBnThreads) {
        const double S0 = sums[c], S1 = sums[C + c];
        if (blockIdx.x == 0) {
            if (dweight) dweight[c] = static_cast<float>(S1);
            if (dbias) dbias[c] = static_cast<float>(S0);
        }
        if (!center) continue;
        const float m0 = static_cast<float>(S0 / count), m1 = static_cast<float>(S1 / count);
        const float mu = __ldg(mean + c), rs = __ldg(rstd + c);
#pragma unroll 8
        for (int r = 0; r < rows; ++r) {
            const long long i = (r0 + r) * C + c;
            const float xh = (prelu(__ldg(pre + i), slope) - mu) * rs;
            dy[i] = dy[i] - m0 - xh * m1;
        }
    }
}
}  // namespace ctn

extern "C" int ctn_bn_tiles(long long R) { return static_cast<int>((R + ctn::kBnRows - 1) / ctn::kBnRows); }

extern "C" int ctn_bn_colsums(const float* dy, const float* pre, const float* slope, const float* mean,
                              const float* rstd, long long R, int C, double scale, float* part, double* sums,
                              void* stream) {
    using namespace ctn;
    CTN_REQUIRE(pre && part && sums, "ctn_bn_colsums: NULL pointer");
    CTN_REQUIRE(R > 0 && C > 0, "ctn_bn_colsums: bad shape");
    CTN_REQUIRE(dy == nullptr || (mean && rstd), "ctn_bn_colsums: the backward sums need (mean, rstd)");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int tiles = ctn_bn_tiles(R);
    if (dy)
        bn_colsums_kernel<1><<<tiles, kBnThreads, 0, st>>>(dy, pre, slope, mean, rstd, R, C, part);
    else
        bn_colsums_kernel<0><<<tiles, kBnThreads, 0, st>>>(nullptr, pre, slope, nullptr, nullptr, R, C, part);
    CTN_LAUNCH_CHECK("bn_colsums_kernel");
    bn_reduce_kernel<<<ceil_div(2 * C, 32), 256, 0, st>>>(part, tiles, 2 * C, scale, sums);
    CTN_LAUNCH_CHECK("bn_reduce_kernel");
    return CTN_OK;
}

extern "C" int ctn_bn_train_fold(const double* sums, double count, const float* weight, const float* bias, float eps,
                                 float momentum, int use_running, float* running_mean, float* running_var,
                                 long long* num_batches_tracked, int C, float* scale, float* shift, float* mean,
                                 float* rstd, void* stream) {
    using namespace ctn;
    CTN_REQUIRE(running_mean && running_var && scale && shift && mean && rstd, "ctn_bn_train_fold: NULL pointer");
    CTN_REQUIRE(use_running || sums, "ctn_bn_train_fold: batch statistics need the column sums");
    CTN_REQUIRE(C > 0 && eps >= 0.f && (use_running || count >= 1.0), "ctn_bn_train_fold: bad arguments");
    bn_train_fold_kernel<<<ceil_div(C, 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(
        sums, count, weight, bias, eps, momentum, use_running, running_mean, running_var, num_batches_tracked, C, scale,
        shift, mean, rstd);
    CTN_LAUNCH_CHECK("bn_train_fold_kernel");
    return CTN_OK;
}

extern "C" int ctn_bn_bwd_center(float* dy, const float* pre, const float* slope, const float* mean, const float* rstd,
                                 const double* sums, double count, int center, long long R, int C, float* dweight,
                                 float* dbias, void* stream) {
    using namespace ctn;
    CTN_REQUIRE(dy && pre && mean && rstd && sums, "ctn_bn_bwd_center: NULL pointer");
    CTN_REQUIRE(R > 0 && C > 0 && count >= 1.0, "ctn_bn_bwd_center: bad arguments");
    const int tiles = center ? ctn_bn_tiles(R) : 1;
    bn_bwd_center_kernel<<<tiles, kBnThreads, 0, static_cast<cudaStream_t>(stream)>>>(dy, pre, slope, mean, rstd, sums,
                                                                                      count, center, R, C, dweight, dbias);
    CTN_LAUNCH_CHECK("bn_bwd_center_kernel");
    return CTN_OK;
}
