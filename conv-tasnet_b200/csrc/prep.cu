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
        sg += __bfloat162float(h) + __bfloat162float(l);
        if (beta) sb = fmaf(w, beta[c], sb);
    }
    sg = warp_sum(sg);
    sb = warp_sum(sb);
    if (lane == 0) {
        if (col_g) col_g[r] = sg;
        if (col_bias) col_bias[r] = sb;
    }
}

}  // namespace ctn

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
