// Tail of the training step (solver.py:182-186): torch.nn.utils.clip_grad_norm_(params, max_norm)
// followed by torch.optim.Adam.step(), as TWO launches over all parameter tensors of the model
// instead of a few dozen multi-tensor launches over 294 tensors.
//
//   kernel 1  sum of squares of every gradient chunk -> one double partial per chunk
//   kernel 2  every CTA re-reduces the partials in a fixed order (bit-reproducible, no atomics,
//             no extra "finalise" launch) -> total norm -> clip coefficient
//                 coef = min(1, max_norm / (norm + 1e-6))          (clip_grad_norm_)
//             then Adam with L2 weight decay on its chunk          (torch.optim.Adam, amsgrad=False)
//                 g = coef * grad + wd * p
//                 m = b1 m + (1 - b1) g ;  v = b2 v + (1 - b2) g^2
//                 p -= step_size * m / (sqrt(v) / sqrt(1 - b2^t) + eps),   step_size = lr / (1 - b1^t)
// HBM-bound: reads p, g, m, v and writes p, m, v once (28 B per parameter + 4 B for the norm).
#include "common.cuh"

namespace ctn {

constexpr int kOptThreads = 256;

__device__ __forceinline__ double block_sum_double(double v, double* red) {
    v = warp_sum(v);
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) red[w] = v;
    __syncthreads();
    double s = 0.0;
    for (int i = 0; i < kOptThreads / 32; ++i) s += red[i];
    __syncthreads();
    return s;
}

__global__ void __launch_bounds__(kOptThreads)
grad_sqnorm_kernel(const ctn_opt_chunk* __restrict__ chunks, double* __restrict__ partials) {
    __shared__ double red[kOptThreads / 32];
    const ctn_opt_chunk c = chunks[blockIdx.x];
    float s = 0.f;
    for (int i = threadIdx.x; i < c.n; i += kOptThreads) {
        const float g = c.grad[i];
        s = fmaf(g, g, s);
    }
    const double t = block_sum_double(static_cast<double>(s), red);
    if (threadIdx.x == 0) partials[blockIdx.x] = t;
}

__global__ void __launch_bounds__(kOptThreads)
clip_adam_kernel(const ctn_opt_chunk* __restrict__ chunks, const double* __restrict__ partials, int n_chunks,
                 float max_norm, float step_size, float beta1, float beta2, float omb1, float omb2, float inv_sqrt_bias2,
                 float eps, float weight_decay, int write_grad, float* __restrict__ norm_out) {
    __shared__ double red[kOptThreads / 32];
    double t = 0.0;
    for (int i = threadIdx.x; i < n_chunks; i += kOptThreads) t += partials[i];
    const float norm = static_cast<float>(sqrt(block_sum_double(t, red)));
    float coef = 1.f;
    if (max_norm > 0.f) coef = fminf(max_norm / (norm + 1e-6f), 1.f);
    if (blockIdx.x == 0 && threadIdx.x == 0 && norm_out != nullptr) *norm_out = norm;
    const ctn_opt_chunk c = chunks[blockIdx.x];
    for (int i = threadIdx.x; i < c.n; i += kOptThreads) {
        const float p = c.param[i];
        const float gc = coef * c.grad[i];
        const float g = fmaf(weight_decay, p, gc);
        const float m = fmaf(beta1, c.exp_avg[i], omb1 * g);
        const float v = fmaf(beta2, c.exp_avg_sq[i], omb2 * g * g);
        const float denom = fmaf(sqrtf(v), inv_sqrt_bias2, eps);
        c.exp_avg[i] = m;
        c.exp_avg_sq[i] = v;
        c.param[i] = p - step_size * (m / denom);
        if (write_grad) c.grad[i] = gc;
    }
}

}  // namespace ctn

using namespace ctn;

extern "C" int ctn_clip_adam(const ctn_opt_chunk* chunks_dev, int n_chunks, double* partials, float max_norm,
                             float step_size, double beta1, double beta2, float inv_sqrt_bias2, float eps,
                             float weight_decay, int write_grad, float* norm_out, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    CTN_REQUIRE(chunks_dev != nullptr && partials != nullptr && n_chunks > 0, "ctn_clip_adam: NULL table or no chunks");
    CTN_REQUIRE(beta1 >= 0.0 && beta1 < 1.0 && beta2 >= 0.0 && beta2 < 1.0 && eps >= 0.f && inv_sqrt_bias2 > 0.f,
                "ctn_clip_adam: bad hyper-parameters");
    grad_sqnorm_kernel<<<n_chunks, kOptThreads, 0, stream>>>(chunks_dev, partials);
    CTN_LAUNCH_CHECK("grad_sqnorm_kernel");
    clip_adam_kernel<<<n_chunks, kOptThreads, 0, stream>>>(chunks_dev, partials, n_chunks, max_norm, step_size,
                                                           static_cast<float>(beta1), static_cast<float>(beta2),
                                                           static_cast<float>(1.0 - beta1), static_cast<float>(1.0 - beta2),
                                                           inv_sqrt_bias2, eps, weight_decay, write_grad, norm_out);
    CTN_LAUNCH_CHECK("clip_adam_kernel");
    return CTN_OK;
}
