// Encoder: w[m, k, n] = relu(sum_l U[n, l] * mixture[m, k*hop + l]),  hop = L/2
// (conv_tasnet.py:97-117), written frames-major as fp32 and as a bf16 (hi, lo)
// pair, together with the per-frame mean / rstd over channels that the leading
// cLN of the separator needs (conv_tasnet.py:167,326-327).  The framing gather is
// done from a shared-memory copy of the tile's samples, so no [M, K, L] frame
// tensor is materialised.  Memory-bound: reads 4*T, writes 8*N*K bytes per utterance.
#include "common.cuh"
#include "ptx.cuh"

namespace ctn {

constexpr int kEncTile = 32;  // frames per CTA
constexpr int kEncWarps = 8;

__global__ void __launch_bounds__(kEncWarps * 32)
encoder_fwd_kernel(const float* __restrict__ mixture, const float* __restrict__ U, int T, int N, int L, int K,
                   float* __restrict__ w_f32, __nv_bfloat16* __restrict__ w_hi, __nv_bfloat16* __restrict__ w_lo,
                   float2* __restrict__ rowstat) {
    ptx::grid_dep_sync();   // PDL: predecessor complete; let the next kernel be scheduled
    extern __shared__ float enc_smem[];
    const int hop = L / 2;
    float* Ut = enc_smem;       // [L][N]
    float* xs = enc_smem + L * N;  // [(tile-1)*hop + L]
    const int m = blockIdx.y;
    const int k0 = blockIdx.x * kEncTile;
    const int nk = min(kEncTile, K - k0);
    const int nsamp = (nk - 1) * hop + L;
    for (int i = threadIdx.x; i < N * L; i += blockDim.x) {
        const int n = i / L, l = i - n * L;
        Ut[l * N + n] = U[i];
    }
    const float* src = mixture + static_cast<long long>(m) * T + static_cast<long long>(k0) * hop;
    for (int i = threadIdx.x; i < nsamp; i += blockDim.x) xs[i] = src[i];
    __syncthreads();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool pair = (N & 1) == 0;
    for (int r = warp; r < nk; r += kEncWarps) {
        const float* x = xs + r * hop;
        const long long grow = static_cast<long long>(m) * K + k0 + r;
        float s = 0.f, ss = 0.f;
        if (pair) {
            for (int n = 2 * lane; n < N; n += 64) {
                float a0 = 0.f, a1 = 0.f;
                for (int l = 0; l < L; ++l) {
                    const float xv = x[l];
                    const float2 u = *reinterpret_cast<const float2*>(Ut + l * N + n);
                    a0 = fmaf(u.x, xv, a0);
                    a1 = fmaf(u.y, xv, a1);
                }
                a0 = fmaxf(a0, 0.f), a1 = fmaxf(a1, 0.f);
                s += a0 + a1;
                ss = fmaf(a0, a0, fmaf(a1, a1, ss));
                if (w_f32) *reinterpret_cast<float2*>(w_f32 + grow * N + n) = make_float2(a0, a1);
                if (w_hi) {
                    uint32_t h, lo;
                    split_pack2(a0, a1, h, lo);
                    *reinterpret_cast<uint32_t*>(w_hi + grow * N + n) = h;
                    *reinterpret_cast<uint32_t*>(w_lo + grow * N + n) = lo;
                }
            }
        } else {
            for (int n = lane; n < N; n += 32) {
                float a0 = 0.f;
                for (int l = 0; l < L; ++l) a0 = fmaf(Ut[l * N + n], x[l], a0);
                a0 = fmaxf(a0, 0.f);
                s += a0;
                ss = fmaf(a0, a0, ss);
                if (w_f32) w_f32[grow * N + n] = a0;
                if (w_hi) {
                    __nv_bfloat16 h, lo;
                    split_bf16(a0, h, lo);
                    w_hi[grow * N + n] = h;
                    w_lo[grow * N + n] = lo;
                }
            }
        }
        if (rowstat) {
            s = warp_sum(s);
            ss = warp_sum(ss);
            if (lane == 0) {
                const float mean = s / N;
                const float var = fmaxf(ss / N - mean * mean, 0.f);
                rowstat[grow] = make_float2(mean, rsqrtf(var + kEps));
            }
        }
    }
}


// Register-blocked variant for the usual filter lengths: a thread keeps the L taps of its two
// channels in registers and walks over the frames of the tile, reading the samples as shared-memory
// broadcasts (L/2 LDS.64 + 2 L FMAs per frame instead of L LDS + L LDS.64 + 2 L FMAs).  N even,
// N / 2 <= 512 threads.  The per-frame channel sums for the cLN statistics go through shared memory
// in a fixed order.
constexpr int kEncTileR = 64;

template <int L>
__global__ void __launch_bounds__(512)
encoder_fwd_reg_kernel(const float* __restrict__ mixture, const float* __restrict__ U, int T, int N, int K,
                       float* __restrict__ w_f32, __nv_bfloat16* __restrict__ w_hi, __nv_bfloat16* __restrict__ w_lo,
                       float2* __restrict__ rowstat) {
    ptx::grid_dep_sync();   // PDL: predecessor complete; let the next kernel be scheduled
    constexpr int hop = L / 2;
    constexpr int kSamp = (kEncTileR - 1) * hop + L;
    __shared__ __align__(16) float xs[kSamp + 2];
    __shared__ float2 red[kEncTileR][16];   // [frame][warp]
    const int m = blockIdx.y;
    const int k0 = blockIdx.x * kEncTileR;
    const int nk = min(kEncTileR, K - k0);
    const int nsamp = (nk - 1) * hop + L;
    const float* src = mixture + static_cast<long long>(m) * T + static_cast<long long>(k0) * hop;
    for (int i = threadIdx.x; i < kSamp + 2; i += blockDim.x) xs[i] = i < nsamp ? src[i] : 0.f;
    const int n = 2 * threadIdx.x;
    const bool act = n < N;
    float u0[L], u1[L];
#pragma unroll
    for (int l = 0; l < L; ++l) {
        u0[l] = act ? __ldg(U + static_cast<long long>(n) * L + l) : 0.f;
        u1[l] = act ? __ldg(U + static_cast<long long>(n + 1) * L + l) : 0.f;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nwarps = (blockDim.x + 31) >> 5;
    for (int r = 0; r < nk; ++r) {
        const float2* x2 = reinterpret_cast<const float2*>(xs + r * hop);   // hop is even: 8-byte aligned
        float a0 = 0.f, a1 = 0.f;
#pragma unroll
        for (int l = 0; l < L / 2; ++l) {
            const float2 xv = x2[l];
            a0 = fmaf(u0[2 * l], xv.x, a0);
            a1 = fmaf(u1[2 * l], xv.x, a1);
            a0 = fmaf(u0[2 * l + 1], xv.y, a0);
            a1 = fmaf(u1[2 * l + 1], xv.y, a1);
        }
        a0 = fmaxf(a0, 0.f), a1 = fmaxf(a1, 0.f);
        const long long o = (static_cast<long long>(m) * K + k0 + r) * N + n;
        if (act) {
            if (w_f32) *reinterpret_cast<float2*>(w_f32 + o) = make_float2(a0, a1);
            if (w_hi) {
                uint32_t h, lo;
                split_pack2(a0, a1, h, lo);
                *reinterpret_cast<uint32_t*>(w_hi + o) = h;
                *reinterpret_cast<uint32_t*>(w_lo + o) = lo;
            }
        }
        if (rowstat) {
            const float s = warp_sum(a0 + a1);
            const float ss = warp_sum(fmaf(a0, a0, a1 * a1));
            if (lane == 0) red[r][warp] = make_float2(s, ss);
        }
    }
    if (rowstat) {
        __syncthreads();
        for (int r = threadIdx.x; r < nk; r += blockDim.x) {
            float s = 0.f, ss = 0.f;
            for (int w = 0; w < nwarps; ++w) s += red[r][w].x, ss += red[r][w].y;
            const float mean = s / N;
            const float var = fmaxf(ss / N - mean * mean, 0.f);
            rowstat[static_cast<long long>(m) * K + k0 + r] = make_float2(mean, rsqrtf(var + kEps));
        }
    }
}

}  // namespace ctn

extern "C" int ctn_encoder_fwd(const float* mixture, const float* U, int M, int T, int N, int L, float* w_f32,
                               void* w_hi, void* w_lo, void* rowstat, void* stream) {
    using namespace ctn;
    CTN_REQUIRE(mixture && U, "ctn_encoder_fwd: NULL input");
    CTN_REQUIRE(M > 0 && N > 0 && L >= 2, "ctn_encoder_fwd: bad shape M=%d N=%d L=%d", M, N, L);
    CTN_REQUIRE(T >= L, "ctn_encoder_fwd: T=%d shorter than the filter length L=%d", T, L);
    CTN_REQUIRE((w_hi == nullptr) == (w_lo == nullptr), "ctn_encoder_fwd: w_hi/w_lo must be given together");
    const int hop = L / 2;
    const int K = (T - L) / hop + 1;
    auto aligned = [](const void* p, int a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; };
    if ((L == 16 || L == 20) && N % 2 == 0 && N <= 1024 && aligned(w_f32, 8) && aligned(w_hi, 4) && aligned(w_lo, 4)) {
        dim3 grid(ceil_div(K, kEncTileR), M);
        const int threads = ceil_div(N / 2, 32) * 32;
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        if (L == 20)
            CTN_LAUNCH_PDL("encoder_fwd_reg_kernel", static_cast<long long>(M) * K, encoder_fwd_reg_kernel<20>, grid, threads, 0, st, mixture, U, T, N, K, w_f32, static_cast<__nv_bfloat16*>(w_hi), static_cast<__nv_bfloat16*>(w_lo), static_cast<float2*>(rowstat));
        else
            CTN_LAUNCH_PDL("encoder_fwd_reg_kernel", static_cast<long long>(M) * K, encoder_fwd_reg_kernel<16>, grid, threads, 0, st, mixture, U, T, N, K, w_f32, static_cast<__nv_bfloat16*>(w_hi), static_cast<__nv_bfloat16*>(w_lo), static_cast<float2*>(rowstat));
        CTN_LAUNCH_CHECK("encoder_fwd_reg_kernel");
        return CTN_OK;
    }
    const size_t smem = (static_cast<size_t>(L) * N + (kEncTile - 1) * hop + L) * sizeof(float);
    if (smem > 48 * 1024) {
        CTN_REQUIRE(smem <= 200 * 1024, "ctn_encoder_fwd: N*L=%d too large for shared memory", N * L);
        cudaError_t e = cudaFuncSetAttribute(encoder_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(smem));
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(encoder_fwd_kernel)");
    }
    dim3 grid(ceil_div(K, kEncTile), M);
    CTN_LAUNCH_PDL("encoder_fwd_kernel", static_cast<long long>(M) * K, encoder_fwd_kernel, grid, kEncWarps * 32, smem, static_cast<cudaStream_t>(stream), mixture, U, T, N, L, K, w_f32, static_cast<__nv_bfloat16*>(w_hi), static_cast<__nv_bfloat16*>(w_lo), static_cast<float2*>(rowstat));
    CTN_LAUNCH_CHECK("encoder_fwd_kernel");
    return CTN_OK;
}
