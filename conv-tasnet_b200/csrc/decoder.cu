// Mask non-linearity + Decoder + overlap-add, fused (conv_tasnet.py:202-209 mask
// relu / softmax over speakers, :137 mask * mixture_w, :138-140 Linear(N -> L),
// :141 + utils.py:9-46 overlap_and_add with step L/2, :56-59 right zero-fill to T).
//
// One CTA owns `kDecTile` consecutive frames of one utterance plus the `halo`
// preceding frames whose tails overlap its first samples.  It computes the
// decoded frames f[k, c, l] into shared memory and then writes each output sample
// exactly once as the sum of the (at most ceil(L/hop)) frames covering it, in
// ascending frame order: deterministic and scatter-free, bit-identical to the
// reference's index_add_ for even L (two commutative addends onto zero).
// Memory-bound: reads 4*(C+1)*N bytes per frame, writes 4*C*hop.
#include "common.cuh"

namespace ctn {

constexpr int kDecTile = 32;
constexpr int kDecWarps = 8;
constexpr int kDecMaxC = 8;

__global__ void __launch_bounds__(kDecWarps * 32)
decoder_fwd_kernel(const float* __restrict__ score, const float* __restrict__ w, const float* __restrict__ V, int K,
                   int N, int L, int C, int T, int softmax, float* __restrict__ est) {
    extern __shared__ float dec_smem[];
    const int hop = L / 2;
    const int halo = (L - 1) / hop;            // frames before the tile that reach into it
    const int CL = C * L;
    float* Vt = dec_smem;                      // [N][L]
    float* fr = Vt + N * L;                    // [halo + kDecTile][C][L]
    float* sw = fr + (halo + kDecTile) * CL;   // [kDecWarps][C][N + 1]
    const int m = blockIdx.y;
    const int k0 = blockIdx.x * kDecTile;
    const int nk = min(kDecTile, K - k0);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < N * L; i += blockDim.x) {
        const int l = i / N, n = i - l * N;
        Vt[n * L + l] = V[i];
    }
    __syncthreads();

    float* my_sw = sw + warp * C * (N + 1);
    for (int r = warp; r < halo + nk; r += kDecWarps) {
        const int k = k0 - halo + r;
        float* f = fr + r * CL;
        if (k < 0) {
            for (int o = lane; o < CL; o += 32) f[o] = 0.f;
            continue;
        }
        const long long grow = static_cast<long long>(m) * K + k;
        const float* srow = score + grow * (static_cast<long long>(C) * N);
        const float* wrow = w + grow * N;
        // masked mixture weights for every speaker
        for (int n = lane; n < N; n += 32) {
            const float wv = wrow[n];
            if (softmax) {
                float sc[kDecMaxC];
                float mx = -INFINITY;
#pragma unroll
                for (int c = 0; c < kDecMaxC; ++c)
                    if (c < C) {
                        sc[c] = srow[c * N + n];
                        mx = fmaxf(mx, sc[c]);
                    }
                float den = 0.f;
#pragma unroll
                for (int c = 0; c < kDecMaxC; ++c)
                    if (c < C) {
                        sc[c] = expf(sc[c] - mx);
                        den += sc[c];
                    }
                const float inv = 1.f / den;
#pragma unroll
                for (int c = 0; c < kDecMaxC; ++c)
                    if (c < C) my_sw[c * (N + 1) + n] = sc[c] * inv * wv;
            } else {
                for (int c = 0; c < C; ++c) my_sw[c * (N + 1) + n] = fmaxf(srow[c * N + n], 0.f) * wv;
            }
        }
        __syncwarp();
        // f[c, l] = sum_n sw[c, n] * V[l, n]
        for (int o = lane; o < CL; o += 32) {
            const int c = o / L, l = o - c * L;
            const float* s = my_sw + c * (N + 1);
            float acc = 0.f;
            for (int n = 0; n < N; ++n) acc = fmaf(s[n], Vt[n * L + l], acc);
            f[o] = acc;
        }
        __syncwarp();
    }
    __syncthreads();

    // samples this tile owns: [k0*hop, (k0+nk)*hop), the last tile runs to T
    const bool last = (k0 + nk == K);
    const int t0 = k0 * hop;
    const int t1 = last ? T : (k0 + nk) * hop;
    for (int i = threadIdx.x; i < (t1 - t0) * C; i += blockDim.x) {
        const int c = i / (t1 - t0), t = t0 + (i - c * (t1 - t0));
        int klo = (t - L + hop) / hop;  // ceil((t - L + 1) / hop) for t - L + 1 >= 0
        if (t - L + 1 <= 0) klo = 0;
        const int khi = min(t / hop, K - 1);
        float acc = 0.f;
        for (int k = klo; k <= khi; ++k) acc += fr[(k - (k0 - halo)) * CL + c * L + (t - k * hop)];
        est[(static_cast<long long>(m) * C + c) * T + t] = acc;
    }
}

__global__ void overlap_add_kernel(const float* __restrict__ frames, int K, int L, int step, int T_out,
                                   float* __restrict__ out) {
    const long long b = blockIdx.y;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T_out) return;
    int klo = t - L + 1 <= 0 ? 0 : (t - L + step) / step;
    const int khi = min(t / step, K - 1);
    const float* f = frames + b * K * L;
    float acc = 0.f;
    for (int k = klo; k <= khi; ++k) acc += f[static_cast<long long>(k) * L + (t - k * step)];
    out[b * T_out + t] = acc;
}

}  // namespace ctn

extern "C" int ctn_decoder_fwd(const float* score, const float* w, const float* V, int M, int K, int N, int L, int C,
                               int T, int mask_nonlinear, float* est, void* stream) {
    using namespace ctn;
    CTN_REQUIRE(score && w && V && est, "ctn_decoder_fwd: NULL pointer");
    CTN_REQUIRE(M > 0 && K > 0 && N > 0 && L >= 2 && C > 0, "ctn_decoder_fwd: bad shape");
    CTN_REQUIRE(C <= kDecMaxC, "ctn_decoder_fwd: C=%d above the supported maximum %d", C, kDecMaxC);
    CTN_REQUIRE(mask_nonlinear == CTN_MASK_RELU || mask_nonlinear == CTN_MASK_SOFTMAX,
                "Unsupported mask non-linear function");
    const int hop = L / 2;
    CTN_REQUIRE(T >= (K - 1) * hop + L, "ctn_decoder_fwd: T=%d shorter than the overlap-added length", T);
    const int halo = (L - 1) / hop;
    const size_t smem =
        (static_cast<size_t>(N) * L + (halo + kDecTile) * C * L + kDecWarps * C * (N + 1)) * sizeof(float);
    CTN_REQUIRE(smem <= 200 * 1024, "ctn_decoder_fwd: shape needs %zu bytes of shared memory", smem);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(decoder_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(smem));
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(decoder_fwd_kernel)");
    }
    dim3 grid(ceil_div(K, kDecTile), M);
    decoder_fwd_kernel<<<grid, kDecWarps * 32, smem, static_cast<cudaStream_t>(stream)>>>(
        score, w, V, K, N, L, C, T, mask_nonlinear == CTN_MASK_SOFTMAX, est);
    CTN_LAUNCH_CHECK("decoder_fwd_kernel");
    return CTN_OK;
}

extern "C" int ctn_overlap_add(const float* frames, int batch, int K, int L, int step, float* out, void* stream) {
    using namespace ctn;
    CTN_REQUIRE(frames && out, "ctn_overlap_add: NULL pointer");
    CTN_REQUIRE(batch > 0 && K > 0 && L > 0 && step > 0 && step <= L, "ctn_overlap_add: bad shape");
    const int T_out = (K - 1) * step + L;
    dim3 grid(ceil_div(T_out, 256), batch);
    overlap_add_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(frames, K, L, step, T_out, out);
    CTN_LAUNCH_CHECK("overlap_add_kernel");
    return CTN_OK;
}
