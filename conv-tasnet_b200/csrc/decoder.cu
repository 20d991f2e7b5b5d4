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
#include "ptx.cuh"

namespace ctn {

constexpr int kDecTile = 32;
constexpr int kDecWarps = 8;
constexpr int kDecMaxC = 8;

__global__ void __launch_bounds__(kDecWarps * 32)
decoder_fwd_kernel(const float* __restrict__ score, const float* __restrict__ w, const float* __restrict__ V, int K,
                   int N, int L, int C, int T, int softmax, float* __restrict__ est) {
    ptx::grid_dep_sync();   // PDL: predecessor complete; let the next kernel be scheduled
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

// Register-blocked variant for the usual shapes (C <= 3, L <= 24, N % 4 == 0).  The CTA stages the
// masked mixture weights sw[r, c, n] of 32 frames (the first `halo` of them belong to the previous
// tile) in shared memory, then thread (r = tid / 8, g = tid % 8) computes the outputs l = g, g + 8,
// g + 16 of all C speakers of frame r with float4 reads along n: C + LPT shared loads (all
// broadcasts or conflict-free) per 4 * C * LPT FMAs.  The generic kernel above does one output per
// lane with scalar loads and 8 of 32 lanes idle in its second pass.
constexpr int kDecRows = 32;     // frames per CTA including the halo
constexpr int kDecPad = 4;       // floats of padding per shared-memory row (bank spread)

template <int C, int LPT>
__global__ void __launch_bounds__(256)
decoder_fwd_blocked_kernel(const float* __restrict__ score, const float* __restrict__ w, const float* __restrict__ V,
                           int K, int N, int L, int T, int softmax, float* __restrict__ est) {
    ptx::grid_dep_sync();   // PDL: predecessor complete; let the next kernel be scheduled
    extern __shared__ float dec_smem[];
    const int hop = L / 2;
    const int halo = (L - 1) / hop;
    const int tile = kDecRows - halo;          // frames this CTA owns
    const int NS = N + kDecPad;
    const int CL = C * L;
    float* Vt = dec_smem;                      // [L][NS]   (V is [L, N] row-major already)
    float* sw = Vt + L * NS;                   // [kDecRows][C][NS]
    float* fr = sw + kDecRows * C * NS;        // [kDecRows][C][L]
    const int m = blockIdx.y;
    const int k0 = blockIdx.x * tile;          // first owned frame
    const int nk = min(tile, K - k0);
    const int rows = halo + nk;                // staged frames: k0 - halo .. k0 + nk - 1
    const int N4 = N >> 2;
    for (int i = threadIdx.x; i < L * N4; i += blockDim.x) {
        const int l = i / N4, n4 = i - l * N4;
        *reinterpret_cast<float4*>(Vt + l * NS + 4 * n4) = *reinterpret_cast<const float4*>(V + l * N + 4 * n4);
    }
    for (int i = threadIdx.x; i < rows * N4; i += blockDim.x) {
        const int r = i / N4, n4 = i - r * N4;
        const int k = k0 - halo + r;
        float4 o[C];
        if (k < 0) {
#pragma unroll
            for (int c = 0; c < C; ++c) o[c] = make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
            const long long grow = static_cast<long long>(m) * K + k;
            const float4 wv = *reinterpret_cast<const float4*>(w + grow * N + 4 * n4);
            float4 sc[C];
#pragma unroll
            for (int c = 0; c < C; ++c)
                sc[c] = *reinterpret_cast<const float4*>(score + grow * (static_cast<long long>(C) * N) + c * N + 4 * n4);
            if (softmax) {
                float* f[C];
#pragma unroll
                for (int c = 0; c < C; ++c) f[c] = reinterpret_cast<float*>(&sc[c]);
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    float mx = -INFINITY;
#pragma unroll
                    for (int c = 0; c < C; ++c) mx = fmaxf(mx, f[c][e]);
                    float den = 0.f;
#pragma unroll
                    for (int c = 0; c < C; ++c) f[c][e] = expf(f[c][e] - mx), den += f[c][e];
                    const float inv = 1.f / den;
#pragma unroll
                    for (int c = 0; c < C; ++c) f[c][e] *= inv;
                }
#pragma unroll
                for (int c = 0; c < C; ++c)
                    o[c] = make_float4(sc[c].x * wv.x, sc[c].y * wv.y, sc[c].z * wv.z, sc[c].w * wv.w);
            } else {
#pragma unroll
                for (int c = 0; c < C; ++c)
                    o[c] = make_float4(fmaxf(sc[c].x, 0.f) * wv.x, fmaxf(sc[c].y, 0.f) * wv.y,
                                       fmaxf(sc[c].z, 0.f) * wv.z, fmaxf(sc[c].w, 0.f) * wv.w);
            }
        }
#pragma unroll
        for (int c = 0; c < C; ++c) *reinterpret_cast<float4*>(sw + (r * C + c) * NS + 4 * n4) = o[c];
    }
    __syncthreads();

    {   // f[r, c, l] = sum_n sw[r, c, n] * V[l, n]
        const int r = threadIdx.x >> 3, g = threadIdx.x & 7;
        if (r < rows) {
            float acc[C][LPT];
#pragma unroll
            for (int c = 0; c < C; ++c)
#pragma unroll
                for (int j = 0; j < LPT; ++j) acc[c][j] = 0.f;
            const float* vrow[LPT];
#pragma unroll
            for (int j = 0; j < LPT; ++j) vrow[j] = Vt + min(g + 8 * j, L - 1) * NS;
            const float* srow = sw + r * C * NS;
            for (int n = 0; n < N; n += 4) {
                float4 sv[C], vv[LPT];
#pragma unroll
                for (int c = 0; c < C; ++c) sv[c] = *reinterpret_cast<const float4*>(srow + c * NS + n);
#pragma unroll
                for (int j = 0; j < LPT; ++j) vv[j] = *reinterpret_cast<const float4*>(vrow[j] + n);
#pragma unroll
                for (int c = 0; c < C; ++c)
#pragma unroll
                    for (int j = 0; j < LPT; ++j) {
                        acc[c][j] = fmaf(sv[c].x, vv[j].x, acc[c][j]);
                        acc[c][j] = fmaf(sv[c].y, vv[j].y, acc[c][j]);
                        acc[c][j] = fmaf(sv[c].z, vv[j].z, acc[c][j]);
                        acc[c][j] = fmaf(sv[c].w, vv[j].w, acc[c][j]);
                    }
            }
#pragma unroll
            for (int c = 0; c < C; ++c)
#pragma unroll
                for (int j = 0; j < LPT; ++j)
                    if (g + 8 * j < L) fr[r * CL + c * L + g + 8 * j] = acc[c][j];
        }
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

template <int C, int LPT>
static int launch_decoder_blocked(const float* score, const float* w, const float* V, int M, int K, int N, int L, int T,
                                  int softmax, float* est, cudaStream_t stream) {
    const int hop = L / 2, halo = (L - 1) / hop;
    const size_t smem = (static_cast<size_t>(L) * (N + kDecPad) + static_cast<size_t>(kDecRows) * C * (N + kDecPad) +
                         static_cast<size_t>(kDecRows) * C * L) * sizeof(float);
    cudaError_t e = cudaFuncSetAttribute(decoder_fwd_blocked_kernel<C, LPT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(decoder_fwd_blocked_kernel)");
    dim3 grid(ceil_div(K, kDecRows - halo), M);
    CTN_LAUNCH_PDL("decoder_fwd_blocked_kernel", static_cast<long long>(M) * K, decoder_fwd_blocked_kernel<C, LPT>, grid, 256, smem, stream, score, w, V, K, N, L, T, softmax, est);
    CTN_LAUNCH_CHECK("decoder_fwd_blocked_kernel");
    return CTN_OK;
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
    {
        const size_t bsmem = (static_cast<size_t>(L) * (N + kDecPad) + static_cast<size_t>(kDecRows) * C * (N + kDecPad) +
                              static_cast<size_t>(kDecRows) * C * L) * sizeof(float);
        auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
        if (C <= 3 && L <= 24 && N % 4 == 0 && halo < 8 && bsmem <= 200 * 1024 && al16(score) && al16(w) && al16(V)) {
            cudaStream_t st = static_cast<cudaStream_t>(stream);
            const int sm = mask_nonlinear == CTN_MASK_SOFTMAX;
            const int lpt = ceil_div(L, 8);
#define CTN_DEC(CC, LL) return launch_decoder_blocked<CC, LL>(score, w, V, M, K, N, L, T, sm, est, st)
            if (C == 1 && lpt == 1) CTN_DEC(1, 1);
            if (C == 1 && lpt == 2) CTN_DEC(1, 2);
            if (C == 1 && lpt == 3) CTN_DEC(1, 3);
            if (C == 2 && lpt == 1) CTN_DEC(2, 1);
            if (C == 2 && lpt == 2) CTN_DEC(2, 2);
            if (C == 2 && lpt == 3) CTN_DEC(2, 3);
            if (C == 3 && lpt == 1) CTN_DEC(3, 1);
            if (C == 3 && lpt == 2) CTN_DEC(3, 2);
            if (C == 3 && lpt == 3) CTN_DEC(3, 3);
#undef CTN_DEC
        }
    }
    const size_t smem =
        (static_cast<size_t>(N) * L + (halo + kDecTile) * C * L + kDecWarps * C * (N + 1)) * sizeof(float);
    CTN_REQUIRE(smem <= 200 * 1024, "ctn_decoder_fwd: shape needs %zu bytes of shared memory", smem);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(decoder_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(smem));
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(decoder_fwd_kernel)");
    }
    dim3 grid(ceil_div(K, kDecTile), M);
    CTN_LAUNCH_PDL("decoder_fwd_kernel", static_cast<long long>(M) * K, decoder_fwd_kernel, grid, kDecWarps * 32, smem, static_cast<cudaStream_t>(stream), score, w, V, K, N, L, C, T, mask_nonlinear == CTN_MASK_SOFTMAX, est);
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
