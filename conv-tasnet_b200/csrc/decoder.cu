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

// Row-blocked variant for the usual shapes (C <= 3, L <= 24, N % 4 == 0).  The CTA stages the masked
// mixture weights sw[r, c, n] of 32 frames (the first `halo` of them belong to the previous tile) and
// the basis V in shared memory.  Lane r of warp q then computes ALL L outputs of ALL C speakers of frame
// r over the q-th eighth of the n range: per 4 n it loads C float4 of its own row (conflict-free: the
// row stride is 4 banks off a multiple of 32) and LT float4 of V that are the same address for the
// whole warp (one-wavefront broadcasts) for 4*C*LT FMAs -- FMA-issue-bound, where the previous layout
// (thread = frame x 3 outputs, C + 3 four-wavefront loads per 24 FMAs) was bound by the shared-memory
// pipe at 0.20 of the HBM roofline.  The 8 partial sums per output are added in a fixed order, so
// a frame's value does not depend on its position in the tile or the batch.
constexpr int kDecRows = 32;     // frames per CTA including the halo (= lanes of a warp)
constexpr int kDecPad = 4;       // floats of padding per shared-memory row (bank spread)
constexpr int kDecSlices = 8;    // n-range slices = warps per CTA

template <int C, int LT>
__global__ void __launch_bounds__(32 * kDecSlices)
decoder_fwd_blocked_kernel(const float* __restrict__ score, const float* __restrict__ w, const float* __restrict__ V,
                           int K, int N, int Np, int L, int T, int softmax, float* __restrict__ est) {
    // Np = channels staged per pass (divides N, multiple of 4): wide models (N = 512, C = 3) walk the n
    // range in several passes so that the staged rows fit in shared memory; accumulators stay in registers
    ptx::grid_dep_sync();   // PDL: predecessor complete; let the next kernel be scheduled
    extern __shared__ float dec_smem[];
    const int hop = L / 2;
    const int halo = (L - 1) / hop;
    const int tile = kDecRows - halo;          // frames this CTA owns
    const int NS = Np + kDecPad;
    float* Vt = dec_smem;                      // [LT][NS]   (V is [L, N] row-major already; rows >= L are zero)
    float* sw = Vt + LT * NS;                  // [kDecRows][C][NS]; later reused as part[kDecSlices][kDecRows][C][LT]
    const int m = blockIdx.y;
    const int k0 = blockIdx.x * tile;          // first owned frame
    const int nk = min(tile, K - k0);
    const int rows = halo + nk;                // staged frames: k0 - halo .. k0 + nk - 1
    const int N4 = Np >> 2;
    const int r = threadIdx.x & 31, q = threadIdx.x >> 5;
    float acc[C][LT];
#pragma unroll
    for (int c = 0; c < C; ++c)
#pragma unroll
        for (int l = 0; l < LT; ++l) acc[c][l] = 0.f;
    for (int nb = 0; nb < N; nb += Np) {       // one pass per Np channels
    if (nb > 0) __syncthreads();               // the previous pass is done with Vt / sw
    // basis -> shared memory with asynchronous copies, in flight together with the row loads below
    for (int i = threadIdx.x; i < LT * N4; i += blockDim.x) {
        const int l = i / N4, n4 = i - l * N4;
        if (l < L)
            ptx::cp_async_16(Vt + l * NS + 4 * n4, V + l * N + nb + 4 * n4);
        else
            *reinterpret_cast<float4*>(Vt + l * NS + 4 * n4) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    // masked mixture weights -> shared memory, kDecUnroll items per thread and step with all global loads
    // of a step issued before the first use (the one-item loop was a chain of 8 DRAM round trips per CTA)
    constexpr int kDecUnroll = 4;
    for (int base = threadIdx.x; base < kDecRows * N4; base += kDecUnroll * blockDim.x) {
        float4 wv[kDecUnroll], sc[kDecUnroll][C];
        int dst[kDecUnroll];
#pragma unroll
        for (int u = 0; u < kDecUnroll; ++u) {
            const int i = base + u * blockDim.x;
            const int r = i / N4, n4 = i - r * N4;
            const int k = k0 - halo + r;
            dst[u] = i < kDecRows * N4 ? r * C * NS + 4 * n4 : -1;
            const bool live = i < kDecRows * N4 && k >= 0 && r < rows;
            wv[u] = make_float4(0.f, 0.f, 0.f, 0.f);     // w = 0 zeroes the row (frames before the utterance / past the tile)
#pragma unroll
            for (int c = 0; c < C; ++c) sc[u][c] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (live) {
                const long long grow = static_cast<long long>(m) * K + k;
                wv[u] = *reinterpret_cast<const float4*>(w + grow * N + nb + 4 * n4);
#pragma unroll
                for (int c = 0; c < C; ++c)
                    sc[u][c] = *reinterpret_cast<const float4*>(score + grow * (static_cast<long long>(C) * N) + c * N + nb + 4 * n4);
            }
        }
#pragma unroll
        for (int u = 0; u < kDecUnroll; ++u) {
            if (dst[u] < 0) continue;
            float4 o[C];
            if (softmax) {
                float* f[C];
#pragma unroll
                for (int c = 0; c < C; ++c) f[c] = reinterpret_cast<float*>(&sc[u][c]);
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
                    o[c] = make_float4(sc[u][c].x * wv[u].x, sc[u][c].y * wv[u].y, sc[u][c].z * wv[u].z, sc[u][c].w * wv[u].w);
            } else {
#pragma unroll
                for (int c = 0; c < C; ++c)
                    o[c] = make_float4(fmaxf(sc[u][c].x, 0.f) * wv[u].x, fmaxf(sc[u][c].y, 0.f) * wv[u].y,
                                       fmaxf(sc[u][c].z, 0.f) * wv[u].z, fmaxf(sc[u][c].w, 0.f) * wv[u].w);
            }
#pragma unroll
            for (int c = 0; c < C; ++c) *reinterpret_cast<float4*>(sw + dst[u] + c * NS) = o[c];
        }
    }
    ptx::cp_async_wait_all();
    __syncthreads();

    // partial f[q][r, c, l] += sum over the q-th slice of this pass's n of sw[r, c, n] * V[l, n]
    {
        const int n4_lo = (q * N4) / kDecSlices, n4_hi = ((q + 1) * N4) / kDecSlices;
        const float* srow = sw + r * C * NS;
        for (int n4 = n4_lo; n4 < n4_hi; ++n4) {
            float4 sv[C];
#pragma unroll
            for (int c = 0; c < C; ++c) sv[c] = *reinterpret_cast<const float4*>(srow + c * NS + 4 * n4);
#pragma unroll
            for (int l = 0; l < LT; ++l) {
                const float4 vv = *reinterpret_cast<const float4*>(Vt + l * NS + 4 * n4);   // warp-wide broadcast
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    acc[c][l] = fmaf(sv[c].x, vv.x, acc[c][l]);
                    acc[c][l] = fmaf(sv[c].y, vv.y, acc[c][l]);
                    acc[c][l] = fmaf(sv[c].z, vv.z, acc[c][l]);
                    acc[c][l] = fmaf(sv[c].w, vv.w, acc[c][l]);
                }
            }
        }
    }
    }                                          // passes
    __syncthreads();                           // every warp is done with sw: reuse it for the partials
    float* part = sw;                          // [kDecSlices][kDecRows][C * LT + 1]  (+1: bank spread over rows)
    constexpr int PR = C * LT + 1;
#pragma unroll
    for (int c = 0; c < C; ++c)
#pragma unroll
        for (int l = 0; l < LT; ++l) part[(q * kDecRows + r) * PR + c * LT + l] = acc[c][l];
    __syncthreads();

    // samples this tile owns: [k0*hop, (k0+nk)*hop), the last tile runs to T.  Each sample is the sum of
    // the frames covering it (ascending), each frame value the sum of its slices (ascending).
    const bool last = (k0 + nk == K);
    const int t0 = k0 * hop;
    const int t1 = last ? T : (k0 + nk) * hop;
    for (int i = threadIdx.x; i < (t1 - t0) * C; i += blockDim.x) {
        const int c = i / (t1 - t0), t = t0 + (i - c * (t1 - t0));
        int klo = (t - L + hop) / hop;  // ceil((t - L + 1) / hop) for t - L + 1 >= 0
        if (t - L + 1 <= 0) klo = 0;
        const int khi = min(t / hop, K - 1);
        float out = 0.f;
        for (int k = klo; k <= khi; ++k) {
            const float* pp = part + (k - (k0 - halo)) * PR + c * LT + (t - k * hop);
            float f = 0.f;
#pragma unroll
            for (int s = 0; s < kDecSlices; ++s) f += pp[s * kDecRows * PR];
            out += f;
        }
        est[(static_cast<long long>(m) * C + c) * T + t] = out;
    }
}

// shared memory of the row-blocked kernel: V, then the staged rows or (later, same region) the partial sums
static size_t decoder_blocked_smem(int C, int LT, int Np) {
    const size_t rows = static_cast<size_t>(kDecRows) * C * (Np + kDecPad);
    const size_t part = static_cast<size_t>(kDecSlices) * kDecRows * (C * LT + 1);
    return (static_cast<size_t>(LT) * (Np + kDecPad) + (rows > part ? rows : part)) * sizeof(float);
}
// channels per pass: the whole row if it fits in ~100 KB (two CTAs per SM), else the largest power-of-two
// fraction of N (>= 32 channels, multiple of 4) that does
static int decoder_blocked_np(int C, int LT, int N) {
    int Np = N;
    while (decoder_blocked_smem(C, LT, Np) > 100 * 1024 && Np % 8 == 0 && Np / 2 >= 32) Np /= 2;
    return Np;
}

template <int C, int LT>
static int launch_decoder_blocked(const float* score, const float* w, const float* V, int M, int K, int N, int L, int T,
                                  int softmax, float* est, cudaStream_t stream) {
    const int hop = L / 2, halo = (L - 1) / hop;
    const int Np = decoder_blocked_np(C, LT, N);
    const size_t smem = decoder_blocked_smem(C, LT, Np);
    static bool attr_done[kMaxDevices] = {};
    bool& attr_set = attr_done[device_slot()];
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(decoder_fwd_blocked_kernel<C, LT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             200 * 1024);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(decoder_fwd_blocked_kernel)");
        attr_set = true;
    }
    dim3 grid(ceil_div(K, kDecRows - halo), M);
    CTN_LAUNCH_PDL("decoder_fwd_blocked_kernel", static_cast<long long>(M) * K, decoder_fwd_blocked_kernel<C, LT>, grid,
                   32 * kDecSlices, smem, stream, score, w, V, K, N, Np, L, T, softmax, est);
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
        const int LT = L <= 16 ? 16 : (L <= 20 ? 20 : 24);
        const size_t bsmem = decoder_blocked_smem(C, LT, decoder_blocked_np(C, LT, N));
        auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
        if (C <= 3 && L <= 24 && N % 4 == 0 && halo < 8 && bsmem <= 200 * 1024 && al16(score) && al16(w) && al16(V)) {
            cudaStream_t st = static_cast<cudaStream_t>(stream);
            const int sm = mask_nonlinear == CTN_MASK_SOFTMAX;
#define CTN_DEC(CC, LL) return launch_decoder_blocked<CC, LL>(score, w, V, M, K, N, L, T, sm, est, st)
            if (C == 1 && LT == 16) CTN_DEC(1, 16);
            if (C == 1 && LT == 20) CTN_DEC(1, 20);
            if (C == 1 && LT == 24) CTN_DEC(1, 24);
            if (C == 2 && LT == 16) CTN_DEC(2, 16);
            if (C == 2 && LT == 20) CTN_DEC(2, 20);
            if (C == 2 && LT == 24) CTN_DEC(2, 24);
            if (C == 3 && LT == 16) CTN_DEC(3, 16);
            if (C == 3 && LT == 20) CTN_DEC(3, 20);
            if (C == 3 && LT == 24) CTN_DEC(3, 24);
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
