// Middle of a TemporalBlock, one pass over the [M*K, H] activation:
//   a --PReLU(a1)--> --norm1 (gLN | cLN) on load--> zero-padded dilated depthwise
//   cross-correlation --PReLU(a2)--> q, plus the statistics the second norm needs.
// Reference: conv_tasnet.py:218-219 (PReLU, norm), :247-252 (depthwise Conv1d,
// groups=H, dilation 2^x), :275-289 (Chomp1d for the causal case), :253 (PReLU).
// Padding semantics: the reference pads the NORMALISED tensor with zeros, so an
// out-of-range tap contributes exactly 0 (not norm(0)).
//   non-causal: d[k] = sum_j wd[j] * n1[k + j*dil - (P-1)*dil/2]
//   causal:     d[k] = sum_j wd[j] * n1[k + (j-(P-1))*dil]          (pad + chomp)
// Layout is frames-major, so one warp owns one frame and its lanes sweep the H
// channels with 16-byte accesses; HBM-bound: reads 4*H, writes 4*H bytes per frame
// (taps at k +- dil re-hit L2).  Inference writes q only as the bf16 (hi, lo) pair the
// pointwise GEMM consumes; training additionally keeps d in fp32 for the backward.
#include "common.cuh"
#include "ptx.cuh"
#include "dwconv_march.cuh"

namespace ctn {

constexpr int kDwWarps = 8;
constexpr int kDwTile = CTN_DW_TILE;  // frames per CTA
constexpr int kDwMaxP = 8;            // largest depthwise kernel size

struct DwParams {
    int K, H, P, dil, causal, norm, in_P;
    const float* in;
    const float* prelu_in;
    const float* gamma_in;
    const float* beta_in;
    const float2* in_stat;
    const float* wd;
    const float* prelu_out;
    float* out_pre;
    __nv_bfloat16* out_hi;
    __nv_bfloat16* out_lo;
    float2* out_stat;
    int out_P;
};

template <int VEC>
struct Vec;
template <>
struct Vec<4> {
    float v[4];
    __device__ __forceinline__ void load(const float* p) {
        float4 t = *reinterpret_cast<const float4*>(p);
        v[0] = t.x, v[1] = t.y, v[2] = t.z, v[3] = t.w;
    }
    __device__ __forceinline__ void store(float* p) const {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    }
};
template <>
struct Vec<1> {
    float v[1];
    __device__ __forceinline__ void load(const float* p) { v[0] = *p; }
    __device__ __forceinline__ void store(float* p) const { *p = v[0]; }
};

template <int VEC>
__device__ __forceinline__ void store_split(__nv_bfloat16* hi, __nv_bfloat16* lo, const float (&q)[VEC]) {
    if (VEC == 4) {
        uint2 h, l;
        split_pack2(q[0], q[1], h.x, l.x);
        split_pack2(q[2 % VEC], q[3 % VEC], h.y, l.y);
        *reinterpret_cast<uint2*>(hi) = h;
        *reinterpret_cast<uint2*>(lo) = l;
    } else {
        __nv_bfloat16 h, l;
        split_bf16(q[0], h, l);
        *hi = h;
        *lo = l;
    }
}

template <int VEC>
__global__ void __launch_bounds__(kDwWarps * 32)
dwconv_fwd_kernel(const DwParams p) {
    ptx::grid_dep_sync();   // PDL: predecessor complete; let the next kernel be scheduled
    extern __shared__ float dw_smem[];
    const int H = p.H, P = p.P;
    float* wd_s = dw_smem;        // [P][H]  tap-major
    float* ga_s = wd_s + P * H;   // gLN: gamma*rstd ; cLN: gamma
    float* be_s = ga_s + H;       // gLN: beta - gamma*rstd*mean ; cLN: beta
    float* red_s = be_s + H;      // [2 * kDwWarps] + 2
    const int m = blockIdx.y;
    const int k0 = blockIdx.x * kDwTile;
    const int nk = min(kDwTile, p.K - k0);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool gln = p.norm == CTN_NORM_GLN;

    if (gln && warp == 0) {
        const float2 st = gln_finalize_warp(p.in_stat + static_cast<long long>(m) * p.in_P, p.in_P,
                                            static_cast<double>(p.K) * H);
        if (lane == 0) red_s[2 * kDwWarps] = st.x, red_s[2 * kDwWarps + 1] = st.y;
    }
    __syncthreads();
    {
        const float mean = gln ? red_s[2 * kDwWarps] : 0.f;
        const float rstd = gln ? red_s[2 * kDwWarps + 1] : 1.f;
        for (int i = threadIdx.x; i < H * P; i += blockDim.x) {
            const int c = i / P, j = i - c * P;
            wd_s[j * H + c] = p.wd[i];
        }
        for (int c = threadIdx.x; c < H; c += blockDim.x) {
            const float g = p.gamma_in[c] * rstd;
            ga_s[c] = g;
            be_s[c] = p.beta_in[c] - g * mean;
        }
    }
    __syncthreads();

    const float a1 = __ldg(p.prelu_in), a2 = __ldg(p.prelu_out);
    const int pad = p.causal ? (P - 1) * p.dil : ((P - 1) * p.dil) / 2;
    const long long ubase = static_cast<long long>(m) * p.K;
    float ls = 0.f, lss = 0.f;  // gLN partial of q over this thread's elements

    for (int r = warp; r < nk; r += kDwWarps) {
        const int k = k0 + r;
        const long long grow = ubase + k;
        float rs = 0.f, rss = 0.f;  // cLN row sums of q
        float tmu[kDwMaxP], trstd[kDwMaxP];
        if (!gln) {
            // per-frame statistics of prelu(a) at each tap, from the GEMM's column-group partials
#pragma unroll
            for (int j = 0; j < kDwMaxP; ++j) {
                tmu[j] = 0.f, trstd[j] = 1.f;
                const int kk = k + j * p.dil - pad;
                if (j < P && kk >= 0 && kk < p.K) {
                    float s = 0.f, ss = 0.f;
                    const float2* rp = p.in_stat + (ubase + kk) * p.in_P;
                    for (int g = 0; g < p.in_P; ++g) {
                        const float2 t = __ldg(rp + g);
                        s += t.x, ss += t.y;
                    }
                    tmu[j] = s / H;
                    trstd[j] = rsqrtf(fmaxf(ss / H - tmu[j] * tmu[j], 0.f) + kEps);
                }
            }
        }
        for (int c = lane * VEC; c < H; c += 32 * VEC) {
            float acc[VEC];
#pragma unroll
            for (int i = 0; i < VEC; ++i) acc[i] = 0.f;
#pragma unroll
            for (int j = 0; j < kDwMaxP; ++j) {
                const int kk = k + j * p.dil - pad;
                if (j >= P || kk < 0 || kk >= p.K) continue;  // zero padding of the normalised input
                Vec<VEC> x, g, b, w;
                x.load(p.in + (ubase + kk) * H + c);
                g.load(ga_s + c);
                b.load(be_s + c);
                w.load(wd_s + j * H + c);
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    float v = prelu(x.v[i], a1);
                    v = gln ? fmaf(g.v[i], v, b.v[i]) : fmaf(g.v[i], (v - tmu[j]) * trstd[j], b.v[i]);
                    acc[i] = fmaf(w.v[i], v, acc[i]);
                }
            }
            if (p.out_pre) {
                Vec<VEC> o;
#pragma unroll
                for (int i = 0; i < VEC; ++i) o.v[i] = acc[i];
                o.store(p.out_pre + grow * H + c);
            }
            float q[VEC];
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                q[i] = prelu(acc[i], a2);
                rs += q[i];
                rss = fmaf(q[i], q[i], rss);
            }
            store_split<VEC>(p.out_hi + grow * H + c, p.out_lo + grow * H + c, q);
        }
        if (gln) {
            ls += rs, lss += rss;
        } else {
            rs = warp_sum(rs);
            rss = warp_sum(rss);
            if (lane == 0) {
                const float mean = rs / H;
                p.out_stat[grow] = make_float2(mean, rsqrtf(fmaxf(rss / H - mean * mean, 0.f) + kEps));
            }
        }
    }
    if (gln) {
        ls = warp_sum(ls);
        lss = warp_sum(lss);
        if (lane == 0) red_s[warp] = ls, red_s[kDwWarps + warp] = lss;
        __syncthreads();
        if (threadIdx.x == 0) {
            float s = 0.f, ss = 0.f;
            for (int w = 0; w < kDwWarps; ++w) s += red_s[w], ss += red_s[kDwWarps + w];
            p.out_stat[static_cast<long long>(m) * p.out_P + blockIdx.x] = make_float2(s, ss);
        }
    }
}

}  // namespace ctn

static bool dw_use_march(int H, int P, int norm) { return norm == CTN_NORM_GLN && P == 3 && H % 4 == 0; }

extern "C" int ctn_dwconv_gln_partials(int K, int H, int P, int dilation) {
    if (dw_use_march(H, P, CTN_NORM_GLN) && dilation > 0) {
        int NS, items;
        ctn::dw_march_geometry(K, H, dilation, &NS, &items);
        return ctn::ceil_div(items, ctn::kDmWarps);
    }
    return ctn::ceil_div(K, CTN_DW_TILE);
}

extern "C" int ctn_dwconv_fwd(const ctn_dwconv_args* args, void* stream) {
    using namespace ctn;
    CTN_REQUIRE(args != nullptr, "ctn_dwconv_fwd: args is NULL");
    const ctn_dwconv_args& a = *args;
    CTN_REQUIRE(a.M > 0 && a.K > 0 && a.H > 0 && a.P > 0 && a.dilation > 0, "ctn_dwconv_fwd: bad shape");
    CTN_REQUIRE(a.norm == CTN_NORM_GLN || a.norm == CTN_NORM_CLN, "ctn_dwconv_fwd: norm must be gLN or cLN");
    CTN_REQUIRE(a.in && a.prelu_in && a.gamma_in && a.beta_in && a.in_stat && a.wd && a.prelu_out && a.out_hi &&
                    a.out_lo && a.out_stat,
                "ctn_dwconv_fwd: NULL pointer");
    CTN_REQUIRE(a.in_P > 0, "ctn_dwconv_fwd: in_P must be positive");
    CTN_REQUIRE(a.P <= kDwMaxP, "ctn_dwconv_fwd: kernel size P=%d above the supported maximum %d", a.P, kDwMaxP);
    // Non-causal output length equals K only when (P-1)*dilation is even; the reference
    // crashes at the residual add otherwise (conv_tasnet.py:236 TODO).
    CTN_REQUIRE(a.causal || ((a.P - 1) * a.dilation) % 2 == 0,
                "ctn_dwconv_fwd: non-causal depthwise with odd (P-1)*dilation changes the length (P=%d, dilation=%d)",
                a.P, a.dilation);
    DwParams p;
    p.K = a.K, p.H = a.H, p.P = a.P, p.dil = a.dilation, p.causal = a.causal, p.norm = a.norm, p.in_P = a.in_P;
    p.in = a.in, p.prelu_in = a.prelu_in, p.gamma_in = a.gamma_in, p.beta_in = a.beta_in;
    p.in_stat = static_cast<const float2*>(a.in_stat);
    p.wd = a.wd, p.prelu_out = a.prelu_out, p.out_pre = a.out_pre;
    p.out_hi = static_cast<__nv_bfloat16*>(a.out_hi), p.out_lo = static_cast<__nv_bfloat16*>(a.out_lo);
    p.out_stat = static_cast<float2*>(a.out_stat);
    p.out_P = ctn_dwconv_gln_partials(a.K, a.H, a.P, a.dilation);
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    if (dw_use_march(a.H, a.P, a.norm) && al16(a.in) && al16(a.out_pre) && al16(a.out_hi) && al16(a.out_lo) &&
        al16(a.gamma_in) && al16(a.beta_in) && al16(a.wd)) {
        DwMarchParams q;
        q.K = a.K, q.H = a.H, q.dil = a.dilation, q.c0 = a.causal ? 2 : 1, q.in_P = a.in_P;
        dw_march_geometry(a.K, a.H, a.dilation, &q.NS, &q.items);
        q.out_P = p.out_P;
        q.in = a.in, q.prelu_in = a.prelu_in, q.gamma_in = a.gamma_in, q.beta_in = a.beta_in;
        q.in_stat = p.in_stat, q.wd = a.wd, q.prelu_out = a.prelu_out, q.out_pre = a.out_pre;
        q.out_hi = p.out_hi, q.out_lo = p.out_lo, q.out_stat = p.out_stat;
        dim3 mgrid(q.out_P, a.M);
        CTN_LAUNCH_PDL("dwconv_march_kernel", static_cast<long long>(a.M) * a.K, dwconv_march_kernel, mgrid, kDmWarps * 32, 0, static_cast<cudaStream_t>(stream), q);
        CTN_LAUNCH_CHECK("dwconv_march_kernel");
        return CTN_OK;
    }
    CTN_REQUIRE(!dw_use_march(a.H, a.P, a.norm), "ctn_dwconv_fwd: pointers must be 16-byte aligned");
    const size_t smem = (static_cast<size_t>(a.P + 2) * a.H + 2 * kDwWarps + 2) * sizeof(float);
    CTN_REQUIRE(smem <= 48 * 1024, "ctn_dwconv_fwd: H*(P+2)=%d too large for shared memory", a.H * (a.P + 2));
    dim3 grid(ceil_div(a.K, kDwTile), a.M);
    const bool vec = a.H % 4 == 0 && al16(a.in) && al16(a.out_pre) && al16(a.out_hi) && al16(a.out_lo);
    if (vec)
        CTN_LAUNCH_PDL("dwconv_fwd_kernel", static_cast<long long>(a.M) * a.K, dwconv_fwd_kernel<4>, grid, kDwWarps * 32, smem, static_cast<cudaStream_t>(stream), p);
    else
        CTN_LAUNCH_PDL("dwconv_fwd_kernel", static_cast<long long>(a.M) * a.K, dwconv_fwd_kernel<1>, grid, kDwWarps * 32, smem, static_cast<cudaStream_t>(stream), p);
    CTN_LAUNCH_CHECK("dwconv_fwd_kernel");
    return CTN_OK;
}
