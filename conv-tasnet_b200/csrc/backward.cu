// Backward kernels of the separator that are not GEMMs (the reference gets all of these
// from autograd over conv_tasnet.py).  Notation per TemporalBlock
// (conv_tasnet.py:212-272), frames-major activations [M*K, H]:
//   a  = conv1x1(x)            p  = prelu(a, a1)      ph = (p - mu1) * r1    n1 = g1*ph + b1
//   d  = depthwise(n1)         q  = prelu(d, a2)      qh = (q - mu2) * r2    u  = g2*qh + b2
//   out = pointwise(u) + x
// gLN backward for y = g*xh + b, xh = (x - mu)*r over the n = H*K elements of an utterance:
//   dg = sum dy*xh,  db = sum dy,  dx = r * (g*dy - mean(g*dy) - xh * mean(g*dy*xh)).
// Every reduction is a fixed-order tree (thread -> warp -> CTA partial -> ctn_reduce_partials),
// so gradients are bit-reproducible run to run.
#include "common.cuh"
#include "bwd_dw_march.cuh"

namespace ctn {

constexpr int kBwThreads = 256;

template <int VEC>
struct V {
    float v[VEC];
    __device__ __forceinline__ void load(const float* p) {
        if (VEC == 4) {
            const float4 t = *reinterpret_cast<const float4*>(p);
            v[0] = t.x, v[1 % VEC] = t.y, v[2 % VEC] = t.z, v[3 % VEC] = t.w;
        } else {
#pragma unroll
            for (int i = 0; i < VEC; ++i) v[i] = p[i];
        }
    }
    __device__ __forceinline__ void store(float* p) const {
        if (VEC == 4) {
            *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1 % VEC], v[2 % VEC], v[3 % VEC]);
        } else {
#pragma unroll
            for (int i = 0; i < VEC; ++i) p[i] = v[i];
        }
    }
};

template <int VEC>
__device__ __forceinline__ void store_hl(__nv_bfloat16* hi, __nv_bfloat16* lo, const float* q) {
    if (VEC == 4) {
        uint2 h, l;
        split_pack2(q[0], q[1 % VEC], h.x, l.x);
        split_pack2(q[2 % VEC], q[3 % VEC], h.y, l.y);
        *reinterpret_cast<uint2*>(hi) = h;
        *reinterpret_cast<uint2*>(lo) = l;
    } else {
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            __nv_bfloat16 h, l;
            split_bf16(q[i], h, l);
            hi[i] = h, lo[i] = l;
        }
    }
}

// block-wide sum of up to NV values per thread; result valid in thread 0.  red: [NV * 8] floats.
template <int NV>
__device__ __forceinline__ void block_sum(float (&v)[NV], float* red) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        v[i] = warp_sum(v[i]);
        if (lane == 0) red[i * 8 + warp] = v[i];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            float s = 0.f;
            for (int w = 0; w < kBwThreads / 32; ++w) s += red[i * 8 + w];
            v[i] = s;
        }
    }
}

// ---------------------------------------------------------------------------------------
// norm_split: out = split_bf16( gamma * (act(in) - mu) * rstd + beta )
// Materialises a normalised activation as the (hi, lo) operand of a weight-gradient GEMM.
// ---------------------------------------------------------------------------------------
struct NormSplitParams {
    int K, C, norm, P;
    double count;
    const float* in;
    const float* slope;
    const float* gamma;
    const float* beta;
    const float2* stat;
    __nv_bfloat16* out_hi;
    __nv_bfloat16* out_lo;
};
// Frames per CTA of the row-wise elementwise kernels below (norm_split, bwd_prelu_norm).  They are latency-bound
// at training batch sizes (20 MB in, 20 MB out per launch): A/B on one box, two alternating rounds, train step
// at batch 3 / batch 8: 32 frames 7.84 / 11.43 ms, 16 frames 7.56 / 11.37 ms, 8 frames 7.54 / 11.53 ms.
#ifndef CTN_NS_TILE
#define CTN_NS_TILE 16
#endif
constexpr int kNsTile = CTN_NS_TILE;
constexpr int kNsPre = 8;   // vectors per thread requested up front: a 16-frame tile of 512 channels in float4s

template <int VEC>
__global__ void __launch_bounds__(kBwThreads) norm_split_kernel(const NormSplitParams p) {
    __shared__ float st_s[2];
    const int m = blockIdx.y, k0 = blockIdx.x * kNsTile;
    const int nk = min(kNsTile, p.K - k0);
    const bool gln = p.norm == CTN_NORM_GLN;
    const int CV = (p.C + VEC - 1) / VEC;
    const long long row0 = static_cast<long long>(m) * p.K + k0;
    // the tile's first kNsPre vectors per thread are requested before the statistics are finalised: the kernel
    // is one round trip to memory long, and the partial sums' round trip would otherwise come in front of it
    V<VEC> xs[kNsPre];
#pragma unroll
    for (int j = 0; j < kNsPre; ++j) {
        const int i = threadIdx.x + j * kBwThreads;
        if (i < nk * CV) {
            const int r = i / CV, c = (i - r * CV) * VEC;
            xs[j].load(p.in + (row0 + r) * p.C + c);
        }
    }
    if (gln && threadIdx.x < 32) {
        const float2 st = gln_finalize_warp(p.stat + static_cast<long long>(m) * p.P, p.P, p.count);
        if (threadIdx.x == 0) st_s[0] = st.x, st_s[1] = st.y;
    }
    __syncthreads();
    const float slope = p.slope ? __ldg(p.slope) : 1.f;
    auto emit = [&](int i, const V<VEC>& x) {
        const int r = i / CV, c = (i - r * CV) * VEC;
        const long long row = row0 + r;
        float mu = st_s[0], rs = st_s[1];
        if (!gln) {
            const float2 st = __ldg(p.stat + row);
            mu = st.x, rs = st.y;
        }
        V<VEC> g, b;
        g.load(p.gamma + c);
        b.load(p.beta + c);
        float o[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) o[j] = fmaf(g.v[j], (prelu(x.v[j], slope) - mu) * rs, b.v[j]);
        store_hl<VEC>(p.out_hi + row * p.C + c, p.out_lo + row * p.C + c, o);
    };
#pragma unroll
    for (int j = 0; j < kNsPre; ++j) {
        const int i = threadIdx.x + j * kBwThreads;
        if (i < nk * CV) emit(i, xs[j]);
    }
    for (int i = threadIdx.x + kNsPre * kBwThreads; i < nk * CV; i += kBwThreads) {
        const int r = i / CV, c = (i - r * CV) * VEC;
        V<VEC> x;
        x.load(p.in + (row0 + r) * p.C + c);
        emit(i, x);
    }
}

// ---------------------------------------------------------------------------------------
// bwd_dw: everything between the pointwise dgrad and the first norm's input gradient.
// In : du = d out / d u (from the pointwise dgrad GEMM, whose epilogue also produced the
//      (S1, S2) = (sum g2*du, sum g2*du*qh) partials), saved a and d, forward statistics.
// Out: dn1 = d/d n1 [M*K, H]; per-channel partial sums (dg2, db2, dwd[0..P), dg1, db1);
//      per-CTA scalars (S1' = sum g1*dn1, S2' = sum g1*dn1*ph, d a2).
// One thread owns VEC channels and walks the frames of its tile, so the per-channel sums
// live in registers.  Partial row layout: [2 + P + 2][H] floats per (CTA, row-group).
// ---------------------------------------------------------------------------------------
constexpr int kBdTile = 128;  // frames per CTA
constexpr int kBdMaxP = 8;

struct BwdDwParams {
    int K, H, P, dil, causal, P1, P2, Pb, cln;   // cln: per-frame statistics (P1 = Pb = column groups, stat2 = (mean, rstd))
    const float* du;
    const float* d;
    const float* a;
    const float2* stat1;   // forward partials of prelu(a)   [M, P1]
    const float2* stat2;   // forward partials of prelu(d)   [M, P2]
    const float2* bstat;   // backward partials (S1, S2)     [M, Pb]
    const float* g1;
    const float* b1;
    const float* g2;
    const float* wd;       // [H, P]
    const float* slope1;
    const float* slope2;
    float* dn1;
    float* chan_part;      // [M * tiles * RG][(4 + P) * H]
    float* scal_part;      // [M * tiles][4]  (S1', S2', d a2, 0)
    int RG;
};

template <int VEC>
__global__ void __launch_bounds__(kBwThreads) bwd_dw_kernel(const BwdDwParams p) {
    __shared__ float sh[8];
    __shared__ float red[3 * 8];
    const int H = p.H, P = p.P;
    const int m = blockIdx.y, tile = blockIdx.x, tiles = gridDim.x;
    const int k0 = tile * kBdTile;
    const int nk = min(kBdTile, p.K - k0);
    const double n = static_cast<double>(p.K) * H;
    if (threadIdx.x < 32 && !p.cln) {
        const float2 s1 = gln_finalize_warp(p.stat1 + static_cast<long long>(m) * p.P1, p.P1, n);
        const float2 s2 = gln_finalize_warp(p.stat2 + static_cast<long long>(m) * p.P2, p.P2, n);
        // (S1, S2) / n
        const int lane = threadIdx.x;
        double a = 0.0, b = 0.0;
        for (int i = lane; i < p.Pb; i += 32) {
            const float2 t = __ldg(p.bstat + static_cast<long long>(m) * p.Pb + i);
            a += t.x, b += t.y;
        }
        a = warp_sum(a), b = warp_sum(b);
        if (lane == 0) {
            sh[0] = s1.x, sh[1] = s1.y, sh[2] = s2.x, sh[3] = s2.y;
            sh[4] = static_cast<float>(a / n), sh[5] = static_cast<float>(b / n);
        }
    }
    __syncthreads();
    const float al1 = __ldg(p.slope1), al2 = __ldg(p.slope2);
    const bool cln = p.cln != 0;
    // statistics of one frame: gLN -> the utterance's scalars, cLN -> finalised from that frame's partials
    auto stats1 = [&](long long row, float& mu1, float& r1) {
        mu1 = sh[0], r1 = sh[1];
        if (cln) {
            float s = 0.f, ss = 0.f;
            for (int g = 0; g < p.P1; ++g) {
                const float2 t = __ldg(p.stat1 + row * p.P1 + g);
                s += t.x, ss += t.y;
            }
            mu1 = s / H;
            r1 = rsqrtf(fmaxf(ss / H - mu1 * mu1, 0.f) + kEps);
        }
    };
    auto stats2 = [&](long long row, float& mu2, float& r2, float& m1, float& m2) {
        mu2 = sh[2], r2 = sh[3], m1 = sh[4], m2 = sh[5];
        if (cln) {
            const float2 t = __ldg(p.stat2 + row);
            mu2 = t.x, r2 = t.y;
            float a = 0.f, b = 0.f;
            for (int g = 0; g < p.Pb; ++g) {
                const float2 u = __ldg(p.bstat + row * p.Pb + g);
                a += u.x, b += u.y;
            }
            m1 = a / H, m2 = b / H;
        }
    };
    const int pad = p.causal ? (P - 1) * p.dil : ((P - 1) * p.dil) / 2;
    const long long ubase = static_cast<long long>(m) * p.K;
    const int CV = (H + VEC - 1) / VEC;
    const int CVt = min(CV, kBwThreads);
    const int RG = p.RG;
    const int rg = threadIdx.x / CVt, cvt = threadIdx.x - rg * CVt;
    const bool active = rg < RG;
    float sc[3] = {0.f, 0.f, 0.f};  // S1', S2', d a2

    // dd at (row kk, channels c..): gradient wrt the depthwise output
    auto dd_at = [&](long long row, int c, const float* g2v, float* out, float* dq_out, float* dval) {
        V<VEC> du, d;
        du.load(p.du + row * H + c);
        d.load(p.d + row * H + c);
        float mu2, r2, m1, m2;
        stats2(row, mu2, r2, m1, m2);
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            const float qh = (prelu(d.v[i], al2) - mu2) * r2;
            const float dq = r2 * (g2v[i] * du.v[i] - m1 - qh * m2);
            out[i] = d.v[i] > 0.f ? dq : al2 * dq;
            if (dq_out) dq_out[i] = dq, dval[i] = d.v[i];
        }
    };

    if (active) {
        for (int cv = cvt; cv < CV; cv += CVt) {
            const int c = cv * VEC;
            V<VEC> g1v, b1v, g2v, wv[kBdMaxP];
            g1v.load(p.g1 + c), b1v.load(p.b1 + c), g2v.load(p.g2 + c);
#pragma unroll
            for (int j = 0; j < kBdMaxP; ++j)
                if (j < P) {
#pragma unroll
                    for (int i = 0; i < VEC; ++i) wv[j].v[i] = (c + i < H) ? p.wd[(c + i) * P + j] : 0.f;
                }
            float acc_dg2[VEC], acc_db2[VEC], acc_dg1[VEC], acc_db1[VEC], acc_dw[kBdMaxP][VEC];
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                acc_dg2[i] = acc_db2[i] = acc_dg1[i] = acc_db1[i] = 0.f;
#pragma unroll
                for (int j = 0; j < kBdMaxP; ++j) acc_dw[j][i] = 0.f;
            }
            for (int r = rg; r < nk; r += RG) {
                const int k = k0 + r;
                const long long row = ubase + k;
                // centre: second-norm parameter grads, PReLU-2 slope grad, dd[k]
                float ddc[VEC], dqc[VEC], dvc[VEC];
                dd_at(row, c, g2v.v, ddc, dqc, dvc);
                {
                    V<VEC> du;
                    du.load(p.du + row * H + c);
                    float mu2, r2, m1u, m2u;
                    stats2(row, mu2, r2, m1u, m2u);
#pragma unroll
                    for (int i = 0; i < VEC; ++i) {
                        const float qh = (prelu(dvc[i], al2) - mu2) * r2;
                        acc_dg2[i] = fmaf(du.v[i], qh, acc_dg2[i]);
                        acc_db2[i] += du.v[i];
                        if (!(dvc[i] > 0.f)) sc[2] = fmaf(dqc[i], dvc[i], sc[2]);
                    }
                }
                float dn[VEC];
#pragma unroll
                for (int i = 0; i < VEC; ++i) dn[i] = 0.f;
#pragma unroll
                for (int j = 0; j < kBdMaxP; ++j) {
                    if (j >= P) continue;
                    const int off = j * p.dil - pad;
                    // depthwise weight grad: dwd[j] += dd[k] * n1[k + off]
                    const int kf = k + off;
                    if (kf >= 0 && kf < p.K) {
                        V<VEC> av;
                        av.load(p.a + (ubase + kf) * H + c);
                        float mu1, r1;
                        stats1(ubase + kf, mu1, r1);
#pragma unroll
                        for (int i = 0; i < VEC; ++i) {
                            const float n1 = fmaf(g1v.v[i], (prelu(av.v[i], al1) - mu1) * r1, b1v.v[i]);
                            acc_dw[j][i] = fmaf(ddc[i], n1, acc_dw[j][i]);
                        }
                    }
                    // input grad: dn1[k] += wd[j] * dd[k - off]
                    const int kb = k - off;
                    if (kb >= 0 && kb < p.K) {
                        float ddb[VEC];
                        if (off == 0) {
#pragma unroll
                            for (int i = 0; i < VEC; ++i) ddb[i] = ddc[i];
                        } else {
                            dd_at(ubase + kb, c, g2v.v, ddb, nullptr, nullptr);
                        }
#pragma unroll
                        for (int i = 0; i < VEC; ++i) dn[i] = fmaf(wv[j].v[i], ddb[i], dn[i]);
                    }
                }
                V<VEC> ac, o;
                ac.load(p.a + row * H + c);
                float mu1, r1;
                stats1(row, mu1, r1);
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    const float ph = (prelu(ac.v[i], al1) - mu1) * r1;
                    acc_dg1[i] = fmaf(dn[i], ph, acc_dg1[i]);
                    acc_db1[i] += dn[i];
                    const float gd = g1v.v[i] * dn[i];
                    sc[0] += gd;
                    sc[1] = fmaf(gd, ph, sc[1]);
                    o.v[i] = dn[i];
                }
                o.store(p.dn1 + row * H + c);
            }
            float* part = p.chan_part + ((static_cast<long long>(m) * tiles + tile) * RG + rg) * (4 + P) * H;
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                if (c + i < H) {
                    part[0 * H + c + i] = acc_dg2[i];
                    part[1 * H + c + i] = acc_db2[i];
                    part[2 * H + c + i] = acc_dg1[i];
                    part[3 * H + c + i] = acc_db1[i];
#pragma unroll
                    for (int j = 0; j < kBdMaxP; ++j)
                        if (j < P) part[(4 + j) * H + c + i] = acc_dw[j][i];
                }
            }
        }
    }
    block_sum<3>(sc, red);
    if (threadIdx.x == 0) {
        float* s = p.scal_part + (static_cast<long long>(m) * tiles + tile) * 4;
        s[0] = sc[0], s[1] = sc[1], s[2] = sc[2], s[3] = 0.f;
    }
}

// ---------------------------------------------------------------------------------------
// bwd_prelu_norm: da = prelu'(a) * r1 * (g1*dn1 - S1'/n - ph*S2'/n)  -> (hi, lo); d a1 partial.
// ---------------------------------------------------------------------------------------
struct BwdPnParams {
    int K, H, P1, tiles_b;
    const float* dn1;
    const float* a;
    const float2* stat1;
    const float* scal_part;  // [M, tiles_b, 4]
    const float* g1;
    const float* slope1;
    __nv_bfloat16* da_hi;
    __nv_bfloat16* da_lo;
    float* alpha_part;       // [M * gridDim.x]
};

constexpr int kPnPre = 2;     // vector pairs (dn1, a) per thread requested ahead of the statistics: what 51 registers hold
// five CTAs per SM: at 3 x 4 s the 600 CTAs of a launch are then resident at once (66 registers made it two waves)
template <int VEC>
__global__ void __launch_bounds__(kBwThreads, 5) bwd_prelu_norm_kernel(const BwdPnParams p) {
    __shared__ float sh[4];
    __shared__ float red[8];
    const int H = p.H;
    const int m = blockIdx.y, k0 = blockIdx.x * kNsTile;
    const int nk = min(kNsTile, p.K - k0);
    const double n = static_cast<double>(p.K) * H;
    const int CV = (H + VEC - 1) / VEC;
    const long long row0 = static_cast<long long>(m) * p.K + k0;
    V<VEC> dns[kPnPre], avs[kPnPre];   // requested before the sums are finalised, as in norm_split
#pragma unroll
    for (int j = 0; j < kPnPre; ++j) {
        const int i = threadIdx.x + j * kBwThreads;
        if (i < nk * CV) {
            const int r = i / CV, c = (i - r * CV) * VEC;
            dns[j].load(p.dn1 + (row0 + r) * H + c);
            avs[j].load(p.a + (row0 + r) * H + c);
        }
    }
    if (threadIdx.x < 32) {
        const float2 s1 = gln_finalize_warp(p.stat1 + static_cast<long long>(m) * p.P1, p.P1, n);
        double a = 0.0, b = 0.0;
        for (int i = threadIdx.x; i < p.tiles_b; i += 32) {
            const float* s = p.scal_part + (static_cast<long long>(m) * p.tiles_b + i) * 4;
            a += s[0], b += s[1];
        }
        a = warp_sum(a), b = warp_sum(b);
        if (threadIdx.x == 0) sh[0] = s1.x, sh[1] = s1.y, sh[2] = static_cast<float>(a / n), sh[3] = static_cast<float>(b / n);
    }
    __syncthreads();
    const float mu1 = sh[0], r1 = sh[1], m1 = sh[2], m2 = sh[3];
    const float al1 = __ldg(p.slope1);
    float da1[1] = {0.f};
    auto emit = [&](int i, const V<VEC>& dn, const V<VEC>& av) {
        const int r = i / CV, c = (i - r * CV) * VEC;
        const long long row = row0 + r;
        V<VEC> g;
        g.load(p.g1 + c);
        float o[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            const float ph = (prelu(av.v[j], al1) - mu1) * r1;
            const float dp = r1 * (g.v[j] * dn.v[j] - m1 - ph * m2);
            o[j] = av.v[j] > 0.f ? dp : al1 * dp;
            if (!(av.v[j] > 0.f)) da1[0] = fmaf(dp, av.v[j], da1[0]);
        }
        store_hl<VEC>(p.da_hi + row * H + c, p.da_lo + row * H + c, o);
    };
#pragma unroll
    for (int j = 0; j < kPnPre; ++j) {
        const int i = threadIdx.x + j * kBwThreads;
        if (i < nk * CV) emit(i, dns[j], avs[j]);
    }
    for (int i = threadIdx.x + (kPnPre) * kBwThreads; i < nk * CV; i += kBwThreads) {
        const int r = i / CV, c = (i - r * CV) * VEC;
        V<VEC> dn, av;
        dn.load(p.dn1 + (row0 + r) * H + c);
        av.load(p.a + (row0 + r) * H + c);
        emit(i, dn, av);
    }
    block_sum<1>(da1, red);
    if (threadIdx.x == 0) p.alpha_part[static_cast<long long>(m) * gridDim.x + blockIdx.x] = da1[0];
}

// ---------------------------------------------------------------------------------------
// bwd_prelu_norm for cLN blocks: the two means of the norm backward are per frame, so one warp
// owns a frame: pass 1 sums g1*dn1 and g1*dn1*ph over the channels, pass 2 writes da (hi, lo).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBwThreads)
bwd_prelu_norm_cln_kernel(const BwdPnParams p) {
    __shared__ float red[8];
    const int H = p.H;
    const int m = blockIdx.y, k0 = blockIdx.x * kNsTile;
    const int nk = min(kNsTile, p.K - k0);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float al1 = __ldg(p.slope1);
    float da1[1] = {0.f};
    for (int r = warp; r < nk; r += kBwThreads / 32) {
        const long long row = static_cast<long long>(m) * p.K + k0 + r;
        float s = 0.f, ss = 0.f;
        for (int g = 0; g < p.P1; ++g) {
            const float2 t = __ldg(p.stat1 + row * p.P1 + g);
            s += t.x, ss += t.y;
        }
        const float mu1 = s / H;
        const float r1 = rsqrtf(fmaxf(ss / H - mu1 * mu1, 0.f) + kEps);
        float s1 = 0.f, s2 = 0.f;
        for (int c = lane; c < H; c += 32) {
            const float ph = (prelu(p.a[row * H + c], al1) - mu1) * r1;
            const float gd = p.g1[c] * p.dn1[row * H + c];
            s1 += gd;
            s2 = fmaf(gd, ph, s2);
        }
        const float m1 = warp_sum(s1) / H, m2 = warp_sum(s2) / H;
        for (int c = lane; c < H; c += 32) {
            const float av = p.a[row * H + c];
            const float ph = (prelu(av, al1) - mu1) * r1;
            const float dp = r1 * (p.g1[c] * p.dn1[row * H + c] - m1 - ph * m2);
            const float o = av > 0.f ? dp : al1 * dp;
            if (!(av > 0.f)) da1[0] = fmaf(dp, av, da1[0]);
            __nv_bfloat16 h, l;
            split_bf16(o, h, l);
            p.da_hi[row * H + c] = h, p.da_lo[row * H + c] = l;
        }
    }
    block_sum<1>(da1, red);
    if (threadIdx.x == 0) p.alpha_part[static_cast<long long>(m) * gridDim.x + blockIdx.x] = da1[0];
}

// ---------------------------------------------------------------------------------------
// reduce_partials: dst[seg][i * stride] (+)= sum_p part[p * part_stride + src_off[seg] + i]
// Fixed summation order over p -> deterministic.
// ---------------------------------------------------------------------------------------
constexpr int kMaxSeg = 12;
struct ReduceParams {
    const float* part[kMaxSeg];
    int n_part[kMaxSeg];
    long long part_stride[kMaxSeg];
    int nseg;
    int accumulate;
    long long src_off[kMaxSeg];
    int len[kMaxSeg];
    int src_stride[kMaxSeg];
    float* dst[kMaxSeg];
    int dst_stride[kMaxSeg];
    int blk_off[kMaxSeg + 1];   // first CTA of each segment (prefix sum of ceil(len / red_span(n_part)))
};

// One launch sums the partials of up to kMaxSeg segments; the grid is ONE dimension over the (segment,
// column block) pairs that exist (blk_off = prefix sum of blocks per segment) -- the 2-D grid (longest
// segment x segments) this replaces launched 45 000 CTAs per block of the paper model, 37 000 of them
// for columns the short segments do not have.  Two layouts of the 256 threads, chosen per segment:
//  * wide (few partials, long rows: weight-gradient splits): 32 columns x 8 slices, four column groups
//    per thread, i.e. 128 outputs per CTA, each warp reading 128 contiguous bytes per partial;
//  * tall (more than kRedTall partials, short rows: per-CTA channel sums of the depthwise backward,
//    up to ~2000 rows of 512): 8 columns x 32 slices, so the serial chain per thread is n_part / 32
//    instead of n_part / 8 and a 512-column segment spreads over 64 CTAs instead of 4 (67 -> 20 us at
//    batch 8 of BASELINE configs[4]).
// Slices are added in a fixed order: deterministic.
constexpr int kRedThreads = 256;
constexpr int kRedCols = 32, kRedSlices = 8, kRedGroups = 4, kRedSpan = kRedCols * kRedGroups;
constexpr int kRedTall = 128, kRedTallCols = 8, kRedTallSlices = kRedThreads / kRedTallCols;
__host__ __device__ inline int red_span(int n_part) { return n_part > kRedTall ? kRedTallCols : kRedSpan; }

__global__ void __launch_bounds__(kRedThreads) reduce_partials_kernel(const ReduceParams p) {
    __shared__ float red[kRedThreads * kRedGroups];
    int seg = 0;
    while (seg + 1 < p.nseg && static_cast<int>(blockIdx.x) >= p.blk_off[seg + 1]) ++seg;
    const int len = p.len[seg];
    const int n_part = p.n_part[seg];
    const long long ps = p.part_stride[seg];
    const int sstr = p.src_stride[seg];
    const float* base = p.part[seg] + p.src_off[seg];
    const int blk = blockIdx.x - p.blk_off[seg];
    if (n_part > kRedTall) {
        const int col = threadIdx.x & (kRedTallCols - 1), slice = threadIdx.x / kRedTallCols;
        const int i = blk * kRedTallCols + col;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
        if (i < len) {
            const float* src = base + static_cast<long long>(i) * sstr;
            int q = slice;
            for (; q + 3 * kRedTallSlices < n_part; q += 4 * kRedTallSlices) {
                s0 += src[(q + 0 * kRedTallSlices) * ps];
                s1 += src[(q + 1 * kRedTallSlices) * ps];
                s2 += src[(q + 2 * kRedTallSlices) * ps];
                s3 += src[(q + 3 * kRedTallSlices) * ps];
            }
            for (; q < n_part; q += kRedTallSlices) s0 += src[q * ps];
        }
        red[slice * kRedTallCols + col] = (s0 + s1) + (s2 + s3);
        __syncthreads();
        if (threadIdx.x < kRedTallCols && i < len) {
            float t = 0.f;
#pragma unroll
            for (int k = 0; k < kRedTallSlices; ++k) t += red[k * kRedTallCols + threadIdx.x];
            float* d = p.dst[seg] + static_cast<long long>(i) * p.dst_stride[seg];
            *d = p.accumulate ? *d + t : t;
        }
        return;
    }
    const int col = threadIdx.x & (kRedCols - 1), slice = threadIdx.x / kRedCols;
    const int i0 = blk * kRedSpan + col;
    float s[kRedGroups];
    const float* src[kRedGroups];
    bool ok[kRedGroups];
#pragma unroll
    for (int g = 0; g < kRedGroups; ++g) {
        const int i = i0 + g * kRedCols;
        ok[g] = i < len;
        src[g] = base + static_cast<long long>(ok[g] ? i : 0) * sstr;
        s[g] = 0.f;
    }
    for (int q = slice; q < n_part; q += kRedSlices) {
        float v[kRedGroups];
#pragma unroll
        for (int g = 0; g < kRedGroups; ++g) v[g] = ok[g] ? src[g][q * ps] : 0.f;
#pragma unroll
        for (int g = 0; g < kRedGroups; ++g) s[g] += v[g];
    }
#pragma unroll
    for (int g = 0; g < kRedGroups; ++g) red[slice * kRedSpan + g * kRedCols + col] = s[g];
    __syncthreads();
    if (threadIdx.x < kRedSpan) {
        const int i = blk * kRedSpan + threadIdx.x;
        if (i < len) {
            float t = 0.f;
#pragma unroll
            for (int k = 0; k < kRedSlices; ++k) t += red[k * kRedSpan + threadIdx.x];
            float* d = p.dst[seg] + static_cast<long long>(i) * p.dst_stride[seg];
            *d = p.accumulate ? *d + t : t;
        }
    }
}

// ---------------------------------------------------------------------------------------
// decoder_bwd: backward of mask non-linearity, mask * w, basis Linear and overlap-add.
//   df[c,l]   = d est[m, c, k*hop + l]                                   (gather = OLA backward)
//   dsw[c,n]  = sum_l df[c,l] V[l,n];   sw[c,n] = mask_c[n] * w[n]
//   dV[l,n]  += sum_c df[c,l] sw[c,n];  dw[n] = sum_c dsw[c,n] mask_c[n]; dmask_c = dsw[c,n] w[n]
//   dscore: relu -> dmask * [score > 0];  softmax -> mask_c (dmask_c - sum_c' mask_c' dmask_c')
// One thread owns channel n and walks the frames of its tile.
// ---------------------------------------------------------------------------------------
constexpr int kDbTile = 32;
constexpr int kDbMaxC = 8;
constexpr int kDbMaxL = 64;

// LT: the filter length as a compile-time constant (0 = run-time L).  With it the per-thread basis column and its
// gradient (2 L floats) live in registers; indexed by a run-time bound they were local-memory arrays, read and
// written in the innermost loop (109 -> 4x us per launch at 3 x 4 s).
template <int LT>
__global__ void __launch_bounds__(kBwThreads)
decoder_bwd_kernel(const float* __restrict__ gest, const float* __restrict__ score, const float* __restrict__ w,
                   const float* __restrict__ V, int K, int N, int L, int C, int T, int softmax,
                   __nv_bfloat16* __restrict__ ds_hi, __nv_bfloat16* __restrict__ ds_lo, float* __restrict__ dw,
                   float* __restrict__ dV_part) {
    extern __shared__ float db_smem[];
    if (LT) L = LT;
    const int hop = L / 2;
    const int CL = C * L;
    constexpr int kLA = LT ? LT : kDbMaxL;
    float* df = db_smem;  // [kDbTile][C][L]
    const int m = blockIdx.y, k0 = blockIdx.x * kDbTile;
    const int nk = min(kDbTile, K - k0);
    for (int i = threadIdx.x; i < nk * CL; i += blockDim.x) {
        const int r = i / CL, cl = i - r * CL;
        const int c = cl / L, l = cl - c * L;
        df[i] = gest[(static_cast<long long>(m) * C + c) * T + static_cast<long long>(k0 + r) * hop + l];
    }
    __syncthreads();
    float* part = dV_part + (static_cast<long long>(m) * gridDim.x + blockIdx.x) * L * N;
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
        float Vn[kLA], dVn[kLA];
#pragma unroll
        for (int l = 0; l < kLA; ++l)
            if (l < L) Vn[l] = V[l * N + n], dVn[l] = 0.f;
        // the next frame's mixture weight and scores are requested before this frame's arithmetic
        float wv_nx = 0.f, sc_nx[kDbMaxC];
        auto fetch = [&](int r) {
            const long long row = static_cast<long long>(m) * K + k0 + r;
            wv_nx = w[row * N + n];
#pragma unroll
            for (int c = 0; c < kDbMaxC; ++c)
                if (c < C) sc_nx[c] = score[row * (static_cast<long long>(C) * N) + c * N + n];
        };
        if (nk > 0) fetch(0);
        for (int r = 0; r < nk; ++r) {
            const long long row = static_cast<long long>(m) * K + k0 + r;
            const float wv = wv_nx;
            float mask[kDbMaxC], dsw[kDbMaxC];
#pragma unroll
            for (int c = 0; c < kDbMaxC; ++c)
                if (c < C) mask[c] = sc_nx[c];   // raw score; relu applied below
            if (r + 1 < nk) fetch(r + 1);
            if (softmax) {
                float mx = -INFINITY;
#pragma unroll
                for (int c = 0; c < kDbMaxC; ++c)
                    if (c < C) mx = fmaxf(mx, mask[c]);
                float den = 0.f;
#pragma unroll
                for (int c = 0; c < kDbMaxC; ++c)
                    if (c < C) mask[c] = expf(mask[c] - mx), den += mask[c];
                const float inv = 1.f / den;
#pragma unroll
                for (int c = 0; c < kDbMaxC; ++c)
                    if (c < C) mask[c] *= inv;
            }
            float dwn = 0.f, dot = 0.f;
#pragma unroll
            for (int c = 0; c < kDbMaxC; ++c) {
                if (c >= C) continue;
                const float* f = df + r * CL + c * L;
                const float mk = softmax ? mask[c] : fmaxf(mask[c], 0.f);
                const float sw = mk * wv;
                float acc = 0.f;
#pragma unroll
                for (int l = 0; l < kLA; ++l) {
                    if (l < L) {
                        acc = fmaf(f[l], Vn[l], acc);
                        dVn[l] = fmaf(f[l], sw, dVn[l]);
                    }
                }
                dwn = fmaf(acc, mk, dwn);
                dsw[c] = acc * wv;              // d mask_c
                dot = fmaf(mk, dsw[c], dot);
            }
            dw[row * N + n] = dwn;
#pragma unroll
            for (int c = 0; c < kDbMaxC; ++c) {
                if (c >= C) continue;
                float g;
                if (softmax) g = mask[c] * (dsw[c] - dot);
                else g = mask[c] > 0.f ? dsw[c] : 0.f;
                __nv_bfloat16 h, lo;
                split_bf16(g, h, lo);
                const long long o = row * (static_cast<long long>(C) * N) + c * N + n;
                ds_hi[o] = h, ds_lo[o] = lo;
            }
        }
#pragma unroll
        for (int l = 0; l < kLA; ++l)
            if (l < L) part[l * N + n] = dVn[l];
    }
}

// ---------------------------------------------------------------------------------------
// cln_bwd: backward of the leading cLN (conv_tasnet.py:167,326-328) + encoder ReLU, merged
// with the decoder's direct gradient on the encoder output:
//   dwh = g0*dn0;  dw_sep = r*(dwh - mean_n(dwh) - wh*mean_n(dwh*wh));  dpre = (dw_sep + dw_dec)*[w>0]
// One warp per frame.  Per-channel (dg0, db0) partials per CTA.
// ---------------------------------------------------------------------------------------
constexpr int kCbTile = 32;
constexpr int kCbMaxPerLane = 32;  // N <= 1024

__global__ void __launch_bounds__(kBwThreads)
cln_bwd_kernel(const float* __restrict__ dn0, const float* __restrict__ w, const float2* __restrict__ rowstat,
               const float* __restrict__ g0, const float* __restrict__ dw_dec, int K, int N,
               float* __restrict__ dpre, float* __restrict__ chan_part) {
    extern __shared__ float cb_smem[];  // [8 warps][2][N]
    const int m = blockIdx.y, k0 = blockIdx.x * kCbTile;
    const int nk = min(kCbTile, K - k0);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* my = cb_smem + warp * 2 * N;
    for (int n = lane; n < N; n += 32) my[n] = 0.f, my[N + n] = 0.f;
    for (int r = warp; r < nk; r += 8) {
        const long long row = static_cast<long long>(m) * K + k0 + r;
        const float2 st = __ldg(rowstat + row);
        float s1 = 0.f, s2 = 0.f;
        for (int n = lane; n < N; n += 32) {
            const float wh = (w[row * N + n] - st.x) * st.y;
            const float dy = dn0[row * N + n];
            const float dwh = g0[n] * dy;
            s1 += dwh;
            s2 = fmaf(dwh, wh, s2);
            my[n] = fmaf(dy, wh, my[n]);
            my[N + n] += dy;
        }
        s1 = warp_sum(s1) / N;
        s2 = warp_sum(s2) / N;
        for (int n = lane; n < N; n += 32) {
            const float wv = w[row * N + n];
            const float wh = (wv - st.x) * st.y;
            const float dwh = g0[n] * dn0[row * N + n];
            const float tot = st.y * (dwh - s1 - wh * s2) + dw_dec[row * N + n];
            dpre[row * N + n] = wv > 0.f ? tot : 0.f;
        }
    }
    __syncthreads();
    float* part = chan_part + (static_cast<long long>(m) * gridDim.x + blockIdx.x) * 2 * N;
    for (int i = threadIdx.x; i < 2 * N; i += blockDim.x) {
        float s = 0.f;
        for (int wi = 0; wi < 8; ++wi) s += cb_smem[wi * 2 * N + i];
        part[i] = s;
    }
}

// ---------------------------------------------------------------------------------------
// encoder_wgrad: dU[n, l] = sum_{m,k} dpre[m,k,n] * mixture[m, k*hop + l]   (conv_tasnet.py:106,116)
// One thread owns channel n; the tile's samples sit in shared memory.
// ---------------------------------------------------------------------------------------
constexpr int kEwTile = 64;

template <int LT>   // filter length at compile time (0 = run-time L): the accumulators stay in registers
__global__ void __launch_bounds__(kBwThreads)
encoder_wgrad_kernel(const float* __restrict__ dpre, const float* __restrict__ mixture, int T, int K, int N, int L,
                     float* __restrict__ dU_part) {
    extern __shared__ float ew_smem[];
    if (LT) L = LT;
    constexpr int kLA = LT ? LT : kDbMaxL;
    const int hop = L / 2;
    const int m = blockIdx.y, k0 = blockIdx.x * kEwTile;
    const int nk = min(kEwTile, K - k0);
    const int nsamp = (nk - 1) * hop + L;
    const float* src = mixture + static_cast<long long>(m) * T + static_cast<long long>(k0) * hop;
    for (int i = threadIdx.x; i < nsamp; i += blockDim.x) ew_smem[i] = src[i];
    __syncthreads();
    float* part = dU_part + (static_cast<long long>(m) * gridDim.x + blockIdx.x) * N * L;
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
        float acc[kLA];
#pragma unroll
        for (int l = 0; l < kLA; ++l) acc[l] = 0.f;
        const float* gp = dpre + (static_cast<long long>(m) * K + k0) * N + n;
        for (int r0 = 0; r0 < nk; r0 += 4) {       // four frames' gradients requested together
            float g[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) g[j] = r0 + j < nk ? gp[static_cast<long long>(r0 + j) * N] : 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (r0 + j < nk) {
                    const float* x = ew_smem + (r0 + j) * hop;
#pragma unroll
                    for (int l = 0; l < kLA; ++l)
                        if (l < L) acc[l] = fmaf(g[j], x[l], acc[l]);
                }
            }
        }
#pragma unroll
        for (int l = 0; l < kLA; ++l)
            if (l < L) part[n * L + l] = acc[l];
    }
}

}  // namespace ctn

// =========================================================================================
using namespace ctn;

static inline bool al16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; }

extern "C" int ctn_norm_split(const float* in, const float* prelu_slope, const float* gamma, const float* beta,
                              int norm, const void* stat, int P, int M, int K, int C, void* out_hi, void* out_lo,
                              void* stream) {
    CTN_REQUIRE(in && gamma && beta && stat && out_hi && out_lo, "ctn_norm_split: NULL pointer");
    CTN_REQUIRE(M > 0 && K > 0 && C > 0, "ctn_norm_split: bad shape");
    CTN_REQUIRE(norm == CTN_NORM_GLN || norm == CTN_NORM_CLN, "ctn_norm_split: norm must be gLN or cLN");
    CTN_REQUIRE(norm != CTN_NORM_GLN || P > 0, "ctn_norm_split: gLN needs P");
    NormSplitParams p;
    p.K = K, p.C = C, p.norm = norm, p.P = P, p.count = static_cast<double>(K) * C;
    p.in = in, p.slope = prelu_slope, p.gamma = gamma, p.beta = beta, p.stat = static_cast<const float2*>(stat);
    p.out_hi = static_cast<__nv_bfloat16*>(out_hi), p.out_lo = static_cast<__nv_bfloat16*>(out_lo);
    dim3 grid(ceil_div(K, kNsTile), M);
    if (C % 4 == 0 && al16(in) && al16(out_hi) && al16(out_lo) && al16(gamma) && al16(beta))
        norm_split_kernel<4><<<grid, kBwThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
    else
        norm_split_kernel<1><<<grid, kBwThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
    CTN_LAUNCH_CHECK("norm_split_kernel");
    return CTN_OK;
}

static bool bwd_use_march(int H, int P, int norm) { return P == 3 && H % 4 == 0 && norm != CTN_NORM_CLN; }

extern "C" int ctn_bwd_dw_emits_u(int H, int P, int norm) { return bwd_use_march(H, P, norm) ? 1 : 0; }

extern "C" int ctn_bwd_dw_layout(int M, int K, int H, int P, int dilation, int norm, int* ctas, int* chan_rows) {
    if (bwd_use_march(H, P, norm) && dilation > 0) {
        int NS, units, groups, n_ctas;
        bwd_march_geometry(M, K, H, dilation, &NS, &units, &groups, &n_ctas);
        if (ctas) *ctas = n_ctas;
        if (chan_rows) *chan_rows = groups;
        return CTN_OK;
    }
    const int vec = H % 4 == 0 ? 4 : 1;
    const int CV = ceil_div(H, vec);
    const int CVt = CV < kBwThreads ? CV : kBwThreads;
    const int tiles = ceil_div(K, kBdTile);
    const int rg = kBwThreads / CVt > 0 ? kBwThreads / CVt : 1;
    if (ctas) *ctas = tiles;
    if (chan_rows) *chan_rows = tiles * rg;
    return CTN_OK;
}

extern "C" int ctn_bwd_dwconv(const ctn_bwd_dw_args* args, void* stream) {
    CTN_REQUIRE(args != nullptr, "ctn_bwd_dwconv: args is NULL");
    const ctn_bwd_dw_args& a = *args;
    CTN_REQUIRE(a.M > 0 && a.K > 0 && a.H > 0 && a.P > 0 && a.P <= kBdMaxP && a.dilation > 0,
                "ctn_bwd_dwconv: bad shape");
    CTN_REQUIRE(a.du && a.d && a.a && a.stat1 && a.stat2 && a.bstat && a.g1 && a.b1 && a.g2 && a.wd && a.slope1 &&
                    a.slope2 && a.dn1 && a.chan_part && a.scal_part,
                "ctn_bwd_dwconv: NULL pointer");
    const bool aligned = al16(a.du) && al16(a.d) && al16(a.a) && al16(a.dn1) && al16(a.g1) && al16(a.b1) &&
                         al16(a.g2) && al16(a.wd) && al16(a.chan_part);
    if (bwd_use_march(a.H, a.P, a.norm)) {
        CTN_REQUIRE(aligned, "ctn_bwd_dwconv: pointers must be 16-byte aligned");
        BwdMarchParams q;
        q.K = a.K, q.H = a.H, q.dil = a.dilation, q.c0 = a.causal ? 2 : 1;
        q.P1 = a.P1, q.P2 = a.P2, q.Pb = a.Pb;
        bwd_march_geometry(a.M, a.K, a.H, a.dilation, &q.NS, &q.units, &q.groups, &q.ctas);
        q.upw = bwd_march_units_per_warp(a.M, q.units, a.H);
        q.chan_rows = q.groups;
        q.du = a.du, q.d = a.d, q.a = a.a;
        q.stat1 = static_cast<const float2*>(a.stat1), q.stat2 = static_cast<const float2*>(a.stat2);
        q.bstat = static_cast<const float2*>(a.bstat);
        q.g1 = a.g1, q.b1 = a.b1, q.g2 = a.g2, q.wd = a.wd, q.slope1 = a.slope1, q.slope2 = a.slope2;
        q.dn1 = a.dn1, q.chan_part = a.chan_part, q.scal_part = a.scal_part;
        dim3 grid(q.ctas, a.M);
        const bool emit_u = a.u_hi != nullptr;
        CTN_REQUIRE(!emit_u || (a.u_lo && a.b2 && al16(a.u_hi) && al16(a.u_lo) && al16(a.b2)),
                    "ctn_bwd_dwconv: u_hi, u_lo and b2 go together, 16-byte aligned");
        q.b2 = a.b2, q.u_hi = static_cast<__nv_bfloat16*>(a.u_hi), q.u_lo = static_cast<__nv_bfloat16*>(a.u_lo);
        void (*kern)(const BwdMarchParams) =
            a.causal ? (emit_u ? bwd_dw_march_kernel<2, true> : bwd_dw_march_kernel<2, false>)
                     : (emit_u ? bwd_dw_march_kernel<1, true> : bwd_dw_march_kernel<1, false>);
        CTN_REQUIRE(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kBmSmemBytes) == cudaSuccess,
                    "ctn_bwd_dwconv: cannot reserve the row ring in shared memory");
        kern<<<grid, kBmWarps * 32, kBmSmemBytes, static_cast<cudaStream_t>(stream)>>>(q);
        CTN_LAUNCH_CHECK("bwd_dw_march_kernel");
        return CTN_OK;
    }
    CTN_REQUIRE(a.u_hi == nullptr && a.u_lo == nullptr, "ctn_bwd_dwconv: u is not written for this shape / norm (ctn_bwd_dw_emits_u)");
    BwdDwParams p;
    p.K = a.K, p.H = a.H, p.P = a.P, p.dil = a.dilation, p.causal = a.causal;
    p.P1 = a.P1, p.P2 = a.P2, p.Pb = a.Pb, p.cln = a.norm == CTN_NORM_CLN;
    p.du = a.du, p.d = a.d, p.a = a.a;
    p.stat1 = static_cast<const float2*>(a.stat1), p.stat2 = static_cast<const float2*>(a.stat2);
    p.bstat = static_cast<const float2*>(a.bstat);
    p.g1 = a.g1, p.b1 = a.b1, p.g2 = a.g2, p.wd = a.wd, p.slope1 = a.slope1, p.slope2 = a.slope2;
    p.dn1 = a.dn1, p.chan_part = a.chan_part, p.scal_part = a.scal_part;
    int tiles, rows;
    ctn_bwd_dw_layout(a.M, a.K, a.H, a.P, a.dilation, a.norm, &tiles, &rows);
    p.RG = rows / tiles;
    dim3 grid(tiles, a.M);
    const bool vec = a.H % 4 == 0 && aligned;
    CTN_REQUIRE(vec || a.H % 4 != 0, "ctn_bwd_dwconv: pointers must be 16-byte aligned");
    if (vec)
        bwd_dw_kernel<4><<<grid, kBwThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
    else
        bwd_dw_kernel<1><<<grid, kBwThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
    CTN_LAUNCH_CHECK("bwd_dw_kernel");
    return CTN_OK;
}

extern "C" int ctn_bwd_prelu_norm_tiles(int K) { return ceil_div(K, kNsTile); }

extern "C" int ctn_bwd_prelu_norm(const float* dn1, const float* a, const void* stat1, int P1, const float* scal_part,
                                  int tiles_b, const float* g1, const float* slope1, int M, int K, int H,
                                  void* da_hi, void* da_lo, float* alpha_part, void* stream) {
    CTN_REQUIRE(dn1 && a && stat1 && scal_part && g1 && slope1 && da_hi && da_lo && alpha_part,
                "ctn_bwd_prelu_norm: NULL pointer");
    CTN_REQUIRE(M > 0 && K > 0 && H > 0 && P1 > 0 && tiles_b > 0, "ctn_bwd_prelu_norm: bad shape");
    BwdPnParams p;
    p.K = K, p.H = H, p.P1 = P1, p.tiles_b = tiles_b;
    p.dn1 = dn1, p.a = a, p.stat1 = static_cast<const float2*>(stat1), p.scal_part = scal_part;
    p.g1 = g1, p.slope1 = slope1;
    p.da_hi = static_cast<__nv_bfloat16*>(da_hi), p.da_lo = static_cast<__nv_bfloat16*>(da_lo);
    p.alpha_part = alpha_part;
    dim3 grid(ceil_div(K, kNsTile), M);
    if (H % 4 == 0 && al16(dn1) && al16(a) && al16(g1) && al16(da_hi) && al16(da_lo))
        bwd_prelu_norm_kernel<4><<<grid, kBwThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
    else
        bwd_prelu_norm_kernel<1><<<grid, kBwThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
    CTN_LAUNCH_CHECK("bwd_prelu_norm_kernel");
    return CTN_OK;
}

extern "C" int ctn_bwd_prelu_norm_cln(const float* dn1, const float* a, const void* stat1, int P1, const float* g1,
                                      const float* slope1, int M, int K, int H, void* da_hi, void* da_lo,
                                      float* alpha_part, void* stream) {
    CTN_REQUIRE(dn1 && a && stat1 && g1 && slope1 && da_hi && da_lo && alpha_part, "ctn_bwd_prelu_norm_cln: NULL pointer");
    CTN_REQUIRE(M > 0 && K > 0 && H > 0 && P1 > 0, "ctn_bwd_prelu_norm_cln: bad shape");
    BwdPnParams p;
    p.K = K, p.H = H, p.P1 = P1, p.tiles_b = 0;
    p.dn1 = dn1, p.a = a, p.stat1 = static_cast<const float2*>(stat1), p.scal_part = nullptr;
    p.g1 = g1, p.slope1 = slope1;
    p.da_hi = static_cast<__nv_bfloat16*>(da_hi), p.da_lo = static_cast<__nv_bfloat16*>(da_lo);
    p.alpha_part = alpha_part;
    dim3 grid(ceil_div(K, kNsTile), M);
    bwd_prelu_norm_cln_kernel<<<grid, kBwThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
    CTN_LAUNCH_CHECK("bwd_prelu_norm_cln_kernel");
    return CTN_OK;
}

extern "C" int ctn_reduce_partials(const ctn_reduce_args* args, void* stream) {
    CTN_REQUIRE(args != nullptr, "ctn_reduce_partials: args is NULL");
    const ctn_reduce_args& a = *args;
    CTN_REQUIRE(a.nseg > 0 && a.nseg <= kMaxSeg, "ctn_reduce_partials: bad arguments");
    ReduceParams p;
    p.nseg = a.nseg, p.accumulate = a.accumulate;
    int blocks = 0;
    for (int s = 0; s < a.nseg; ++s) {
        CTN_REQUIRE(a.dst[s] != nullptr && a.len[s] > 0, "ctn_reduce_partials: bad segment %d", s);
        p.part[s] = a.seg_part[s] ? a.seg_part[s] : a.part;
        p.n_part[s] = a.seg_part[s] ? a.seg_n_part[s] : a.n_part;
        p.part_stride[s] = a.seg_part[s] ? a.seg_part_stride[s] : a.part_stride;
        CTN_REQUIRE(p.part[s] != nullptr && p.n_part[s] > 0, "ctn_reduce_partials: segment %d has no partials", s);
        p.src_off[s] = a.src_off[s], p.len[s] = a.len[s], p.src_stride[s] = a.src_stride[s] ? a.src_stride[s] : 1;
        p.dst[s] = a.dst[s], p.dst_stride[s] = a.dst_stride[s] ? a.dst_stride[s] : 1;
        p.blk_off[s] = blocks;
        blocks += ceil_div(a.len[s], red_span(p.n_part[s]));
    }
    p.blk_off[a.nseg] = blocks;
    reduce_partials_kernel<<<blocks, kRedThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
    CTN_LAUNCH_CHECK("reduce_partials_kernel");
    return CTN_OK;
}

extern "C" int ctn_decoder_bwd_tiles(int K) { return ceil_div(K, kDbTile); }

extern "C" int ctn_decoder_bwd(const float* grad_est, const float* score, const float* w, const float* V, int M,
                               int K, int N, int L, int C, int T, int mask_nonlinear, void* dscore_hi,
                               void* dscore_lo, float* dw, float* dV_part, void* stream) {
    CTN_REQUIRE(grad_est && score && w && V && dscore_hi && dscore_lo && dw && dV_part, "ctn_decoder_bwd: NULL pointer");
    CTN_REQUIRE(M > 0 && K > 0 && N > 0 && L >= 2 && L <= kDbMaxL && C > 0 && C <= kDbMaxC,
                "ctn_decoder_bwd: bad shape (L <= %d, C <= %d)", kDbMaxL, kDbMaxC);
    CTN_REQUIRE(mask_nonlinear == CTN_MASK_RELU || mask_nonlinear == CTN_MASK_SOFTMAX,
                "Unsupported mask non-linear function");
    CTN_REQUIRE(T >= (K - 1) * (L / 2) + L, "ctn_decoder_bwd: T too short");
    const size_t smem = static_cast<size_t>(kDbTile) * C * L * sizeof(float);
    CTN_REQUIRE(smem <= 48 * 1024, "ctn_decoder_bwd: C*L too large");
    dim3 grid(ceil_div(K, kDbTile), M);
    auto kern = L == 20 ? decoder_bwd_kernel<20> : L == 16 ? decoder_bwd_kernel<16> : L == 40 ? decoder_bwd_kernel<40>
                                                                                                 : decoder_bwd_kernel<0>;
    kern<<<grid, kBwThreads, smem, static_cast<cudaStream_t>(stream)>>>(
        grad_est, score, w, V, K, N, L, C, T, mask_nonlinear == CTN_MASK_SOFTMAX,
        static_cast<__nv_bfloat16*>(dscore_hi), static_cast<__nv_bfloat16*>(dscore_lo), dw, dV_part);
    CTN_LAUNCH_CHECK("decoder_bwd_kernel");
    return CTN_OK;
}

extern "C" int ctn_cln_bwd_tiles(int K) { return ceil_div(K, kCbTile); }

extern "C" int ctn_cln_bwd(const float* dn0, const float* w, const void* rowstat, const float* gamma0,
                           const float* dw_dec, int M, int K, int N, float* dpre, float* chan_part, void* stream) {
    CTN_REQUIRE(dn0 && w && rowstat && gamma0 && dw_dec && dpre && chan_part, "ctn_cln_bwd: NULL pointer");
    CTN_REQUIRE(M > 0 && K > 0 && N > 0, "ctn_cln_bwd: bad shape");
    const size_t smem = static_cast<size_t>(8) * 2 * N * sizeof(float);
    CTN_REQUIRE(smem <= 200 * 1024, "ctn_cln_bwd: N too large");
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(cln_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(smem));
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(cln_bwd_kernel)");
    }
    dim3 grid(ceil_div(K, kCbTile), M);
    cln_bwd_kernel<<<grid, kBwThreads, smem, static_cast<cudaStream_t>(stream)>>>(
        dn0, w, static_cast<const float2*>(rowstat), gamma0, dw_dec, K, N, dpre, chan_part);
    CTN_LAUNCH_CHECK("cln_bwd_kernel");
    return CTN_OK;
}

extern "C" int ctn_encoder_wgrad_tiles(int K) { return ceil_div(K, kEwTile); }

extern "C" int ctn_encoder_wgrad(const float* dpre, const float* mixture, int M, int T, int N, int L, float* dU_part,
                                 void* stream) {
    CTN_REQUIRE(dpre && mixture && dU_part, "ctn_encoder_wgrad: NULL pointer");
    CTN_REQUIRE(M > 0 && N > 0 && L >= 2 && L <= kDbMaxL && T >= L, "ctn_encoder_wgrad: bad shape");
    const int hop = L / 2;
    const int K = (T - L) / hop + 1;
    const size_t smem = static_cast<size_t>((kEwTile - 1) * hop + L) * sizeof(float);
    dim3 grid(ceil_div(K, kEwTile), M);
    auto kern = L == 20 ? encoder_wgrad_kernel<20> : L == 16 ? encoder_wgrad_kernel<16> : L == 40 ? encoder_wgrad_kernel<40>
                                                                                                       : encoder_wgrad_kernel<0>;
    kern<<<grid, kBwThreads, smem, static_cast<cudaStream_t>(stream)>>>(dpre, mixture, T, K, N, L,
                                                                                        dU_part);
    CTN_LAUNCH_CHECK("encoder_wgrad_kernel");
    return CTN_OK;
}
