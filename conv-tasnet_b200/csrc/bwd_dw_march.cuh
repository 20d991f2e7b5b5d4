// Fast path of ctn_bwd_dwconv for P = 3, H % 4 == 0 (the paper configurations): the backward
// counterpart of dwconv_march.cuh.  A warp owns 128 channels of one dilation phase and marches
// along its frames, so du, d and a are each read ONCE (12 B / element) and dn1 written once; the
// three-tap windows of dd (gradient at the depthwise output) and of n1 (its input) live in
// registers, as do all per-channel gradient sums.  The row-per-thread kernel it replaces re-read
// every neighbour and walked 64 rows serially per thread: 40 % of a training step.
//
// Per loaded row i of the phase (sequence index; rows outside [0, J) are zero padding):
//   q  = prelu(d, a2)        qh = (q - mu2) r2        dq = r2 (g2 du - m1 - qh m2)
//   dd = d > 0 ? dq : a2 dq  (grad at the depthwise output)          da2 += dq * min(d, 0)
//   ph = (prelu(a, a1) - mu1) r1                      n1 = g1 ph + b1
//   dwd[t] += dd[j] n1[j - c0 + t]        with j = i + c0 - 2   (all three taps are in the window)
//   dn1[j'] = w0 dd[i] + w1 dd[i-1] + w2 dd[i-2]   with j' = i - c0
//   dg1 += dn1 ph,  db1 += dn1,  S1' += g1 dn1,  S2' += g1 dn1 ph      (row j')
//   dg2 += du qh,   db2 += du                                          (row i)
// c0 = 1 non-causal, 2 causal (conv_tasnet.py:175-176, 247-252, 289).
#pragma once
#include "common.cuh"
#include "f32x2.cuh"
#include "ptx.cuh"

namespace ctn {

constexpr int kBmUnits = 4;   // most (phase, 32-frame segment) units a warp accumulates before writing partials
#ifndef CTN_BM_RING
#define CTN_BM_RING 8
#endif
// Rows in flight per thread.  They wait in a shared-memory ring filled by 16-byte asynchronous copies (LDGSTS), each
// thread copying and reading back only its own slots: no barrier, no registers held by loads in flight.  The 4-row
// register ring this replaces (48 registers; every first use of a loaded row was a scoreboard stall, 45 % of the
// kernel's samples at batch 3) could not be deepened under the 168-register bound of three CTAs per SM.
constexpr int kBmAhead = CTN_BM_RING;
static_assert((kBmAhead & (kBmAhead - 1)) == 0 && kBmAhead >= 4, "ring depth: a power of two");
constexpr int kBmSmemBytes = kBmAhead * 3 * 4 * 32 * 16;   // [slot][du, d, a][thread] float4
template <int N>
__device__ __forceinline__ void bm_wait_group() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bm_commit_group() { asm volatile("cp.async.commit_group;" ::: "memory"); }
// CTA shape: 4 warps, 3 CTAs per SM (168 registers per thread under that bound, a few spilled values).
// The 8-warp, one-per-SM shape (204 registers) it replaces left a third of the register file idle and,
// worse, quantised badly: at the reference's batch of 3 x 4 s the 1200 warps of work made 150 CTAs for
// 148 SMs, i.e. two rounds (51 us per launch for 25 us of work).  Small CTAs fill the 1776 warp slots
// evenly: backward 6.0 -> 5.2 ms at batch 3, 18.6 -> 16.5 ms at batch 16.
constexpr int kBmWarps = 4;
#ifndef CTN_BM_CTAS
#define CTN_BM_CTAS 3
#endif
constexpr int kBmCtasPerSm = CTN_BM_CTAS;

struct BwdMarchParams {
    int K, H, dil, c0, P1, P2, Pb, NS, units, groups, chan_rows, ctas, upw;
    const float* du;
    const float* d;
    const float* a;
    const float2* stat1;
    const float2* stat2;
    const float2* bstat;
    const float* g1;
    const float* b1;
    const float* g2;
    const float* wd;
    const float* slope1;
    const float* slope2;
    float* dn1;
    float* chan_part;   // [M * chan_rows][7 * H]: dg2, db2, dg1, db1, dwd[0..3)
    float* scal_part;   // [M * ctas][4]: S1', S2', d a2, 0
    const float* b2;    // EMIT_U: u = g2 qh + b2 as bf16 (hi, lo), rows of the unit's centre only
    __nv_bfloat16* u_hi;
    __nv_bfloat16* u_lo;
};

// units per utterance, warp groups (= partial rows) per utterance, CTAs per utterance
// units per warp: as many as kBmUnits (fewer partial rows to reduce) while the grid still fills the GPU
inline int bwd_march_units_per_warp(int M, int units, int H) {
    const long long warps1 = static_cast<long long>(M) * units * ((H + 127) / 128);   // warps at one unit each
    const long long want = 2ll * 148 * kBmWarps;
    long long u = warps1 / want;
    return u < 1 ? 1 : (u > kBmUnits ? kBmUnits : static_cast<int>(u));
}
inline void bwd_march_geometry(int M, int K, int H, int dil, int* NS, int* units, int* groups, int* ctas) {
    const int Jmax = (K + dil - 1) / dil;
    *NS = (Jmax + kDmSeg - 1) / kDmSeg;
    *units = dil * (*NS);
    const int upw = bwd_march_units_per_warp(M, *units, H);
    *groups = (*units + upw - 1) / upw;
    const int CG = (H + 127) / 128;
    *ctas = ((*groups) * CG + kBmWarps - 1) / kBmWarps;
}

// block-wide sum of two doubles in a fixed order; every thread gets the result.  red: 2 * warps doubles.
__device__ __forceinline__ void block_sum2_double(double& s, double& ss, double* red) {
    s = warp_sum(s), ss = warp_sum(ss);
    const int nw = blockDim.x >> 5, w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) red[w] = s, red[nw + w] = ss;
    __syncthreads();
    s = 0.0, ss = 0.0;
    for (int i = 0; i < nw; ++i) s += red[i], ss += red[nw + i];
    __syncthreads();   // red may be reused
}

template <int C0, bool EMIT_U>
__global__ void __launch_bounds__(kBmWarps * 32, kBmCtasPerSm)
bwd_dw_march_kernel(const BwdMarchParams p) {
    __shared__ double fin_s[6 * kBmWarps];
    __shared__ float red_s[3 * kBmWarps];
    extern __shared__ __align__(16) float4 ring_s[];   // [kBmAhead][3][kBmWarps * 32]
    const int m = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int H = p.H, d = p.dil;
    const double cnt = static_cast<double>(p.K) * H;
    // The three sets of partials (norm-1 and norm-2 statistics, the dgrad epilogue's backward sums) are fetched
    // together and reduced with one barrier: three block reductions in a row were a tenth of the kernel at batch 3.
    // Same summation order as gln_finalize_block.
    double acc[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    {
        const float2* src[3] = {p.stat1 + static_cast<long long>(m) * p.P1, p.stat2 + static_cast<long long>(m) * p.P2,
                                p.bstat + static_cast<long long>(m) * p.Pb};
        const int cnts[3] = {p.P1, p.P2, p.Pb};
        constexpr int kT = kBmWarps * 32;
        float2 v[3][3];
#pragma unroll
        for (int q = 0; q < 3; ++q)
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const int i = threadIdx.x + r * kT;
                v[q][r] = i < cnts[q] ? __ldg(src[q] + i) : make_float2(0.f, 0.f);
            }
#pragma unroll
        for (int q = 0; q < 3; ++q) {
#pragma unroll
            for (int r = 0; r < 3; ++r) acc[2 * q] += static_cast<double>(v[q][r].x), acc[2 * q + 1] += static_cast<double>(v[q][r].y);
            for (int i = threadIdx.x + 3 * kT; i < cnts[q]; i += kT) {
                const float2 t = __ldg(src[q] + i);
                acc[2 * q] += static_cast<double>(t.x), acc[2 * q + 1] += static_cast<double>(t.y);
            }
        }
#pragma unroll
        for (int q = 0; q < 6; ++q) {
            acc[q] = warp_sum(acc[q]);
            if ((threadIdx.x & 31) == 0) fin_s[q * kBmWarps + (threadIdx.x >> 5)] = acc[q];
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < 6; ++q) {
            acc[q] = 0.0;
            for (int i = 0; i < kBmWarps; ++i) acc[q] += fin_s[q * kBmWarps + i];
        }
    }
    auto mean_rstd = [&](double sum, double sumsq) {
        const double mean = sum / cnt;
        double var = sumsq / cnt - mean * mean;
        var = var > 0.0 ? var : 0.0;
        return make_float2(static_cast<float>(mean), static_cast<float>(1.0 / sqrt(var + 1e-8)));
    };
    const float2 st1 = mean_rstd(acc[0], acc[1]), st2 = mean_rstd(acc[2], acc[3]);
    const double bs = acc[4], bss = acc[5];
    const float mu1 = st1.x, r1 = st1.y, mu2 = st2.x, r2 = st2.y;
    const float m1 = static_cast<float>(bs / cnt), m2 = static_cast<float>(bss / cnt);
    const float al1 = __ldg(p.slope1), al2 = __ldg(p.slope2);

    const int item = blockIdx.x * kBmWarps + warp;
    const int CG = (H + 127) / 128;
    const int cg = item % CG;
    const int grp = item / CG;
    const int c = cg * 128 + lane * 4;
    const bool active = grp < p.groups && c < H;
    const f2_t zero = 0ull;
    f2_t dg2[2] = {zero, zero}, db2[2] = {zero, zero}, dg1[2] = {zero, zero}, db1[2] = {zero, zero};
    f2_t dw0[2] = {zero, zero}, dw1[2] = {zero, zero}, dw2[2] = {zero, zero};
    f2_t sS1[2] = {zero, zero}, sS2[2] = {zero, zero}, sA2[2] = {zero, zero};

    if (active) {
        const f2_t neg1 = f2_pack(-1.f, -1.f);
        const f2_t a1p = f2_pack(al1, al1), a2p = f2_pack(al2, al2);
        const f2_t r1p = f2_pack(r1, r1), nmu1r1 = f2_pack(-mu1 * r1, -mu1 * r1);
        const f2_t r2p = f2_pack(r2, r2), nmu2r2 = f2_pack(-mu2 * r2, -mu2 * r2);
        const f2_t nm1r2 = f2_pack(-m1 * r2, -m1 * r2), nm2r2 = f2_pack(-m2 * r2, -m2 * r2);
        f2_t g1p[2], b1p[2], g2p[2], w0[2], w1[2], w2[2];
        f2_t b2p[2] = {zero, zero};
        if (EMIT_U) {
            const float4 b2v = *reinterpret_cast<const float4*>(p.b2 + c);
            b2p[0] = f2_pack(b2v.x, b2v.y), b2p[1] = f2_pack(b2v.z, b2v.w);
        }
        {
            const float4 g1v = *reinterpret_cast<const float4*>(p.g1 + c);
            const float4 b1v = *reinterpret_cast<const float4*>(p.b1 + c);
            const float4 g2v = *reinterpret_cast<const float4*>(p.g2 + c);
            g1p[0] = f2_pack(g1v.x, g1v.y), g1p[1] = f2_pack(g1v.z, g1v.w);
            b1p[0] = f2_pack(b1v.x, b1v.y), b1p[1] = f2_pack(b1v.z, b1v.w);
            g2p[0] = f2_pack(g2v.x, g2v.y), g2p[1] = f2_pack(g2v.z, g2v.w);
            const float4 t0 = *reinterpret_cast<const float4*>(p.wd + c * 3);
            const float4 t1 = *reinterpret_cast<const float4*>(p.wd + c * 3 + 4);
            const float4 t2 = *reinterpret_cast<const float4*>(p.wd + c * 3 + 8);
            const float f[12] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w, t2.x, t2.y, t2.z, t2.w};
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                w0[i] = f2_pack(f[6 * i], f[6 * i + 3]);
                w1[i] = f2_pack(f[6 * i + 1], f[6 * i + 4]);
                w2[i] = f2_pack(f[6 * i + 2], f[6 * i + 5]);
            }
        }
        const long long ubase = static_cast<long long>(m) * p.K;
        const long long strd = static_cast<long long>(d) * H;
        for (int uu = 0; uu < p.upw; ++uu) {
            const int unit = grp * p.upw + uu;
            if (unit >= p.units) break;
            const int ph = unit % d, seg = unit / d;
            const int J = ph < p.K ? (p.K - ph + d - 1) / d : 0;
            const int j0 = seg * kDmSeg;
            const int n = min(kDmSeg, J - j0);
            if (n <= 0) continue;
            const int i0 = j0 - C0;              // sequence index of loaded row s = 0
            const int L = n + 2 * C0;            // loaded rows
            const long long off0 = (ubase + ph + static_cast<long long>(i0) * d) * H + c;   // may precede the buffer
            auto valid = [&](int s) { return s < L && i0 + s >= 0 && i0 + s < J; };
            // one commit group per row (empty for rows that do not exist), so that the group count is uniform
            float4* mine = ring_s + threadIdx.x;
            auto load = [&](int s) {
                if (valid(s)) {
                    const long long o = off0 + s * strd;
                    float4* dst = mine + (s & (kBmAhead - 1)) * (3 * kBmWarps * 32);
                    ptx::cp_async_16(dst, p.du + o);
                    ptx::cp_async_16(dst + kBmWarps * 32, p.d + o);
                    ptx::cp_async_16(dst + 2 * kBmWarps * 32, p.a + o);
                }
                bm_commit_group();
            };
#pragma unroll
            for (int s = 0; s < kBmAhead; ++s) load(s);
            f2_t dd0[2] = {zero, zero}, dd1[2] = {zero, zero}, dd2[2];
            f2_t n0[2] = {zero, zero}, n1[2] = {zero, zero}, n2[2];
            f2_t h0[2] = {zero, zero}, h1[2] = {zero, zero}, h2[2];
            constexpr int kUnroll = 3;      // the three-row windows rotate by renaming
            for (int sb = 0; sb < L; sb += kUnroll) {
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) {
                    const int s = sb + u;
                    if (s < L) {
                        const bool ok = valid(s);
                        const bool centre = s >= C0 && s < C0 + n;
                        bm_wait_group<kBmAhead - 1>();          // row s has landed
                        const float4* src = mine + (s & (kBmAhead - 1)) * (3 * kBmWarps * 32);
                        float4 vdu, vd, va;
                        if (ok) vdu = src[0], vd = src[kBmWarps * 32], va = src[2 * kBmWarps * 32];
                        load(s + kBmAhead);                     // refills the slot just read (same thread, program order)
                        if (ok) {
                            const f2_t du2[2] = {f2_pack(vdu.x, vdu.y), f2_pack(vdu.z, vdu.w)};
                            const f2_t d2[2] = {f2_pack(vd.x, vd.y), f2_pack(vd.z, vd.w)};
                            const f2_t a2v[2] = {f2_pack(va.x, va.y), f2_pack(va.z, va.w)};
                            const float dsc[4] = {vd.x, vd.y, vd.z, vd.w};
                            uint32_t uh[2], ul[2];
#pragma unroll
                            for (int i = 0; i < 2; ++i) {
                                f2_t mx, mn;
                                f2_split_sign(d2[i], neg1, mx, mn);
                                const f2_t q = f2_fma(a2p, mn, mx);
                                const f2_t qh = f2_fma(q, r2p, nmu2r2);
                                const f2_t gdu = f2_mul(g2p[i], du2[i]);
                                const f2_t dq = f2_fma(qh, nm2r2, f2_fma(gdu, r2p, nm1r2));
                                if (centre) {
                                    dg2[i] = f2_fma(du2[i], qh, dg2[i]);
                                    db2[i] = f2_add(db2[i], du2[i]);
                                    sA2[i] = f2_fma(dq, mn, sA2[i]);
                                    if (EMIT_U) {
                                        float u0, u1;
                                        f2_unpack(f2_fma(g2p[i], qh, b2p[i]), u0, u1);
                                        split_pack2(u0, u1, uh[i], ul[i]);
                                    }
                                }
                                const f2_t fac = f2_pack(dsc[2 * i] > 0.f ? 1.f : al2, dsc[2 * i + 1] > 0.f ? 1.f : al2);
                                dd2[i] = f2_mul(dq, fac);
                                f2_t amx, amn;
                                f2_split_sign(a2v[i], neg1, amx, amn);
                                const f2_t pa = f2_fma(a1p, amn, amx);
                                h2[i] = f2_fma(pa, r1p, nmu1r1);
                                n2[i] = f2_fma(g1p[i], h2[i], b1p[i]);
                            }
                            if (EMIT_U && centre) {
                                const long long o = off0 + s * strd;
                                *reinterpret_cast<uint2*>(p.u_hi + o) = make_uint2(uh[0], uh[1]);
                                *reinterpret_cast<uint2*>(p.u_lo + o) = make_uint2(ul[0], ul[1]);
                            }
                        } else {
                            dd2[0] = zero, dd2[1] = zero, n2[0] = zero, n2[1] = zero, h2[0] = zero, h2[1] = zero;
                        }
                        if (s >= 2 && s < n + 2) {           // depthwise weight gradient, dd row j = i + C0 - 2
#pragma unroll
                            for (int i = 0; i < 2; ++i) {
                                const f2_t ddj = C0 == 1 ? dd1[i] : dd2[i];
                                dw0[i] = f2_fma(ddj, n0[i], dw0[i]);
                                dw1[i] = f2_fma(ddj, n1[i], dw1[i]);
                                dw2[i] = f2_fma(ddj, n2[i], dw2[i]);
                            }
                        }
                        if (s >= 2 * C0 && s < n + 2 * C0) {  // input gradient of row j' = i - C0
                            float4 outv;
                            f2_t dn[2];
#pragma unroll
                            for (int i = 0; i < 2; ++i) {
                                dn[i] = f2_fma(w2[i], dd0[i], f2_fma(w1[i], dd1[i], f2_mul(w0[i], dd2[i])));
                                const f2_t phj = C0 == 1 ? h1[i] : h0[i];
                                dg1[i] = f2_fma(dn[i], phj, dg1[i]);
                                db1[i] = f2_add(db1[i], dn[i]);
                                const f2_t gd = f2_mul(g1p[i], dn[i]);
                                sS1[i] = f2_add(sS1[i], gd);
                                sS2[i] = f2_fma(gd, phj, sS2[i]);
                            }
                            f2_unpack(dn[0], outv.x, outv.y);
                            f2_unpack(dn[1], outv.z, outv.w);
                            *reinterpret_cast<float4*>(p.dn1 + off0 + (s - C0) * strd) = outv;
                        }
#pragma unroll
                        for (int i = 0; i < 2; ++i) {
                            dd0[i] = dd1[i], dd1[i] = dd2[i];
                            n0[i] = n1[i], n1[i] = n2[i];
                            h0[i] = h1[i], h1[i] = h2[i];
                        }
                    }
                }
            }
        }
        float* part = p.chan_part + (static_cast<long long>(m) * p.chan_rows + grp) * 7 * H + c;
        const f2_t* vals[7] = {dg2, db2, dg1, db1, dw0, dw1, dw2};
#pragma unroll
        for (int v = 0; v < 7; ++v) {
            float4 o;
            f2_unpack(vals[v][0], o.x, o.y);
            f2_unpack(vals[v][1], o.z, o.w);
            *reinterpret_cast<float4*>(part + static_cast<long long>(v) * H) = o;
        }
    }
    float sc[3];
    {
        float a, b, c2, d2;
        f2_unpack(sS1[0], a, b), f2_unpack(sS1[1], c2, d2);
        sc[0] = warp_sum((a + b) + (c2 + d2));
        f2_unpack(sS2[0], a, b), f2_unpack(sS2[1], c2, d2);
        sc[1] = warp_sum((a + b) + (c2 + d2));
        f2_unpack(sA2[0], a, b), f2_unpack(sA2[1], c2, d2);
        sc[2] = warp_sum((a + b) + (c2 + d2));
    }
    if (lane == 0) red_s[warp] = sc[0], red_s[kBmWarps + warp] = sc[1], red_s[2 * kBmWarps + warp] = sc[2];
    __syncthreads();
    if (threadIdx.x == 0) {
        float t0 = 0.f, t1 = 0.f, t2 = 0.f;
        for (int i = 0; i < kBmWarps; ++i) t0 += red_s[i], t1 += red_s[kBmWarps + i], t2 += red_s[2 * kBmWarps + i];
        float* s = p.scal_part + (static_cast<long long>(m) * p.ctas + blockIdx.x) * 4;
        s[0] = t0, s[1] = t1, s[2] = t2, s[3] = 0.f;
    }
}

}  // namespace ctn
