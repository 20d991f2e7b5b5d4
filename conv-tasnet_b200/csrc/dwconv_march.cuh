// Fast path of ctn_dwconv_fwd for gLN, P = 3, H % 4 == 0 (the paper configurations).
// A warp owns 128 channels (one float4 per lane) of one dilation phase and marches along the
// frames k = ph, ph + dil, ph + 2 dil, ...  The three taps of consecutive outputs are then
// consecutive elements of that sequence, so every input element is loaded, PReLU'd and
// normalised ONCE and reused from a 3-deep register window (the row-per-warp kernel does
// that work three times).  Loads run 8 rows ahead of use.  HBM-bound by design: 4*H bytes
// read and 4*H written per frame (+4*H for d in training), 2/32 halo re-reads per segment --
// so the arithmetic is written with the packed fp32x2 instructions of sm_100 (fma.rn.f32x2
// ...): the scalar version of this loop issued 34 instructions per element and was
// issue-bound at half the memory roofline.
#pragma once
#include "common.cuh"
#include "ptx.cuh"
#include "f32x2.cuh"

namespace ctn {

#ifndef CTN_DM_AHEAD
#define CTN_DM_AHEAD 8
#endif
constexpr int kDmAhead = CTN_DM_AHEAD;   // rows loaded ahead of use (a power of two)

struct DwMarchParams {
    int K, H, dil, c0, in_P, NS, items, out_P;
    const float* in;
    const float* prelu_in;
    const float* gamma_in;
    const float* beta_in;
    const float2* in_stat;
    const float* wd;   // [H, 3]
    const float* prelu_out;
    float* out_pre;
    __nv_bfloat16* out_hi;
    __nv_bfloat16* out_lo;
    float2* out_stat;
};

__global__ void __launch_bounds__(kDmWarps * 32)
dwconv_march_kernel(const DwMarchParams p) {
    ptx::grid_dep_sync();   // PDL: predecessor complete; let the next kernel be scheduled
    __shared__ double fin_s[2 * kDmWarps];
    __shared__ float red_s[2 * kDmWarps];
    const int m = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int H = p.H, d = p.dil;
    const float2 st = gln_finalize_block(p.in_stat + static_cast<long long>(m) * p.in_P, p.in_P,
                                         static_cast<double>(p.K) * H, fin_s);
    const float mean = st.x, rstd = st.y;
    const float a1 = __ldg(p.prelu_in), a2 = __ldg(p.prelu_out);
    f2_t ls[2] = {0ull, 0ull}, lss[2] = {0ull, 0ull};   // packed (0.f, 0.f)

    const int item = blockIdx.x * kDmWarps + warp;
    const int CG = (H + 127) / 128;
    const int cg = item % CG;
    const int rest = item / CG;
    const int ph = rest % d, seg = rest / d;
    const int c = cg * 128 + lane * 4;
    const int J = ph < p.K ? (p.K - ph + d - 1) / d : 0;   // frames of this phase
    const int j0 = seg * kDmSeg;
    const int n = min(kDmSeg, J - j0);                      // outputs of this segment
    if (item < p.items && c < H && n > 0) {
        const f2_t neg1 = f2_pack(-1.f, -1.f);
        const f2_t a2p = f2_pack(a2, a2);
        // per-channel constants: n1 = A * max(a, 0) + (A * a1) * min(a, 0) + B   (PReLU + gLN)
        f2_t A[2], Aa[2], Bc[2], w0[2], w1[2], w2[2];
        {
            const float4 g = *reinterpret_cast<const float4*>(p.gamma_in + c);
            const float4 b = *reinterpret_cast<const float4*>(p.beta_in + c);
            const float Af[4] = {g.x * rstd, g.y * rstd, g.z * rstd, g.w * rstd};
            const float Bf[4] = {b.x - Af[0] * mean, b.y - Af[1] * mean, b.z - Af[2] * mean, b.w - Af[3] * mean};
            const float4 t0 = *reinterpret_cast<const float4*>(p.wd + c * 3);
            const float4 t1 = *reinterpret_cast<const float4*>(p.wd + c * 3 + 4);
            const float4 t2 = *reinterpret_cast<const float4*>(p.wd + c * 3 + 8);
            const float f[12] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w, t2.x, t2.y, t2.z, t2.w};
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                A[i] = f2_pack(Af[2 * i], Af[2 * i + 1]);
                Aa[i] = f2_pack(Af[2 * i] * a1, Af[2 * i + 1] * a1);
                Bc[i] = f2_pack(Bf[2 * i], Bf[2 * i + 1]);
                w0[i] = f2_pack(f[6 * i], f[6 * i + 3]);
                w1[i] = f2_pack(f[6 * i + 1], f[6 * i + 4]);
                w2[i] = f2_pack(f[6 * i + 2], f[6 * i + 5]);
            }
        }
        const long long ubase = static_cast<long long>(m) * p.K;
        const int r0 = j0 - p.c0;   // sequence index of the first window element
        const int nrows = n + 2;
        const long long strd = static_cast<long long>(d) * H;            // elements between rows of the phase
        const long long in0 = (ubase + ph + static_cast<long long>(r0) * d) * H + c;   // may point before the buffer
        auto valid = [&](int s) { return s < nrows && r0 + s >= 0 && r0 + s < J; };
        auto load = [&](int s) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (valid(s)) v = *reinterpret_cast<const float4*>(p.in + in0 + s * strd);
            return v;
        };
        // PReLU + norm of one loaded row; rows outside the phase are the convolution's zero padding
        auto xform = [&](const float4& v, bool ok, f2_t (&o)[2]) {
            if (ok) {
                const f2_t in2[2] = {f2_pack(v.x, v.y), f2_pack(v.z, v.w)};
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    f2_t mx, mn;
                    f2_split_sign(in2[i], neg1, mx, mn);
                    o[i] = f2_fma(A[i], mx, f2_fma(Aa[i], mn, Bc[i]));
                }
            } else {
                o[0] = 0ull, o[1] = 0ull;
            }
        };
        float4 q[kDmAhead];
#pragma unroll
        for (int s = 0; s < kDmAhead; ++s) q[s] = load(s);
        f2_t x0[2], x1[2], x2[2];
        xform(q[0], valid(0), x0);
        q[0] = load(kDmAhead);
        xform(q[1], valid(1), x1);
        q[1] = load(kDmAhead + 1);
        long long orow = (ubase + ph + static_cast<long long>(j0) * d) * H + c;
        for (int o = 0; o < n; o += kDmAhead) {
#pragma unroll
            for (int u = 0; u < kDmAhead; ++u) {
                if (o + u < n) {
                    const int s = o + u + 2;
                    xform(q[(u + 2) & (kDmAhead - 1)], valid(s), x2);
                    q[(u + 2) & (kDmAhead - 1)] = load(s + kDmAhead);
                    uint2 h, l;
                    f2_t acc[2];
#pragma unroll
                    for (int i = 0; i < 2; ++i) {
                        acc[i] = f2_fma(w2[i], x2[i], f2_fma(w1[i], x1[i], f2_mul(w0[i], x0[i])));
                        f2_t mx, mn;
                        f2_split_sign(acc[i], neg1, mx, mn);
                        const f2_t qq = f2_fma(a2p, mn, mx);          // PReLU
                        ls[i] = f2_add(ls[i], qq);
                        lss[i] = f2_fma(qq, qq, lss[i]);
                        f2_split_bf16(qq, neg1, i == 0 ? h.x : h.y, i == 0 ? l.x : l.y);
                        x0[i] = x1[i], x1[i] = x2[i];
                    }
                    if (p.out_pre) {
                        float4 dv;
                        f2_unpack(acc[0], dv.x, dv.y);
                        f2_unpack(acc[1], dv.z, dv.w);
                        *reinterpret_cast<float4*>(p.out_pre + orow) = dv;
                    }
                    *reinterpret_cast<uint2*>(p.out_hi + orow) = h;
                    *reinterpret_cast<uint2*>(p.out_lo + orow) = l;
                    orow += strd;
                }
            }
        }
    }
    float s0, s1, s2, s3, t0, t1, t2, t3;
    f2_unpack(ls[0], s0, s1), f2_unpack(ls[1], s2, s3);
    f2_unpack(lss[0], t0, t1), f2_unpack(lss[1], t2, t3);
    const float wls = warp_sum((s0 + s1) + (s2 + s3));
    const float wlss = warp_sum((t0 + t1) + (t2 + t3));
    if (lane == 0) red_s[warp] = wls, red_s[kDmWarps + warp] = wlss;
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.f, ss = 0.f;
        for (int i = 0; i < kDmWarps; ++i) s += red_s[i], ss += red_s[kDmWarps + i];
        p.out_stat[static_cast<long long>(m) * p.out_P + blockIdx.x] = make_float2(s, ss);
    }
}

// segments per phase and warp work items per utterance
inline void dw_march_geometry(int K, int H, int dil, int* NS, int* items) {
    const int Jmax = (K + dil - 1) / dil;
    *NS = (Jmax + kDmSeg - 1) / kDmSeg;
    *items = ((H + 127) / 128) * dil * (*NS);
}

}  // namespace ctn
