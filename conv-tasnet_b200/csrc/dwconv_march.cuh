// Fast path of ctn_dwconv_fwd for gLN, P = 3, H % 4 == 0 (the paper configurations).
// A warp owns 128 channels (one float4 per lane) of one dilation phase and marches along the
// frames k = ph, ph + dil, ph + 2 dil, ...  The three taps of consecutive outputs are then
// consecutive elements of that sequence, so every input element is loaded, PReLU'd and
// normalised ONCE and reused from a 3-deep register window (the row-per-warp kernel does
// that work three times).  Loads run 4 rows ahead of use.  HBM-bound: 4*H bytes read and
// 4*H written per frame (+4*H for d in training), 2/32 halo re-reads per segment.
#pragma once
#include "common.cuh"

namespace ctn {

constexpr int kDmSeg = 32;    // outputs per warp segment
constexpr int kDmWarps = 8;
constexpr int kDmAhead = 4;   // rows loaded ahead of use

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
    __shared__ float st_s[2];
    __shared__ float red_s[2 * kDmWarps];
    const int m = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int H = p.H, d = p.dil;
    if (warp == 0) {
        const float2 st = gln_finalize_warp(p.in_stat + static_cast<long long>(m) * p.in_P, p.in_P,
                                            static_cast<double>(p.K) * H);
        if (lane == 0) st_s[0] = st.x, st_s[1] = st.y;
    }
    __syncthreads();
    const float mean = st_s[0], rstd = st_s[1];
    const float a1 = __ldg(p.prelu_in), a2 = __ldg(p.prelu_out);
    float ls = 0.f, lss = 0.f;

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
        // per-channel constants: n1 = A * prelu(a) + B
        const float4 g = *reinterpret_cast<const float4*>(p.gamma_in + c);
        const float4 b = *reinterpret_cast<const float4*>(p.beta_in + c);
        const float A[4] = {g.x * rstd, g.y * rstd, g.z * rstd, g.w * rstd};
        const float Bc[4] = {b.x - A[0] * mean, b.y - A[1] * mean, b.z - A[2] * mean, b.w - A[3] * mean};
        float w[4][3];
        {
            const float4 w0 = *reinterpret_cast<const float4*>(p.wd + c * 3);
            const float4 w1 = *reinterpret_cast<const float4*>(p.wd + c * 3 + 4);
            const float4 w2 = *reinterpret_cast<const float4*>(p.wd + c * 3 + 8);
            const float f[12] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w, w2.x, w2.y, w2.z, w2.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int t = 0; t < 3; ++t) w[i][t] = f[i * 3 + t];
        }
        const long long ubase = static_cast<long long>(m) * p.K;
        const int r0 = j0 - p.c0;   // sequence index of the first window element
        const int nrows = n + 2;
        auto valid = [&](int s) { return s < nrows && r0 + s >= 0 && r0 + s < J; };
        auto load = [&](int s) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (valid(s)) v = *reinterpret_cast<const float4*>(p.in + (ubase + ph + static_cast<long long>(r0 + s) * d) * H + c);
            return v;
        };
        auto xform = [&](const float4& v, bool ok, float (&o)[4]) {
            const float r[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) o[i] = ok ? fmaf(A[i], prelu(r[i], a1), Bc[i]) : 0.f;  // zero padding after the norm
        };
        float4 q[kDmAhead];
#pragma unroll
        for (int s = 0; s < kDmAhead; ++s) q[s] = load(s);
        float x0[4], x1[4], x2[4];
        xform(q[0], valid(0), x0);
        q[0] = load(4);
        xform(q[1], valid(1), x1);
        q[1] = load(5);
        for (int o = 0; o < n; o += kDmAhead) {
#pragma unroll
            for (int u = 0; u < kDmAhead; ++u) {
                if (o + u < n) {
                    const int s = o + u + 2;
                    xform(q[(u + 2) & 3], valid(s), x2);
                    q[(u + 2) & 3] = load(s + kDmAhead);
                    const long long row = ubase + ph + static_cast<long long>(j0 + o + u) * d;
                    float acc[4], qq[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        acc[i] = fmaf(w[i][2], x2[i], fmaf(w[i][1], x1[i], w[i][0] * x0[i]));
                        qq[i] = prelu(acc[i], a2);
                        ls += qq[i];
                        lss = fmaf(qq[i], qq[i], lss);
                        x0[i] = x1[i], x1[i] = x2[i];
                    }
                    if (p.out_pre) *reinterpret_cast<float4*>(p.out_pre + row * H + c) = make_float4(acc[0], acc[1], acc[2], acc[3]);
                    uint2 h, l;
                    split_pack2(qq[0], qq[1], h.x, l.x);
                    split_pack2(qq[2], qq[3], h.y, l.y);
                    *reinterpret_cast<uint2*>(p.out_hi + row * H + c) = h;
                    *reinterpret_cast<uint2*>(p.out_lo + row * H + c) = l;
                }
            }
        }
    }
    ls = warp_sum(ls);
    lss = warp_sum(lss);
    if (lane == 0) red_s[warp] = ls, red_s[kDmWarps + warp] = lss;
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
