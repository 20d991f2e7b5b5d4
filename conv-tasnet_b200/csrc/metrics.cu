// SI-SNR improvement of a batch on the device (evaluate.py:108-144, called per utterance at
// evaluate.py:66-80 on host numpy arrays after remove_pad).
//
// For utterance b with n = length_b valid samples, source c:
//   sisnr(ref, out): zero-mean both over the n samples (:134-135), ref_energy = sum ref~^2 + eps (:136),
//   proj = <ref~, out~> ref~ / ref_energy (:137), noise = out~ - proj (:138),
//   10 log10(sum proj^2 / (sum noise^2 + eps) + eps) (:139-140)
//   sisnri_b = mean_c [ sisnr(ref_c, est_c) - sisnr(ref_c, mix) ] (:114-124; the reference hard-codes c = 0, 1)
// Everything reduces to eight sums per (b, c) taken in ONE pass over ref_c, est_c and mix
// (HBM-bound, reads 4*(2C+C)*n bytes per utterance; the mixture is re-read per source and
// stays in L2).  Products and sums are formed in fp64 (exact products of fp32 inputs, fixed
// reduction order), so the closed forms below do not lose the noise energy to cancellation
// even at high SNR:
//   R = sum ref~^2 = Srr - n mr^2,  O = sum out~^2 = Soo - n mo^2,  D = <ref~, out~> = Sro - n mr mo
//   a = D / (R + eps);  sum proj^2 = a^2 R;  sum noise^2 = O - 2 a D + a^2 R
#include "common.cuh"

namespace ctn {

constexpr int kMetThreads = 512;

__device__ __forceinline__ double sisnr_closed(double n, double Sr, double Srr, double So, double Soo, double Sro) {
    const double eps = 1e-8;
    const double mr = Sr / n, mo = So / n;
    const double R = Srr - n * mr * mr;
    const double O = Soo - n * mo * mo;
    const double D = Sro - n * mr * mo;
    const double a = D / (R + eps);
    const double proj = a * a * R;
    double noise = O - 2.0 * a * D + a * a * R;
    noise = noise > 0.0 ? noise : 0.0;
    return 10.0 * log10(proj / (noise + eps) + eps);
}

// grid (C, B); one CTA per (utterance, source)
__global__ void __launch_bounds__(kMetThreads)
sisnri_pair_kernel(const float* __restrict__ ref, const float* __restrict__ est, const float* __restrict__ mix,
                   const long long* __restrict__ lengths, int C, int T, float* __restrict__ sisnr) {
    const int c = blockIdx.x, b = blockIdx.y;
    long long len = lengths ? lengths[b] : T;
    len = len < 0 ? 0 : (len > T ? T : len);
    const float* r = ref + (static_cast<long long>(b) * C + c) * T;
    const float* e = est + (static_cast<long long>(b) * C + c) * T;
    const float* m = mix + static_cast<long long>(b) * T;
    double acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.0;
    for (int t = threadIdx.x; t < len; t += kMetThreads) {
        const double rv = r[t], ev = e[t], mv = m[t];
        acc[0] += rv;
        acc[1] += ev;
        acc[2] += mv;
        acc[3] = fma(rv, rv, acc[3]);
        acc[4] = fma(ev, ev, acc[4]);
        acc[5] = fma(mv, mv, acc[5]);
        acc[6] = fma(rv, ev, acc[6]);
        acc[7] = fma(rv, mv, acc[7]);
    }
    __shared__ double red[kMetThreads / 32][8];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const double v = warp_sum(acc[i]);
        if (lane == 0) red[warp][i] = v;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double s[8];
        for (int i = 0; i < 8; ++i) {
            s[i] = 0.0;
            for (int w = 0; w < kMetThreads / 32; ++w) s[i] += red[w][i];
        }
        const double n = static_cast<double>(len);
        float* o = sisnr + (static_cast<long long>(b) * C + c) * 2;
        o[0] = static_cast<float>(sisnr_closed(n, s[0], s[3], s[1], s[4], s[6]));  // estimate
        o[1] = static_cast<float>(sisnr_closed(n, s[0], s[3], s[2], s[5], s[7]));  // mixture (the baseline)
    }
}

__global__ void sisnri_mean_kernel(const float* __restrict__ sisnr, int B, int C, float* __restrict__ sisnri) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    double s = 0.0;
    for (int c = 0; c < C; ++c)
        s += static_cast<double>(sisnr[(static_cast<long long>(b) * C + c) * 2]) -
             static_cast<double>(sisnr[(static_cast<long long>(b) * C + c) * 2 + 1]);
    sisnri[b] = static_cast<float>(s / C);
}

}  // namespace ctn

extern "C" int ctn_sisnri(const float* ref, const float* est, const float* mix, const long long* lengths, int B, int C,
                          int T, float* sisnr, float* sisnri, void* stream) {
    using namespace ctn;
    CTN_REQUIRE(ref && est && mix && sisnr, "ctn_sisnri: NULL pointer");
    CTN_REQUIRE(B > 0 && C > 0 && T > 0, "ctn_sisnri: bad shape B=%d C=%d T=%d", B, C, T);
    CTN_REQUIRE(B <= 65535, "ctn_sisnri: B=%d above the grid limit", B);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    sisnri_pair_kernel<<<dim3(C, B), kMetThreads, 0, st>>>(ref, est, mix, lengths, C, T, sisnr);
    CTN_LAUNCH_CHECK("sisnri_pair_kernel");
    if (sisnri) {
        sisnri_mean_kernel<<<ceil_div(B, 128), 128, 0, st>>>(sisnr, B, C, sisnri);
        CTN_LAUNCH_CHECK("sisnri_mean_kernel");
    }
    return CTN_OK;
}
