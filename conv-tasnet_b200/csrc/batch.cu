// Batch assembly and its inverse on the device.
//
// The reference pads ragged utterances on the host (data.py:295-301 pad_list, called from the
// collate functions at data.py:146-152 and :222-225), permutes the sources [B, T, C] -> [B, C, T]
// (:153-154), uploads the padded tensors (solver.py:173-176) and, after separation, slices the padding
// off again with one device-to-host copy per utterance (utils.py:49-66 remove_pad, called at
// evaluate.py:62-65 and separate.py:75-76).
//
// Here the valid samples travel packed (one upload / one download of exactly the valid bytes) and the
// padded layout the model needs - zeros included, since gLN statistics count padded frames
// (conv_tasnet.py:352-353) - is produced and undone by two copy kernels:
//   pad_batch:   mix_pad[b, t]    = t < len_b ? mix_packed[off_b + t]           : pad
//                src_pad[b, c, t] = t < len_b ? src_packed[(off_b + t) * C + c] : pad   (T x C per utterance)
//   remove_pad:  packed[C * off_b + c * len_b + t] = in[b, c, t],  t < len_b
// off = exclusive prefix sum of the lengths (offsets[B] = total).  Pure HBM-bound copies:
// 4 * (1 + C) * (sum len + B * T) bytes for pad_batch, 8 * C * sum len for remove_pad.
#include "common.cuh"

namespace ctn {

constexpr int kBatchThreads = 256;
constexpr int kBatchPerThread = 4;
constexpr int kBatchSpan = kBatchThreads * kBatchPerThread;

// grid (ceil(T / span), B, 1 + C): z = 0 is the mixture row, z = 1 + c source c
__global__ void __launch_bounds__(kBatchThreads)
pad_batch_kernel(const float* __restrict__ mix_packed, const float* __restrict__ src_packed,
                 const long long* __restrict__ offsets, int C, int T, float pad, float* __restrict__ mix_pad,
                 float* __restrict__ src_pad) {
    const int b = blockIdx.y, z = blockIdx.z;
    const long long off = offsets[b];
    long long len = offsets[b + 1] - off;
    len = len > T ? T : len;
    const int t0 = blockIdx.x * kBatchSpan + threadIdx.x;
    if (z == 0) {
        if (!mix_pad) return;
        float* dst = mix_pad + static_cast<long long>(b) * T;
#pragma unroll
        for (int i = 0; i < kBatchPerThread; ++i) {
            const int t = t0 + i * kBatchThreads;
            if (t < T) dst[t] = t < len ? mix_packed[off + t] : pad;
        }
    } else {
        const int c = z - 1;
        float* dst = src_pad + (static_cast<long long>(b) * C + c) * T;
#pragma unroll
        for (int i = 0; i < kBatchPerThread; ++i) {
            const int t = t0 + i * kBatchThreads;
            // stride-C reads: the C planes of one (b, t) span are read by C CTAs back to back and hit in L2
            if (t < T) dst[t] = t < len ? src_packed[(off + t) * C + c] : pad;
        }
    }
}

// grid (ceil(T / span), B, C)
__global__ void __launch_bounds__(kBatchThreads)
remove_pad_kernel(const float* __restrict__ in, const long long* __restrict__ offsets, int C, int T,
                  float* __restrict__ packed) {
    const int b = blockIdx.y, c = blockIdx.z;
    const long long off = offsets[b];
    long long len = offsets[b + 1] - off;
    len = len > T ? T : len;
    const float* src = in + (static_cast<long long>(b) * C + c) * T;
    float* dst = packed + off * C + c * len;
    const int t0 = blockIdx.x * kBatchSpan + threadIdx.x;
#pragma unroll
    for (int i = 0; i < kBatchPerThread; ++i) {
        const int t = t0 + i * kBatchThreads;
        if (t < len) dst[t] = src[t];
    }
}

}  // namespace ctn

extern "C" int ctn_pad_batch(const float* mix_packed, const float* src_packed, const long long* offsets, int B, int C,
                             int T, float pad_value, float* mix_pad, float* src_pad, void* stream) {
    using namespace ctn;
    CTN_REQUIRE(offsets, "ctn_pad_batch: NULL offsets");
    CTN_REQUIRE((mix_packed == nullptr) == (mix_pad == nullptr), "ctn_pad_batch: mixture in/out must be given together");
    CTN_REQUIRE((src_packed == nullptr) == (src_pad == nullptr), "ctn_pad_batch: sources in/out must be given together");
    CTN_REQUIRE(mix_pad || src_pad, "ctn_pad_batch: nothing to do");
    CTN_REQUIRE(B >= 0 && T >= 0 && C >= 0 && C <= 65534, "ctn_pad_batch: bad shape B=%d C=%d T=%d", B, C, T);
    CTN_REQUIRE(B <= 65535, "ctn_pad_batch: at most 65535 utterances per call (B=%d)", B);
    if (B == 0 || T == 0) return CTN_OK;
    const int planes = src_pad ? 1 + C : 1;
    dim3 grid(ceil_div(T, kBatchSpan), B, planes);
    pad_batch_kernel<<<grid, kBatchThreads, 0, static_cast<cudaStream_t>(stream)>>>(mix_packed, src_packed, offsets, C, T,
                                                                                   pad_value, mix_pad, src_pad);
    CTN_LAUNCH_CHECK("pad_batch_kernel");
    return CTN_OK;
}

extern "C" int ctn_remove_pad(const float* in, const long long* offsets, int B, int C, int T, float* packed,
                              void* stream) {
    using namespace ctn;
    CTN_REQUIRE(in && offsets && packed, "ctn_remove_pad: NULL argument");
    CTN_REQUIRE(B >= 0 && T >= 0 && C >= 1 && C <= 65535, "ctn_remove_pad: bad shape B=%d C=%d T=%d", B, C, T);
    CTN_REQUIRE(B <= 65535, "ctn_remove_pad: at most 65535 utterances per call (B=%d)", B);
    if (B == 0 || T == 0) return CTN_OK;
    dim3 grid(ceil_div(T, kBatchSpan), B, C);
    remove_pad_kernel<<<grid, kBatchThreads, 0, static_cast<cudaStream_t>(stream)>>>(in, offsets, C, T, packed);
    CTN_LAUNCH_CHECK("remove_pad_kernel");
    return CTN_OK;
}
