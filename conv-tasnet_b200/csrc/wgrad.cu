// Weight gradients of the 1x1 convolutions: G[i, j] = sum_r x[r, i] * y[r, j], the reduction
// running over all M*K frames of the batch (autograd's convolution_backward for the weights
// of conv_tasnet.py:169,185,217,256).
//
// Both operands are frames-major [R, C] bf16 (hi, lo) pairs, i.e. MN-major for the tensor
// core (the reduction dimension is the strided one).  tcgen05 reads them straight from
// 128B-swizzled TMA tiles through MN-major shared-memory descriptors, so no transposed copy of
// an activation is ever made.  The row range is split across CTAs (split-K); every CTA writes
// its fp32 [128 x BN] partial tile, and ctn_reduce_partials sums the splits in a fixed order.
// One k-step (16 frames) issues hi*hi + lo*hi + hi*lo like the forward GEMM.
#include <cuda.h>

#include "common.cuh"
#include "ptx.cuh"

namespace ctn {

int make_tmap_bf16_box(CUtensorMap* tm, const void* ptr, long long rows, int cols, int box_cols, int box_rows);

constexpr int kWgBM = 128;   // output rows (i) per tile = TMEM lanes
constexpr int kWgBR = 64;    // frames per k-block
constexpr int kWgThreads = 192;
constexpr int kAtomBytes = kWgBR * 128;  // one MN atom: 64 channels (128 B) x 64 frames

template <int BN>
struct WgCfg {
    static constexpr int kStages = (BN == 256) ? 2 : 3;
    static constexpr int kXBytes = 2 * kAtomBytes;            // 128 channels, one of hi / lo
    static constexpr int kYBytes = (BN / 64) * kAtomBytes;
    static constexpr int kStageBytes = 2 * kXBytes + 2 * kYBytes;
    static constexpr int kSmemBytes = kStages * kStageBytes + 256;
};

template <int BN>
__global__ void __launch_bounds__(kWgThreads, 1)
wgrad_tc_kernel(const __grid_constant__ CUtensorMap tm_x_hi, const __grid_constant__ CUtensorMap tm_x_lo,
                const __grid_constant__ CUtensorMap tm_y_hi, const __grid_constant__ CUtensorMap tm_y_lo,
                float* __restrict__ partials, int Ci, int Cj, int j_tiles, int rows_per_split, long long R) {
    using Cfg = WgCfg<BN>;
    extern __shared__ __align__(1024) uint8_t smem[];
    if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
    uint64_t* empty_bar = full_bar + Cfg::kStages;
    uint64_t* done_bar = empty_bar + Cfg::kStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x, split = blockIdx.y;
    const int it = tile / j_tiles, jt = tile % j_tiles;
    const int i0 = it * kWgBM, j0 = jt * BN;
    const long long r_begin = static_cast<long long>(split) * rows_per_split;
    long long r_end = r_begin + rows_per_split;
    if (r_end > R) r_end = R;
    const int nkb = r_end > r_begin ? static_cast<int>((r_end - r_begin + kWgBR - 1) / kWgBR) : 0;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tm_x_hi);
        ptx::prefetch_tmap(&tm_x_lo);
        ptx::prefetch_tmap(&tm_y_hi);
        ptx::prefetch_tmap(&tm_y_lo);
        for (int i = 0; i < Cfg::kStages; ++i) {
            ptx::mbar_init(full_bar + i, 1);
            ptx::mbar_init(empty_bar + i, 1);
        }
        ptx::mbar_init(done_bar, 1);
        ptx::fence_mbar_init();
    } else if (warp == 1) {
        ptx::tmem_alloc(tmem_slot, BN);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = 0; kb < nkb; ++kb) {
                ptx::mbar_wait(empty_bar + stage, phase ^ 1);
                uint8_t* st = smem + stage * Cfg::kStageBytes;
                const int r0 = static_cast<int>(r_begin) + kb * kWgBR;
                // Rows at or beyond r_end that belong to the NEXT split must not be counted twice:
                // splits are multiples of 64 rows, so a k-block never straddles two splits; rows
                // beyond R are zero-filled by TMA.
                ptx::mbar_expect_tx(full_bar + stage, Cfg::kStageBytes);
#pragma unroll
                for (int a = 0; a < 2; ++a) {
                    ptx::tma_load_2d(st + a * kAtomBytes, &tm_x_hi, full_bar + stage, i0 + a * 64, r0);
                    ptx::tma_load_2d(st + Cfg::kXBytes + a * kAtomBytes, &tm_x_lo, full_bar + stage, i0 + a * 64, r0);
                }
#pragma unroll
                for (int a = 0; a < BN / 64; ++a) {
                    ptx::tma_load_2d(st + 2 * Cfg::kXBytes + a * kAtomBytes, &tm_y_hi, full_bar + stage, j0 + a * 64, r0);
                    ptx::tma_load_2d(st + 2 * Cfg::kXBytes + Cfg::kYBytes + a * kAtomBytes, &tm_y_lo, full_bar + stage,
                                     j0 + a * 64, r0);
                }
                if (++stage == Cfg::kStages) stage = 0, phase ^= 1;
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = ptx::make_idesc_bf16(kWgBM, BN, 1, 1);  // both operands MN-major
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = 0; kb < nkb; ++kb) {
                ptx::mbar_wait(full_bar + stage, phase);
                ptx::tc_fence_after();
                const uint32_t sa = ptx::smem_u32(smem + stage * Cfg::kStageBytes);
                // MN-major, 128B swizzle: LBO = stride between 64-channel atoms, SBO = 8 frames x 128 B
                const uint64_t x_hi = ptx::make_smem_desc(sa, kAtomBytes, 1024, ptx::kLayoutSwizzle128B);
                const uint64_t x_lo = ptx::make_smem_desc(sa + Cfg::kXBytes, kAtomBytes, 1024, ptx::kLayoutSwizzle128B);
                const uint64_t y_hi = ptx::make_smem_desc(sa + 2 * Cfg::kXBytes, kAtomBytes, 1024, ptx::kLayoutSwizzle128B);
                const uint64_t y_lo =
                    ptx::make_smem_desc(sa + 2 * Cfg::kXBytes + Cfg::kYBytes, kAtomBytes, 1024, ptx::kLayoutSwizzle128B);
#pragma unroll
                for (int k = 0; k < kWgBR / 16; ++k) {
                    const uint64_t adv = static_cast<uint64_t>((k * 16 * 128) >> 4);  // 16 frames x 128 B
                    ptx::umma_bf16(tmem_base, x_hi + adv, y_hi + adv, idesc, (kb | k) != 0);
                    ptx::umma_bf16(tmem_base, x_lo + adv, y_hi + adv, idesc, 1);
                    ptx::umma_bf16(tmem_base, x_hi + adv, y_lo + adv, idesc, 1);
                }
                ptx::umma_commit(empty_bar + stage);
                if (kb == nkb - 1) ptx::umma_commit(done_bar);
                if (++stage == Cfg::kStages) stage = 0, phase ^= 1;
            }
        }
    } else {
        const int ew = warp & 3;
        const int i = i0 + ew * 32 + lane;
        float* dst = partials + (static_cast<long long>(split) * Ci + i) * Cj + j0;
        if (nkb > 0) {
            ptx::mbar_wait(done_bar, 0);
            ptx::tc_fence_after();
        }
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
            if (j0 + c * 32 >= Cj) break;
            float v[32];
            if (nkb > 0) {
                uint32_t r[32];
                ptx::tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + c * 32, r);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = 0.f;
            }
            if (i < Ci) {
                const int ncols = min(32, Cj - (j0 + c * 32));  // multiple of 8
#pragma unroll
                for (int j = 0; j < 32; j += 4)
                    if (j < ncols)
                        *reinterpret_cast<float4*>(dst + c * 32 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
            }
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, BN);
    }
}

// any shape: one thread per output element and split
__global__ void wgrad_simt_kernel(const __nv_bfloat16* __restrict__ x_hi, const __nv_bfloat16* __restrict__ x_lo,
                                  const __nv_bfloat16* __restrict__ y_hi, const __nv_bfloat16* __restrict__ y_lo,
                                  float* __restrict__ partials, int Ci, int Cj, int rows_per_split, long long R) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= Ci * Cj) return;
    const int i = e / Cj, j = e - i * Cj;
    const long long r0 = static_cast<long long>(blockIdx.y) * rows_per_split;
    long long r1 = r0 + rows_per_split;
    if (r1 > R) r1 = R;
    float acc = 0.f;
    for (long long r = r0; r < r1; ++r) {
#if CTN_OPERAND_FP16
        const float xv = op16_value(x_hi[r * Ci + i]) + op16_value(x_lo[r * Ci + i]);   // the weight gradients keep 3 terms
        const float yv = op16_value(y_hi[r * Cj + j]) + op16_value(y_lo[r * Cj + j]);
#else
        const float xv = __bfloat162float(x_hi[r * Ci + i]) + __bfloat162float(x_lo[r * Ci + i]);
        const float yv = __bfloat162float(y_hi[r * Cj + j]) + __bfloat162float(y_lo[r * Cj + j]);
#endif
        acc = fmaf(xv, yv, acc);
    }
    partials[static_cast<long long>(blockIdx.y) * Ci * Cj + e] = acc;
}

static int wgrad_bn(int Cj) { return Cj % 256 == 0 ? 256 : 128; }

static void wgrad_geometry(long long R, int Ci, int Cj, int* splits, int* rows_per_split) {
    const int tiles = ceil_div(Ci, kWgBM) * ceil_div(Cj, wgrad_bn(Cj));
    long long kblocks = (R + kWgBR - 1) / kWgBR;
    int S = sm_count() / (tiles > 0 ? tiles : 1);
    if (S < 1) S = 1;
    if (S > kblocks) S = static_cast<int>(kblocks);
    const long long kb_per = (kblocks + S - 1) / S;
    S = static_cast<int>((kblocks + kb_per - 1) / kb_per);
    *splits = S;
    *rows_per_split = static_cast<int>(kb_per * kWgBR);
}

template <int BN>
static int launch_wgrad(const void* x_hi, const void* x_lo, const void* y_hi, const void* y_lo, long long R, int Ci,
                        int Cj, float* partials, int S, int rps, cudaStream_t stream) {
    using Cfg = WgCfg<BN>;
    CUtensorMap tx_hi, tx_lo, ty_hi, ty_lo;
    int rc;
    if ((rc = make_tmap_bf16_box(&tx_hi, x_hi, R, Ci, 64, kWgBR)) != CTN_OK) return rc;
    if ((rc = make_tmap_bf16_box(&tx_lo, x_lo, R, Ci, 64, kWgBR)) != CTN_OK) return rc;
    if ((rc = make_tmap_bf16_box(&ty_hi, y_hi, R, Cj, 64, kWgBR)) != CTN_OK) return rc;
    if ((rc = make_tmap_bf16_box(&ty_lo, y_lo, R, Cj, 64, kWgBR)) != CTN_OK) return rc;
    static bool attr_done[kMaxDevices] = {};
    bool& attr_set = attr_done[device_slot()];
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(wgrad_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Cfg::kSmemBytes);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(wgrad_tc_kernel)");
        attr_set = true;
    }
    const int j_tiles = ceil_div(Cj, BN);
    dim3 grid(ceil_div(Ci, kWgBM) * j_tiles, S);
    wgrad_tc_kernel<BN><<<grid, kWgThreads, Cfg::kSmemBytes, stream>>>(tx_hi, tx_lo, ty_hi, ty_lo, partials, Ci, Cj,
                                                                       j_tiles, rps, R);
    CTN_LAUNCH_CHECK("wgrad_tc_kernel");
    return CTN_OK;
}

}  // namespace ctn

using namespace ctn;

extern "C" int ctn_wgrad_splits(long long R, int Ci, int Cj) {
    if (R <= 0 || Ci <= 0 || Cj <= 0) return 0;
    int S, rps;
    wgrad_geometry(R, Ci, Cj, &S, &rps);
    return S;
}

extern "C" int ctn_wgrad(const void* x_hi, const void* x_lo, const void* y_hi, const void* y_lo, long long R, int Ci,
                         int Cj, float* partials, int force_simt, void* stream_) {
    CTN_REQUIRE(x_hi && x_lo && y_hi && y_lo && partials, "ctn_wgrad: NULL pointer");
    CTN_REQUIRE(R > 0 && R < (1ll << 31) && Ci > 0 && Cj > 0, "ctn_wgrad: bad shape");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    int S, rps;
    wgrad_geometry(R, Ci, Cj, &S, &rps);
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    const bool tc = !force_simt && Ci % 8 == 0 && Cj % 8 == 0 && al16(x_hi) && al16(x_lo) && al16(y_hi) &&
                    al16(y_lo) && al16(partials);
    if (tc) {
        if (wgrad_bn(Cj) == 256) return launch_wgrad<256>(x_hi, x_lo, y_hi, y_lo, R, Ci, Cj, partials, S, rps, stream);
        return launch_wgrad<128>(x_hi, x_lo, y_hi, y_lo, R, Ci, Cj, partials, S, rps, stream);
    }
    dim3 grid(ceil_div(Ci * Cj, 256), S);
    wgrad_simt_kernel<<<grid, 256, 0, stream>>>(
        static_cast<const __nv_bfloat16*>(x_hi), static_cast<const __nv_bfloat16*>(x_lo),
        static_cast<const __nv_bfloat16*>(y_hi), static_cast<const __nv_bfloat16*>(y_lo), partials, Ci, Cj, rps, R);
    CTN_LAUNCH_CHECK("wgrad_simt_kernel");
    return CTN_OK;
}
