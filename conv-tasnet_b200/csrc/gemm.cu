// 1x1 convolutions of the separator as one GEMM over all frames of the batch:
//     out[r, n] = sum_k a[r, k] * w[n, k]      r = m*K + frame, both operands K-major.
// Reference: nn.Conv1d(.., 1, bias=False) at conv_tasnet.py:169 (bottleneck), :217
// (TemporalBlock conv1x1), :256 (pointwise, + residual :237), :185 (mask conv).
//
// Tensor-core path (sm_100a): persistent, warp-specialised CTA of 6 warps
//   warp 0    TMA producer   (cp.async.bulk.tensor, 128B-swizzled K-major tiles)
//   warp 1    MMA issuer     (tcgen05.mma kind::f16, 128 x BN x 16, accumulators in TMEM)
//   warps 2-5 epilogue       (tcgen05.ld 32x32b -> registers -> fused epilogue -> global)
// Operands are bf16 (hi, lo) pairs; each k-step issues hi*hi + lo*hi + hi*lo so the
// product carries 16 mantissa bits per operand with fp32 accumulation (plain bf16
// fails the 1e-3 waveform tolerance).  Two TMEM accumulator
// stages let the MMAs of tile i+1 run under the epilogue of tile i.
//
// The epilogue fuses: the folded input norm (scale / shift / per-channel bias, see
// ctn_weight_prep), the residual add, the (hi, lo) split of the output and the
// (sum, sumsq) statistics of the next gLN / cLN (after an optional PReLU).
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "ptx.cuh"

namespace ctn {

__device__ __forceinline__ uint32_t r2u(float f) { return __float_as_uint(f); }

struct EpiParams {
    int K, Cout, groups, tiles_per_utt;
    float* out_f32;
    __nv_bfloat16* out_hi;
    __nv_bfloat16* out_lo;
    const float* resid;
    const __nv_bfloat16* resid_hi;   // residual as a (hi, lo) pair instead of fp32
    const __nv_bfloat16* resid_lo;
    const float* col_g;
    const float* col_bias;
    int in_norm;
    const float2* in_stat;
    int in_P;
    double in_count;
    int out_stat;
    float2* out_stat_buf;
    int out_P;
    const float* stat_prelu;
    // backward statistics (pointwise dgrad): see ctn_gemm_args.bw_*
    const float* bw_pre;
    const float* bw_gamma;
    const float* bw_slope;
    const float2* bw_stat;
    int bw_P;
    int store_prelu;   // the stored output is PReLU(out) (slope *stat_prelu), not only the statistics
    int dbg;   // CTN_GEMM_DBG experiment bits (0 in production): 1 skip epilogue, 2 skip stores, 4 skip MMAs
};

// One thread owns output row `grow` (valid or not) and 32 consecutive columns
// starting at n0 whose accumulators are in v[].  VEC: Cout % 8 == 0 so 4-wide
// groups are entirely inside or outside [0, Cout).
template <bool VEC>
__device__ __forceinline__ void epilogue_chunk(const EpiParams& ep, float (&v)[32], long long grow, int n0,
                                               bool valid, float s, float t, float slope, bool use_slope,
                                               float bw_mean, float bw_rstd, float& ls, float& lss) {
    if (n0 >= ep.Cout) return;
    const int ncols = min(32, ep.Cout - n0);
    // folded norm + bias
    if (ep.col_g != nullptr || ep.col_bias != nullptr || ep.in_norm != CTN_NORM_NONE) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            if (j < ncols) {
                float g = ep.col_g ? __ldg(ep.col_g + n0 + j) : 0.f;
                float b = ep.col_bias ? __ldg(ep.col_bias + n0 + j) : 0.f;
                v[j] = fmaf(s, v[j], fmaf(t, g, b));
            }
        }
    }
    if (!valid) return;
    const long long base = grow * ep.Cout + n0;
    if (ep.resid_hi != nullptr) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
#if CTN_OPERAND_FP16
            if (j < ncols) v[j] += op16_value(ep.resid_hi[base + j]) + op16_value(ep.resid_lo[base + j]);
#else
            if (j < ncols) v[j] += __bfloat162float(ep.resid_hi[base + j]) + __bfloat162float(ep.resid_lo[base + j]);
#endif
    }
    if (ep.resid != nullptr) {
        if (VEC) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                if (j < ncols) {
                    float4 r = *reinterpret_cast<const float4*>(ep.resid + base + j);
                    v[j] += r.x, v[j + 1] += r.y, v[j + 2] += r.z, v[j + 3] += r.w;
                }
            }
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
                if (j < ncols) v[j] += ep.resid[base + j];
        }
    }
    if (ep.bw_pre != nullptr) {
        // gLN backward sums: slope = PReLU slope of the saved pre-activation, (s, t) reused as (mean, rstd)
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            if (j < ncols) {
                const float qh = (prelu(ep.bw_pre[base + j], slope) - bw_mean) * bw_rstd;
                const float gd = __ldg(ep.bw_gamma + n0 + j) * v[j];
                ls += gd;
                lss = fmaf(gd, qh, lss);
            }
        }
    } else if (ep.out_stat != CTN_NORM_NONE) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            if (j < ncols) {
                float x = use_slope ? prelu(v[j], slope) : v[j];
                ls += x;
                lss = fmaf(x, x, lss);
                if (ep.store_prelu) v[j] = x;
            }
        }
    }
    if (ep.out_f32 != nullptr) {
        if (VEC) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
                if (j < ncols)
                    *reinterpret_cast<float4*>(ep.out_f32 + base + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
                if (j < ncols) ep.out_f32[base + j] = v[j];
        }
    }
    if (ep.out_hi != nullptr) {
        if (VEC) {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
                if (j < ncols) {
                    uint4 h, l;
                    split_pack2(v[j], v[j + 1], h.x, l.x);
                    split_pack2(v[j + 2], v[j + 3], h.y, l.y);
                    split_pack2(v[j + 4], v[j + 5], h.z, l.z);
                    split_pack2(v[j + 6], v[j + 7], h.w, l.w);
                    *reinterpret_cast<uint4*>(ep.out_hi + base + j) = h;
                    *reinterpret_cast<uint4*>(ep.out_lo + base + j) = l;
                }
            }
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                if (j < ncols) {
                    __nv_bfloat16 h, l;
                    split_bf16(v[j], h, l);
                    ep.out_hi[base + j] = h;
                    ep.out_lo[base + j] = l;
                }
            }
        }
    }
}

// Per-tile scale/shift of the folded input norm.  Warp-collective (gLN path).
__device__ __forceinline__ void epilogue_row_affine(const EpiParams& ep, int m, long long grow, bool valid, float& s,
                                                    float& t) {
    s = 1.f, t = 0.f;
    if (ep.in_norm == CTN_NORM_GLN) {
        float2 st = gln_finalize_warp(ep.in_stat + static_cast<long long>(m) * ep.in_P, ep.in_P, ep.in_count);
        s = st.y, t = -st.y * st.x;
    } else if (ep.in_norm == CTN_NORM_CLN) {
        if (valid) {
            float2 st = __ldg(ep.in_stat + grow);
            s = st.y, t = -st.y * st.x;
        }
    }
}

// (mean, rstd) of the forward norm whose backward sums the epilogue accumulates.  Warp-collective.
// gLN (out_stat = GLN): per utterance, from bw_P partials.  cLN (out_stat = CLN): per frame, bw_stat[row] holds
// (mean, rstd) directly.
__device__ __forceinline__ void epilogue_bw_stats(const EpiParams& ep, int m, long long grow, bool valid, float& mean,
                                                  float& rstd) {
    mean = 0.f, rstd = 1.f;
    if (ep.bw_pre == nullptr) return;
    if (ep.out_stat == CTN_NORM_GLN) {
        const float2 st = gln_finalize_warp(ep.bw_stat + static_cast<long long>(m) * ep.bw_P, ep.bw_P,
                                            static_cast<double>(ep.K) * ep.Cout);
        mean = st.x, rstd = st.y;
    } else if (valid) {
        const float2 st = __ldg(ep.bw_stat + grow);
        mean = st.x, rstd = st.y;
    }
}

// Flush the statistics of one 128-column group.  Warp-collective (gLN path).
__device__ __forceinline__ void epilogue_flush_stats(const EpiParams& ep, int m, int tk, int grp, int ew,
                                                     long long grow, bool valid, float& ls, float& lss) {
    if (ep.out_stat == CTN_NORM_GLN) {
        float ws = warp_sum(ls), wss = warp_sum(lss);
        if ((threadIdx.x & 31) == 0)
            ep.out_stat_buf[static_cast<long long>(m) * ep.out_P + (tk * ep.groups + grp) * 4 + ew] =
                make_float2(ws, wss);
    } else if (ep.out_stat == CTN_NORM_CLN) {
        if (valid) ep.out_stat_buf[grow * ep.groups + grp] = make_float2(ls, lss);
    }
    ls = 0.f, lss = 0.f;
}

// ======================================================================
// tcgen05 kernel
// ======================================================================
constexpr int kBM = CTN_ROW_TILE;  // 128 rows = TMEM lanes
constexpr int kBK = 64;            // bf16 elements per k-block = one 128B swizzle atom
// threads = TMA warp + MMA warp + epilogue warps (4 per 128 accumulator columns: a lone warp per
// scheduler cannot hide its own instruction latencies, which made the epilogue the bottleneck)

// PAIR: two CTAs of a cluster (cta_group::2) share one 256-column weight tile -- each loads half
// of it -- and each owns a 128-row tile of its own, so the weight bytes crossing L2 -> SM halve.
template <int BN, bool PAIR>
struct TcCfg {
    static constexpr int kWRows = PAIR ? BN / 2 : BN;   // weight rows this CTA loads per k-block
    static constexpr int kEpiWarps = 4 * (BN / 128);    // warp w: TMEM lane quarter w & 3, column group w >> 2
    static constexpr int kThreads = 64 + 32 * kEpiWarps;
    static constexpr int kStages = 3;
    static constexpr int kABytes = kBM * kBK * 2;  // one of hi / lo
    static constexpr int kWBytes = kWRows * kBK * 2;
    static constexpr int kStageBytes = 2 * kABytes + 2 * kWBytes;
    // no epilogue staging: outputs go straight from registers to global memory
    static constexpr int kStgWarpBytes = 0;
    static constexpr int kStgBytes = kEpiWarps * kStgWarpBytes;
    static constexpr int kColBytes = 2 * BN * 4;   // col_g, col_bias of the current column tile
    static constexpr int kBarBytes = 256;
    static constexpr int kSmemBytes = kStages * kStageBytes + kStgBytes + kColBytes + kBarBytes;
    static constexpr int kTmemCols = 2 * BN;
};

constexpr int kEpiPlain = 0;    // optional PReLU statistics of the output
constexpr int kEpiAffine = 1;   // folded input norm (scale / shift / bias) + optional residual
constexpr int kEpiBw = 2;       // gLN backward sums against a saved pre-activation

template <int BN, bool PAIR, int EPI>
__global__ void __launch_bounds__((TcCfg<BN, PAIR>::kThreads), 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tm_a_hi, const __grid_constant__ CUtensorMap tm_a_lo,
               const __grid_constant__ CUtensorMap tm_w_hi, const __grid_constant__ CUtensorMap tm_w_lo,
               const __grid_constant__ CUtensorMap tm_o_f32, const __grid_constant__ CUtensorMap tm_o_hi,
               const __grid_constant__ CUtensorMap tm_o_lo, const EpiParams ep, int Cin, int n_tiles,
               int total_tiles, int row_tiles) {
    // Work items: `total_tiles` = (row tile or row-tile PAIR) x column tile, strided over the CTAs
    // (or CTA pairs).  In PAIR mode CTA `rank` of the pair owns row tile 2*item_row + rank.
    using Cfg = TcCfg<BN, PAIR>;
    const uint32_t rank = PAIR ? ptx::cluster_ctarank() : 0u;
    const int worker = PAIR ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
    const int n_workers = PAIR ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
    extern __shared__ __align__(1024) uint8_t smem[];
    if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();  // 128B-swizzled tiles need 1024 B alignment
    uint8_t* stg = smem + Cfg::kStages * Cfg::kStageBytes;
    float* col_s = reinterpret_cast<float*>(stg + Cfg::kStgBytes);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(stg + Cfg::kStgBytes + Cfg::kColBytes);
    uint64_t* empty_bar = full_bar + Cfg::kStages;
    uint64_t* tfull_bar = empty_bar + Cfg::kStages;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int nkb = (Cin + kBK - 1) / kBK;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tm_a_hi);
        ptx::prefetch_tmap(&tm_a_lo);
        ptx::prefetch_tmap(&tm_w_hi);
        ptx::prefetch_tmap(&tm_w_lo);
        for (int i = 0; i < Cfg::kStages; ++i) {
            ptx::mbar_init(full_bar + i, 1);
            ptx::mbar_init(empty_bar + i, 1);
        }
        for (int i = 0; i < 2; ++i) {
            ptx::mbar_init(tfull_bar + i, 1);
            ptx::mbar_init(tempty_bar + i, (PAIR ? 2 : 1) * Cfg::kEpiWarps);   // one arrival per epilogue warp (of both CTAs)
        }
        ptx::fence_mbar_init();
    } else if (warp == 1) {
        if (PAIR) {
            ptx::tmem_alloc_pair(tmem_slot, Cfg::kTmemCols);
            ptx::tmem_relinquish_pair();
        } else {
            ptx::tmem_alloc(tmem_slot, Cfg::kTmemCols);
            ptx::tmem_relinquish();
        }
    }
    ptx::tc_fence_before();
    if (PAIR) ptx::cluster_sync(); else __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // PDL: barrier init, TMEM allocation and descriptor prefetch above overlapped the predecessor's tail;
    // from here on the kernel reads what the predecessor wrote
    ptx::grid_dep_sync();

    if (warp == 0) {
        // ------------------------------------------------ TMA producer
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            constexpr bool kLoTerms = CTN_MMA_TERMS == 3;   // the (a_lo, w_lo) tiles are only read by the correction MMAs
            constexpr uint32_t kStageTx = kLoTerms ? Cfg::kStageBytes : Cfg::kStageBytes / 2;
            for (int tile = worker; tile < total_tiles; tile += n_workers) {
                const int tn = tile % n_tiles;
                const int rt = PAIR ? 2 * (tile / n_tiles) + static_cast<int>(rank) : tile / n_tiles;
                const int m = rt / ep.tiles_per_utt, tk = rt % ep.tiles_per_utt;
                // a pair's odd leftover row tile does not exist: load rows past the end (zero fill)
                const int row0 = rt < row_tiles ? m * ep.K + tk * kBM : 0x7fffff00;
                const int n0 = tn * BN + static_cast<int>(rank) * Cfg::kWRows;
                for (int kb = 0; kb < nkb; ++kb) {
                    ptx::mbar_wait(empty_bar + stage, phase ^ 1);
                    uint8_t* st = smem + stage * Cfg::kStageBytes;
                    if (PAIR) {
                        // both CTAs' bytes complete on the leader's barrier, which the MMA issuer waits on
                        const uint32_t fb = ptx::mapa(ptx::smem_u32(full_bar + stage), 0);
                        if (rank == 0) ptx::mbar_expect_tx(full_bar + stage, 2 * kStageTx);
                        ptx::tma_load_2d_pair(st, &tm_a_hi, fb, kb * kBK, row0);
                        if (kLoTerms) ptx::tma_load_2d_pair(st + Cfg::kABytes, &tm_a_lo, fb, kb * kBK, row0);
                        ptx::tma_load_2d_pair(st + 2 * Cfg::kABytes, &tm_w_hi, fb, kb * kBK, n0);
                        if (kLoTerms) ptx::tma_load_2d_pair(st + 2 * Cfg::kABytes + Cfg::kWBytes, &tm_w_lo, fb, kb * kBK, n0);
                    } else {
                        ptx::mbar_expect_tx(full_bar + stage, kStageTx);
                        ptx::tma_load_2d(st, &tm_a_hi, full_bar + stage, kb * kBK, row0);
                        if (kLoTerms) ptx::tma_load_2d(st + Cfg::kABytes, &tm_a_lo, full_bar + stage, kb * kBK, row0);
                        ptx::tma_load_2d(st + 2 * Cfg::kABytes, &tm_w_hi, full_bar + stage, kb * kBK, n0);
                        if (kLoTerms) ptx::tma_load_2d(st + 2 * Cfg::kABytes + Cfg::kWBytes, &tm_w_lo, full_bar + stage, kb * kBK, n0);
                    }
                    if (++stage == Cfg::kStages) stage = 0, phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        // -------------------------------------------------- MMA issuer
        if (lane == 0 && rank == 0) {
            constexpr uint32_t idesc = ptx::make_idesc_bf16(PAIR ? 2 * kBM : kBM, BN, 0, 0);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int tile = worker; tile < total_tiles; tile += n_workers, ++it) {
                const int as = it & 1;
                const uint32_t aphase = (it >> 1) & 1;
                ptx::mbar_wait(tempty_bar + as, aphase ^ 1);
                ptx::tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * BN;
                for (int kb = 0; kb < nkb; ++kb) {
                    ptx::mbar_wait(full_bar + stage, phase);
                    ptx::tc_fence_after();
                    const uint32_t sa = ptx::smem_u32(smem + stage * Cfg::kStageBytes);
                    // K-major, 128B swizzle: 8-row groups are 1024 B apart (SBO); LBO unused.
                    const uint64_t a_hi = ptx::make_smem_desc(sa, 0, 1024, ptx::kLayoutSwizzle128B);
                    const uint64_t a_lo = ptx::make_smem_desc(sa + Cfg::kABytes, 0, 1024, ptx::kLayoutSwizzle128B);
                    const uint64_t w_hi = ptx::make_smem_desc(sa + 2 * Cfg::kABytes, 0, 1024, ptx::kLayoutSwizzle128B);
                    const uint64_t w_lo =
                        ptx::make_smem_desc(sa + 2 * Cfg::kABytes + Cfg::kWBytes, 0, 1024, ptx::kLayoutSwizzle128B);
#pragma unroll
                    for (int k = 0; k < ((ep.dbg & 4) ? 0 : kBK / 16); ++k) {
                        const uint64_t adv = static_cast<uint64_t>((k * 16 * 2) >> 4);  // +32 B per k-step
                        if (PAIR) {
                            ptx::umma_bf16_pair(d_tmem, a_hi + adv, w_hi + adv, idesc, (kb | k) != 0);
                            if (CTN_MMA_TERMS == 3) {
                                ptx::umma_bf16_pair(d_tmem, a_lo + adv, w_hi + adv, idesc, 1);
                                ptx::umma_bf16_pair(d_tmem, a_hi + adv, w_lo + adv, idesc, 1);
                            }
                        } else {
                            ptx::umma_bf16(d_tmem, a_hi + adv, w_hi + adv, idesc, (kb | k) != 0);
                            if (CTN_MMA_TERMS == 3) {
                                ptx::umma_bf16(d_tmem, a_lo + adv, w_hi + adv, idesc, 1);
                                ptx::umma_bf16(d_tmem, a_hi + adv, w_lo + adv, idesc, 1);
                            }
                        }
                    }
                    if (PAIR) {
                        ptx::umma_commit_pair(empty_bar + stage, 3);  // frees the stage in BOTH CTAs
                        if (kb == nkb - 1) ptx::umma_commit_pair(tfull_bar + as, 3);
                    } else {
                        ptx::umma_commit(empty_bar + stage);  // frees the smem stage when these MMAs retire
                        if (kb == nkb - 1) ptx::umma_commit(tfull_bar + as);
                    }
                    if (++stage == Cfg::kStages) stage = 0, phase ^= 1;
                }
            }
        }
    } else {
        // ---------------------------------------------------- epilogue
        // TMEM -> registers -> fused math -> swizzled smem staging -> TMA store.  A warp owns 32
        // rows x 128 columns (4 chunks of 32, fully unrolled) and its own 8 KB staging area, so the
        // stores need no cross-warp barrier.  EPI selects the fused math at compile time: a lone
        // generic epilogue spent 3/4 of its issue slots on predicates and register shuffles and,
        // not the MMAs, set the kernel's duration.
        const int ew = warp & 3;          // TMEM lane quarter this warp may read
        const int wq = warp - 2;          // 0 .. kEpiWarps-1: staging area
        const int cgrp = wq >> 2;         // 128-column group of the tile this warp drains
        const int et = threadIdx.x - 64;  // index among the epilogue threads
        constexpr int kEpiThreads = 32 * Cfg::kEpiWarps;
        constexpr bool kAffine = EPI == kEpiAffine;
        constexpr bool kBw = EPI == kEpiBw;
        const bool use_slope = ep.stat_prelu != nullptr;
        const float slope = use_slope ? __ldg(ep.stat_prelu) : 1.f;   // PReLU with slope 1 = identity
        const float bw_slope = kBw ? __ldg(ep.bw_slope) : 0.f;
        const float* aux = kBw ? ep.bw_pre : (kAffine ? ep.resid : nullptr);  // per-element side input
        const bool aux_hl = kAffine && ep.resid_hi != nullptr;                // ... or a (hi, lo) residual
        const bool want_stat = ep.out_stat != CTN_NORM_NONE;
        int it = 0, cur_tn = -1, cur_m = -1;
        float s = 1.f, t = 0.f, bw_mean = 0.f, bw_rstd = 1.f;
        const uint32_t tempty_remote = PAIR ? ptx::mapa(ptx::smem_u32(tempty_bar), 0) : 0u;
        for (int tile = worker; tile < total_tiles; tile += n_workers, ++it) {
            const int as = it & 1;
            const uint32_t aphase = (it >> 1) & 1;
            const int tn = tile % n_tiles;
            const int rt = PAIR ? 2 * (tile / n_tiles) + static_cast<int>(rank) : tile / n_tiles;
            const bool tile_ok = rt < row_tiles;   // false only for the leftover half of an odd pair
            const int m = tile_ok ? rt / ep.tiles_per_utt : 0, tk = tile_ok ? rt % ep.tiles_per_utt : 0;
            const int k = tk * kBM + ew * 32 + lane;
            const bool valid = tile_ok && k < ep.K;
            const long long grow = static_cast<long long>(m) * ep.K + k;
            // gLN statistics depend on the utterance only: re-reduce the partials when m changes
            if (kAffine && (ep.in_norm != CTN_NORM_GLN || m != cur_m)) epilogue_row_affine(ep, m, grow, valid, s, t);
            if (kBw && (ep.out_stat != CTN_NORM_GLN || m != cur_m)) epilogue_bw_stats(ep, m, grow, valid, bw_mean, bw_rstd);
            cur_m = m;
            if ((kAffine || kBw) && tn != cur_tn) {  // per-column constants of this column tile -> smem
                ptx::named_bar_sync(1, kEpiThreads);
                for (int i = et; i < BN; i += kEpiThreads) {
                    const int n = tn * BN + i;
                    const float* cg = kBw ? ep.bw_gamma : ep.col_g;
                    col_s[i] = (cg && n < ep.Cout) ? __ldg(cg + n) : 0.f;
                    col_s[BN + i] = (ep.col_bias && n < ep.Cout) ? __ldg(ep.col_bias + n) : 0.f;
                }
                ptx::named_bar_sync(1, kEpiThreads);
                cur_tn = tn;
            }
            const int col0 = tn * BN + cgrp * 128;           // first column of this warp's group
            const bool grp_ok = col0 < ep.Cout;              // Cout % 32 == 0: chunks are whole or absent
            const int krow0 = tk * kBM + ew * 32;            // first frame of this warp's 32 rows
            const bool store_ok = tile_ok && krow0 < ep.K && !(ep.dbg & 2);  // rows past K: clipped by the map
            float ra[4][8], rb[4][8];                        // side input of even / odd chunks
            float rc3[4][8];                                 // kDeep: a third buffer, the prefetch runs two chunks ahead
            constexpr bool kDeep = kBw && !PAIR;             // (single-CTA kernel: small problems, latency-bound epilogue)
            auto load_aux = [&](float (&dst)[4][8], int ci) {
                if (!valid || col0 + ci * 32 >= ep.Cout || (ep.dbg & 32)) return;
                if (aux != nullptr) {
                    const float* rp = aux + grow * ep.Cout + col0 + ci * 32;
#pragma unroll
                    for (int j = 0; j < 4; ++j) ptx::ld_global_v8(rp + 8 * j, dst[j]);
                } else if (aux_hl) {   // 32 bf16 hi in dst[0..1], 32 bf16 lo in dst[2..3] (raw bits)
                    const long long o = grow * ep.Cout + col0 + ci * 32;
                    ptx::ld_global_v8(ep.resid_hi + o, dst[0]);
                    ptx::ld_global_v8(ep.resid_hi + o + 16, dst[1]);
                    ptx::ld_global_v8(ep.resid_lo + o, dst[2]);
                    ptx::ld_global_v8(ep.resid_lo + o + 16, dst[3]);
                }
            };
            if (kAffine || kBw) load_aux(ra, 0);
            if constexpr (kDeep) load_aux(rb, 1);
            if (kBw && valid && grp_ok) {
                // the saved pre-activation comes from DRAM (written a whole forward pass ago): bring the lane's other
                // three lines of this tile into L2 while the MMAs run; the register prefetch is only one chunk deep
#pragma unroll
                for (int ci = 1; ci < 4; ++ci)
                    if (col0 + ci * 32 < ep.Cout) ptx::prefetch_l2(aux + grow * ep.Cout + col0 + ci * 32);
            }
            ptx::mbar_wait(tfull_bar + as, aphase);
            ptx::tc_fence_after();
            float ls[4] = {0.f, 0.f, 0.f, 0.f}, lss[4] = {0.f, 0.f, 0.f, 0.f};
            if (!(ep.dbg & 1) && grp_ok) {
#pragma unroll
                for (int ci = 0; ci < 4; ++ci) {
                    const int n0 = col0 + ci * 32;
                    if (n0 < ep.Cout) {
                        uint32_t r[32];
                        if (ep.dbg & 8) {
#pragma unroll
                            for (int j = 0; j < 32; ++j) r[j] = j + lane;
                        } else
                        ptx::tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + as * BN + cgrp * 128 + ci * 32, r);
                        if constexpr (kDeep) {          // buffers ra, rb, rc3 in turn: chunk ci + 2 -> buffer (ci + 2) % 3
                            if (ci == 0) load_aux(rc3, 2);
                            if (ci == 1) load_aux(ra, 3);
                        } else if ((kAffine || kBw) && ci < 3) {
                            if (ci & 1) load_aux(ra, ci + 1); else load_aux(rb, ci + 1);
                        }
                        ptx::tmem_ld_wait();
                        float v[32];
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                        const float (&rc)[4][8] = kDeep ? ((ci == 1) ? rb : ((ci == 2) ? rc3 : ra)) : ((ci & 1) ? rb : ra);
                        const float* cs = col_s + cgrp * 128 + ci * 32;
                        if (kAffine) {
#pragma unroll
                            for (int j = 0; j < 32; j += 4) {
                                const float4 g = *reinterpret_cast<const float4*>(cs + j);
                                const float4 b = *reinterpret_cast<const float4*>(cs + BN + j);
                                v[j] = fmaf(s, v[j], fmaf(t, g.x, b.x));
                                v[j + 1] = fmaf(s, v[j + 1], fmaf(t, g.y, b.y));
                                v[j + 2] = fmaf(s, v[j + 2], fmaf(t, g.z, b.z));
                                v[j + 3] = fmaf(s, v[j + 3], fmaf(t, g.w, b.w));
                            }
                            if (aux != nullptr && valid) {
#pragma unroll
                                for (int j = 0; j < 32; ++j) v[j] += rc[j >> 3][j & 7];
                            } else if (aux_hl && valid) {
#pragma unroll
                                for (int j = 0; j < 16; ++j) {   // word j holds columns 2j (low half) and 2j + 1
                                    const uint32_t h = __float_as_uint(rc[j >> 3][j & 7]);
                                    const uint32_t l = __float_as_uint(rc[2 + (j >> 3)][j & 7]);
#if CTN_OPERAND_FP16
                                    float h0, h1, l0, l1;
                                    op16_value2(h, h0, h1), op16_value2(l, l0, l1);
                                    v[2 * j] += h0 + l0;
                                    v[2 * j + 1] += h1 + l1;
#else
                                    v[2 * j] += __uint_as_float(h << 16) + __uint_as_float(l << 16);
                                    v[2 * j + 1] += __uint_as_float(h & 0xffff0000u) + __uint_as_float(l & 0xffff0000u);
#endif
                                }
                            }
                        }
                        if (kBw) {
                            if (valid) {
#pragma unroll
                                for (int j = 0; j < 8; ++j) {
                                    const float dv[4] = {rc[j >> 1][(j & 1) * 4], rc[j >> 1][(j & 1) * 4 + 1],
                                                         rc[j >> 1][(j & 1) * 4 + 2], rc[j >> 1][(j & 1) * 4 + 3]};
                                    const float4 g = *reinterpret_cast<const float4*>(cs + 4 * j);
                                    const float gv[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
                                    for (int q = 0; q < 4; ++q) {
                                        const float qh = (prelu(dv[q], bw_slope) - bw_mean) * bw_rstd;
                                        const float gd = gv[q] * v[4 * j + q];
                                        ls[q] += gd;
                                        lss[q] = fmaf(gd, qh, lss[q]);
                                    }
                                }
                            }
                        } else if (want_stat && valid && !(ep.dbg & 16)) {
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                const float x = fmaf(slope, fminf(v[j], 0.f), fmaxf(v[j], 0.f));   // prelu
                                ls[j & 3] += x;
                                lss[j & 3] = fmaf(x, x, lss[j & 3]);
                                if (ep.store_prelu) v[j] = x;
                            }
                        }
                        // Straight from registers: a thread owns 32 consecutive columns of its row, i.e.
                        // four whole 32 B sectors of fp32 (two of bf16), written with 256-bit stores.  The
                        // earlier smem-staged TMA stores cost 2 x 8 KB of shared-memory bandwidth per chunk,
                        // which the MMAs' operand reads and the TMA fills already saturate.
                        if (valid && !(ep.dbg & 2)) {
                            if (ep.out_f32 != nullptr) {
                                float* op = ep.out_f32 + grow * ep.Cout + n0;
#pragma unroll
                                for (int j = 0; j < 4; ++j)
                                    ptx::st_global_v8(op + 8 * j, r2u(v[8 * j]), r2u(v[8 * j + 1]), r2u(v[8 * j + 2]),
                                                      r2u(v[8 * j + 3]), r2u(v[8 * j + 4]), r2u(v[8 * j + 5]),
                                                      r2u(v[8 * j + 6]), r2u(v[8 * j + 7]));
                            }
                            if (ep.out_hi != nullptr) {
                                uint32_t h[16], l[16];
#pragma unroll
                                for (int j = 0; j < 16; ++j) split_pack2(v[2 * j], v[2 * j + 1], h[j], l[j]);
                                __nv_bfloat16* hp = ep.out_hi + grow * ep.Cout + n0;
                                __nv_bfloat16* lp = ep.out_lo + grow * ep.Cout + n0;
#pragma unroll
                                for (int j = 0; j < 2; ++j) {
                                    ptx::st_global_v8(hp + 16 * j, h[8 * j], h[8 * j + 1], h[8 * j + 2], h[8 * j + 3],
                                                      h[8 * j + 4], h[8 * j + 5], h[8 * j + 6], h[8 * j + 7]);
                                    ptx::st_global_v8(lp + 16 * j, l[8 * j], l[8 * j + 1], l[8 * j + 2], l[8 * j + 3],
                                                      l[8 * j + 4], l[8 * j + 5], l[8 * j + 6], l[8 * j + 7]);
                                }
                            }
                        }
                    }
                }
                if ((want_stat || kBw) && tile_ok) {
                    float a0 = (ls[0] + ls[1]) + (ls[2] + ls[3]), a1 = (lss[0] + lss[1]) + (lss[2] + lss[3]);
                    epilogue_flush_stats(ep, m, tk, col0 / CTN_COL_TILE, ew, grow, valid, a0, a1);
                }
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (PAIR) ptx::mbar_arrive_cluster(tempty_remote + as * 8);
                else ptx::mbar_arrive(tempty_bar + as);
            }
        }
    }

    ptx::tc_fence_before();
    if (PAIR) ptx::cluster_sync(); else __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        if (PAIR) ptx::tmem_dealloc_pair(tmem_base, Cfg::kTmemCols);
        else ptx::tmem_dealloc(tmem_base, Cfg::kTmemCols);
    }
}

// ======================================================================
// SIMT kernel: same tiling, statistics layout and epilogue; any Cin / Cout.
// One thread per output row, 32 columns at a time.  Used for shapes the
// tensor-core path does not take (channel counts not multiples of 8).
// ======================================================================
__global__ void __launch_bounds__(128)
gemm_simt_kernel(const __nv_bfloat16* __restrict__ a_hi, const __nv_bfloat16* __restrict__ a_lo,
                 const __nv_bfloat16* __restrict__ w_hi, const __nv_bfloat16* __restrict__ w_lo, const EpiParams ep,
                 int Cin) {
    ptx::grid_dep_sync();   // PDL: predecessor complete; let the next kernel be scheduled
    const int rt = blockIdx.x;
    const int m = rt / ep.tiles_per_utt, tk = rt % ep.tiles_per_utt;
    const int ew = threadIdx.x >> 5;
    const int k = tk * kBM + threadIdx.x;
    const bool valid = k < ep.K;
    const long long grow = static_cast<long long>(m) * ep.K + k;
    float slope = 0.f;
    const bool use_slope = ep.stat_prelu != nullptr;
    if (use_slope) slope = __ldg(ep.stat_prelu);
    float s, t, bw_mean, bw_rstd;
    epilogue_row_affine(ep, m, grow, valid, s, t);
    epilogue_bw_stats(ep, m, grow, valid, bw_mean, bw_rstd);
    if (ep.bw_pre != nullptr) slope = __ldg(ep.bw_slope);
    float ls = 0.f, lss = 0.f;
    for (int grp = 0; grp < ep.groups; ++grp) {
        for (int c = 0; c < CTN_COL_TILE / 32; ++c) {
            const int n0 = grp * CTN_COL_TILE + c * 32;
            if (n0 >= ep.Cout) break;
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = 0.f;
            if (valid) {
                for (int kk = 0; kk < Cin; ++kk) {
#if CTN_OPERAND_FP16
                    const float a = op16_value(a_hi[grow * Cin + kk]);     // as the tensor-core path: hi only
#else
                    const float a = __bfloat162float(a_hi[grow * Cin + kk]) + __bfloat162float(a_lo[grow * Cin + kk]);
#endif
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        if (n0 + j < ep.Cout) {
                            const long long wi = static_cast<long long>(n0 + j) * Cin + kk;
#if CTN_OPERAND_FP16
                            v[j] = fmaf(a, op16_value(w_hi[wi]), v[j]);
#else
                            v[j] = fmaf(a, __bfloat162float(w_hi[wi]) + __bfloat162float(w_lo[wi]), v[j]);
#endif
                        }
                    }
                }
            }
            epilogue_chunk<false>(ep, v, grow, n0, valid, s, t, slope, use_slope, bw_mean, bw_rstd, ls, lss);
        }
        epilogue_flush_stats(ep, m, tk, grp, ew, grow, valid, ls, lss);
    }
}

// ======================================================================
// host side
// ======================================================================
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
    static PFN_encodeTiled fn = nullptr;  // benign race: every thread resolves the same pointer
    if (fn == nullptr) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_encodeTiled>(p);
    }
    return fn;
}

int make_tmap_bf16_kmajor(CUtensorMap* tm, const void* ptr, long long rows, int cols, int box_rows);
// same map, named for its use by the weight-gradient kernel (MN-major reads of the same tiles)
int make_tmap_bf16_box(CUtensorMap* tm, const void* ptr, long long rows, int cols, int box_cols, int box_rows) {
    if (box_cols != kBK) {
        set_error("make_tmap_bf16_box: only 64-column boxes are supported");
        return CTN_ERR_INVALID;
    }
    return make_tmap_bf16_kmajor(tm, ptr, rows, cols, box_rows);
}

// bf16 [rows, cols] row-major, box = 64 columns x box_rows rows, 128B swizzle, zero OOB fill
int make_tmap_bf16_kmajor(CUtensorMap* tm, const void* ptr, long long rows, int cols, int box_rows) {
    PFN_encodeTiled enc = get_encode_fn();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled not available from the driver");
        return CTN_ERR_CUDA;
    }
    cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
    cuuint64_t strides[1] = {static_cast<cuuint64_t>(cols) * 2};
    cuuint32_t box[2] = {static_cast<cuuint32_t>(kBK), static_cast<cuuint32_t>(box_rows)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (%d) ptr=%p rows=%lld cols=%d", static_cast<int>(r), ptr, rows, cols);
        return CTN_ERR_CUDA;
    }
    return CTN_OK;
}

// [M, K, C] row-major output, box = 32 columns x 32 frames of one utterance; rows past K are
// clipped by the map, so a partial last row tile never touches the next utterance.
int make_tmap_out(CUtensorMap* tm, const void* ptr, int M, int K, int C, bool f32) {
    PFN_encodeTiled enc = get_encode_fn();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled not available from the driver");
        return CTN_ERR_CUDA;
    }
    const cuuint64_t es = f32 ? 4 : 2;
    cuuint64_t dims[3] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(K), static_cast<cuuint64_t>(M)};
    cuuint64_t strides[2] = {static_cast<cuuint64_t>(C) * es, static_cast<cuuint64_t>(C) * es * K};
    cuuint32_t box[3] = {32, 32, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(tm, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3,
                     const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     f32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled(out) failed (%d) ptr=%p M=%d K=%d C=%d", static_cast<int>(r), ptr, M, K, C);
        return CTN_ERR_CUDA;
    }
    return CTN_OK;
}

// bf16 [M, K, C] row-major, box = 64 channels x box_rows frames of ONE utterance, 128B swizzle: loads zero-fill
// frames past K, stores clip them (a row tile never touches the next utterance).
int make_tmap_bf16_3d(CUtensorMap* tm, const void* ptr, int M, int K, int C, int box_rows) {
    PFN_encodeTiled enc = get_encode_fn();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled not available from the driver");
        return CTN_ERR_CUDA;
    }
    cuuint64_t dims[3] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(K), static_cast<cuuint64_t>(M)};
    cuuint64_t strides[2] = {static_cast<cuuint64_t>(C) * 2, static_cast<cuuint64_t>(C) * 2 * K};
    cuuint32_t box[3] = {static_cast<cuuint32_t>(kBK), static_cast<cuuint32_t>(box_rows), 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled(3d) failed (%d) ptr=%p M=%d K=%d C=%d", static_cast<int>(r), ptr, M, K, C);
        return CTN_ERR_CUDA;
    }
    return CTN_OK;
}

// fp32 [M, K, C] seen by dilation phase: dims (C, d, ceil(K / d), M), element (c, ph, j, m) = frame j d + ph of
// utterance m; box = 32 columns x 1 phase x box_j frames of that phase, 128B swizzle.  Stores only, and only for boxes
// whose frames all exist: a (ph, j) with j d + ph >= K aliases the next utterance and is NOT clipped by the map.
int make_tmap_phase_f32(CUtensorMap* tm, const void* ptr, int M, int K, int C, int d, int box_j) {
    PFN_encodeTiled enc = get_encode_fn();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled not available from the driver");
        return CTN_ERR_CUDA;
    }
    const cuuint64_t row = static_cast<cuuint64_t>(C) * 4;
    cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(d), static_cast<cuuint64_t>((K + d - 1) / d),
                          static_cast<cuuint64_t>(M)};
    cuuint64_t strides[3] = {row, row * d, row * K};
    cuuint32_t box[4] = {32, 1, static_cast<cuuint32_t>(box_j), 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled(phase) failed (%d) ptr=%p M=%d K=%d C=%d d=%d", static_cast<int>(r), ptr, M, K, C, d);
        return CTN_ERR_CUDA;
    }
    return CTN_OK;
}

template <int BN, bool PAIR, int EPI>
static int launch_tc(const ctn_gemm_args& a, const EpiParams& ep, cudaStream_t stream) {
    using Cfg = TcCfg<BN, PAIR>;
    CUtensorMap ta_hi, ta_lo, tw_hi, tw_lo;
    const long long R = static_cast<long long>(a.M) * a.K;
    int rc;
    if ((rc = make_tmap_bf16_kmajor(&ta_hi, a.a_hi, R, a.Cin, kBM)) != CTN_OK) return rc;
    if ((rc = make_tmap_bf16_kmajor(&ta_lo, a.a_lo, R, a.Cin, kBM)) != CTN_OK) return rc;
    if ((rc = make_tmap_bf16_kmajor(&tw_hi, a.w_hi, a.Cout, a.Cin, Cfg::kWRows)) != CTN_OK) return rc;
    if ((rc = make_tmap_bf16_kmajor(&tw_lo, a.w_lo, a.Cout, a.Cin, Cfg::kWRows)) != CTN_OK) return rc;
    CUtensorMap to_f32, to_hi, to_lo;
    // unused maps still have to be valid objects: point them at an operand
    if ((rc = make_tmap_out(&to_f32, a.out_f32 ? a.out_f32 : a.out_hi, a.M, a.K, a.Cout, a.out_f32 != nullptr)) != CTN_OK)
        return rc;
    if ((rc = make_tmap_out(&to_hi, a.out_hi ? a.out_hi : a.out_f32, a.M, a.K, a.Cout, a.out_hi == nullptr)) != CTN_OK)
        return rc;
    if ((rc = make_tmap_out(&to_lo, a.out_lo ? a.out_lo : a.out_f32, a.M, a.K, a.Cout, a.out_lo == nullptr)) != CTN_OK)
        return rc;
    static bool attr_done[kMaxDevices] = {};  // idempotent; set once per device per instantiation
    bool& attr_set = attr_done[device_slot()];
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<BN, PAIR, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Cfg::kSmemBytes);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(gemm_tc_kernel)");
        attr_set = true;
    }
    const int n_tiles = ceil_div(a.Cout, BN);
    const int row_tiles = a.M * ep.tiles_per_utt;
    const int items = (PAIR ? ceil_div(row_tiles, 2) : row_tiles) * n_tiles;
    const int workers_max = PAIR ? sm_count() / 2 : sm_count();
    const int workers = items < workers_max ? items : workers_max;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(PAIR ? 2 * workers : workers, 1, 1);
    cfg.blockDim = dim3(Cfg::kThreads, 1, 1);
    cfg.dynamicSmemBytes = Cfg::kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = PAIR ? 2 : 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = pdl_allowed(stream, static_cast<long long>(a.M) * a.K);
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_tc_kernel<BN, PAIR, EPI>, ta_hi, ta_lo, tw_hi, tw_lo, to_f32, to_hi, to_lo,
                                       ep, a.Cin, n_tiles, items, row_tiles);
    if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx(gemm_tc_kernel)");
    return CTN_OK;
}

}  // namespace ctn

using namespace ctn;

extern "C" int ctn_gemm_gln_partials(int K, int Cout) {
    return ceil_div(K, CTN_ROW_TILE) * ceil_div(Cout, CTN_COL_TILE) * 4;
}
extern "C" int ctn_gemm_cln_groups(int Cout) { return ceil_div(Cout, CTN_COL_TILE); }

extern "C" int ctn_gemm1x1(const ctn_gemm_args* args, void* stream_) {
    CTN_REQUIRE(args != nullptr, "ctn_gemm1x1: args is NULL");
    const ctn_gemm_args& a = *args;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    CTN_REQUIRE(a.M > 0 && a.K > 0 && a.Cin > 0 && a.Cout > 0, "ctn_gemm1x1: non-positive shape M=%d K=%d Cin=%d Cout=%d",
                a.M, a.K, a.Cin, a.Cout);
    CTN_REQUIRE(static_cast<long long>(a.M) * a.K < (1ll << 31), "ctn_gemm1x1: M*K exceeds 2^31 rows");
    CTN_REQUIRE(a.a_hi && a.a_lo && a.w_hi && a.w_lo, "ctn_gemm1x1: NULL operand");
    CTN_REQUIRE((a.out_hi == nullptr) == (a.out_lo == nullptr), "ctn_gemm1x1: out_hi/out_lo must be given together");
    CTN_REQUIRE(a.out_f32 || a.out_hi, "ctn_gemm1x1: no output requested");
    CTN_REQUIRE(a.in_norm >= 0 && a.in_norm <= 2 && a.out_stat >= 0 && a.out_stat <= 2, "ctn_gemm1x1: bad norm mode");
    CTN_REQUIRE(a.in_norm == CTN_NORM_NONE || (a.in_stat && a.col_g), "ctn_gemm1x1: in_norm needs in_stat and col_g");
    CTN_REQUIRE(a.in_norm != CTN_NORM_GLN || (a.in_P > 0 && a.in_count > 0), "ctn_gemm1x1: gLN needs in_P, in_count");
    CTN_REQUIRE(a.out_stat == CTN_NORM_NONE || a.out_stat_buf, "ctn_gemm1x1: out_stat needs out_stat_buf");

    EpiParams ep;
    ep.K = a.K, ep.Cout = a.Cout;
    ep.groups = ceil_div(a.Cout, CTN_COL_TILE);
    ep.tiles_per_utt = ceil_div(a.K, kBM);
    ep.out_f32 = a.out_f32;
    ep.out_hi = static_cast<__nv_bfloat16*>(a.out_hi);
    ep.out_lo = static_cast<__nv_bfloat16*>(a.out_lo);
    ep.resid = a.resid, ep.col_g = a.col_g, ep.col_bias = a.col_bias;
    ep.resid_hi = static_cast<const __nv_bfloat16*>(a.resid_hi), ep.resid_lo = static_cast<const __nv_bfloat16*>(a.resid_lo);
    CTN_REQUIRE((a.resid_hi == nullptr) == (a.resid_lo == nullptr) && !(a.resid && a.resid_hi),
                "ctn_gemm1x1: resid_hi/resid_lo go together and exclude resid");
    ep.in_norm = a.in_norm, ep.in_stat = static_cast<const float2*>(a.in_stat), ep.in_P = a.in_P;
    ep.in_count = a.in_count;
    ep.out_stat = a.out_stat, ep.out_stat_buf = static_cast<float2*>(a.out_stat_buf);
    ep.out_P = ctn_gemm_gln_partials(a.K, a.Cout);
    ep.stat_prelu = a.stat_prelu;
    ep.store_prelu = a.store_prelu;
    CTN_REQUIRE(!a.store_prelu || (a.stat_prelu != nullptr && a.out_stat != CTN_NORM_NONE && a.bw_pre == nullptr),
                "ctn_gemm1x1: store_prelu needs stat_prelu and an out_stat mode");
    ep.bw_pre = a.bw_pre, ep.bw_gamma = a.bw_gamma, ep.bw_slope = a.bw_slope;
    ep.bw_stat = static_cast<const float2*>(a.bw_stat), ep.bw_P = a.bw_P;
    {
        const char* d = getenv("CTN_GEMM_DBG");
        ep.dbg = d ? atoi(d) : 0;
    }
    CTN_REQUIRE(a.bw_pre == nullptr || (a.bw_gamma && a.bw_slope && a.bw_stat && a.out_stat != CTN_NORM_NONE &&
                                        (a.out_stat != CTN_NORM_GLN || a.bw_P > 0) && a.resid == nullptr && a.resid_hi == nullptr &&
                                        a.in_norm == CTN_NORM_NONE),
                "ctn_gemm1x1: bw_pre needs bw_gamma/bw_slope/bw_stat, an out_stat mode and no resid / in_norm");

    auto aligned16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    const bool tc_ok = !(a.force_simt & 1) && a.Cin % 8 == 0 && a.Cout % 32 == 0 && aligned16(a.a_hi) && aligned16(a.a_lo) &&
                       aligned16(a.w_hi) && aligned16(a.w_lo) && aligned16(a.out_f32) && aligned16(a.out_hi) &&
                       aligned16(a.out_lo) && aligned16(a.resid) && aligned16(a.bw_pre) && aligned16(a.resid_hi) &&
                       aligned16(a.resid_lo);
    if (tc_ok) {
        // 256-wide tiles halve the re-reads of the activation tile but need enough tiles to fill the SMs
        const long long tiles256 = static_cast<long long>(a.M) * ep.tiles_per_utt * ceil_div(a.Cout, 256);
        const int mode = a.force_simt >> 1;   // test hook: 1 = single-CTA kernels only, 2 = pair kernel
        // ... or few enough that every CTA pair gets at most one 256 x 256 item where 128 x 128 tiles would need a
        // second round for a handful of tiles (the reference's batch of 3 x 4 s: 150 tiles of a 256-column output
        // for 148 SMs; training step 3.92 -> 3.83 ms backward)
        const long long row_tiles_ = static_cast<long long>(a.M) * ep.tiles_per_utt;
        const bool one_round = (row_tiles_ + 1) / 2 * ceil_div(a.Cout, 256) <= sm_count() / 2 &&
                               row_tiles_ * ceil_div(a.Cout, 128) > sm_count();
        const bool pair = a.Cout % 256 == 0 && mode != 1 && (mode == 2 || tiles256 >= 2ll * sm_count() || one_round);
        const bool affine = a.col_g != nullptr || a.col_bias != nullptr || a.in_norm != CTN_NORM_NONE ||
                            a.resid != nullptr || a.resid_hi != nullptr;
        if (a.bw_pre != nullptr)
            return pair ? launch_tc<256, true, kEpiBw>(a, ep, stream) : launch_tc<128, false, kEpiBw>(a, ep, stream);
        if (affine)
            return pair ? launch_tc<256, true, kEpiAffine>(a, ep, stream)
                        : launch_tc<128, false, kEpiAffine>(a, ep, stream);
        return pair ? launch_tc<256, true, kEpiPlain>(a, ep, stream) : launch_tc<128, false, kEpiPlain>(a, ep, stream);
    }
    const int grid = a.M * ep.tiles_per_utt;
    CTN_LAUNCH_PDL("gemm_simt_kernel", static_cast<long long>(a.M) * a.K, gemm_simt_kernel, grid, 128, 0, stream, static_cast<const __nv_bfloat16*>(a.a_hi), static_cast<const __nv_bfloat16*>(a.a_lo), static_cast<const __nv_bfloat16*>(a.w_hi), static_cast<const __nv_bfloat16*>(a.w_lo), ep, a.Cin);
    CTN_LAUNCH_CHECK("gemm_simt_kernel");
    return CTN_OK;
}
