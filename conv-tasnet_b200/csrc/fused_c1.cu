// Deferred residual update fused into the next block's conv1x1 (inference):
//     x <- x + r y - r mu rowsum + W2 beta2         (gLN-2 of the PREVIOUS block, conv_tasnet.py:254-256, :237)
//     a  = PReLU(W1 x),  (sum, sumsq) of a          (conv1x1 of THIS block, :217-218, statistics for :219)
// ctn_dwpw_fwd leaves the previous block's pointwise product y un-normalised because gLN-2 needs the
// statistics of the whole utterance; here they are complete.  SIMT producer warps read x (hi, lo)
// and y, form the new x, write it back (the residual stream, 1 KB per frame) and store it
// 128B-swizzled, K-major into the shared-memory slot the tcgen05 MMAs read as their A operand.  The
// A tile (128 frames x 256 channels) stays resident while BOTH 256-column tiles of Cout = 512 are
// multiplied, so it is produced once per row tile; only the weights stream (TMA, 3 slots).
// Replaces ctn_resid_update + ctn_gemm1x1: 6 KB -> 5 KB per frame and block, one launch fewer.
//
// CTA pair (cta_group::2), warps: 0 TMA (weights), 1 MMA issuer, 2-5 drain the TMEM accumulators
// (PReLU, statistics, 256-bit stores), 6-13 producers.  The producers are latency-bound, not
// FMA-bound like fused_block.cu's, so they never leave their load stream to drain.
#include <cuda.h>

#include "common.cuh"
#include "f32x2.cuh"
#include "ptx.cuh"

namespace ctn {

int make_tmap_bf16_kmajor(CUtensorMap* tm, const void* ptr, long long rows, int cols, int box_rows);  // gemm.cu

namespace c1 {

constexpr int kBM = 128;
constexpr int kBN = 256;             // output columns per accumulator
constexpr int kBK = 64;
constexpr int kCin = 256;            // B: the whole reduction dimension is resident
constexpr int kKB = kCin / kBK;      // 4 k-blocks = 4 resident A slots
constexpr int kWSlots = 3;
constexpr int kABytes = kBM * kBK * 2;           // one of hi / lo
constexpr int kWBytes = (kBN / 2) * kBK * 2;     // this CTA's half of a weight k-block, one of hi / lo
constexpr int kASlotBytes = 2 * kABytes;
constexpr int kWSlotBytes = 2 * kWBytes;
constexpr int kTileBytes = kKB * kASlotBytes + kWSlots * kWSlotBytes;
constexpr int kBarBytes = 256;
constexpr int kColBytes = 2 * kCin * 4;          // col_g, col_bias of the previous pointwise conv
constexpr int kMaxTiles = 32;                    // row tiles per CTA whose gLN-2 scale / shift are precomputed
constexpr int kSmemBytes = kTileBytes + kBarBytes + kColBytes + kMaxTiles * 8;
constexpr int kDrainWarps = 4;
constexpr int kProdWarps = 8;
constexpr int kThreads = 32 * (2 + kDrainWarps + kProdWarps);
constexpr int kRows = kBM / (kProdWarps * 2);   // frames a producer thread handles per unit (a warp spans 2 frame groups)

struct Params {
    int K, Cout, n_tn, tiles_per_utt, row_tiles, items, groups;
    __nv_bfloat16* x_hi;
    __nv_bfloat16* x_lo;
    const float* y;
    const float2* in_stat;   // gLN-2 partials of the previous block
    int in_P;
    double in_count;
    const float* col_g;
    const float* col_bias;
    float* a;                // out: PReLU(W1 x), fp32 [M*K, Cout]
    float2* out_stat;        // out: gLN-1 partials, ctn_gemm1x1 layout
    int out_P;
    const float* slope;
};

struct Row {   // one frame's worth of a thread's inputs: 4 channels
    uint2 hi, lo;
    float4 y;
};

__global__ void __launch_bounds__(kThreads, 1)
updc1_kernel(const __grid_constant__ CUtensorMap tm_w_hi, const __grid_constant__ CUtensorMap tm_w_lo, const Params p) {
    const uint32_t rank = ptx::cluster_ctarank();
    const int worker = static_cast<int>(blockIdx.x >> 1);
    const int n_workers = static_cast<int>(gridDim.x >> 1);
    extern __shared__ __align__(1024) uint8_t smem[];
    if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
    uint8_t* wsm = smem + kKB * kASlotBytes;
    uint64_t* afull_bar = reinterpret_cast<uint64_t*>(smem + kTileBytes);
    uint64_t* aempty_bar = afull_bar + kKB;
    uint64_t* wfull_bar = aempty_bar + kKB;
    uint64_t* wempty_bar = wfull_bar + kWSlots;
    uint64_t* tfull_bar = wempty_bar + kWSlots;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
    float* colc = reinterpret_cast<float*>(smem + kTileBytes + kBarBytes);   // [2][kCin]
    float2* tstat = reinterpret_cast<float2*>(smem + kTileBytes + kBarBytes + kColBytes);   // [kMaxTiles] (rstd, -rstd mean)

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < kCin; i += kThreads) {
        colc[i] = __ldg(p.col_g + i);
        colc[kCin + i] = __ldg(p.col_bias + i);
    }
    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tm_w_hi);
        ptx::prefetch_tmap(&tm_w_lo);
        for (int i = 0; i < kKB; ++i) {
            ptx::mbar_init(afull_bar + i, 2 * kProdWarps);   // one arrival per producer warp, both CTAs
            ptx::mbar_init(aempty_bar + i, 1);
        }
        for (int i = 0; i < kWSlots; ++i) {
            ptx::mbar_init(wfull_bar + i, 1);
            ptx::mbar_init(wempty_bar + i, 1);
        }
        for (int i = 0; i < 2; ++i) {
            ptx::mbar_init(tfull_bar + i, 1);
            ptx::mbar_init(tempty_bar + i, 2 * kDrainWarps);
        }
        ptx::fence_mbar_init();
    } else if (warp == 1) {
        ptx::tmem_alloc_pair(tmem_slot, 2 * kBN);
        ptx::tmem_relinquish_pair();
    }
    ptx::tc_fence_before();
    ptx::cluster_sync();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int n_tiles = worker < p.items ? (p.items - worker + n_workers - 1) / n_workers : 0;   // of this CTA pair

    if (warp == 0) {
        // ------------------------------------------------ TMA producer (weights)
        if (lane == 0) {
            int slot = 0;
            uint32_t phase = 0;
            const uint32_t fb0 = ptx::mapa(ptx::smem_u32(wfull_bar), 0);
            for (int it = 0; it < n_tiles; ++it) {
                for (int tn = 0; tn < p.n_tn; ++tn) {
                    for (int kb = 0; kb < kKB; ++kb) {
                        ptx::mbar_wait(wempty_bar + slot, phase ^ 1);
                        uint8_t* st = wsm + slot * kWSlotBytes;
                        if (rank == 0) ptx::mbar_expect_tx(wfull_bar + slot, 4 * kWBytes);
                        const int n0 = tn * kBN + static_cast<int>(rank) * (kBN / 2);
                        ptx::tma_load_2d_pair(st, &tm_w_hi, fb0 + slot * 8, kb * kBK, n0);
                        ptx::tma_load_2d_pair(st + kWBytes, &tm_w_lo, fb0 + slot * 8, kb * kBK, n0);
                        if (++slot == kWSlots) slot = 0, phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // -------------------------------------------------- MMA issuer (leader CTA)
        if (lane == 0 && rank == 0) {
            constexpr uint32_t idesc = ptx::make_idesc_bf16(2 * kBM, kBN, 0, 0);
            int wslot = 0;
            uint32_t wph = 0;
            int d = 0;   // accumulator counter
            for (int it = 0; it < n_tiles; ++it) {
                const uint32_t aph = it & 1;
                for (int tn = 0; tn < p.n_tn; ++tn, ++d) {
                    const int as = d & 1;
                    ptx::mbar_wait(tempty_bar + as, ((d >> 1) & 1) ^ 1);
                    ptx::tc_fence_after();
                    const uint32_t d_tmem = tmem_base + as * kBN;
                    for (int kb = 0; kb < kKB; ++kb) {
                        ptx::mbar_wait(wfull_bar + wslot, wph);
                        if (tn == 0) ptx::mbar_wait(afull_bar + kb, aph);
                        ptx::tc_fence_after();
                        const uint32_t sa = ptx::smem_u32(smem + kb * kASlotBytes);
                        const uint32_t sw = ptx::smem_u32(wsm + wslot * kWSlotBytes);
                        const uint64_t a_hi = ptx::make_smem_desc(sa, 0, 1024, ptx::kLayoutSwizzle128B);
                        const uint64_t a_lo = ptx::make_smem_desc(sa + kABytes, 0, 1024, ptx::kLayoutSwizzle128B);
                        const uint64_t w_hi = ptx::make_smem_desc(sw, 0, 1024, ptx::kLayoutSwizzle128B);
                        const uint64_t w_lo = ptx::make_smem_desc(sw + kWBytes, 0, 1024, ptx::kLayoutSwizzle128B);
#pragma unroll
                        for (int k = 0; k < kBK / 16; ++k) {
                            const uint64_t adv = static_cast<uint64_t>((k * 16 * 2) >> 4);
                            ptx::umma_bf16_pair(d_tmem, a_hi + adv, w_hi + adv, idesc, (kb | k) != 0);
                            ptx::umma_bf16_pair(d_tmem, a_lo + adv, w_hi + adv, idesc, 1);
                            ptx::umma_bf16_pair(d_tmem, a_hi + adv, w_lo + adv, idesc, 1);
                        }
                        ptx::umma_commit_pair(wempty_bar + wslot, 3);
                        if (tn == p.n_tn - 1) ptx::umma_commit_pair(aempty_bar + kb, 3);   // last use of the A slot
                        if (kb == kKB - 1) ptx::umma_commit_pair(tfull_bar + as, 3);
                        if (++wslot == kWSlots) wslot = 0, wph ^= 1;
                    }
                }
            }
        }
    } else if (warp < 2 + kDrainWarps) {
        // ------------------------------------------------------------- drainers
        // accumulator d (row tile it = d / n_tn, column tile tn = d % n_tn) -> PReLU -> statistics + a;
        // strictly in order, so the parity wait on a TMEM stage can never see a later accumulator
        const int ew = warp & 3;                               // TMEM lane quarter this warp may read
        const uint32_t tempty_remote = ptx::mapa(ptx::smem_u32(tempty_bar), 0);
        const float slope = __ldg(p.slope);
        const int n_acc = n_tiles * p.n_tn;
        int it = 0, tn = 0;
        for (int d = 0; d < n_acc; ++d) {
            const int as = d & 1;
            const int rt = 2 * (worker + it * n_workers) + static_cast<int>(rank);
            const bool tile_ok = rt < p.row_tiles;
            const int m = tile_ok ? rt / p.tiles_per_utt : 0, tk = tile_ok ? rt % p.tiles_per_utt : 0;
            const int k = tk * kBM + ew * 32 + lane;
            const bool valid = tile_ok && k < p.K;
            float* arow = p.a + (static_cast<long long>(m) * p.K + k) * p.Cout + tn * kBN;
            ptx::mbar_wait(tfull_bar + as, (d >> 1) & 1);
            ptx::tc_fence_after();
            float ls = 0.f, lss = 0.f;
#pragma unroll 1
            for (int ci = 0; ci < kBN / 32; ++ci) {
                uint32_t r[32];
                ptx::tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + as * kBN + ci * 32, r);
                ptx::tmem_ld_wait();
                if (valid) {
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float t = __uint_as_float(r[j]);
                        v[j] = fmaf(slope, fminf(t, 0.f), fmaxf(t, 0.f));   // PReLU
                    }
                    float l0 = 0.f, l1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
                    for (int j = 0; j < 32; j += 2) {
                        l0 += v[j], l1 += v[j + 1];
                        q0 = fmaf(v[j], v[j], q0), q1 = fmaf(v[j + 1], v[j + 1], q1);
                    }
                    ls += l0 + l1, lss += q0 + q1;
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        ptx::st_global_v8(arow + ci * 32 + 8 * j, __float_as_uint(v[8 * j]), __float_as_uint(v[8 * j + 1]),
                                          __float_as_uint(v[8 * j + 2]), __float_as_uint(v[8 * j + 3]),
                                          __float_as_uint(v[8 * j + 4]), __float_as_uint(v[8 * j + 5]),
                                          __float_as_uint(v[8 * j + 6]), __float_as_uint(v[8 * j + 7]));
                }
                if ((ci & 3) == 3) {   // end of a 128-column group: one partial per (row tile, group, quarter)
                    const float ws = warp_sum(ls), wss = warp_sum(lss);
                    if (lane == 0 && tile_ok)
                        p.out_stat[static_cast<long long>(m) * p.out_P + (tk * p.groups + tn * 2 + (ci >> 2)) * 4 + ew] =
                            make_float2(ws, wss);
                    ls = 0.f, lss = 0.f;
                }
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive_cluster(tempty_remote + as * 8);
            if (++tn == p.n_tn) tn = 0, ++it;
        }
    } else {
        // ------------------------------------------------------------ producers
        // All eight warps fill one unit (row tile it, k-block kb: 128 frames x 64 channels) at a time;
        // thread (quad, s) owns 4 channels of frames s*8 .. s*8+7.  While a row of the current unit is
        // consumed, the same row of the NEXT unit is requested, so 8 rows (256 B) per thread are always
        // in flight, also while the warp waits for its A slot to be released.
        const int pt = (warp - 2 - kDrainWarps) * 32 + lane;   // 0 .. 255
        const int quad = pt & 15, s = pt >> 4;
        const uint32_t full_remote = ptx::mapa(ptx::smem_u32(afull_bar), 0);
        const uint32_t row_off0 = static_cast<uint32_t>(s * kRows) * 128u;
        const uint32_t cs = static_cast<uint32_t>(quad >> 1) * 16u + static_cast<uint32_t>(quad & 1) * 8u;
        const f2_t neg1 = f2_pack(-1.f, -1.f);

        auto ld_row = [&](long long off, bool ok) {
            Row r;
            r.hi = make_uint2(0u, 0u), r.lo = make_uint2(0u, 0u), r.y = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ok) {
                asm volatile("ld.volatile.global.v2.b32 {%0, %1}, [%2];" : "=r"(r.hi.x), "=r"(r.hi.y) : "l"(p.x_hi + off) : "memory");
                asm volatile("ld.volatile.global.v2.b32 {%0, %1}, [%2];" : "=r"(r.lo.x), "=r"(r.lo.y) : "l"(p.x_lo + off) : "memory");
                asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                             : "=f"(r.y.x), "=f"(r.y.y), "=f"(r.y.z), "=f"(r.y.w)
                             : "l"(p.y + off)
                             : "memory");
            }
            return r;
        };
        struct Unit {
            int m, nrows;
            long long ox;      // element offset of (first row, first channel); x and y are both [M*K, 256]
        };
        auto unit = [&](int u) {
            Unit g;
            const int it = u >> 2, kb = u & 3;
            const int rt = 2 * (worker + it * n_workers) + static_cast<int>(rank);
            const bool ok = it < n_tiles && rt < p.row_tiles;
            g.m = ok ? rt / p.tiles_per_utt : 0;
            const int tk = ok ? rt % p.tiles_per_utt : 0;
            const int k0 = tk * kBM + s * kRows;
            g.nrows = ok ? max(0, min(kRows, p.K - k0)) : 0;
            g.ox = (static_cast<long long>(g.m) * p.K + k0) * kCin + kb * kBK + quad * 4;
            return g;
        };

        const int n_units = n_tiles * kKB;
        // Consecutive row tiles of a CTA lie in different utterances (stride 2 * n_workers tiles), so the
        // statistics of the previous block change with every tile: finalise them up front, one tile per
        // warp, instead of a chain of dependent loads and fp64 math in front of every tile.
        for (int t = warp - 2 - kDrainWarps; t < min(n_tiles, kMaxTiles); t += kProdWarps) {
            const int rt = 2 * (worker + t * n_workers) + static_cast<int>(rank);
            const int m = rt < p.row_tiles ? rt / p.tiles_per_utt : 0;
            const float2 st = gln_finalize_warp(p.in_stat + static_cast<long long>(m) * p.in_P, p.in_P, p.in_count);
            if (lane == 0) tstat[t] = make_float2(st.y, -st.y * st.x);
        }
        ptx::named_bar_sync(1, 32 * kProdWarps);
        Unit nx = unit(0);
        Row q[kRows];
#pragma unroll
        for (int t = 0; t < kRows; ++t) q[t] = ld_row(nx.ox + static_cast<long long>(t) * kCin, t < nx.nrows);
        for (int u = 0; u < n_units; ++u) {
            const Unit g = nx;
            nx = unit(u + 1);
            const int it = u >> 2, kb = u & 3;
            float sc, sh;
            if (it < kMaxTiles) {
                const float2 st = tstat[it];
                sc = st.x, sh = st.y;
            } else {
                const float2 st = gln_finalize_warp(p.in_stat + static_cast<long long>(g.m) * p.in_P, p.in_P, p.in_count);
                sc = st.y, sh = -st.y * st.x;
            }
            const int c = kb * kBK + quad * 4;
            const float4 cg = *reinterpret_cast<const float4*>(colc + c);
            const float4 cb = *reinterpret_cast<const float4*>(colc + kCin + c);
            const f2_t cc0 = f2_pack(fmaf(sh, cg.x, cb.x), fmaf(sh, cg.y, cb.y));
            const f2_t cc1 = f2_pack(fmaf(sh, cg.z, cb.z), fmaf(sh, cg.w, cb.w));
            const f2_t sc2 = f2_pack(sc, sc);
            ptx::mbar_wait(aempty_bar + kb, (static_cast<uint32_t>(it) & 1u) ^ 1u);
            const uint32_t hi_base = ptx::smem_u32(smem + kb * kASlotBytes) + row_off0;
#pragma unroll
            for (int o = 0; o < kRows; ++o) {
                const Row r = q[o];
                q[o] = ld_row(nx.ox + static_cast<long long>(o) * kCin, o < nx.nrows);
                // x = hi + lo (exact in fp32), x' = x + sc * y + cc
#if CTN_OPERAND_FP16
                float e0, e1, e2, e3, e4, e5, e6, e7;
                op16_value2(r.hi.x, e0, e1), op16_value2(r.hi.y, e2, e3), op16_value2(r.lo.x, e4, e5), op16_value2(r.lo.y, e6, e7);
                const f2_t h0 = f2_pack(e0, e1), h1 = f2_pack(e2, e3), l0 = f2_pack(e4, e5), l1 = f2_pack(e6, e7);
#else
                const f2_t h0 = f2_pack(__uint_as_float(r.hi.x << 16), __uint_as_float(r.hi.x & 0xffff0000u));
                const f2_t h1 = f2_pack(__uint_as_float(r.hi.y << 16), __uint_as_float(r.hi.y & 0xffff0000u));
                const f2_t l0 = f2_pack(__uint_as_float(r.lo.x << 16), __uint_as_float(r.lo.x & 0xffff0000u));
                const f2_t l1 = f2_pack(__uint_as_float(r.lo.y << 16), __uint_as_float(r.lo.y & 0xffff0000u));
#endif
                f2_t v0 = f2_add(f2_fma(sc2, f2_pack(r.y.x, r.y.y), cc0), f2_add(h0, l0));
                f2_t v1 = f2_add(f2_fma(sc2, f2_pack(r.y.z, r.y.w), cc1), f2_add(h1, l1));
                uint32_t nh[2] = {0u, 0u}, nl[2] = {0u, 0u};
                if (o < g.nrows) {
                    f2_split_bf16(v0, neg1, nh[0], nl[0]);
                    f2_split_bf16(v1, neg1, nh[1], nl[1]);
                    const long long off = g.ox + static_cast<long long>(o) * kCin;
                    asm volatile("st.global.v2.b32 [%0], {%1, %2};" ::"l"(p.x_hi + off), "r"(nh[0]), "r"(nh[1]) : "memory");
                    asm volatile("st.global.v2.b32 [%0], {%1, %2};" ::"l"(p.x_lo + off), "r"(nl[0]), "r"(nl[1]) : "memory");
                }
                const uint32_t ad = ((hi_base + cs) ^ (static_cast<uint32_t>(o & 7) << 4)) + static_cast<uint32_t>(o) * 128u;
                asm volatile("st.volatile.shared.v2.b32 [%0], {%1, %2};" ::"r"(ad), "r"(nh[0]), "r"(nh[1]) : "memory");
                asm volatile("st.volatile.shared.v2.b32 [%0], {%1, %2};" ::"r"(ad + kABytes), "r"(nl[0]), "r"(nl[1]) : "memory");
            }
            ptx::fence_async_smem();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive_cluster(full_remote + kb * 8);
        }
    }

    ptx::tc_fence_before();
    ptx::cluster_sync();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc_pair(tmem_base, 2 * kBN);
    }
}

}  // namespace c1
}  // namespace ctn

using namespace ctn;

extern "C" int ctn_updc1_supported(int B, int H, int norm) {
    return norm == CTN_NORM_GLN && B == c1::kCin && H >= c1::kBN && H <= 4 * c1::kBN && H % c1::kBN == 0;
}

extern "C" int ctn_updc1_fwd(const ctn_updc1_args* args, void* stream_) {
    CTN_REQUIRE(args != nullptr, "ctn_updc1_fwd: args is NULL");
    const ctn_updc1_args& a = *args;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    CTN_REQUIRE(a.M > 0 && a.K > 0, "ctn_updc1_fwd: non-positive shape M=%d K=%d", a.M, a.K);
    CTN_REQUIRE(ctn_updc1_supported(a.B, a.H, CTN_NORM_GLN), "ctn_updc1_fwd: needs B == 256 and H in {256, 512, 768, 1024} (B=%d H=%d)", a.B,
                a.H);
    CTN_REQUIRE(static_cast<long long>(a.M) * a.K < (1ll << 31), "ctn_updc1_fwd: M*K exceeds 2^31 rows");
    CTN_REQUIRE(a.x_hi && a.x_lo && a.y && a.in_stat && a.in_P > 0 && a.in_count > 0 && a.col_g && a.col_bias && a.w_hi &&
                    a.w_lo && a.a && a.out_stat && a.slope,
                "ctn_updc1_fwd: NULL argument");
    auto aligned = [](const void* p, int n) { return (reinterpret_cast<uintptr_t>(p) & (n - 1)) == 0; };
    CTN_REQUIRE(aligned(a.x_hi, 16) && aligned(a.x_lo, 16) && aligned(a.y, 16) && aligned(a.a, 32) && aligned(a.w_hi, 16) &&
                    aligned(a.w_lo, 16) && aligned(a.col_g, 16) && aligned(a.col_bias, 16),
                "ctn_updc1_fwd: buffers must be 16-byte aligned (a: 32)");
    c1::Params p;
    p.K = a.K, p.Cout = a.H, p.n_tn = a.H / c1::kBN;
    p.tiles_per_utt = ceil_div(a.K, c1::kBM);
    p.row_tiles = a.M * p.tiles_per_utt;
    p.items = ceil_div(p.row_tiles, 2);
    p.groups = ceil_div(a.H, CTN_COL_TILE);
    p.x_hi = static_cast<__nv_bfloat16*>(a.x_hi), p.x_lo = static_cast<__nv_bfloat16*>(a.x_lo);
    p.y = a.y, p.in_stat = static_cast<const float2*>(a.in_stat), p.in_P = a.in_P, p.in_count = a.in_count;
    p.col_g = a.col_g, p.col_bias = a.col_bias, p.a = a.a, p.out_stat = static_cast<float2*>(a.out_stat);
    p.out_P = ctn_gemm_gln_partials(a.K, a.H);
    p.slope = a.slope;
    CUtensorMap tw_hi, tw_lo;
    int rc;
    if ((rc = make_tmap_bf16_kmajor(&tw_hi, a.w_hi, a.H, a.B, c1::kBN / 2)) != CTN_OK) return rc;
    if ((rc = make_tmap_bf16_kmajor(&tw_lo, a.w_lo, a.H, a.B, c1::kBN / 2)) != CTN_OK) return rc;
    static bool attr_done[kMaxDevices] = {};
    bool& attr_set = attr_done[device_slot()];
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(c1::updc1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, c1::kSmemBytes);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(updc1_kernel)");
        attr_set = true;
    }
    const int workers_max = sm_count() / 2;
    const int workers = p.items < workers_max ? p.items : workers_max;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * workers, 1, 1);
    cfg.blockDim = dim3(c1::kThreads, 1, 1);
    cfg.dynamicSmemBytes = c1::kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, c1::updc1_kernel, tw_hi, tw_lo, p);
    if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx(updc1_kernel)");
    return CTN_OK;
}
