// Deferred residual update fused into the next block's conv1x1 (inference):
//     x <- x + r y - r mu rowsum + W2 beta2         (gLN-2 of the PREVIOUS block, conv_tasnet.py:254-256, :237)
//     a  = PReLU(W1 x),  (sum, sumsq) of a          (conv1x1 of THIS block, :217-218, statistics for :219)
// ctn_dwpw_fwd leaves the previous block's pointwise product y un-normalised because gLN-2 needs the
// statistics of the whole utterance; here they are complete.  Replaces ctn_resid_update + ctn_gemm1x1:
// 6 KB -> 5 KB per frame and block, one launch fewer.
//
// CTA pair (cta_group::2); per CTA a row tile of 128 frames whose whole reduction dimension (B = 256: four
// k-block slots of (hi, lo), 128 KB) stays resident while BOTH 256-column tiles of Cout = 512 are multiplied.
// Nothing on the residual stream passes through the LSU as a global access:
//   x in   one thread TMA-loads the (hi, lo) tiles of a k-block straight into its A slot (128B-swizzled,
//          K-major: the tcgen05 operand layout) as soon as the MMAs of the previous row tile released it;
//   fix-up eight warps rewrite the slot in place: x += sc y + cc, re-split; y streams through a 16-row
//          register ring (two units ahead);
//   x out  one thread TMA-stores the rewritten slot back to the residual stream (3-D maps clip at K);
//   MMA    one thread, three MMAs per k-step, two TMEM accumulator stages; weights by TMA (2 slots);
//   drain  four warps: TMEM -> PReLU -> statistics -> 4 KB staging buffer -> TMA store of `a`.
// Measured per 64 utterances (tools/bench_updc1.py): 203 us against 99 + 155..162 us for the two kernels it
// replaces; register-ring loads of x and 256-bit global stores (the first version) gave 243 us: the loads
// of a latency-bound ring and the scattered stores queue behind each other in the LSU, and every load's
// latency stretches to several microseconds once the stores saturate the memory system.
// Warps: 0 TMA (weights), 1 MMA issuer, 2-5 drain, 6-13 fix-up, 14 x in, 15 x out.
#include <cuda.h>
#include <stdlib.h>

#include <type_traits>

#include "common.cuh"
#include "f32x2.cuh"
#include "ptx.cuh"

namespace ctn {

int make_tmap_bf16_kmajor(CUtensorMap* tm, const void* ptr, long long rows, int cols, int box_rows);  // gemm.cu
int make_tmap_out(CUtensorMap* tm, const void* ptr, int M, int K, int C, bool f32);                     // gemm.cu
int make_tmap_bf16_3d(CUtensorMap* tm, const void* ptr, int M, int K, int C, int box_rows);             // gemm.cu

namespace c1 {

// -DCTN_TRACE=1 (experiment builds only): CTA 0 records clock64() at the pipeline events of each of its row tiles
// in a device array that ctn_debug_trace_read() copies out (tools/bench_updc1.py prints the timeline).
#if CTN_TRACE
__device__ long long g_trace[64 * 32];
#define C1_TRACE(it, ev)                                                              \
    do {                                                                              \
        if (blockIdx.x == 0 && (it) < 64) g_trace[(it) * 32 + (ev)] = clock64();      \
    } while (0)
#else
#define C1_TRACE(it, ev) \
    do {                 \
    } while (0)
#endif

constexpr int kBM = 128;
constexpr int kBN = 256;             // output columns per accumulator
constexpr int kBK = 64;
constexpr int kCin = 256;            // B: the whole reduction dimension is resident
constexpr int kKB = kCin / kBK;      // 4 k-blocks = 4 resident A slots
constexpr int kWSlots = 2;
constexpr int kABytes = kBM * kBK * 2;           // one of hi / lo
constexpr int kWBytes = (kBN / 2) * kBK * 2;     // this CTA's half of a weight k-block, one of hi / lo
constexpr int kASlotBytes = 2 * kABytes;
constexpr int kWSlotBytes = 2 * kWBytes;
constexpr int kTileBytes = kKB * kASlotBytes + kWSlots * kWSlotBytes;
constexpr int kBarBytes = 256;
constexpr int kColBytes = 2 * kCin * 4;          // col_g, col_bias of the previous pointwise conv
constexpr int kMaxTiles = 32;                    // row tiles per CTA whose gLN-2 scale / shift are precomputed
constexpr int kDrainWarps = 4;
constexpr int kStgChunk = 32 * 32 * 4;           // one drained chunk: 32 frames x 32 columns of fp32
constexpr int kStgBytes = kDrainWarps * 2 * kStgChunk;   // two per drain warp: the TMA store of one reads while the next fills
constexpr int kSmemBytes = kTileBytes + kStgBytes + kBarBytes + kColBytes + kMaxTiles * 8;
constexpr int kProdWarps = 8;
constexpr int kFirstProd = 2 + kDrainWarps;      // warps 0 TMA (weights), 1 MMA, 2-5 drain, 6-13 fix-up, 14 x in, 15 x out
constexpr int kInWarp = kFirstProd + kProdWarps;
constexpr int kOutWarp = kInWarp + 1;
constexpr int kThreads = 32 * (kOutWarp + 1);
constexpr int kRows = kBM / (kProdWarps * 2);   // frames a fix-up thread handles per unit (a warp spans 2 frame groups)

struct Params {
    int K, Cout, n_tn, tiles_per_utt, row_tiles, items, groups;
    uint32_t tpu_magic;      // ceil(2^32 / tiles_per_utt) (0: one tile per utterance): no integer division per unit
    const float* y;
    const float2* in_stat;   // gLN-2 partials of the previous block
    int in_P;
    double in_count;
    const float* col_g;
    const float* col_bias;
    float2* out_stat;        // out: gLN-1 partials, ctn_gemm1x1 layout
    int out_P;
    const float* slope;
    int store_raw;           // `a` leaves before the PReLU (training); the statistics are those of PReLU(a)
    int dbg;   // CTN_C1_DBG experiment bits (0 in production): 1 no `a` stores, 2 no x stores, 4 no MMAs, 8 no y loads
};

// The residual stream never passes through registers on its way in or out: the x-io thread TMA-loads the
// (hi, lo) tiles of a k-block straight into the A slot (128B-swizzled, K-major: the tcgen05 operand layout) as
// soon as the MMAs of the previous row tile have released it - up to a whole row tile ahead of its use -,
// the fix-up warps rewrite the slot in place (x += sc y + cc, re-split), and the same thread TMA-stores the
// slot back to the residual stream.  Only y goes through a register ring (16 rows = two units ahead).  The
// LSU sees no global stores at all: `a` leaves through per-warp staging buffers and TMA stores as well.
__global__ void __launch_bounds__(kThreads, 1)
updc1_kernel(const __grid_constant__ CUtensorMap tm_w_hi, const __grid_constant__ CUtensorMap tm_w_lo,
             const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_x_hi,
             const __grid_constant__ CUtensorMap tm_x_lo, const __grid_constant__ CUtensorMap tm_xo_hi,
             const __grid_constant__ CUtensorMap tm_xo_lo, const Params p) {
    const uint32_t rank = ptx::cluster_ctarank();
    const int worker = static_cast<int>(blockIdx.x >> 1);
    const int n_workers = static_cast<int>(gridDim.x >> 1);
    extern __shared__ __align__(1024) uint8_t smem[];
    if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
    uint8_t* wsm = smem + kKB * kASlotBytes;
    uint8_t* stg = smem + kTileBytes;
    uint64_t* afull_bar = reinterpret_cast<uint64_t*>(smem + kTileBytes + kStgBytes);   // fixed-up A slot (leader's: both CTAs)
    uint64_t* aempty_bar = afull_bar + kKB;     // slot released: MMAs of the last column tile + the x store have read it
    uint64_t* araw_bar = aempty_bar + kKB;      // raw (hi, lo) tiles have landed
    uint64_t* xdone_bar = araw_bar + kKB;       // all fix-up warps of this CTA have rewritten the slot
    uint64_t* wfull_bar = xdone_bar + kKB;
    uint64_t* wempty_bar = wfull_bar + kWSlots;
    uint64_t* tfull_bar = wempty_bar + kWSlots;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
    float* colc = reinterpret_cast<float*>(smem + kTileBytes + kStgBytes + kBarBytes);   // [2][kCin]
    float2* tstat = reinterpret_cast<float2*>(smem + kTileBytes + kStgBytes + kBarBytes + kColBytes);   // [kMaxTiles] (rstd, -rstd mean)

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    if (threadIdx.x == 0) C1_TRACE(0, 28);     // kernel entry
    for (int i = threadIdx.x; i < kCin; i += kThreads) {
        colc[i] = __ldg(p.col_g + i);
        colc[kCin + i] = __ldg(p.col_bias + i);
    }
    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tm_w_hi);
        ptx::prefetch_tmap(&tm_w_lo);
        ptx::prefetch_tmap(&tm_a);
        ptx::prefetch_tmap(&tm_x_hi);
        ptx::prefetch_tmap(&tm_x_lo);
        ptx::prefetch_tmap(&tm_xo_hi);
        ptx::prefetch_tmap(&tm_xo_lo);
        for (int i = 0; i < kKB; ++i) {
            ptx::mbar_init(afull_bar + i, 2 * kProdWarps);   // one arrival per fix-up warp, both CTAs
            ptx::mbar_init(aempty_bar + i, 2);               // MMA commit + x store
            ptx::mbar_init(araw_bar + i, 1);
            ptx::mbar_init(xdone_bar + i, kProdWarps);
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
                for (int step = 0; step < p.n_tn * kKB; ++step) {
                    const int tn = step / kKB, kb = step % kKB;
                    ptx::mbar_wait(wempty_bar + slot, phase ^ 1);
                    uint8_t* st = wsm + slot * kWSlotBytes;
                    // (the lo tiles are read by the correction MMAs only: not loaded by one-MMA builds)
                    if (rank == 0) ptx::mbar_expect_tx(wfull_bar + slot, (CTN_MMA_TERMS == 3 ? 4 : 2) * kWBytes);
                    const int n0 = tn * kBN + static_cast<int>(rank) * (kBN / 2);
                    ptx::tma_load_2d_pair(st, &tm_w_hi, fb0 + slot * 8, kb * kBK, n0);
                    if (CTN_MMA_TERMS == 3) ptx::tma_load_2d_pair(st + kWBytes, &tm_w_lo, fb0 + slot * 8, kb * kBK, n0);
                    if (++slot == kWSlots) slot = 0, phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        // -------------------------------------------------- MMA issuer (leader CTA)
        if (lane == 0 && rank == 0) {
            constexpr uint32_t idesc = ptx::make_idesc_bf16(2 * kBM, kBN, 0, 0);
            int wslot = 0;
            uint32_t wph = 0;
            for (int it = 0; it < n_tiles; ++it) {
                const uint32_t aph = it & 1;
                for (int step = 0; step < p.n_tn * kKB; ++step) {
                    const int tn = step / kKB, kb = step % kKB;
                    const int d = it * p.n_tn + tn;   // accumulator number: TMEM stage d & 1
                    const int as = d & 1;
                    if (kb == 0) {                    // first MMA into this stage: its previous contents must be drained
                        ptx::mbar_wait(tempty_bar + as, ((d >> 1) & 1) ^ 1);
                        ptx::tc_fence_after();
                    }
                    const uint32_t d_tmem = tmem_base + as * kBN;
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
                    for (int k = 0; k < ((p.dbg & 4) ? 0 : kBK / 16); ++k) {
                        const uint64_t adv = static_cast<uint64_t>((k * 16 * 2) >> 4);
                        ptx::umma_bf16_pair(d_tmem, a_hi + adv, w_hi + adv, idesc, (kb | k) != 0);
                        if (CTN_MMA_TERMS == 3) {
                            ptx::umma_bf16_pair(d_tmem, a_lo + adv, w_hi + adv, idesc, 1);
                            ptx::umma_bf16_pair(d_tmem, a_hi + adv, w_lo + adv, idesc, 1);
                        }
                    }
                    C1_TRACE(it, 16 + step);
                    ptx::umma_commit_pair(wempty_bar + wslot, 3);
                    if (tn == p.n_tn - 1) ptx::umma_commit_pair(aempty_bar + kb, 3);   // last use of the A slot
                    if (kb == kKB - 1) ptx::umma_commit_pair(tfull_bar + as, 3);
                    if (++wslot == kWSlots) wslot = 0, wph ^= 1;
                }
            }
        }
    } else if (warp == kInWarp) {
        // ------------------------------------------------ x in (TMA): the raw (hi, lo) tiles of a k-block, as soon
        // as its slot is free.  Rows past the end of the utterance are zero-filled (3-D map [M, K, C]).
        if (lane == 0) {
            for (int it = 0; it < n_tiles; ++it) {
                const int rt = 2 * (worker + it * n_workers) + static_cast<int>(rank);
                const bool tile_ok = rt < p.row_tiles;
                // a pair's odd leftover row tile does not exist: coordinates past the last utterance (zero fill)
                const int m = tile_ok ? rt / p.tiles_per_utt : 0x7fffff, tk = tile_ok ? rt % p.tiles_per_utt : 0;
                const uint32_t ph = static_cast<uint32_t>(it) & 1u;
                for (int kb = 0; kb < kKB; ++kb) {
                    ptx::mbar_wait(aempty_bar + kb, ph ^ 1u);
                    uint8_t* sl = smem + kb * kASlotBytes;
                    ptx::mbar_expect_tx(araw_bar + kb, 2 * kABytes);
                    ptx::tma_load_3d(sl, &tm_x_hi, araw_bar + kb, kb * kBK, tk * kBM, m);
                    ptx::tma_load_3d(sl + kABytes, &tm_x_lo, araw_bar + kb, kb * kBK, tk * kBM, m);
                    C1_TRACE(it, kb);
                }
            }
        }
    } else if (warp == kOutWarp) {
        // ------------------------------------------------ x out (TMA): the rewritten slots, as soon as the fix-up
        // warps are done with them; frames past the end of the utterance are clipped by the map.  A slot is
        // released (together with the MMAs' commit) once the store has read it.
        if (lane == 0) {
            for (int it = 0; it < n_tiles; ++it) {
                const int rt = 2 * (worker + it * n_workers) + static_cast<int>(rank);
                const bool tile_ok = rt < p.row_tiles;
                const int m = tile_ok ? rt / p.tiles_per_utt : 0, tk = tile_ok ? rt % p.tiles_per_utt : 0;
                const uint32_t ph = static_cast<uint32_t>(it) & 1u;
                for (int kb = 0; kb < kKB; ++kb) {
                    ptx::mbar_wait(xdone_bar + kb, ph);
                    if (tile_ok && !(p.dbg & 2)) {
                        const uint32_t sl = ptx::smem_u32(smem + kb * kASlotBytes);
                        ptx::tma_store_3d_addr(&tm_xo_hi, sl, kb * kBK, tk * kBM, m);
                        ptx::tma_store_3d_addr(&tm_xo_lo, sl + kABytes, kb * kBK, tk * kBM, m);
                        ptx::tma_store_commit();
                    }
                    C1_TRACE(it, 12 + kb);
                    if (kb & 1) {
                        ptx::tma_store_wait_read<0>();
                        ptx::mbar_arrive(aempty_bar + kb - 1);
                        ptx::mbar_arrive(aempty_bar + kb);
                    }
                }
            }
            ptx::tma_store_wait_all<0>();
        }
    } else if (warp < kFirstProd) {
        // ------------------------------------------------------------- drainers
        // accumulator d (row tile it = d / n_tn, column tile tn = d % n_tn) -> PReLU -> statistics + a;
        // strictly in order, so the parity wait on a TMEM stage can never see a later accumulator
        const int ew = warp & 3;                               // TMEM lane quarter this warp may read
        const uint32_t tempty_remote = ptx::mapa(ptx::smem_u32(tempty_bar), 0);
        const float slope = __ldg(p.slope);
        const int n_acc = n_tiles * p.n_tn;
        const uint32_t stg_base = ptx::smem_u32(stg) + static_cast<uint32_t>(warp - 2) * (2 * kStgChunk);
        int it = 0, tn = 0, chunk = 0;
        for (int d = 0; d < n_acc; ++d) {
            const int as = d & 1;
            const int rt = 2 * (worker + it * n_workers) + static_cast<int>(rank);
            const bool tile_ok = rt < p.row_tiles;
            const int m = tile_ok ? rt / p.tiles_per_utt : 0, tk = tile_ok ? rt % p.tiles_per_utt : 0;
            const int k = tk * kBM + ew * 32 + lane;
            const bool valid = tile_ok && k < p.K;
            ptx::mbar_wait(tfull_bar + as, (d >> 1) & 1);
            ptx::tc_fence_after();
            if (warp == 2 && lane == 0) C1_TRACE(it, 24 + 2 * tn);
            float ls = 0.f, lss = 0.f;
            const bool rows_ok = tile_ok && tk * kBM + ew * 32 < p.K;   // any of this warp's 32 frames exists
#pragma unroll 1
            for (int ci = 0; ci < kBN / 32; ++ci, ++chunk) {
                uint32_t r[32];
                ptx::tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + as * kBN + ci * 32, r);
                // the staging buffer about to be refilled: its previous TMA store (two chunks ago) must have read it
                if (lane == 0) ptx::tma_store_wait_read<1>();
                __syncwarp();
                ptx::tmem_ld_wait();
                float v[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const float t = __uint_as_float(r[j]);
                    v[j] = fmaf(slope, fminf(t, 0.f), fmaxf(t, 0.f));   // PReLU
                }
                const bool raw = p.store_raw != 0;
                if (valid) {
                    float l0 = 0.f, l1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
                    for (int j = 0; j < 32; j += 2) {
                        l0 += v[j], l1 += v[j + 1];
                        q0 = fmaf(v[j], v[j], q0), q1 = fmaf(v[j + 1], v[j + 1], q1);
                    }
                    ls += l0 + l1, lss += q0 + q1;
                }
                // lane = frame: its 128 B row of the chunk, 16 B pieces XOR-swizzled with (row & 7) as the
                // 128B-swizzled tensor map expects (also makes the warp's stores conflict-free)
                const uint32_t sb = stg_base + static_cast<uint32_t>(chunk & 1) * kStgChunk + static_cast<uint32_t>(lane) * 128u;
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    ptx::st_shared_v4(sb + ((static_cast<uint32_t>(j) ^ (static_cast<uint32_t>(lane) & 7u)) << 4),
                                      raw ? r[4 * j] : __float_as_uint(v[4 * j]), raw ? r[4 * j + 1] : __float_as_uint(v[4 * j + 1]),
                                      raw ? r[4 * j + 2] : __float_as_uint(v[4 * j + 2]),
                                      raw ? r[4 * j + 3] : __float_as_uint(v[4 * j + 3]));
                ptx::fence_async_smem();
                __syncwarp();
                if (lane == 0 && rows_ok && !(p.dbg & 1)) {   // frames past K are clipped by the map
                    ptx::tma_store_3d_addr(&tm_a, stg_base + static_cast<uint32_t>(chunk & 1) * kStgChunk, tn * kBN + ci * 32,
                                           tk * kBM + ew * 32, m);
                    ptx::tma_store_commit();
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
            if (warp == 2 && lane == 0) C1_TRACE(it, 25 + 2 * tn);
            if (++tn == p.n_tn) tn = 0, ++it;
        }
        if (lane == 0) ptx::tma_store_wait_all<0>();   // the staging buffers must outlive their stores
    } else {
        // ------------------------------------------------------------ fix-up warps
        // All eight warps rewrite one unit (row tile it, k-block kb: 128 frames x 64 channels) at a time;
        // thread (quad, s) owns 4 channels of frames s*8 .. s*8+7.  x (hi, lo) is read from and written back
        // to the A slot; y streams through a 16-row register ring: while a row of unit u is consumed, the
        // same row of unit u + 2 is requested.
        const int pt = (warp - kFirstProd) * 32 + lane;   // 0 .. 255
        const int quad = pt & 15, s = pt >> 4;
        const uint32_t full_remote = ptx::mapa(ptx::smem_u32(afull_bar), 0);
        const uint32_t row_off0 = static_cast<uint32_t>(s * kRows) * 128u;
        const uint32_t cs = static_cast<uint32_t>(quad >> 1) * 16u + static_cast<uint32_t>(quad & 1) * 8u;
        const f2_t neg1 = f2_pack(-1.f, -1.f);

        auto ld_y = [&](long long off, bool ok) {
            float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ok && !(p.dbg & 8))
                asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                             : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                             : "l"(p.y + off)
                             : "memory");
            return r;
        };
        struct Unit {
            int nrows;
            long long oy;      // element offset of (first row, first channel) in y [M*K, 256]
        };
        auto unit = [&](int u) {
            Unit g;
            const int it = u >> 2, kb = u & 3;
            const int rt = 2 * (worker + it * n_workers) + static_cast<int>(rank);
            const bool ok = it < n_tiles && rt < p.row_tiles;
            const int mq = p.tpu_magic == 0u ? rt : static_cast<int>(__umulhi(static_cast<uint32_t>(rt), p.tpu_magic));
            const int m = ok ? mq : 0;
            const int tk = ok ? rt - mq * p.tiles_per_utt : 0;
            const int k0 = tk * kBM + s * kRows;
            g.nrows = ok ? max(0, min(kRows, p.K - k0)) : 0;
            g.oy = (static_cast<long long>(m) * p.K + k0) * kCin + kb * kBK + quad * 4;
            return g;
        };

        const int n_units = n_tiles * kKB;
        // Consecutive row tiles of a CTA lie in different utterances (stride 2 * n_workers tiles), so the
        // statistics of the previous block change with every tile: finalise them up front, one tile per
        // warp, instead of a chain of dependent loads and fp64 math in front of every tile.
        for (int t = warp - kFirstProd; t < min(n_tiles, kMaxTiles); t += kProdWarps) {
            const int rt = 2 * (worker + t * n_workers) + static_cast<int>(rank);
            const int m = rt < p.row_tiles ? rt / p.tiles_per_utt : 0;
            const float2 st = gln_finalize_warp(p.in_stat + static_cast<long long>(m) * p.in_P, p.in_P, p.in_count);
            if (lane == 0) tstat[t] = make_float2(st.y, -st.y * st.x);
        }
        ptx::named_bar_sync(1, 32 * kProdWarps);
        float4 q[2][kRows];
        Unit g0 = unit(0), g1 = unit(1);
#pragma unroll
        for (int t = 0; t < kRows; ++t) q[0][t] = ld_y(g0.oy + static_cast<long long>(t) * kCin, t < g0.nrows);
#pragma unroll
        for (int t = 0; t < kRows; ++t) q[1][t] = ld_y(g1.oy + static_cast<long long>(t) * kCin, t < g1.nrows);

        // unit u with its y rows in ring half `par` (a compile-time index: the ring stays in registers)
        auto do_unit = [&](int u, auto par_tag) {
            constexpr int par = decltype(par_tag)::value;
            const Unit nx = unit(u + 2);
            const int it = u >> 2, kb = u & 3;
            float sc, sh;
            if (it < kMaxTiles) {
                const float2 st = tstat[it];
                sc = st.x, sh = st.y;
            } else {
                const int rt = 2 * (worker + it * n_workers) + static_cast<int>(rank);
                const int m = rt < p.row_tiles ? rt / p.tiles_per_utt : 0;
                const float2 st = gln_finalize_warp(p.in_stat + static_cast<long long>(m) * p.in_P, p.in_P, p.in_count);
                sc = st.y, sh = -st.y * st.x;
            }
            const int c = kb * kBK + quad * 4;
            const float4 cg = *reinterpret_cast<const float4*>(colc + c);
            const float4 cb = *reinterpret_cast<const float4*>(colc + kCin + c);
            const f2_t cc0 = f2_pack(fmaf(sh, cg.x, cb.x), fmaf(sh, cg.y, cb.y));
            const f2_t cc1 = f2_pack(fmaf(sh, cg.z, cb.z), fmaf(sh, cg.w, cb.w));
            const f2_t sc2 = f2_pack(sc, sc);
            ptx::mbar_wait(araw_bar + kb, static_cast<uint32_t>(it) & 1u);
            if (pt == 0) C1_TRACE(it, 4 + kb);
            const uint32_t hi_base = ptx::smem_u32(smem + kb * kASlotBytes) + row_off0;
#pragma unroll
            for (int o = 0; o < kRows; ++o) {
                const uint32_t ad = ((hi_base + cs) ^ (static_cast<uint32_t>(o & 7) << 4)) + static_cast<uint32_t>(o) * 128u;
                uint32_t xh0, xh1, xl0, xl1;
                asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(xh0), "=r"(xh1) : "r"(ad) : "memory");
                asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(xl0), "=r"(xl1) : "r"(ad + kABytes) : "memory");
                const float4 yv = q[par][o];
                // x = hi + lo (exact in fp32), x' = x + sc * y + cc
#if CTN_OPERAND_FP16
                float e0, e1, e2, e3, e4, e5, e6, e7;
                op16_value2(xh0, e0, e1), op16_value2(xh1, e2, e3), op16_value2(xl0, e4, e5), op16_value2(xl1, e6, e7);
                const f2_t h0 = f2_pack(e0, e1), h1 = f2_pack(e2, e3), l0 = f2_pack(e4, e5), l1 = f2_pack(e6, e7);
#else
                const f2_t h0 = f2_pack(__uint_as_float(xh0 << 16), __uint_as_float(xh0 & 0xffff0000u));
                const f2_t h1 = f2_pack(__uint_as_float(xh1 << 16), __uint_as_float(xh1 & 0xffff0000u));
                const f2_t l0 = f2_pack(__uint_as_float(xl0 << 16), __uint_as_float(xl0 & 0xffff0000u));
                const f2_t l1 = f2_pack(__uint_as_float(xl1 << 16), __uint_as_float(xl1 & 0xffff0000u));
#endif
                const f2_t v0 = f2_add(f2_fma(sc2, f2_pack(yv.x, yv.y), cc0), f2_add(h0, l0));
                const f2_t v1 = f2_add(f2_fma(sc2, f2_pack(yv.z, yv.w), cc1), f2_add(h1, l1));
                // refill AFTER the last use of the ring slot, so that the load targets the registers it frees
                q[par][o] = ld_y(nx.oy + static_cast<long long>(o) * kCin, o < nx.nrows);
                uint32_t nh0, nl0, nh1, nl1;
                f2_split_bf16(v0, neg1, nh0, nl0);
                f2_split_bf16(v1, neg1, nh1, nl1);
                asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(ad), "r"(nh0), "r"(nh1) : "memory");
                asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(ad + kABytes), "r"(nl0), "r"(nl1) : "memory");
            }
            ptx::fence_async_smem();      // generic-proxy stores -> visible to the tensor core and the TMA store
            __syncwarp();
            if (lane == 0) {
                ptx::mbar_arrive_cluster(full_remote + kb * 8);
                ptx::mbar_arrive(xdone_bar + kb);
            }
            if (pt == 0) C1_TRACE(it, 8 + kb);
        };
        for (int u = 0; u < n_units; u += 2) {
            do_unit(u, std::integral_constant<int, 0>{});
            if (u + 1 < n_units) do_unit(u + 1, std::integral_constant<int, 1>{});
        }
    }

    ptx::tc_fence_before();
    ptx::cluster_sync();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc_pair(tmem_base, 2 * kBN);
        if (lane == 0) C1_TRACE(0, 29);        // kernel exit
    }
}

}  // namespace c1
}  // namespace ctn

using namespace ctn;

#if CTN_TRACE
extern "C" int ctn_debug_trace_read(long long* host, int n) {
    return cudaMemcpyFromSymbol(host, c1::g_trace, sizeof(long long) * n) == cudaSuccess ? CTN_OK : CTN_ERR_CUDA;
}
#endif

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
    // exact for row tile numbers < 2^20 and at most 2^12 tiles per utterance (see fused_block.cu: div_magic)
    CTN_REQUIRE(p.tiles_per_utt <= 4096 && p.row_tiles + 2 * sm_count() < (1 << 20),
                "ctn_updc1_fwd: batch too large for the tile index arithmetic (M=%d K=%d)", a.M, a.K);
    p.tpu_magic = p.tiles_per_utt == 1 ? 0u : static_cast<uint32_t>(((1ull << 32) + p.tiles_per_utt - 1) / p.tiles_per_utt);
    p.groups = ceil_div(a.H, CTN_COL_TILE);
    p.y = a.y, p.in_stat = static_cast<const float2*>(a.in_stat), p.in_P = a.in_P, p.in_count = a.in_count;
    p.col_g = a.col_g, p.col_bias = a.col_bias, p.out_stat = static_cast<float2*>(a.out_stat);
    p.out_P = ctn_gemm_gln_partials(a.K, a.H);
    p.slope = a.slope;
    p.store_raw = a.store_raw;
    {
        const char* d = getenv("CTN_C1_DBG");
        p.dbg = d ? atoi(d) : 0;
    }
    CUtensorMap tw_hi, tw_lo;
    int rc;
    if ((rc = make_tmap_bf16_kmajor(&tw_hi, a.w_hi, a.H, a.B, c1::kBN / 2)) != CTN_OK) return rc;
    if ((rc = make_tmap_bf16_kmajor(&tw_lo, a.w_lo, a.H, a.B, c1::kBN / 2)) != CTN_OK) return rc;
    CUtensorMap ta;   // [M, K, H] fp32, box 32 columns x 32 frames of one utterance, 128B swizzle
    if ((rc = make_tmap_out(&ta, a.a, a.M, a.K, a.H, true)) != CTN_OK) return rc;
    CUtensorMap tx_hi, tx_lo;   // [M, K, B] bf16, box 64 channels x 128 frames of one utterance, 128B swizzle (in and out)
    if ((rc = make_tmap_bf16_3d(&tx_hi, a.x_hi, a.M, a.K, a.B, c1::kBM)) != CTN_OK) return rc;
    if ((rc = make_tmap_bf16_3d(&tx_lo, a.x_lo, a.M, a.K, a.B, c1::kBM)) != CTN_OK) return rc;
    CTN_REQUIRE((a.x_out_hi == nullptr) == (a.x_out_lo == nullptr), "ctn_updc1_fwd: x_out_hi / x_out_lo go together");
    CUtensorMap txo_hi = tx_hi, txo_lo = tx_lo;   // in place unless an output pair is given
    if (a.x_out_hi != nullptr) {
        CTN_REQUIRE(aligned(a.x_out_hi, 16) && aligned(a.x_out_lo, 16), "ctn_updc1_fwd: x_out must be 16-byte aligned");
        if ((rc = make_tmap_bf16_3d(&txo_hi, a.x_out_hi, a.M, a.K, a.B, c1::kBM)) != CTN_OK) return rc;
        if ((rc = make_tmap_bf16_3d(&txo_lo, a.x_out_lo, a.M, a.K, a.B, c1::kBM)) != CTN_OK) return rc;
    }

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
    cudaError_t e = cudaLaunchKernelEx(&cfg, c1::updc1_kernel, tw_hi, tw_lo, ta, tx_hi, tx_lo, txo_hi, txo_lo, p);
    if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx(updc1_kernel)");
    return CTN_OK;
}
