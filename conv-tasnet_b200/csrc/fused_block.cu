// Depthwise stage of a TemporalBlock fused into the pointwise 1x1 convolution (inference):
//     a -> PReLU -> gLN-1 -> depthwise conv (P = 3, dilation d) -> PReLU -> q     conv_tasnet.py:218-253
//     y = (W2 diag(gamma2)) q                                                    conv_tasnet.py:254-256
// q never reaches HBM: SIMT producer warps compute 128 frames x 64 channels of q, split it
// into the bf16 (hi, lo) pair and store it 128B-swizzled, K-major into the shared-memory
// stage the tcgen05 MMAs read as their A operand (the layout a TMA load would have written).
//
// gLN-2 normalises q with the statistics of the WHOLE utterance, which exist only when every
// tile of q has been produced.  The normalisation is already an epilogue affine of the folded
// weight (ctn_weight_prep: W(gamma (q - mu) r + beta) = r (W gamma) q - r mu rowsum + W beta), so
// it is deferred: this kernel writes the raw product y and the (sum, sumsq) partials of q, and
// ctn_resid_update applies  x += r y - r mu rowsum + W beta  once the partials are complete.
//
// Tiles are not runs of consecutive frames.  Thread (quad, s) of a producer group owns four
// channels and tile rows 16 s .. 16 s + 15, which are the frames  ph + d (j0 + jj)  of ONE
// dilation phase: the three taps of consecutive outputs are consecutive elements of that
// sequence, so each input row is loaded and activated once and reused from a 3-deep register
// window (18 loads per 16 outputs, for every dilation).  A GEMM row may be any frame: the
// epilogue thread that owns TMEM lane r simply stores to the row of frame(r).
//
// CTA pair (cta_group::2) as in gemm.cu: each CTA owns a 128-row tile and loads half of the
// 256-column weight tile.  Warps: 0 TMA (weights), 1 MMA issuer, 2-13 producers (three groups of
// four warps; group g fills stage g and drains the TMEM accumulators of tiles g, g + 3, ...: there
// are no dedicated epilogue warps).  14 warps = at most 4 per scheduler partition = 128 registers
// per thread, which the producers' load ring needs.
#include <cuda.h>
#include <stdlib.h>

#include <type_traits>

#include "common.cuh"
#include "f32x2.cuh"
#include "ptx.cuh"

namespace ctn {

int make_tmap_bf16_kmajor(CUtensorMap* tm, const void* ptr, long long rows, int cols, int box_rows);  // gemm.cu
int make_tmap_phase_f32(CUtensorMap* tm, const void* ptr, int M, int K, int C, int d, int box_j);       // gemm.cu

namespace fb {

constexpr int kBM = 128;             // tile rows = TMEM lanes
constexpr int kBN = 256;             // output channels (B) per tile = whole pointwise output
constexpr int kBK = 64;              // channels per k-block = one 128 B swizzle atom of bf16
#ifndef CTN_FB_TMA_DRAIN
#define CTN_FB_TMA_DRAIN 1
#endif
constexpr bool kTmaDrain = CTN_FB_TMA_DRAIN != 0;   // y of full row groups leaves through staging + TMA tensor stores
#ifndef CTN_FB_CST_GLOBAL
#define CTN_FB_CST_GLOBAL 1
#endif
// per-channel constants (gamma, beta, taps) read through L1 at the start of a unit instead of from a shared-memory copy:
// together with staging buffers per (TMEM stage, lane quarter) this leaves room for the fourth A slot next to the TMA drain
constexpr bool kCstGlobal = CTN_FB_CST_GLOBAL != 0;
constexpr int kASlots = (kTmaDrain && !kCstGlobal) ? 3 : 4;           // produced A units in flight: more than producer groups, so a group never
                                     // waits for the MMAs to consume its own previous unit
constexpr int kWSlots = 2;           // weight k-blocks in flight (TMA, L2-resident)
constexpr int kABytes = kBM * kBK * 2;           // one of hi / lo
constexpr int kWBytes = (kBN / 2) * kBK * 2;     // this CTA's half of the weight tile, one of hi / lo
constexpr int kASlotBytes = 2 * kABytes;
constexpr int kWSlotBytes = 2 * kWBytes;
constexpr int kTileBytes = kASlots * kASlotBytes + kWSlots * kWSlotBytes;
constexpr int kProdWarps = 12;       // three groups of four warps; they also drain TMEM
constexpr int kCtrlWarps = 4;        // one warpgroup: 0 TMA, 1 MMA, 2-3 idle (they only hand their registers over)
constexpr int kThreads = 32 * (kCtrlWarps + kProdWarps);
constexpr int kCtrlRegs = 56, kProdRegs = 152;   // per scheduler partition: 1 control + 3 producer warps <= 512 x 32 registers
constexpr int kBarBytes = 256;
constexpr int kStgWarp = 32 * 128;      // one 32-column chunk of a warp's 32 rows
constexpr int kStgBufs = kCstGlobal ? 8 : 12;   // per (TMEM stage, lane quarter): two tiles drain at once at most / per warp
constexpr int kStgBytes = kTmaDrain ? kStgBufs * kStgWarp : 0;
constexpr int kMaxTiles = 32;        // row tiles per CTA whose gLN-1 (mean, rstd) are precomputed
constexpr int kSmemFixed = kTileBytes + kBarBytes + kStgBytes + kMaxTiles * 8;   // + 20 H bytes of per-channel constants
constexpr int kSmemMax = 232448;
constexpr int kRowsPerThread = 16;   // outputs a producer thread marches per k-block
constexpr int kPartialsPerTile = kProdWarps;

struct Params {
    int K, H, d, c0;                 // frames, channels, dilation, taps before the output frame
    int w, w_log2, jspan, n_ig, tiles_per_utt, row_tiles, items;
    // the producers' index arithmetic runs once per unit on the critical path of a 3-warp-per-scheduler kernel:
    // no integer division there (20-30 dependent instructions each) - powers of two by shift, the rest by
    // multiplication with a rounded-up reciprocal (exact for the operand ranges asserted on the host)
    int d_log2, n_ig_log2, nkb;
    uint32_t tpu_magic, nkb_magic;
    int in_P, out_P;
    const float* a;
    const float* prelu_in;
    const float* gamma_in;
    const float* beta_in;
    const float2* in_stat;
    const float* wd;
    const float* prelu_out;
    float* y;
    float2* out_stat;
    float* d_out;   // training: depthwise output before the second PReLU (kStoreD instantiations)
};

// x / divisor for 0 <= x < 2^20 with magic = ceil(2^32 / divisor), 2 <= divisor <= 2^12; magic = 0 stands for divisor 1
__device__ __forceinline__ int div_magic(int x, uint32_t magic) {
    return magic == 0u ? x : static_cast<int>(__umulhi(static_cast<uint32_t>(x), magic));
}
// frame of tile row r of tile tt (may be >= K: the row does not exist)
__device__ __forceinline__ int tile_phase(const Params& p, int tt, int s) {
    return (tt & (p.n_ig - 1)) * p.w + (s & (p.w - 1));
}
__device__ __forceinline__ int tile_j0(const Params& p, int tt, int s) {
    return (tt >> p.n_ig_log2) * p.jspan + (s >> p.w_log2) * kRowsPerThread;
}

// kPreact: the input is already PReLU(conv1x1 output) (ctn_gemm1x1 with store_prelu), prelu_in is unused
// kStoreD: the depthwise output before the second PReLU is also written to d_out (the training forward)
template <bool kPreact, bool kStoreD = false>
__global__ void __launch_bounds__(kThreads, 1)
dwpw_kernel(const __grid_constant__ CUtensorMap tm_w_hi, const __grid_constant__ CUtensorMap tm_w_lo,
            const __grid_constant__ CUtensorMap tm_y, const Params p) {
    const uint32_t rank = ptx::cluster_ctarank();
    const int worker = static_cast<int>(blockIdx.x >> 1);
    const int n_workers = static_cast<int>(gridDim.x >> 1);
    extern __shared__ __align__(1024) uint8_t smem[];
    if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
    uint8_t* wsm = smem + kASlots * kASlotBytes;                // weight slots follow the A slots
    uint64_t* afull_bar = reinterpret_cast<uint64_t*>(smem + kTileBytes + kStgBytes);
    uint64_t* aempty_bar = afull_bar + kASlots;
    uint64_t* wfull_bar = aempty_bar + kASlots;
    uint64_t* wempty_bar = wfull_bar + kWSlots;
    uint64_t* tfull_bar = wempty_bar + kWSlots;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int nkb = p.nkb;
    // per-channel constants of the producers: gamma, beta and the three taps, [5][H]
    uint8_t* stg_smem = smem + kTileBytes;                      // [12 warps][32 rows][128 B] (TMA drain), 1024-aligned
    float2* tstat = reinterpret_cast<float2*>(smem + kTileBytes + kBarBytes + kStgBytes);   // [kMaxTiles] (mean, rstd)
    float* cst = reinterpret_cast<float*>(smem + kTileBytes + kBarBytes + kStgBytes + kMaxTiles * 8);
    for (int i = threadIdx.x; i < (kCstGlobal ? 0 : p.H); i += kThreads) {
        cst[i] = __ldg(p.gamma_in + i);
        cst[p.H + i] = __ldg(p.beta_in + i);
        cst[2 * p.H + i] = __ldg(p.wd + 3 * i);
        cst[3 * p.H + i] = __ldg(p.wd + 3 * i + 1);
        cst[4 * p.H + i] = __ldg(p.wd + 3 * i + 2);
    }

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tm_w_hi);
        ptx::prefetch_tmap(&tm_w_lo);
        ptx::prefetch_tmap(&tm_y);
        for (int i = 0; i < kASlots; ++i) {
            ptx::mbar_init(afull_bar + i, 2 * 4);   // one arrival per producer warp of the filling group, both CTAs
            ptx::mbar_init(aempty_bar + i, 1);
        }
        for (int i = 0; i < kWSlots; ++i) {
            ptx::mbar_init(wfull_bar + i, 1);       // leader: weight bytes of both CTAs (one expect_tx arrival)
            ptx::mbar_init(wempty_bar + i, 1);
        }
        for (int i = 0; i < 2; ++i) {
            ptx::mbar_init(tfull_bar + i, 1);
            ptx::mbar_init(tempty_bar + i, 2 * 4);
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
    // PDL: everything above touched shared memory, TMEM and model parameters only (the predecessor, a
    // conv1x1 GEMM, writes activations and statistics); wait for it before reading those
    ptx::grid_dep_sync();

    if (warp < kCtrlWarps) {
    ptx::setmaxnreg_dec<kCtrlRegs>();
    if (warp == 0) {
        // ------------------------------------------------ TMA producer (weights only)
        if (lane == 0) {
            int slot = 0;
            uint32_t phase = 0;
            const uint32_t fb0 = ptx::mapa(ptx::smem_u32(wfull_bar), 0);
            for (int item = worker; item < p.items; item += n_workers) {
                for (int kb = 0; kb < nkb; ++kb) {
                    ptx::mbar_wait(wempty_bar + slot, phase ^ 1);
                    uint8_t* st = wsm + slot * kWSlotBytes;
                    if (rank == 0) ptx::mbar_expect_tx(wfull_bar + slot, (CTN_MMA_TERMS == 3 ? 4 : 2) * kWBytes);
                    ptx::tma_load_2d_pair(st, &tm_w_hi, fb0 + slot * 8, kb * kBK, static_cast<int>(rank) * (kBN / 2));
                    if (CTN_MMA_TERMS == 3)
                        ptx::tma_load_2d_pair(st + kWBytes, &tm_w_lo, fb0 + slot * 8, kb * kBK,
                                              static_cast<int>(rank) * (kBN / 2));
                    if (++slot == kWSlots) slot = 0, phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        // -------------------------------------------------- MMA issuer (leader CTA)
        if (lane == 0 && rank == 0) {
            constexpr uint32_t idesc = ptx::make_idesc_bf16(2 * kBM, kBN, 0, 0);
            int aslot = 0, wslot = 0;
            uint32_t aph = 0, wph = 0;
            int it = 0;
            for (int item = worker; item < p.items; item += n_workers, ++it) {
                const int as = it & 1;
                const uint32_t aphase = (it >> 1) & 1;
                ptx::mbar_wait(tempty_bar + as, aphase ^ 1);
                ptx::tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * kBN;
                for (int kb = 0; kb < nkb; ++kb) {
                    ptx::mbar_wait(wfull_bar + wslot, wph);
                    ptx::mbar_wait(afull_bar + aslot, aph);
                    ptx::tc_fence_after();
                    const uint32_t sa = ptx::smem_u32(smem + aslot * kASlotBytes);
                    const uint32_t sw = ptx::smem_u32(wsm + wslot * kWSlotBytes);
                    const uint64_t a_hi = ptx::make_smem_desc(sa, 0, 1024, ptx::kLayoutSwizzle128B);
                    const uint64_t a_lo = ptx::make_smem_desc(sa + kABytes, 0, 1024, ptx::kLayoutSwizzle128B);
                    const uint64_t w_hi = ptx::make_smem_desc(sw, 0, 1024, ptx::kLayoutSwizzle128B);
                    const uint64_t w_lo = ptx::make_smem_desc(sw + kWBytes, 0, 1024, ptx::kLayoutSwizzle128B);
#pragma unroll
                    for (int k = 0; k < kBK / 16; ++k) {
                        const uint64_t adv = static_cast<uint64_t>((k * 16 * 2) >> 4);
                        ptx::umma_bf16_pair(d_tmem, a_hi + adv, w_hi + adv, idesc, (kb | k) != 0);
                        if (CTN_MMA_TERMS == 3) {
                            ptx::umma_bf16_pair(d_tmem, a_lo + adv, w_hi + adv, idesc, 1);
                            ptx::umma_bf16_pair(d_tmem, a_hi + adv, w_lo + adv, idesc, 1);
                        }
                    }
                    ptx::umma_commit_pair(aempty_bar + aslot, 3);   // frees the slots in BOTH CTAs
                    ptx::umma_commit_pair(wempty_bar + wslot, 3);
                    if (kb == nkb - 1) ptx::umma_commit_pair(tfull_bar + as, 3);
                    if (++aslot == kASlots) aslot = 0, aph ^= 1;
                    if (++wslot == kWSlots) wslot = 0, wph ^= 1;
                }
            }
        }
    }
    } else {
        ptx::setmaxnreg_inc<kProdRegs>();
        // ---------------------------------------------------- q producers (which also drain TMEM)
        // Three groups of four warps; the CTA's units (tile, k-block) are dealt round-robin, unit u to
        // group u % 3, which therefore always fills stage u % 3.  Thread (quad, s) of a group owns four
        // channels and tile rows 16 s .. 16 s + 15 of the unit: 18 input rows, 16 outputs.  Input rows
        // stream through a 9-deep register ring that runs ahead across unit and tile boundaries, so
        // loads are in flight all the time.  Interior units (all 18 rows inside the utterance) take a
        // path without any predicate.  There are no epilogue warps: the four warps of group d % 3
        // (one per TMEM lane quarter) drain the accumulator of tile d between two of their units.
        const int pw = warp - kCtrlWarps;                      // 0 .. 11
        const int grp = pw >> 2;
        const int pt = (pw & 3) * 32 + lane;                   // 0 .. 127 within the group
        const int quad = pt & 15, s = pt >> 4;
        const int ew = warp & 3;                               // TMEM lane quarter this warp may read
        const uint32_t full_remote = ptx::mapa(ptx::smem_u32(afull_bar), 0);
        const uint32_t tempty_remote = ptx::mapa(ptx::smem_u32(tempty_bar), 0);
        const float a1 = kPreact ? 1.f : __ldg(p.prelu_in), a2 = __ldg(p.prelu_out);
        const f2_t a1m1 = f2_pack(a1 - 1.f, a1 - 1.f), a2m1 = f2_pack(a2 - 1.f, a2 - 1.f);
        const f2_t neg1 = f2_pack(-1.f, -1.f), one2 = f2_pack(1.f, 1.f);
        // byte offset of this thread's 8 bytes inside a [128 x 64] bf16 tile, for row 16 s + jj:
        // 128 B rows, 16 B chunks XOR-swizzled with (row & 7)
        const uint32_t row_off0 = static_cast<uint32_t>(s * kRowsPerThread) * 128u;
        const uint32_t cs = static_cast<uint32_t>(quad >> 1) * 16u + static_cast<uint32_t>(quad & 1) * 8u;
        const long long strd = static_cast<long long>(p.d) * p.H;
        constexpr int kRing = (kRowsPerThread + 2) / 2;        // 9

        struct Tile {          // this thread's view of one tile
            const float* base; // row t = 0 of k-block 0 (may point before the buffer)
            int lo, hi;        // rows t in [lo, hi) exist (inside the utterance and needed)
            int n;             // outputs
        };
        auto tile_setup = [&](int item, int& m, int& tt) {
            Tile T;
            const int rt = 2 * item + static_cast<int>(rank);
            const bool ok = item < p.items && rt < p.row_tiles;
            const int mq = div_magic(rt, p.tpu_magic);
            m = ok ? mq : 0;
            tt = ok ? rt - mq * p.tiles_per_utt : -1;
            const int ph = tile_phase(p, max(tt, 0), s);
            const int j0 = tile_j0(p, max(tt, 0), s);
            const int J = ph < p.K ? (p.K - ph + p.d - 1) >> p.d_log2 : 0;   // frames of this phase
            T.n = ok ? max(0, min(kRowsPerThread, J - j0)) : 0;
            const int r0 = j0 - p.c0;                                  // sequence index of window row t = 0
            T.lo = max(0, -r0);
            T.hi = T.n > 0 ? min(T.n + 2, J - r0) : 0;
            T.base = p.a + (static_cast<long long>(m) * p.K + ph + static_cast<long long>(r0) * p.d) * p.H + quad * 4;
            return T;
        };
        // volatile: ptxas otherwise sinks the ring refills down to their first use to save registers,
        // which turns the prefetch ring into synchronous loads
        auto ld_row = [](const float* ptr) {
            float4 v;
            asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                         : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                         : "l"(ptr)
                         : "memory");
            return v;
        };
        auto is_full = [](const Tile& T) { return T.lo == 0 && T.hi == kRowsPerThread + 2; };

        // TMEM accumulator of the CTA's tile number d -> y.  A lane owns one output row and writes its 32
        // columns of a chunk with four 256-bit stores straight from registers.  (Staging the chunk in
        // shared memory and letting the async proxy copy it out relieves the LSU, which paid off while
        // dedicated epilogue warps could wait for the copies; a producer group must not wait.)
        auto drain = [&](int d) {
            const int as = d & 1;
            const uint32_t aphase = (d >> 1) & 1;
            const int rt = 2 * (worker + d * n_workers) + static_cast<int>(rank);
            const bool tile_ok = rt < p.row_tiles;
            const int mq = div_magic(rt, p.tpu_magic);
            const int m = tile_ok ? mq : 0, tt = tile_ok ? rt - mq * p.tiles_per_utt : 0;
            const int r = ew * 32 + lane;
            const int sr = r >> 4;
            const int frame = tile_phase(p, tt, sr) + p.d * (tile_j0(p, tt, sr) + (r & 15));
            const bool valid = tile_ok && frame < p.K;
            float* yrow = p.y + (static_cast<long long>(m) * p.K + frame) * kBN;
            ptx::mbar_wait(tfull_bar + as, aphase);
            ptx::tc_fence_after();
            const uint32_t tsrc = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + as * kBN;
            const bool full = __all_sync(0xffffffffu, valid);
            if (kTmaDrain && full) {
                // All 32 rows exist: the warp stages a 32-column chunk (lane = row, 16 B pieces XOR-swizzled with the
                // row as the 128B-swizzled map expects: conflict-free) and lane 0 hands it to the async proxy as two
                // tensor stores, one per 16-frame run of a dilation phase.  Shared-memory stores release their
                // operands at once; the register path's 256-bit global stores held theirs until the LSU had taken
                // them, and the next TMEM load's write-after-read wait was the producers' largest single stall.
                const uint32_t buf = ptx::smem_u32(stg_smem) + static_cast<uint32_t>(kCstGlobal ? as * 4 + ew : pw) * kStgWarp;
                const uint32_t srow = buf + static_cast<uint32_t>(lane) * 128u;
                const int ph_l = tile_phase(p, tt, sr), j_l = tile_j0(p, tt, sr);
                const int ph1 = __shfl_sync(0xffffffffu, ph_l, 16), j1 = __shfl_sync(0xffffffffu, j_l, 16);
#pragma unroll 1
                for (int ci = 0; ci < kBN / 32; ++ci) {
                    uint32_t v[32];
                    ptx::tmem_ld_32x32(tsrc + ci * 32, v);
                    if (lane == 0) ptx::tma_store_wait_read<0>();   // the previous chunk's stores have read the buffer
                    __syncwarp();
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        ptx::st_shared_v4(srow + ((static_cast<uint32_t>(j) ^ (static_cast<uint32_t>(lane) & 7u)) << 4), v[4 * j],
                                          v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                    ptx::fence_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        ptx::tma_store_4d_addr(&tm_y, buf, ci * 32, ph_l, j_l, m);
                        ptx::tma_store_4d_addr(&tm_y, buf + 2048u, ci * 32, ph1, j1, m);
                        ptx::tma_store_commit();
                    }
                }
                // the buffer belongs to (TMEM stage, lane quarter): the warp that drains the stage next is another
                // group's, and the release of the stage below is what orders its writes after these reads
                if (kCstGlobal && lane == 0) ptx::tma_store_wait_read<0>();
            } else {
            // 16-column pieces in two alternating register sets: the TMEM load of piece h + 1 is in flight while
            // piece h is stored, and a set is reloaded only one piece after its stores were issued (the loads'
            // write-after-read wait on the stores' operands was the drain's main stall)
            const uint32_t tsrc = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + as * kBN;
            uint32_t va[16], vb[16];
            ptx::tmem_ld_32x16(tsrc, va);
#pragma unroll 1
            for (int h = 0; h < kBN / 16; h += 2) {
                ptx::tmem_ld_wait();
                ptx::tmem_ld_32x16(tsrc + (h + 1) * 16, vb);
                if (valid) {
                    ptx::st_global_v8(yrow + h * 16, va[0], va[1], va[2], va[3], va[4], va[5], va[6], va[7]);
                    ptx::st_global_v8(yrow + h * 16 + 8, va[8], va[9], va[10], va[11], va[12], va[13], va[14], va[15]);
                }
                ptx::tmem_ld_wait();
                if (h + 2 < kBN / 16) ptx::tmem_ld_32x16(tsrc + (h + 2) * 16, va);
                if (valid) {
                    ptx::st_global_v8(yrow + (h + 1) * 16, vb[0], vb[1], vb[2], vb[3], vb[4], vb[5], vb[6], vb[7]);
                    ptx::st_global_v8(yrow + (h + 1) * 16 + 8, vb[8], vb[9], vb[10], vb[11], vb[12], vb[13], vb[14], vb[15]);
                }
            }
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive_cluster(tempty_remote + as * 8);
        };

        const int n_tiles = worker < p.items ? (p.items - worker + n_workers - 1) / n_workers : 0;   // of this CTA
        const int n_units = n_tiles * nkb;
        int next_drain = grp;                                  // this group drains tiles grp, grp + 3, ...
        int cur_it = -1, cur_m = -1, m = 0, tt = -1;
        float mean = 0.f, rstd = 1.f;
        Tile T = {nullptr, 0, 0, 0};
        f2_t ls[2] = {0ull, 0ull}, lss[2] = {0ull, 0ull};
        // every (tile, warp) slot of the partials is written exactly once: tiles in which the group had
        // no unit (only possible when H / 64 < 3) get zeros
        auto zero_stats = [&](int it_lo, int it_hi) {
            for (int t = it_lo; t < it_hi; ++t) {
                const int rt = 2 * (worker + t * n_workers) + static_cast<int>(rank);
                const int mq = div_magic(rt, p.tpu_magic);
                if (lane == 0 && rt < p.row_tiles)
                    p.out_stat[static_cast<long long>(mq) * p.out_P + (rt - mq * p.tiles_per_utt) * kPartialsPerTile + pw] =
                        make_float2(0.f, 0.f);
            }
        };
        auto flush_stats = [&]() {   // (sum, sumsq) of this warp's share of the tile it has just left
            float s0, s1, s2, s3, t0, t1, t2, t3;
            f2_unpack(ls[0], s0, s1), f2_unpack(ls[1], s2, s3);
            f2_unpack(lss[0], t0, t1), f2_unpack(lss[1], t2, t3);
            const float wls = warp_sum((s0 + s1) + (s2 + s3));
            const float wlss = warp_sum((t0 + t1) + (t2 + t3));
            if (lane == 0 && tt >= 0)
                p.out_stat[static_cast<long long>(m) * p.out_P + tt * kPartialsPerTile + pw] = make_float2(wls, wlss);
            ls[0] = ls[1] = lss[0] = lss[1] = 0ull;
        };
        // Consecutive row tiles of a CTA lie in different utterances (stride 2 * n_workers tiles): finalise
        // the statistics of all of them up front, one tile per warp, instead of a chain of dependent
        // loads and fp64 math in front of every tile of every group.
        for (int t = pw; t < min(n_tiles, kMaxTiles); t += kProdWarps) {
            const int rt = 2 * (worker + t * n_workers) + static_cast<int>(rank);
            const int mt = rt < p.row_tiles ? div_magic(rt, p.tpu_magic) : 0;
            const float2 st = gln_finalize_warp(p.in_stat + static_cast<long long>(mt) * p.in_P, p.in_P,
                                                static_cast<double>(p.K) * p.H);
            if (lane == 0) tstat[t] = st;
        }
        ptx::named_bar_sync(1, 32 * kProdWarps);
        float4 q[kRing];
        if (grp < n_units) {   // rows 0 .. 8 of the group's first unit
            int m0, tt0;
            const int it0 = div_magic(grp, p.nkb_magic);
            const Tile T0 = tile_setup(worker + it0 * n_workers, m0, tt0);
            const float* b0 = T0.base + (grp - it0 * nkb) * kBK;
#pragma unroll
            for (int t = 0; t < kRing; ++t) {
                q[t] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (t >= T0.lo && t < T0.hi) q[t] = ld_row(b0 + t * strd);
            }
        }
        for (int u = grp; u < n_units; u += 3) {
            const int it = div_magic(u, p.nkb_magic), kb = u - it * nkb;
            if (it != cur_it) {
                if (cur_it >= 0) flush_stats();
                zero_stats(cur_it + 1, it);
                T = tile_setup(worker + it * n_workers, m, tt);
                cur_it = it;
                if (it < kMaxTiles) {
                    const float2 st = tstat[it];
                    mean = st.x, rstd = st.y;
                } else if (m != cur_m) {
                    const float2 st = gln_finalize_warp(p.in_stat + static_cast<long long>(m) * p.in_P, p.in_P,
                                                        static_cast<double>(p.K) * p.H);
                    mean = st.x, rstd = st.y;
                }
                cur_m = m;
            }
            constexpr int nslots = kASlots;
            const int aslot = u % nslots;                                 // units are consumed in order
            const uint32_t phase = static_cast<uint32_t>(u / nslots) & 1u;
            // the unit whose first rows replace this one's in the ring as they are consumed
            const int un = u + 3;
            const int itn = div_magic(un, p.nkb_magic), kbn = un - itn * nkb;
            Tile Tn = T;
            if (itn != it) {
                int mn_, ttn;
                Tn = tile_setup(un < n_units ? worker + itn * n_workers : p.items, mn_, ttn);
            }
            // rows kRing .. 17 of this unit, then rows 0 .. kRing-1 of the next one
            const float* pc = T.base + kb * kBK + kRing * strd;
            const float* pn = Tn.base + kbn * kBK;
            const int c = kb * kBK + quad * 4;
            const bool fast = is_full(T) && is_full(Tn);
            // per-channel constants.  Interior units fold the norm into the taps:
            //   d = sum_t w_t (A pv_t + B) = sum_t (w_t A) pv_t + B (w_0 + w_1 + w_2),  pv = PReLU(v)
            // boundary units keep n1 = A pv + B explicit, because the zero padding applies to n1.
            f2_t A[2], Bc[2], w0[2], w1[2], w2[2];
            {
                float4 g, b, u0, u1, u2;
                if (kCstGlobal) {
                    g = __ldg(reinterpret_cast<const float4*>(p.gamma_in + c));
                    b = __ldg(reinterpret_cast<const float4*>(p.beta_in + c));
                    // wd is [H, 3]: the taps of channels c .. c + 3 are twelve consecutive floats
                    const float4 t0 = __ldg(reinterpret_cast<const float4*>(p.wd + 3 * c));
                    const float4 t1 = __ldg(reinterpret_cast<const float4*>(p.wd + 3 * c + 4));
                    const float4 t2 = __ldg(reinterpret_cast<const float4*>(p.wd + 3 * c + 8));
                    u0 = make_float4(t0.x, t0.w, t1.z, t2.y);
                    u1 = make_float4(t0.y, t1.x, t1.w, t2.z);
                    u2 = make_float4(t0.z, t1.y, t2.x, t2.w);
                } else {
                    g = *reinterpret_cast<const float4*>(cst + c);
                    b = *reinterpret_cast<const float4*>(cst + p.H + c);
                    u0 = *reinterpret_cast<const float4*>(cst + 2 * p.H + c);
                    u1 = *reinterpret_cast<const float4*>(cst + 3 * p.H + c);
                    u2 = *reinterpret_cast<const float4*>(cst + 4 * p.H + c);
                }
                const float Af[4] = {g.x * rstd, g.y * rstd, g.z * rstd, g.w * rstd};
                const float Bf[4] = {b.x - Af[0] * mean, b.y - Af[1] * mean, b.z - Af[2] * mean, b.w - Af[3] * mean};
                // f[3 ch + t]: tap t of channel ch
                const float f[12] = {u0.x, u1.x, u2.x, u0.y, u1.y, u2.y, u0.z, u1.z, u2.z, u0.w, u1.w, u2.w};
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    A[i] = f2_pack(Af[2 * i], Af[2 * i + 1]);
                    const f2_t Bp = f2_pack(Bf[2 * i], Bf[2 * i + 1]);
                    const f2_t r0 = f2_pack(f[6 * i], f[6 * i + 3]), r1 = f2_pack(f[6 * i + 1], f[6 * i + 4]);
                    const f2_t r2 = f2_pack(f[6 * i + 2], f[6 * i + 5]);
                    // products of packed multiplies are born as register pairs (packing loaded scalars at
                    // every use costs two moves per operand)
                    const f2_t sc = fast ? A[i] : one2;
                    w0[i] = f2_mul(r0, sc), w1[i] = f2_mul(r1, sc), w2[i] = f2_mul(r2, sc);
                    Bc[i] = fast ? f2_mul(Bp, f2_add(f2_add(r0, r1), r2)) : f2_mul(Bp, one2);
                }
            }
            ptx::mbar_wait(aempty_bar + aslot, phase ^ 1);
            const uint32_t hi_base = ptx::smem_u32(smem + aslot * kASlotBytes) + row_off0;

            float* dptr = kStoreD ? p.d_out + (T.base - p.a) + kb * kBK : nullptr;
            auto unit = [&](auto fast_tag) {
                constexpr bool kFast = decltype(fast_tag)::value;
                // row t -> pv = PReLU(v) (interior) or n1 = A pv + B, zero outside the utterance
                // (boundary); then the ring slot is refilled
                auto xform = [&](int t, f2_t (&o)[2]) {
                    const float4 v = q[t % kRing];
                    const f2_t v2[2] = {f2_pack(v.x, v.y), f2_pack(v.z, v.w)};
#pragma unroll
                    for (int i = 0; i < 2; ++i) {
                        if (kPreact) {
                            o[i] = v2[i];
                        } else {
                            const f2_t mn = i == 0 ? f2_pack(fminf(v.x, 0.f), fminf(v.y, 0.f))
                                                   : f2_pack(fminf(v.z, 0.f), fminf(v.w, 0.f));
                            o[i] = f2_fma(a1m1, mn, v2[i]);
                        }
                        if (!kFast) o[i] = (t >= T.lo && t < T.hi) ? f2_fma(A[i], o[i], Bc[i]) : 0ull;
                    }
                    if (t < kRing) {
                        if (kFast || (t + kRing >= T.lo && t + kRing < T.hi)) q[t % kRing] = ld_row(pc);
                        pc += strd;
                    } else {
                        if (kFast || (t - kRing >= Tn.lo && t - kRing < Tn.hi))
                            q[t % kRing] = ld_row(pn);
                        else
                            q[t % kRing] = make_float4(0.f, 0.f, 0.f, 0.f);
                        pn += strd;
                    }
                };
                f2_t x0[2], x1[2], x2[2];
                xform(0, x0);
                xform(1, x1);
#pragma unroll
                for (int o = 0; o < kRowsPerThread; ++o) {
                    xform(o + 2, x2);
                    uint32_t h[2], l[2];
                    float dv[4];
#pragma unroll
                    for (int i = 0; i < 2; ++i) {
                        const f2_t acc = kFast ? f2_fma(w2[i], x2[i], f2_fma(w1[i], x1[i], f2_fma(w0[i], x0[i], Bc[i])))
                                               : f2_fma(w2[i], x2[i], f2_fma(w1[i], x1[i], f2_mul(w0[i], x0[i])));
                        float c0, c1;
                        f2_unpack(acc, c0, c1);
                        if (kStoreD) dv[2 * i] = c0, dv[2 * i + 1] = c1;
                        f2_t qq = f2_fma(a2m1, f2_pack(fminf(c0, 0.f), fminf(c1, 0.f)), acc);   // PReLU
                        if (!kFast && o >= T.n) qq = 0ull;
                        ls[i] = f2_add(ls[i], qq);
                        lss[i] = f2_fma(qq, qq, lss[i]);
                        f2_split_bf16(qq, neg1, h[i], l[i]);
                        x0[i] = x1[i], x1[i] = x2[i];
                    }
                    if (kStoreD && (kFast || o < T.n))   // output o = sequence row c0 + o of the window that starts at T.base
                        *reinterpret_cast<float4*>(dptr + (p.c0 + o) * strd) = make_float4(dv[0], dv[1], dv[2], dv[3]);
                    // row 16 s + o of the tile: 128 B rows, the 16 B chunk index XOR (row & 7) = XOR (o & 7).
                    // hi_base is 1024-aligned, so the XOR can be applied to the whole address.
                    const uint32_t ad = ((hi_base + cs) ^ (static_cast<uint32_t>(o & 7) << 4)) + static_cast<uint32_t>(o) * 128u;
                    asm volatile("st.volatile.shared.v2.b32 [%0], {%1, %2};" ::"r"(ad), "r"(h[0]), "r"(h[1]) : "memory");
                    if (CTN_MMA_TERMS == 3)
                        asm volatile("st.volatile.shared.v2.b32 [%0], {%1, %2};" ::"r"(ad + kABytes), "r"(l[0]), "r"(l[1]) : "memory");
                }
            };
            if (fast) unit(std::true_type{});
            else unit(std::false_type{});
            ptx::fence_async_smem();      // generic-proxy stores -> visible to the tensor core (async proxy)
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive_cluster(full_remote + aslot * 8);
            // drain duty: tiles before the current one whose accumulators are this group's
            while (next_drain < it) {
                drain(next_drain);
                next_drain += 3;
            }
        }
        if (cur_it >= 0) flush_stats();
        zero_stats(cur_it + 1, n_tiles);
        while (next_drain < n_tiles) {
            drain(next_drain);
            next_drain += 3;
        }
        if (kTmaDrain && lane == 0) ptx::tma_store_wait_all<0>();   // the staging buffers must outlive their stores
    }

    ptx::tc_fence_before();
    ptx::cluster_sync();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc_pair(tmem_base, 2 * kBN);
    }
}

// x += r y - r mu rowsum + W beta  (the deferred gLN-2 + residual, conv_tasnet.py:254-256, :237) on
// the (hi, lo) residual stream, in place.  One thread owns 8 channels of a row.
constexpr int kUpdThreads = 256;
constexpr int kUpdRows = 64;   // rows per CTA

__global__ void __launch_bounds__(kUpdThreads)
resid_update_kernel(__nv_bfloat16* __restrict__ x_hi, __nv_bfloat16* __restrict__ x_lo, const float* __restrict__ y,
                    const float2* __restrict__ stat, int P, double count, const float* __restrict__ col_g,
                    const float* __restrict__ col_bias, int K, int B) {
    ptx::grid_dep_sync();   // PDL: predecessor complete; let the next kernel be scheduled
    __shared__ double red[2 * (kUpdThreads / 32)];
    const int m = blockIdx.y;
    const float2 st = gln_finalize_block(stat + static_cast<long long>(m) * P, P, count, red);
    const float sc = st.y, sh = -st.y * st.x;
    const int tpr = B / 8;                       // threads per row
    const int rpp = kUpdThreads / tpr;           // rows per pass
    const int c = (threadIdx.x % tpr) * 8;
    float cc[8];
    {
        const float4 g0 = __ldg(reinterpret_cast<const float4*>(col_g + c)), g1 = __ldg(reinterpret_cast<const float4*>(col_g + c + 4));
        const float4 b0 = __ldg(reinterpret_cast<const float4*>(col_bias + c)), b1 = __ldg(reinterpret_cast<const float4*>(col_bias + c + 4));
        cc[0] = fmaf(sh, g0.x, b0.x), cc[1] = fmaf(sh, g0.y, b0.y), cc[2] = fmaf(sh, g0.z, b0.z), cc[3] = fmaf(sh, g0.w, b0.w);
        cc[4] = fmaf(sh, g1.x, b1.x), cc[5] = fmaf(sh, g1.y, b1.y), cc[6] = fmaf(sh, g1.z, b1.z), cc[7] = fmaf(sh, g1.w, b1.w);
    }
    const int k0 = blockIdx.x * kUpdRows;
    const int k1 = min(K, k0 + kUpdRows);
    for (int k = k0 + threadIdx.x / tpr; k < k1; k += rpp) {
        const long long o = (static_cast<long long>(m) * K + k) * B + c;
        const uint4 h = *reinterpret_cast<const uint4*>(x_hi + o);
        const uint4 l = *reinterpret_cast<const uint4*>(x_lo + o);
        float yv[8];
        ptx::ld_global_v8(y + o, yv);
        const uint32_t hw[4] = {h.x, h.y, h.z, h.w}, lw[4] = {l.x, l.y, l.z, l.w};
        uint32_t oh[4], ol[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
#if CTN_OPERAND_FP16
            float xh0, xh1, xl0, xl1;
            op16_value2(hw[j], xh0, xh1), op16_value2(lw[j], xl0, xl1);
            const float x0 = xh0 + xl0, x1 = xh1 + xl1;
#else
            const float x0 = __uint_as_float(hw[j] << 16) + __uint_as_float(lw[j] << 16);
            const float x1 = __uint_as_float(hw[j] & 0xffff0000u) + __uint_as_float(lw[j] & 0xffff0000u);
#endif
            const float v0 = fmaf(sc, yv[2 * j], cc[2 * j]) + x0;
            const float v1 = fmaf(sc, yv[2 * j + 1], cc[2 * j + 1]) + x1;
            split_pack2(v0, v1, oh[j], ol[j]);
        }
        *reinterpret_cast<uint4*>(x_hi + o) = make_uint4(oh[0], oh[1], oh[2], oh[3]);
        *reinterpret_cast<uint4*>(x_lo + o) = make_uint4(ol[0], ol[1], ol[2], ol[3]);
    }
}

struct Geometry {
    int w, w_log2, jspan, n_ig, tiles_per_utt;
};

static Geometry geometry(int K, int d) {
    Geometry g;
    g.w = d < 8 ? d : 8;
    g.w_log2 = 0;
    while ((1 << g.w_log2) < g.w) ++g.w_log2;
    g.jspan = kRowsPerThread * (8 / g.w);
    g.n_ig = d / g.w;
    const int Jtot = ceil_div(K, d);
    g.tiles_per_utt = g.n_ig * ceil_div(Jtot, g.jspan);
    return g;
}

}  // namespace fb
}  // namespace ctn

using namespace ctn;

extern "C" int ctn_dwpw_partials(int K, int dilation) {
    if (K <= 0 || dilation <= 0 || (dilation & (dilation - 1)) != 0) return 0;
    return fb::geometry(K, dilation).tiles_per_utt * fb::kPartialsPerTile;
}

extern "C" int ctn_dwpw_supported(int H, int B, int P, int dilation, int norm) {
    return norm == CTN_NORM_GLN && P == 3 && B == fb::kBN && H % fb::kBK == 0 && H >= fb::kBK &&
           fb::kSmemFixed + (fb::kCstGlobal ? 0 : 20 * H) <= fb::kSmemMax && dilation > 0 &&
           (dilation & (dilation - 1)) == 0;
}

extern "C" int ctn_dwpw_fwd(const ctn_dwpw_args* args, void* stream_) {
    CTN_REQUIRE(args != nullptr, "ctn_dwpw_fwd: args is NULL");
    const ctn_dwpw_args& a = *args;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    CTN_REQUIRE(a.M > 0 && a.K > 0, "ctn_dwpw_fwd: non-positive shape M=%d K=%d", a.M, a.K);
    CTN_REQUIRE(ctn_dwpw_supported(a.H, a.B, 3, a.dilation, CTN_NORM_GLN),
                "ctn_dwpw_fwd: needs B == 256, H %% 64 == 0 and a power-of-two dilation (H=%d B=%d dilation=%d)", a.H, a.B,
                a.dilation);
    CTN_REQUIRE(static_cast<long long>(a.M) * a.K < (1ll << 31), "ctn_dwpw_fwd: M*K exceeds 2^31 rows");
    CTN_REQUIRE(a.a && a.gamma_in && a.beta_in && a.in_stat && a.in_P > 0 && a.wd && a.prelu_out && a.w_hi &&
                    a.w_lo && a.y && a.out_stat,
                "ctn_dwpw_fwd: NULL argument");
    auto aligned16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    CTN_REQUIRE(aligned16(a.a) && aligned16(a.gamma_in) && aligned16(a.beta_in) && aligned16(a.wd) && aligned16(a.w_hi) &&
                    aligned16(a.w_lo) && (reinterpret_cast<uintptr_t>(a.y) & 31) == 0,
                "ctn_dwpw_fwd: buffers must be 16-byte aligned (y: 32)");
    const fb::Geometry g = fb::geometry(a.K, a.dilation);
    fb::Params p;
    p.K = a.K, p.H = a.H, p.d = a.dilation, p.c0 = a.causal ? 2 : 1;
    p.w = g.w, p.w_log2 = g.w_log2, p.jspan = g.jspan, p.n_ig = g.n_ig, p.tiles_per_utt = g.tiles_per_utt;
    p.row_tiles = a.M * g.tiles_per_utt;
    p.items = ceil_div(p.row_tiles, 2);
    p.nkb = a.H / fb::kBK;
    p.d_log2 = p.n_ig_log2 = 0;
    while ((1 << p.d_log2) < a.dilation) ++p.d_log2;
    while ((1 << p.n_ig_log2) < g.n_ig) ++p.n_ig_log2;
    p.tpu_magic = g.tiles_per_utt == 1 ? 0u : static_cast<uint32_t>(((1ull << 32) + g.tiles_per_utt - 1) / g.tiles_per_utt);
    p.nkb_magic = p.nkb == 1 ? 0u : static_cast<uint32_t>(((1ull << 32) + p.nkb - 1) / p.nkb);
    // div_magic is exact while x * (magic * divisor - 2^32) < 2^32; here divisors <= 2^12 and x < 2^20: row tile
    // numbers, and unit numbers of one CTA (< row_tiles / sm_count * nkb, nkb <= 18 by the shared-memory bound on H)
    CTN_REQUIRE(g.tiles_per_utt <= 4096 && p.nkb <= 64 && p.row_tiles + 2 * sm_count() < (1 << 20),
                "ctn_dwpw_fwd: batch too large for the tile index arithmetic (M=%d K=%d)", a.M, a.K);
    p.in_P = a.in_P, p.out_P = g.tiles_per_utt * fb::kPartialsPerTile;
    p.a = a.a, p.prelu_in = a.prelu_in, p.gamma_in = a.gamma_in, p.beta_in = a.beta_in;
    p.in_stat = static_cast<const float2*>(a.in_stat);
    p.wd = a.wd, p.prelu_out = a.prelu_out, p.y = a.y, p.out_stat = static_cast<float2*>(a.out_stat);
    CUtensorMap tw_hi, tw_lo;
    int rc;
    if ((rc = make_tmap_bf16_kmajor(&tw_hi, a.w_hi, a.B, a.H, fb::kBN / 2)) != CTN_OK) return rc;
    if ((rc = make_tmap_bf16_kmajor(&tw_lo, a.w_lo, a.B, a.H, fb::kBN / 2)) != CTN_OK) return rc;
    CUtensorMap ty;
    if ((rc = make_tmap_phase_f32(&ty, a.y, a.M, a.K, a.B, a.dilation, fb::kRowsPerThread)) != CTN_OK) return rc;
    const bool preact = a.prelu_in == nullptr;
    p.d_out = a.d_out;
    CTN_REQUIRE(a.d_out == nullptr || (!preact && aligned16(a.d_out)),
                "ctn_dwpw_fwd: d_out (training) goes with prelu_in and must be 16-byte aligned");
    auto kern = a.d_out ? fb::dwpw_kernel<false, true> : preact ? fb::dwpw_kernel<true> : fb::dwpw_kernel<false>;
    static bool attr_done[kMaxDevices][3] = {};
    bool* attr_set = attr_done[device_slot()] + (a.d_out ? 2 : 0);
    if (!attr_set[a.d_out ? 0 : preact]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, fb::kSmemMax);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(dwpw_kernel)");
        attr_set[a.d_out ? 0 : preact] = true;
    }
    const int workers_max = sm_count() / 2;
    const int workers = p.items < workers_max ? p.items : workers_max;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * workers, 1, 1);
    cfg.blockDim = dim3(fb::kThreads, 1, 1);
    cfg.dynamicSmemBytes = fb::kSmemFixed + (fb::kCstGlobal ? 0 : 20 * a.H);
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = pdl_allowed(stream, static_cast<long long>(a.M) * a.K);
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tw_hi, tw_lo, ty, p);
    if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx(dwpw_kernel)");
    return CTN_OK;
}

extern "C" int ctn_resid_update(void* x_hi, void* x_lo, const float* y, const void* stat, int P, double count,
                                const float* col_g, const float* col_bias, int M, int K, int B, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    CTN_REQUIRE(M > 0 && K > 0 && B > 0 && B % 8 == 0 && B <= 8 * fb::kUpdThreads && (8 * fb::kUpdThreads) % B == 0,
                "ctn_resid_update: bad shape M=%d K=%d B=%d (B must divide 2048 and be a multiple of 8)", M, K, B);
    CTN_REQUIRE(x_hi && x_lo && y && stat && P > 0 && count > 0 && col_g && col_bias, "ctn_resid_update: NULL argument");
    auto aligned = [](const void* p, int n) { return (reinterpret_cast<uintptr_t>(p) & (n - 1)) == 0; };
    CTN_REQUIRE(aligned(x_hi, 16) && aligned(x_lo, 16) && aligned(y, 32) && aligned(col_g, 16) && aligned(col_bias, 16),
                "ctn_resid_update: buffers must be 16-byte aligned (y: 32)");
    dim3 grid(ceil_div(K, fb::kUpdRows), M);
    CTN_LAUNCH_PDL("resid_update_kernel", static_cast<long long>(M) * K, fb::resid_update_kernel, grid, fb::kUpdThreads, 0, stream, static_cast<__nv_bfloat16*>(x_hi), static_cast<__nv_bfloat16*>(x_lo), y, static_cast<const float2*>(stat), P, count, col_g, col_bias, K, B);
    CTN_LAUNCH_CHECK("resid_update_kernel");
    return CTN_OK;
}
