// glowtts_b200/csrc/wgrad_group.cu -- ALL weight and bias gradients of one coupling block in one persistent,
// stream-K tcgen05 launch + one reduce launch.
//
//   dW_q[m][c][j] = sum_{b,t} dY_q[b, t, m] * X_q[b, t + (j - K_q/2) dil_q, c]          for every conv q of the block
//   db_q[m]       = sum_{b,t} dY_q[b, t, m]                                             (a column of ones appended to X_q)
//
// Work = tiles x (utterance, 64-frame chunk) k-steps.  A tile is
//   * (conv, 128-row M tile, tap) with all X columns: the default.  One tile per tap, each fetching dY and a shifted
//     copy of X: ~490 MB of L2 -> SM traffic per coupling block, which is what bounds it (profiles/wgrad_overlap_r02.txt);
//   * opt-in (GTK_WGRAD_ALL_TAPS=1), k > 1 convs: "all-taps" tiles (conv, 128-row M tile, 64-column block of X) with EVERY
//     tap inside: the X rows [t0 - halo, t0 + 64 + halo) are staged ONCE per k-step and tap j is a row-shifted MN-major
//     descriptor on them (the 128-byte swizzle is a function of the shared-memory address only), one TMEM accumulator of
//     64 columns per tap.  Less than half the traffic, but N = 64 MMAs run at ~74 cycles each: no faster overall
//     (profiles/mma_issue_r02.txt).
// All tiles of all convs are laid
// end to end (weighted by the bytes one k-step stages) and cut into `grid` equal contiguous shares, one per SM
// ("stream-K"): no wave quantisation, no per-conv launch, and each CTA writes at most a few partial tiles.  A CTA
// walks its segments with one TMA producer thread, one MMA thread and four epilogue warps; two 256-column TMEM
// accumulators take alternate k16 steps (independent MMA chains) and are added when a segment is drained.
//
// Operands are frame-major bf16 ([B, T, channels]; reduction dim = rows), i.e. MN-major UMMA operands fetched as
// {64 channels x 64 frames} SWIZZLE_128B boxes, the conv tap being a shifted (zero-filled) box of X -- as in
// wgrad_tc.cu.  The bias gradient rides along as X block 3 = [1, 0, 0, ...] written once to shared memory.
// The reduce kernel sums a tile's partial slots and finishes weight norm (dg, dv) and the bias gradient per row.
#include <stdlib.h>
#include <string.h>

#include "tc.cuh"
#include "ops.h"

namespace gtk {
namespace tc {

constexpr int GG_TK = 64;                        // frames per k-step
constexpr int GG_BOX = GG_TK * 128;              // bytes of one {64 ch, 64 frames} box
constexpr int GG_STAGES = 4;
constexpr int GG_PAIR_ROWS = 8;                  // extra X rows a stage can hold: a tap PAIR shares one staged X tile when dil <= 8
constexpr int GG_XSTRIDE = (GG_TK + GG_PAIR_ROWS) * 128;   // bytes between the 64-column X blocks of a stage (9 swizzle atoms)
constexpr int GG_STAGE_BYTES = 2 * GG_BOX + 4 * GG_XSTRIDE;   // 2 dY boxes (128 rows of M) + up to 4 X blocks (N <= 256)
constexpr int GG_ONES_BLK = 3;                   // X block that holds the ones column (never written by TMA)
constexpr int GG_SLOT_COLS = 512;                // one partial tile: 128 rows x up to 512 accumulator columns
constexpr int GG_SLOT_FLOATS = 128 * GG_SLOT_COLS;
constexpr int GG_ALL_TAPS = 0xFF;                // GgTile::tap of an all-taps tile
constexpr int GG_CB = 64;                        // X columns of an all-taps tile = accumulator columns per tap
// Stream-K share weights = measured SM cycles of one k-step (profiles/gg_trace_r02.py, B200): a single-tap tile is bound
// by the latency of its operand stage under full-chip load (~820 cycles for 40 KB, the MMA warp waits 45 % of the time),
// a tap pair by its eight MMAs on two accumulator chains (~1060 cycles, it never waits for operands)
constexpr int GG_W_SINGLE = 100, GG_W_PAIR = 130;

struct GgJobDev {
    int taps, dil, n_blk, ones;                  // n_blk = real 64-column blocks of X; ones: bias gradient wanted
    int halo;                                    // (taps / 2) * dil: rows staged either side of a k-step (all-taps tiles)
};
// tap == GG_ALL_TAPS: every tap, X columns [64 cblk, 64 cblk + 64); otherwise taps [tap, tap + cblk] (cblk = 0 or 1) with all
// X columns: a PAIR of taps shares the dY tile and ONE staged X tile of 64 + dil rows (tap + 1 is a row-shifted MN-major
// descriptor on it) and owns one 256-column accumulator per tap -- the same two independent MMA chains as a single-tap
// tile (which alternates its k16 steps between the two accumulators), for half the L2 -> SM operand traffic per MMA
struct GgTile { unsigned char job, m_tile, tap, cblk; };
struct GgSeg { unsigned short tile, s_lo, s_hi, slot; };
struct GgCta { unsigned short seg0, nseg; };

struct GgParams {
    int B, T, chunks_t;
    const int* n_rows;                           // varlen packing (B == 1): only rows [0, *n_rows) hold data, or null
    float* part;                                 // [slots][128][512] fp32
    int red;                                     // 1: ONE zeroed slot per tile, every share adds into it (red.global.add.v4.f32)
    GgJobDev job[WGG_MAX_JOBS];
    GgTile tile[WGG_MAX_TILES];
    GgSeg seg[WGG_MAX_SEGS];
    GgCta cta[WGG_MAX_CTAS];
    CUtensorMap tmY[WGG_MAX_JOBS], tmX[WGG_MAX_JOBS];
    CUtensorMap tmXh[WGG_MAX_JOBS];              // X with a {64 columns, 64 + 2 halo rows} box (all-taps tiles),
                                                 // or {64 columns, 64 + dil rows} (tap pairs)
};

struct GgCtl {
    uint64_t full[GG_STAGES], empty[GG_STAGES], acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};

// Instrumented builds (-DGTK_INSTRUMENT): SM cycles per CTA of the last launch: [0] MMA warp waiting for operand stages,
// [1] waiting for the drain of the previous segment, [2] MMA warp total, [3] k-steps, [4] segments, [5] kernel entry ->
// MMA warp start (pdl wait + setup), [6] producer waiting for free stages, [7] drain warps total drain cycles (warp 4)
#ifdef GTK_INSTRUMENT
__device__ long long g_gg_wait[WGG_MAX_CTAS * 8];
#define GG_CLK() clock64()
#define GG_STORE(i, v) do { g_gg_wait[blockIdx.x * 8 + (i)] = (v); } while (0)
int gg_wait_read(long long* out, int n_ctas) {
    GTK_REQUIRE(out && n_ctas >= 0 && n_ctas <= WGG_MAX_CTAS, "n_ctas");
    GTK_CUDA(cudaMemcpyFromSymbol(out, g_gg_wait, size_t(n_ctas) * 8 * sizeof(long long)));
    return GTK_OK;
}
#else
#define GG_CLK() 0ll
#define GG_STORE(i, v) do { } while (0)
#endif

__global__ void __launch_bounds__(NTHREADS, 1) wgrad_group_kernel(const __grid_constant__ GgParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    GgCtl* ctl = reinterpret_cast<GgCtl*>(smem + (size_t)GG_STAGES * GG_STAGE_BYTES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const GgCta me = p.cta[blockIdx.x];
    const long long k_t0 = GG_CLK();
    pdl_trigger();
    // The host cut every tile's k-steps [0, S_max) into contiguous shares.  With packed rows only the first S_val steps
    // hold data: every segment boundary is rescaled by S_val / S_max (monotone, so the segments still tile [0, S_val)
    // and the shares stay balanced).  A segment that becomes empty writes a zero partial tile.
    const long long S_max = (long long)p.B * p.chunks_t;
    const long long S_val = p.n_rows ? min(S_max, (long long)ceil_div(*p.n_rows, GG_TK)) : S_max;
#define GG_SCALE(s) ((int)((long long)(s) * S_val / S_max))

    if (threadIdx.x == 0) {
        for (int s = 0; s < GG_STAGES; ++s) { mbar_init(&ctl->full[s], 1); mbar_init(&ctl->empty[s], 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(&ctl->acc_full[a], 1); mbar_init(&ctl->acc_empty[a], 4); }
        mbar_fence_init();
    }
    if (warp == 2) tmem_alloc<512>(&ctl->tmem_base);
    // ones block of every stage: row r (frame) is 128 B = 64 channels, channel 0 sits in 16-byte chunk (0 ^ (r & 7))
    for (int i = threadIdx.x; i < GG_STAGES * GG_TK * 8; i += NTHREADS) {
        const int s = i / (GG_TK * 8), r = (i / 8) % GG_TK, ch = i % 8;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (ch == (r & 7)) v.x = 0x3F80u;        // bf16 1.0 in the low half-word = channel 0
        *reinterpret_cast<uint4*>(smem + (size_t)s * GG_STAGE_BYTES + 2 * GG_BOX + GG_ONES_BLK * GG_XSTRIDE + r * 128 + ch * 16) = v;
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    pdl_wait();                                          // dY / X of the data-gradient chain are complete and visible
    const uint32_t tmem = ctl->tmem_base;

    if (warp == 0) {
        // TMA producer: whole warp in the uniform loops, one elected lane issues (see the MMA warp)
        int it = 0;
        long long w_free = 0;
        for (int sg = 0; sg < me.nseg; ++sg) {
            const GgSeg S = p.seg[me.seg0 + sg];
            const GgTile Tl = p.tile[S.tile];
            const GgJobDev J = p.job[Tl.job];
            const CUtensorMap* tmX = &p.tmX[Tl.job];
            const CUtensorMap* tmY = &p.tmY[Tl.job];
            const int m0 = Tl.m_tile * 128;
            const bool all_taps = Tl.tap == GG_ALL_TAPS;
            const int shift = all_taps ? -J.halo : ((int)Tl.tap - J.taps / 2) * J.dil;
            const bool pair = !all_taps && Tl.cblk != 0;
            const uint32_t bytes = all_taps ? (uint32_t)(2 * GG_BOX + (GG_TK + 2 * J.halo) * 128)
                                 : pair ? (uint32_t)(2 * GG_BOX + J.n_blk * (GG_TK + J.dil) * 128) : (uint32_t)(2 + J.n_blk) * GG_BOX;
            const CUtensorMap* tmXs = pair ? &p.tmXh[Tl.job] : tmX;
            const int s_lo = GG_SCALE(S.s_lo), s_hi = GG_SCALE(S.s_hi);
            for (int step = s_lo; step < s_hi; ++step, ++it) {
                const int s = it % GG_STAGES, use = it / GG_STAGES;
                const long long c0 = GG_CLK();
                if (use > 0) mbar_wait(&ctl->empty[s], (use - 1) & 1);
                w_free += GG_CLK() - c0;
                const int b = step / p.chunks_t, t0 = (step % p.chunks_t) * GG_TK;
                unsigned char* sa = smem + (size_t)s * GG_STAGE_BYTES;
                unsigned char* sb = sa + 2 * GG_BOX;
                if (elect_one()) {
                    mbar_arrive_expect_tx(&ctl->full[s], bytes);
                    tma_load_3d(sa, tmY, &ctl->full[s], m0, t0, b);
                    tma_load_3d(sa + GG_BOX, tmY, &ctl->full[s], m0 + 64, t0, b);
                    if (all_taps) {
                        tma_load_3d(sb, &p.tmXh[Tl.job], &ctl->full[s], Tl.cblk * GG_CB, t0 + shift, b);
                    } else {
                        for (int nb = 0; nb < J.n_blk; ++nb)
                            tma_load_3d(sb + (size_t)nb * GG_XSTRIDE, tmXs, &ctl->full[s], nb * 64, t0 + shift, b);
                    }
                }
                __syncwarp();
            }
        }
        GG_STORE(6, w_free);
    } else if (warp == 1) {
        // MMA issue.  The WHOLE warp walks the (warp-uniform) loops and one elected lane issues: the descriptors then live
        // in uniform registers.  Under `if (lane == 0)` the compiler cannot prove uniformity and wraps every tcgen05.mma in
        // an R2UR waterfall loop (~15 instructions, >100 cycles per MMA: profiles/mma_issue_r02.txt).
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
        int it = 0, ne = 0;                            // ne: non-empty segments so far (accumulator hand-shake parity)
        long long w_full = 0, w_drain = 0;
        const long long m_t0 = GG_CLK();
        GG_STORE(5, m_t0 - k_t0);
        for (int sg = 0; sg < me.nseg; ++sg) {
            const GgSeg S = p.seg[me.seg0 + sg];
            const GgTile Tl = p.tile[S.tile];
            const GgJobDev J = p.job[Tl.job];
            const int s_lo = GG_SCALE(S.s_lo), s_hi = GG_SCALE(S.s_hi);
            if (s_hi <= s_lo) continue;
            { const long long c0 = GG_CLK();
              if (ne > 0) { mbar_wait(&ctl->acc_empty[0], (ne - 1) & 1); tc_fence_after(); }
              w_drain += GG_CLK() - c0; }
            if (Tl.tap == GG_ALL_TAPS) {
                // one accumulator per tap (columns [64 j, 64 j + 64)): independent MMA chains by construction; the bias
                // gradient (ones column) takes 16 more columns on the tile of X block 0
                const uint32_t idesc_t = umma_idesc_bf16(128, GG_CB, 1, 1), idesc_1 = umma_idesc_bf16(128, 16, 1, 1);
                const bool ones = J.ones && Tl.cblk == 0;
                for (int step = s_lo; step < s_hi; ++step, ++it) {
                    const int s = it % GG_STAGES, use = it / GG_STAGES;
                    mbar_wait(&ctl->full[s], use & 1);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + (size_t)s * GG_STAGE_BYTES);
                    const uint32_t sb = sa + 2 * GG_BOX;
                    if (elect_one()) {
#pragma unroll
                        for (int k = 0; k < GG_TK / 16; ++k) {
                            const uint64_t adesc = umma_desc_mnmajor_sw128(sa + k * 2048, GG_BOX, 1024);
                            const bool acc = (step != s_lo) || k > 0;
                            for (int j = 0; j < J.taps; ++j) {
                                const uint64_t bdesc = umma_desc_mnmajor_sw128(sb + (uint32_t)(j * J.dil) * 128u + k * 2048, GG_XSTRIDE, 1024);
                                umma_bf16(tm + j * GG_CB, adesc, bdesc, idesc_t, acc);
                            }
                            if (ones) {
                                const uint64_t bdesc = umma_desc_mnmajor_sw128(sb + GG_ONES_BLK * GG_XSTRIDE + k * 2048, GG_XSTRIDE, 1024);
                                umma_bf16(tm + J.taps * GG_CB, adesc, bdesc, idesc_1, acc);
                            }
                        }
                        umma_commit(&ctl->empty[s]);
                    }
                    __syncwarp();
                }
            } else if (Tl.cblk != 0) {
                // tap pair: accumulator 0 = tap `tap` (with the ones column when tap == 0), accumulator 1 = tap + 1 read
                // dil rows further down the same staged X tile
                const int N0 = (J.ones && Tl.tap == 0) ? (GG_ONES_BLK + 1) * 64 : J.n_blk * 64;
                const uint32_t idesc0 = umma_idesc_bf16(128, N0, 1, 1), idesc1 = umma_idesc_bf16(128, J.n_blk * 64, 1, 1);
                for (int step = s_lo; step < s_hi; ++step, ++it) {
                    const int s = it % GG_STAGES, use = it / GG_STAGES;
                    mbar_wait(&ctl->full[s], use & 1);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + (size_t)s * GG_STAGE_BYTES);
                    const uint32_t sb = sa + 2 * GG_BOX;
                    if (elect_one()) {
#pragma unroll
                        for (int k = 0; k < GG_TK / 16; ++k) {
                            const uint64_t adesc = umma_desc_mnmajor_sw128(sa + k * 2048, GG_BOX, 1024);
                            const uint64_t b0 = umma_desc_mnmajor_sw128(sb + k * 2048, GG_XSTRIDE, 1024);
                            const uint64_t b1 = umma_desc_mnmajor_sw128(sb + (uint32_t)J.dil * 128u + k * 2048, GG_XSTRIDE, 1024);
                            const bool acc = (step != s_lo) || k > 0;
                            umma_bf16(tm, adesc, b0, idesc0, acc);
                            umma_bf16(tm + 256, adesc, b1, idesc1, acc);
                        }
                        umma_commit(&ctl->empty[s]);
                    }
                    __syncwarp();
                }
            } else {
                const int N = (J.ones && Tl.tap == 0) ? (GG_ONES_BLK + 1) * 64 : J.n_blk * 64;
                const uint32_t idesc = umma_idesc_bf16(128, N, 1, 1);
                // Two accumulators (TMEM columns 0.. and 256..) take alternate k16 steps: back-to-back MMAs into one
                // accumulator serialise on its read-modify-write latency.  The epilogue adds them; the next segment
                // starts when it has drained both (segments are ~100 k-steps long, the drain is a few per cent).
                for (int step = s_lo; step < s_hi; ++step, ++it) {
                    const int s = it % GG_STAGES, use = it / GG_STAGES;
                    const long long c0 = GG_CLK();
                    mbar_wait(&ctl->full[s], use & 1);
                    w_full += GG_CLK() - c0;
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + (size_t)s * GG_STAGE_BYTES);
                    const uint32_t sb = sa + 2 * GG_BOX;
                    if (elect_one()) {
#pragma unroll
                        for (int k = 0; k < GG_TK / 16; ++k) {
                            const uint64_t adesc = umma_desc_mnmajor_sw128(sa + k * 2048, GG_BOX, 1024);
                            const uint64_t bdesc = umma_desc_mnmajor_sw128(sb + k * 2048, GG_XSTRIDE, 1024);
                            umma_bf16(tm + (k & 1) * 256, adesc, bdesc, idesc, (step != s_lo) || k >= 2);
                        }
                        umma_commit(&ctl->empty[s]);
                    }
                    __syncwarp();
                }
            }
            if (elect_one()) umma_commit(&ctl->acc_full[0]);
            __syncwarp();
            ++ne;
        }
        GG_STORE(0, w_full); GG_STORE(1, w_drain); GG_STORE(2, GG_CLK() - m_t0); GG_STORE(3, (long long)it); GG_STORE(4, (long long)ne);
    } else if (warp >= 4) {
        const int q = warp & 3;
        int ne = 0;
        long long t_drain = 0;
        for (int sg = 0; sg < me.nseg; ++sg) {
            const GgSeg S = p.seg[me.seg0 + sg];
            const GgTile Tl = p.tile[S.tile];
            const GgJobDev J = p.job[Tl.job];
            float* prow = p.part + (size_t)S.slot * GG_SLOT_FLOATS + (size_t)(q * 32 + lane) * GG_SLOT_COLS;
            const bool all_taps = Tl.tap == GG_ALL_TAPS;
            const bool pair = !all_taps && Tl.cblk != 0;
            const int n_cols = all_taps ? J.taps * GG_CB : J.n_blk * 64;             // weight-gradient columns of the tile (per tap)
            const bool ones = J.ones && (all_taps ? Tl.cblk == 0 : Tl.tap == 0);
            const int ones_col = all_taps ? J.taps * GG_CB : GG_ONES_BLK * 64;
            if (GG_SCALE(S.s_hi) <= GG_SCALE(S.s_lo)) {              // empty share: its partial tile is zero
                if (p.red) continue;                                 // ... and adds nothing to the tile's one slot
                for (int c0 = 0; c0 < n_cols; c0 += 4) *reinterpret_cast<float4*>(prow + c0) = make_float4(0.f, 0.f, 0.f, 0.f);
                if (pair)
                    for (int c0 = 0; c0 < n_cols; c0 += 4) *reinterpret_cast<float4*>(prow + 256 + c0) = make_float4(0.f, 0.f, 0.f, 0.f);
                if (ones) prow[ones_col] = 0.f;
                continue;
            }
            mbar_wait(&ctl->acc_full[0], ne & 1);
            ++ne;
            tc_fence_after();
            const long long d0 = GG_CLK();
            const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
            for (int half = 0; half < (pair ? 2 : 1); ++half)      // a pair keeps tap + 1 in accumulator columns 256..
                for (int c0 = 0; c0 < n_cols; c0 += 16) {
                    float v[16];
                    tmem_ld16_dual(trow + half * 256 + c0, !all_taps && !pair, v);
                    float4* o4 = reinterpret_cast<float4*>(prow + half * 256 + c0);
                    if (p.red) {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" :: "l"(o4 + i), "f"(v[4 * i]),
                                         "f"(v[4 * i + 1]), "f"(v[4 * i + 2]), "f"(v[4 * i + 3]) : "memory");
                    } else {
#pragma unroll
                        for (int i = 0; i < 4; ++i) o4[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                    }
                }
            if (ones) {
                float v[16];
                tmem_ld16_dual(trow + ones_col, !all_taps && !pair, v);
                if (p.red) atomicAdd(prow + ones_col, v[0]);
                else prow[ones_col] = v[0];
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl->acc_empty[0]);
            t_drain += GG_CLK() - d0;
        }
        if (warp == 4 && lane == 0) GG_STORE(7, t_drain);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc<512>(tmem);
    }
}

// ------------------------------------------------------------------------------------------ reduce
struct GgRedJob {
    const float* g;            // weight_g [M] or null (plain weight)
    const float* v;            // weight_v / weight [M][Cin][taps]
    float* dg;
    float* dv;
    float* db;                 // bias gradient [M] or null
    int M, Cin, taps, tile0, row0, ones;       // row0: first block of this conv in the block-per-row launch (k > 1 convs)
    int row0w;                 // first block in the warp-per-row launch (1x1 convs; four rows per block); a conv has rows in one
    int n_cblk;                // > 0: all-taps tiles, tile = tile0 + m_tile * n_cblk + c / 64, column = 64 tap + c % 64
    int n_grp;                 // > 0: tap pairs, n_grp tiles per m_tile: tile = tile0 + m_tile * n_grp + tap / 2, column = 256 (tap % 2) + c
};
struct GgRedParams {
    const float* part;
    int n_jobs;
    GgRedJob job[WGG_MAX_JOBS];
    unsigned short slot0[WGG_MAX_TILES], nslot[WGG_MAX_TILES];
};

// One block per output row m of one conv.  Plain weight: dv[m][c][j] += sum.  Weight-normed (torch.nn.utils.weight_norm,
// modules.py:124,131,141):  dg[m] += <dW_m, v_m>/||v_m||,  dv_m += g_m/||v_m|| (dW_m - v_m <dW_m, v_m>/||v_m||^2).
// The kernel is latency-bound (a few dependent L2 round trips per block), so every thread issues all the loads of its
// (up to four) row elements -- partial slots four at a time, plus v -- before consuming any of them.
constexpr int GG_RED_NT = 128;
struct GgRedMulti {
    int n;
    int row_end[WGG_MAX_DEFER];                    // cumulative blocks of stream-K launch k in the block-per-row launch
    int row_end_w[WGG_MAX_DEFER];                  // ... in the warp-per-row launch
    GgRedParams r[WGG_MAX_DEFER];
};
struct WgradReduceBatch {
    GgRedMulti M;
    float* part_ws;
    int capacity;
    size_t max_row;
    int zeroed[WGG_MAX_DEFER];                     // adding shares: tile slots of launch k that were zeroed ahead of time
};
// 1x1 convs (res_skip, start, end: half of a step's 38 784 rows, 15 % of its bytes): a WARP per output row, the row (<= 256
// elements, already in torch order) stays in registers -- no shared memory, no block barrier, a quarter of the blocks.  Its own
// kernel: inside the block-per-row kernel its 48 registers cost that kernel a third of its occupancy (4.066 against 4.054 ms).
__global__ void __launch_bounds__(GG_RED_NT, 10) wgrad_group_reduce_rows_kernel(const __grid_constant__ GgRedMulti pm) {
    pdl_trigger();
    pdl_wait();
    int k = 0;
    while (k + 1 < pm.n && (int)blockIdx.x >= pm.row_end_w[k]) ++k;
    const GgRedParams& p = pm.r[k];
    const int brow = (int)blockIdx.x - (k ? pm.row_end_w[k - 1] : 0);
    int q = 0;
    while (q + 1 < p.n_jobs && brow >= p.job[q + 1].row0w) ++q;
    const GgRedJob& J = p.job[q];
    const int lane = threadIdx.x & 31;
    const int m = (brow - J.row0w) * (GG_RED_NT / 32) + (threadIdx.x >> 5);
    if (m >= J.M) return;
    const int Cin = J.Cin, tile = J.tile0 + (m >> 7);
    const float* src = p.part + (size_t)p.slot0[tile] * GG_SLOT_FLOATS + (size_t)(m & 127) * GG_SLOT_COLS;
    const int ns = p.nslot[tile];
    const float* __restrict__ vrow = J.v + (size_t)m * Cin;
    float* __restrict__ dvrow = J.dv + (size_t)m * Cin;
    const bool wn = J.g != nullptr;
    float a[8], v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
        const int c = lane + 32 * u;
        const bool ok = c < Cin;
        a[u] = ok ? __ldg(src + c) : 0.f;
        v[u] = (wn && ok) ? __ldg(vrow + c) : 0.f;
    }
    for (int sl = 1; sl < ns; ++sl) {                          // adding shares: one slot per tile, never taken
#pragma unroll
        for (int u = 0; u < 8; ++u)
            if (lane + 32 * u < Cin) a[u] += __ldg(src + (size_t)sl * GG_SLOT_FLOATS + lane + 32 * u);
    }
    if (J.db && J.ones && lane == 31) {
        float acc = 0.f;
        for (int sl = 0; sl < ns; ++sl) acc += src[(size_t)sl * GG_SLOT_FLOATS + GG_ONES_BLK * 64];
        J.db[m] += acc;
    }
    if (!wn) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
            if (lane + 32 * u < Cin) dvrow[lane + 32 * u] += a[u];
        return;
    }
    float ss = 0.f, dd = 0.f;
#pragma unroll
    for (int u = 0; u < 8; ++u) { ss += v[u] * v[u]; dd += v[u] * a[u]; }
    ss = warp_sum(ss);
    dd = warp_sum(dd);
    ss = __shfl_sync(0xffffffffu, ss, 0);
    dd = __shfl_sync(0xffffffffu, dd, 0);
    const float nrm = sqrtf(ss);
    const float sc = J.g[m] / nrm, proj = dd / ss;
#pragma unroll
    for (int u = 0; u < 8; ++u)
        if (lane + 32 * u < Cin) dvrow[lane + 32 * u] += sc * (a[u] - v[u] * proj);
    if (lane == 0) J.dg[m] += dd / nrm;
}

__global__ void __launch_bounds__(GG_RED_NT) wgrad_group_reduce_kernel(const __grid_constant__ GgRedMulti pm) {
    extern __shared__ float row[];                 // [Cin * taps] reduced dW row, torch order (c, j)
    __shared__ float red[2][GG_RED_NT / 32];
    pdl_trigger();
    pdl_wait();                                    // the stream-K partial tiles are complete and visible
    int k = 0;
    while (k + 1 < pm.n && (int)blockIdx.x >= pm.row_end[k]) ++k;
    const GgRedParams& p = pm.r[k];
    const int brow = (int)blockIdx.x - (k ? pm.row_end[k - 1] : 0);
    int q = 0;
    while (q + 1 < p.n_jobs && brow >= p.job[q + 1].row0) ++q;
    const GgRedJob& J = p.job[q];
    const int m = brow - J.row0, Cin = J.Cin, taps = J.taps, R = Cin * taps;
    const int m_tile = m >> 7, r = m & 127;
    const float* __restrict__ vrow = J.v + (size_t)m * R;
    float* __restrict__ dvrow = J.dv + (size_t)m * R;
    const bool wn = J.g != nullptr;
    const float* part_row = p.part + (size_t)r * GG_SLOT_COLS;
    float ss = 0.f, dd = 0.f;
    // pass 1: partial slots -> reduced row in shared memory, torch order.  ncu (round 2): this kernel is bound by issue slots
    // (71 % busy, XU pipe saturated by the per-element e / Cin; DRAM at 17 % of peak).  Per-tap and tap-pair tiles: taps in the
    // outer loop (tile, slot base and slot count are uniform per tap), the threads cover c (Cin <= 256: two columns each),
    // coalesced read-only loads of the tile rows.  v is NOT touched here (see pass 2).
    if (!J.n_cblk) {
        const int c0 = threadIdx.x, c1 = threadIdx.x + GG_RED_NT;
        const bool ok0 = c0 < Cin, ok1 = c1 < Cin;
        const int tbase = J.tile0 + m_tile * (J.n_grp ? J.n_grp : taps);
        constexpr int JB = 5;                      // taps whose loads are issued back to back (one round trip for a k = 5 row)
        for (int j0 = 0; j0 < taps; j0 += JB) {
            float a0[JB], a1[JB];
#pragma unroll
            for (int u = 0; u < JB; ++u) {
                const int j = j0 + u;
                a0[u] = 0.f; a1[u] = 0.f;
                if (j < taps) {
                    const int tile = tbase + (J.n_grp ? (j >> 1) : j);
                    const float* src = part_row + (size_t)p.slot0[tile] * GG_SLOT_FLOATS + (J.n_grp ? (j & 1) * 256 : 0);
                    if (ok0) a0[u] = __ldg(src + c0);
                    if (ok1) a1[u] = __ldg(src + c1);
                    for (int sl = 1; sl < p.nslot[tile]; ++sl) {      // adding shares: one slot per tile, never taken
                        src += GG_SLOT_FLOATS;
                        if (ok0) a0[u] += __ldg(src + c0);
                        if (ok1) a1[u] += __ldg(src + c1);
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < JB; ++u) {
                const int j = j0 + u;
                if (j < taps && ok0) row[c0 * taps + j] = a0[u];
                if (j < taps && ok1) row[c1 * taps + j] = a1[u];
            }
        }
    } else {
        // all-taps tiles (opt-in): element by element (the partial tiles are read c-fastest = coalesced;
        // v is NOT touched here: indexed c * taps + j it would be a stride-taps read, 5x the sectors for the k = 5 convs)
        for (int e0 = threadIdx.x; e0 < R; e0 += 4 * GG_RED_NT) {
            float acc[4];
            int idx[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int e = e0 + u * GG_RED_NT;
                acc[u] = 0.f; idx[u] = -1;
                if (e < R) {
                    const int j = e / Cin, c = e - j * Cin;        // c fastest: coalesced reads of the partial tiles
                    const int tile = J.n_cblk ? J.tile0 + m_tile * J.n_cblk + c / GG_CB
                                   : J.n_grp ? J.tile0 + m_tile * J.n_grp + (j >> 1) : J.tile0 + m_tile * taps + j;
                    const float* src = part_row + (size_t)p.slot0[tile] * GG_SLOT_FLOATS +
                                       (J.n_cblk ? j * GG_CB + c % GG_CB : J.n_grp ? (j & 1) * 256 + c : c);
                    const int ns = p.nslot[tile];
                    idx[u] = c * taps + j;
                    for (int s = 0; s < ns; s += 4) {
                        const float a0 = src[(size_t)s * GG_SLOT_FLOATS];
                        const float a1 = (s + 1 < ns) ? src[(size_t)(s + 1) * GG_SLOT_FLOATS] : 0.f;
                        const float a2 = (s + 2 < ns) ? src[(size_t)(s + 2) * GG_SLOT_FLOATS] : 0.f;
                        const float a3 = (s + 3 < ns) ? src[(size_t)(s + 3) * GG_SLOT_FLOATS] : 0.f;
                        acc[u] += (a0 + a1) + (a2 + a3);
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (idx[u] >= 0) row[idx[u]] = acc[u];
        }
    }
    if (J.db && J.ones && threadIdx.x == GG_RED_NT - 1) {
        const int tile = J.tile0 + m_tile * (J.n_cblk ? J.n_cblk : J.n_grp ? J.n_grp : taps);
        const float* src = part_row + (size_t)p.slot0[tile] * GG_SLOT_FLOATS + (J.n_cblk ? taps * GG_CB : GG_ONES_BLK * 64);
        float acc = 0.f;
        for (int s = 0; s < p.nslot[tile]; ++s) acc += src[(size_t)s * GG_SLOT_FLOATS];
        J.db[m] += acc;
    }
    __syncthreads();
    if (!wn) {
        for (int e = threadIdx.x; e < R; e += GG_RED_NT) dvrow[e] += row[e];
        return;
    }
    // pass 2 (coalesced, torch order): <v, v> and <dW, v> of the row
    for (int e = threadIdx.x; e < R; e += GG_RED_NT) {
        const float vv = vrow[e];
        ss += vv * vv;
        dd += vv * row[e];
    }
    ss = warp_sum(ss);
    dd = warp_sum(dd);
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = ss; red[1][threadIdx.x >> 5] = dd; }
    __syncthreads();
    ss = 0.f; dd = 0.f;
#pragma unroll
    for (int w = 0; w < GG_RED_NT / 32; ++w) { ss += red[0][w]; dd += red[1][w]; }
    const float nrm = sqrtf(ss);
    const float sc = J.g[m] / nrm, proj = dd / ss;
    for (int e = threadIdx.x; e < R; e += GG_RED_NT) dvrow[e] += sc * (row[e] - vrow[e] * proj);
    if (threadIdx.x == 0) J.dg[m] += dd / nrm;
}

// ------------------------------------------------------------------------------------------ host
static bool wgrad_reduce_warp_rows() { const char* e = getenv("GTK_WGRAD_REDUCE_ROWS"); return !(e && e[0] == '0'); }
bool wgrad_red_enabled() { const char* e = getenv("GTK_WGRAD_RED"); return !(e && e[0] == '0'); }

size_t wgrad_group_workspace_bytes() { return (size_t)(WGG_MAX_CTAS + WGG_MAX_TILES) * GG_SLOT_FLOATS * sizeof(float); }

static int tmap_frames(CUtensorMap* out, const __nv_bfloat16* base, int ld, int cols, int T, int B, int box_rows = GG_TK) {
    const uint64_t dims[3] = {(uint64_t)cols, (uint64_t)T, (uint64_t)B};
    const uint64_t strides[2] = {(uint64_t)ld * 2, (uint64_t)ld * 2 * T};
    const uint32_t box[3] = {64, (uint32_t)box_rows, 1};
    return make_tmap_bf16(out, base, 3, dims, strides, box, true);
}

WgradReduceBatch* wgrad_reduce_batch(float* part_ws, int capacity) {
    static thread_local WgradReduceBatch rb;
    rb.M.n = 0; rb.part_ws = part_ws; rb.capacity = max(1, min(capacity, WGG_MAX_DEFER)); rb.max_row = 0;
    memset(rb.zeroed, 0, sizeof(rb.zeroed));
    return &rb;
}

int wgrad_group_reduce_flush(WgradReduceBatch* rb, cudaStream_t st) {
    if (!rb || rb->M.n == 0) return GTK_OK;
    const int rows = rb->M.row_end[rb->M.n - 1], rows_w = rb->M.row_end_w[rb->M.n - 1];
    ProfScope prof("wgrad_group_reduce", st);
    if (rows > 0) {
        GTK_CUDA(launch_pdl(wgrad_group_reduce_kernel, dim3(rows), dim3(GG_RED_NT), rb->max_row, st, rb->M));
        if (rows_w > 0) { int rc = check_launch("wgrad_group_reduce"); if (rc) return rc; }
    }
    if (rows_w > 0) GTK_CUDA(launch_pdl(wgrad_group_reduce_rows_kernel, dim3(rows_w), dim3(GG_RED_NT), 0, st, rb->M));
    rb->M.n = 0; rb->max_row = 0;
    memset(rb->zeroed, 0, sizeof(rb->zeroed));
    return check_launch("wgrad_group_reduce");
}

int wgrad_group_launch(const WgradGroupJob* jobs, int n_jobs, int B, int T, float* part_ws, const int* n_rows, cudaStream_t st,
                       int max_ctas, WgradReduceBatch* defer) {
    GTK_REQUIRE(jobs && n_jobs >= 1 && n_jobs <= WGG_MAX_JOBS && part_ws, "bad job list");
    if ((long long)B * T == 0) return GTK_OK;
    static thread_local GgParams P;                 // ~12 KB of kernel parameters, rebuilt per launch
    static thread_local GgRedMulti R1;              // immediate mode: a batch of one
    if (defer && defer->M.n == defer->capacity) { int rc = wgrad_group_reduce_flush(defer, st); if (rc) return rc; }
    if (defer) part_ws = defer->part_ws + (size_t)defer->M.n * (wgrad_group_workspace_bytes() / sizeof(float));
    GgRedMulti& RM = defer ? defer->M : R1;
    if (!defer) RM.n = 0;
    GgRedParams& R = RM.r[RM.n];
    memset(&P, 0, sizeof(P));
    memset(&R, 0, sizeof(R));
    GTK_REQUIRE(!n_rows || B == 1, "packed rows are one long sequence");
    P.B = B; P.T = T; P.chunks_t = ceil_div(T, GG_TK); P.part = part_ws; P.n_rows = n_rows;
    R.part = part_ws; R.n_jobs = n_jobs;
    const int S = B * P.chunks_t;
    GTK_REQUIRE(S <= 65535, "too many k-steps for the segment table");

    // all-taps tiles (opt-in, GTK_WGRAD_ALL_TAPS=1): k > 1, the per-tap accumulators (+ 16 bias columns) fit the 512 TMEM
    // columns, the staged X rows fit the three X boxes of a stage.  Measured on the bench workload (B200, round 2,
    // profiles/mma_issue_r02.txt): L2 -> SM traffic 490 -> 223 MB per coupling block, but an N = 64 MMA costs ~74 cycles
    // (its tensor-pipe time is 32), so the kernel ends up at the same time as the per-tap tiles (0.74 vs 0.70 ms per
    // step).  Per-tap stays the default; N = 96 column blocks (two boxes per stage) are the next thing to try.
    static const bool all_taps_on = getenv("GTK_WGRAD_ALL_TAPS") && atoi(getenv("GTK_WGRAD_ALL_TAPS")) != 0;
    auto all_taps = [&](const WgradGroupJob& J) {
        return all_taps_on && J.K > 1 && J.K * GG_CB + 16 <= GG_SLOT_COLS && GG_TK + 2 * (J.K / 2) * J.dil <= 3 * GG_TK;
    };
    // tap pairs (default, GTK_WGRAD_TAP_PAIRS=0 disables): k > 1 and the dil extra rows fit the staged X block
    static const bool pairs_on = !(getenv("GTK_WGRAD_TAP_PAIRS") && atoi(getenv("GTK_WGRAD_TAP_PAIRS")) == 0);
    auto tap_pairs = [&](const WgradGroupJob& J) { return pairs_on && !all_taps(J) && J.K > 1 && J.dil <= GG_PAIR_ROWS; };
    bool ones_ok = true;
    for (int q = 0; q < n_jobs; ++q) ones_ok = ones_ok && (all_taps(jobs[q]) || ceil_div(jobs[q].Cin, 64) <= GG_ONES_BLK);
    int n_tiles = 0, rows = 0, rows_w = 0, rc_at = 0;
    size_t max_row = 0;
    long long w_tile[WGG_MAX_TILES];
    for (int q = 0; q < n_jobs; ++q) {
        const WgradGroupJob& J = jobs[q];
        GTK_REQUIRE(J.dy && J.x && J.dv && (!J.g || (J.v && J.dg)), "null");
        GTK_REQUIRE(J.K >= 1 && (J.K & 1) && J.dil >= 1 && J.M >= 1 && J.Cin >= 1 && J.Cin <= 256, "bad conv shape");
        GTK_REQUIRE(J.ld_y % 8 == 0 && J.ld_x % 8 == 0 && J.cols_x >= J.Cin, "frame pitch");
        GTK_REQUIRE(J.cols_y >= J.M, "dY columns");
        GgJobDev& D = P.job[q];
        D.taps = J.K; D.dil = J.dil; D.n_blk = ceil_div(J.Cin, 64); D.halo = (J.K / 2) * J.dil;
        D.ones = (J.db && ones_ok) ? 1 : 0;
        const bool at = all_taps(J), tp = tap_pairs(J);
        if (at || tp) {
            rc_at = tmap_frames(&P.tmXh[q], J.x, J.ld_x, J.cols_x, T, B, at ? GG_TK + 2 * D.halo : GG_TK + J.dil);
            if (rc_at) return rc_at;
        }
        int rc = tmap_frames(&P.tmY[q], J.dy, J.ld_y, J.cols_y, T, B);
        if (rc) return rc;
        rc = tmap_frames(&P.tmX[q], J.x, J.ld_x, J.cols_x, T, B);
        if (rc) return rc;
        GgRedJob& Q = R.job[q];
        Q.g = J.g; Q.v = J.v; Q.dg = J.dg; Q.dv = J.dv; Q.db = J.db;
        Q.M = J.M; Q.Cin = J.Cin; Q.taps = J.K; Q.tile0 = n_tiles; Q.row0 = rows; Q.row0w = rows_w; Q.ones = D.ones;
        const bool warp_rows = J.K == 1 && !at && wgrad_reduce_warp_rows();
        Q.n_cblk = at ? D.n_blk : 0;
        Q.n_grp = tp ? (J.K + 1) / 2 : 0;
        if (warp_rows) rows_w += ceil_div(J.M, GG_RED_NT / 32);
        else rows += J.M;
        max_row = max(max_row, (size_t)J.Cin * J.K * sizeof(float));
        const int m_tiles = ceil_div(J.M, 128);
        const int per_m = at ? D.n_blk : tp ? (J.K + 1) / 2 : J.K;
        GTK_REQUIRE(n_tiles + m_tiles * per_m <= WGG_MAX_TILES && m_tiles <= 255 && J.K < GG_ALL_TAPS, "too many tiles in one group");
        // tile weight = measured cycles of one k16 step: with two accumulator chains an MMA of ANY N <= 256 is accepted every
        // ~128 cycles (profiles/mma_issue_r02.txt), so a single-tap tile costs 128 per k16 step whatever its width and a tap
        // pair 256; five chains of N = 64 (all-taps) ~74 each.  (Round-2 ncu capture of the kernel with weights N / 2:
        // sm cycles active 68 k of 99 k elapsed -- CTAs whose share was made of narrow tiles got up to twice the k-steps.)
        for (int mt = 0; mt < m_tiles; ++mt)
            for (int j = 0; j < per_m; ++j) {
                GgTile& Tl = P.tile[n_tiles];
                Tl.job = (unsigned char)q; Tl.m_tile = (unsigned char)mt;
                if (at) {
                    Tl.tap = (unsigned char)GG_ALL_TAPS; Tl.cblk = (unsigned char)j;
                    w_tile[n_tiles] = (long long)J.K * 76 + ((D.ones && j == 0) ? 76 : 0);
                } else if (tp) {
                    Tl.tap = (unsigned char)(2 * j); Tl.cblk = (unsigned char)((2 * j + 1 < J.K) ? 1 : 0);
                    w_tile[n_tiles] = Tl.cblk ? GG_W_PAIR : GG_W_SINGLE;
                } else {
                    Tl.tap = (unsigned char)j; Tl.cblk = 0;
                    w_tile[n_tiles] = GG_W_SINGLE;
                }
                ++n_tiles;
            }
    }
    GTK_REQUIRE(max_row <= 48 * 1024, "weight row too long for the reduce kernel");

    // stream-K cut: tile i covers [pos_i, pos_i + w_i S) on a line of length U; CTA c owns [c U / G, (c+1) U / G)
    long long U = 0;
    for (int i = 0; i < n_tiles; ++i) U += w_tile[i] * S;
    int G = (int)min((long long)min(sm_count(), WGG_MAX_CTAS), (long long)n_tiles * S);
    // Overlapped launch (side stream): leave the other SMs to the data-gradient chain -- and take them as whole TPCs
    // (clusters of 2): the fused layer kernels launch as CTA pairs, which need both SMs of a TPC free.
    const bool paired = max_ctas > 0;
    if (paired && G > max_ctas) G = max_ctas;
    if (paired && G > 1) G &= ~1;
    // Adding shares (default, GTK_WGRAD_RED=0 restores one partial tile per share): every share of a tile adds its
    // accumulator into the tile's ONE zeroed slot with vector reds, which merge in L2 -- the reduce kernel reads one tile
    // instead of ~4 partial tiles per tile (its DRAM reads were 534 MB per step, profiles/wgrad_group_trace_r02.txt).  The
    // price is the summation order: weight gradients differ in their last bits from run to run (like the bias gradients).
    const bool red = wgrad_red_enabled();
    P.red = red ? 1 : 0;
    int n_seg = 0, slot = 0, c = 0;
    long long pos = 0;
    for (int i = 0; i < n_tiles; ++i) {
        const long long w = w_tile[i];
        R.slot0[i] = (unsigned short)slot;
        int s = 0;
        while (s < S) {
            while (c + 1 < G && pos + w * s >= (long long)(c + 1) * U / G) ++c;       // CTA that owns k-step s
            const long long hi = (c + 1 < G) ? (long long)(c + 1) * U / G : U + w;     // first position it does not own
            int s_hi = (int)min((long long)S, (hi - pos + w - 1) / w);
            if (s_hi <= s) s_hi = s + 1;
            GTK_REQUIRE(n_seg < WGG_MAX_SEGS, "segment table overflow");
            GgSeg& Sg = P.seg[n_seg];
            Sg.tile = (unsigned short)i; Sg.s_lo = (unsigned short)s; Sg.s_hi = (unsigned short)s_hi; Sg.slot = (unsigned short)slot;
            if (P.cta[c].nseg == 0) P.cta[c].seg0 = (unsigned short)n_seg;
            ++P.cta[c].nseg;
            ++n_seg;
            if (!red) ++slot;
            s = s_hi;
        }
        if (red) ++slot;
        R.nslot[i] = (unsigned short)(slot - R.slot0[i]);
        pos += w * S;
    }
    GTK_REQUIRE((size_t)slot * GG_SLOT_FLOATS * sizeof(float) <= wgrad_group_workspace_bytes(), "partial workspace");

    const size_t smem = (size_t)GG_STAGES * GG_STAGE_BYTES + sizeof(GgCtl) + 1024;
    GTK_CUDA(cudaFuncSetAttribute(wgrad_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (red) {
        // the tile slots must be zero when the first share lands.  A batch zeroes the slots of its NEXT launch right behind
        // this one (the stream is idle then: the memset stays off the path of the launch that needs it)
        const size_t slot_bytes = (size_t)GG_SLOT_FLOATS * sizeof(float);
        const int have = defer ? defer->zeroed[defer->M.n] : 0;
        if (n_tiles > have)
            GTK_CUDA(cudaMemsetAsync(part_ws + (size_t)have * GG_SLOT_FLOATS, 0, (size_t)(n_tiles - have) * slot_bytes, st));
    }
    {
        ProfScope prof("wgrad_group", st);
        if (paired && G > 1) {
            cudaLaunchConfig_t cfg;
            memset(&cfg, 0, sizeof(cfg));
            cfg.gridDim = dim3(G); cfg.blockDim = dim3(NTHREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
            cudaLaunchAttribute attr[2];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
            attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
            cfg.attrs = attr; cfg.numAttrs = 2;
            GTK_CUDA(cudaLaunchKernelEx(&cfg, wgrad_group_kernel, P));
        } else
        GTK_CUDA(launch_pdl(wgrad_group_kernel, dim3(G), dim3(NTHREADS), smem, st, P));
        int rc = check_launch("wgrad_group");
        if (rc) return rc;
    }
    RM.row_end[RM.n] = (RM.n ? RM.row_end[RM.n - 1] : 0) + rows;
    RM.row_end_w[RM.n] = (RM.n ? RM.row_end_w[RM.n - 1] : 0) + rows_w;
    ++RM.n;
    if (defer) {
        if (red && defer->M.n < defer->capacity) {
            GTK_CUDA(cudaMemsetAsync(defer->part_ws + (size_t)defer->M.n * (wgrad_group_workspace_bytes() / sizeof(float)), 0,
                                     (size_t)n_tiles * GG_SLOT_FLOATS * sizeof(float), st));
            defer->zeroed[defer->M.n] = n_tiles;
        }
        defer->max_row = max(defer->max_row, max_row);
    } else {
        ProfScope prof("wgrad_group_reduce", st);
        if (rows > 0) {
            GTK_CUDA(launch_pdl(wgrad_group_reduce_kernel, dim3(rows), dim3(GG_RED_NT), (size_t)max_row, st, RM));
            if (rows_w > 0) { int rc = check_launch("wgrad_group_reduce"); if (rc) return rc; }
        }
        if (rows_w > 0) GTK_CUDA(launch_pdl(wgrad_group_reduce_rows_kernel, dim3(rows_w), dim3(GG_RED_NT), 0, st, RM));
        int rc = check_launch("wgrad_group_reduce");
        if (rc) return rc;
    }
    // bias gradients that could not ride along (Cin > 192): plain column sums
    for (int q = 0; q < n_jobs; ++q) {
        const WgradGroupJob& J = jobs[q];
        if (!J.db || P.job[q].ones) continue;
        int rc = n_rows ? colsum_frames_seg(J.dy, J.ld_y, J.M, 1, n_rows - 1, J.db, 0, st)     // hdr = {0, n_rows}
                        : colsum_frames(J.dy, J.ld_y, J.M, 1, (long long)B * T, J.db, 0, st);
        if (rc) return rc;
    }
    return GTK_OK;
}

}  // namespace tc
}  // namespace gtk
