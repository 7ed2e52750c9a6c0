// glowtts_b200/csrc/wn_tc.cu -- one WaveNet layer (modules.py:151-170) as ONE tcgen05 kernel per 128-frame tile:
//
//   GEMM 1   xin[t, 0:2H]  = sum_{j,c} h[t + (j-K/2) dil, c] * Win[j][n][c]          (dilated conv)
//   epilogue acts[t, 0:H]  = tanh(drop(xin[:, :H] + b) + g) * sigmoid(drop(xin[:, H:] + b) + g)   (commons.py:61-68)
//            -> written as bf16 straight into the shared-memory A operand of GEMM 2 (canonical SWIZZLE_128B layout)
//   GEMM 2   rs[t, 0:RS]   = acts[t, :] * Wrs[n][c]                                   (1x1 res/skip conv)
//   epilogue h' = (h + rs[:, :H] + b) * mask ;  skip (+)= rs[:, H:] + b               (modules.py:166-170)
//
// The 2H-channel pre-activation and the gate output never touch HBM (except the bf16 copies the backward needs,
// written only when training).  Thread <-> frame ownership (TMEM lane = frame) makes every epilogue step thread-local.
//
// Round-2 structure (the round-1 kernel ran GEMM 1 -> epilogue 1 -> GEMM 2 -> epilogue 2 strictly in sequence, re-read
// the activation tile from L2 once per tap and streamed all 884 KB of weights into every CTA: per launch it was bound by
// one CTA's latency and by L2 bytes, profiles/wn_layer_phase_trace_r01.txt, profiles/wgrad_overlap_r02.txt):
//   * the activation tile (128 frames + halo) is loaded ONCE and stays in shared memory; a tap is a row-shifted UMMA
//     descriptor on it (the 128-byte swizzle is a function of the address only, profiles/shift_probe.py);
//   * CTAs run as PAIRS (cluster of 2 neighbouring tiles): every weight tile is fetched half by each CTA and TMA-multicast
//     into both shared memories -- half the L2 -> SM weight traffic;
//   * GEMM 1 is cut into N-chunks of 64 gate channels (64 tanh + 64 sigmoid columns, two TMEM accumulators taking alternate
//     k16 steps).  Chunks alternate between two TMEM regions, so the gate epilogue of chunk j -- and its bf16 tanh|sigmoid
//     stores -- runs under the MMAs of chunk j + 1 instead of after the whole GEMM;
//   * warp roles: warp 0 = TMA producer, warp 1 = MMA issuer, warp 2 = TMEM allocator, warps 4..19 = epilogue.
#include <string.h>

#include "tc.cuh"
#include "ops.h"

namespace gtk {
namespace tc {

constexpr int WN_STAGES = 6;                         // weight ring
constexpr int WN_STAGE_BYTES = 128 * KC * 2;         // one stage: [128 weight rows x 64 k] bf16 = 16 KB
constexpr int WN_CL = 2;                             // CTAs per cluster
constexpr int WN_EPI_WARPS = 16, WN_EPI_NT = WN_EPI_WARPS * 32;
constexpr int WN_NT = 128 + WN_EPI_NT;               // 4 role warps + 16 epilogue warps
constexpr int WN_MAX_CHUNKS = 4;

struct WnCtl {
    uint64_t full[WN_STAGES], empty[WN_STAGES];      // weight ring (empty counts the pair: both CTAs read every stage)
    uint64_t a_full;                                 // activation tile landed
    uint64_t g1_full[WN_MAX_CHUNKS];                 // GEMM-1 chunk j complete in TMEM
    uint64_t reg_empty[2];                           // TMEM region (chunk parity) drained by the epilogue
    uint64_t acts_ready, g2_full;
    uint32_t tmem_base;
    alignas(16) float bias_in[512];                  // [2H]
    alignas(16) float cw[512];                       // [2H]  WNP conditioner weight / bias per channel (zeros when there is none)
    alignas(16) float cb[512];
    alignas(16) float bias_rs[512];                  // [RS]
};

// Instrumented builds only (-DGTK_INSTRUMENT): phase timestamps (globaltimer, ns) of every CTA: g_wn_trace[cta][start,
// setup done, first GEMM-1 chunk done, last gate epilogue done, GEMM 2 done, end], cta = blockIdx.y * gridDim.x +
// blockIdx.x (first WN_TRACE_CTAS only); read with gtk_debug_wn_trace for the per-phase breakdown in profiles/.
#ifdef GTK_INSTRUMENT
constexpr int WN_TRACE_CTAS = 512;
__device__ unsigned long long g_wn_trace[WN_TRACE_CTAS * 6];
__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#define WN_TRACE(i)                                                                     \
    do {                                                                                \
        const int cta_ = blockIdx.y * gridDim.x + blockIdx.x;                           \
        if (threadIdx.x == 128 && cta_ < WN_TRACE_CTAS) g_wn_trace[cta_ * 6 + (i)] = gtimer(); \
    } while (0)
#else
#define WN_TRACE(i) do { } while (0)
#endif

// One item of the MMA stream = one 16 KB weight stage.
//   it <  n1: GEMM 1, chunk j, tap, k-chunk kc: rows 0..63 = tanh channels [64 j, 64 j + 64), rows 64..127 = the sigmoid ones
//   it >= n1: GEMM 2, k-chunk kc, output block m: rows = res/skip channels [128 m, 128 m + nrows)
struct WnItem { int g2, j, tap, kc, m, nrows; };
__device__ __forceinline__ WnItem wn_item(int it, int n1, int taps, int n_kh, int m2, int RS) {
    WnItem I;
    if (it < n1) {
        I.g2 = 0; I.j = it / (taps * n_kh);
        const int rem = it - I.j * taps * n_kh;
        I.tap = rem / n_kh; I.kc = rem - I.tap * n_kh; I.m = 0; I.nrows = 128;
    } else {
        const int i2 = it - n1;
        I.g2 = 1; I.kc = i2 / m2; I.m = i2 - I.kc * m2; I.j = 0; I.tap = 0;
        I.nrows = min(128, RS - 128 * I.m);
    }
    return I;
}

__global__ void __launch_bounds__(WN_NT, 1)
wn_layer_kernel(const __grid_constant__ CUtensorMap tmH, const __grid_constant__ CUtensorMap tmWin,
                const __grid_constant__ CUtensorMap tmWrs, const __grid_constant__ CUtensorMap tmActs,
                const __grid_constant__ CUtensorMap tmOut, const WnLayerParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int H = p.H, H2 = 2 * p.H, RS = p.RS;
    const int n_kh = H / KC;                             // k-chunks of the hidden dimension == GEMM-1 chunks (<= 3)
    const int halo = (p.taps / 2) * p.dil;
    const int R = (BT + 2 * halo + 7) & ~7;              // rows of the resident activation tile
    const int a_tile = R * 128;                          // bytes of one k-chunk of it (a multiple of 1024)
    unsigned char* a_res = smem;                         // [n_kh][R][64] bf16: frames t0 - halo .. ; later the output staging
    unsigned char* acts = a_res + (size_t)n_kh * a_tile; // [n_kh][128][64] bf16: gate output = A operand of GEMM 2
    unsigned char* ring = acts + (size_t)n_kh * A_BYTES;
    WnCtl* ctl = reinterpret_cast<WnCtl*>(ring + (size_t)WN_STAGES * WN_STAGE_BYTES);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t0 = blockIdx.x * BT, b = blockIdx.y;
    pdl_trigger();
    // varlen packing: a pair of tiles with nothing alive leaves at once (pair-uniform: the peer's barriers are involved)
    if (p.n_rows && (int)(blockIdx.x & ~(unsigned)(WN_CL - 1)) * BT >= *p.n_rows) return;
    WN_TRACE(0);
    const int n1 = n_kh * p.taps * n_kh;                 // GEMM-1 items
    const int m2 = (RS + 127) / 128;                     // output blocks of GEMM 2
    const int n_all = n1 + n_kh * m2;
    const uint32_t rank = cluster_ctarank();
    const uint16_t cl_mask = (uint16_t)((1u << WN_CL) - 1);

    if (threadIdx.x == 0) {
        for (int s = 0; s < WN_STAGES; ++s) { mbar_init(&ctl->full[s], 1); mbar_init(&ctl->empty[s], WN_CL); }
        mbar_init(&ctl->a_full, 1);
        for (int j = 0; j < WN_MAX_CHUNKS; ++j) mbar_init(&ctl->g1_full[j], 1);
        mbar_init(&ctl->reg_empty[0], WN_EPI_NT);
        mbar_init(&ctl->reg_empty[1], WN_EPI_NT);
        mbar_init(&ctl->acts_ready, WN_EPI_NT);
        mbar_init(&ctl->g2_full, 1);
        mbar_fence_init();
    }
    for (int i = threadIdx.x; i < H2; i += WN_NT) {
        ctl->bias_in[i] = p.bias_in[i];
        ctl->cw[i] = p.trk ? p.cw[i] : 0.f;
        ctl->cb[i] = p.trk ? p.cb[i] : 0.f;
    }
    for (int i = threadIdx.x; i < RS; i += WN_NT) ctl->bias_rs[i] = p.bias_rs[i];
    if (warp == 2) tmem_alloc<512>(&ctl->tmem_base);
    tc_fence_before();
    cluster_sync_all();                                  // the peer's barriers exist before anything is multicast
    tc_fence_after();
    // PDL: everything the previous kernel of the stream wrote (h, skip) may be touched only after pdl_wait().  The TMA
    // producer first requests the WEIGHT tiles of its first stages (packed long before the previous kernel started), so
    // they are in flight while the previous kernel drains, and waits afterwards.
    const bool is_producer = (warp == 0 && lane == 0);
    if (!is_producer) pdl_wait();
    const uint32_t tmem = ctl->tmem_base;
    WN_TRACE(1);

    if (warp == 0) {
        // ------------------------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            tma_prefetch_desc(&tmH);
            tma_prefetch_desc(&tmWin);
            tma_prefetch_desc(&tmWrs);
            // this CTA's half of weight stage `it` (the peer delivers the other half), multicast into both CTAs
            auto load_w = [&](int it, int s) {
                const WnItem I = wn_item(it, n1, p.taps, n_kh, m2, RS);
                unsigned char* sb = ring + (size_t)s * WN_STAGE_BYTES;
                mbar_arrive_expect_tx(&ctl->full[s], (uint32_t)(I.nrows * KC * 2));
                if (!I.g2) {
                    // rank 0: the 64 tanh rows, rank 1: the 64 sigmoid rows
                    tma_load_3d_multicast(sb + (size_t)rank * 64 * 128, &tmWin, &ctl->full[s], I.kc * KC, (int)rank * H + 64 * I.j,
                                          I.tap, cl_mask);
                } else if (I.nrows == 128) {
                    tma_load_2d_multicast(sb + (size_t)rank * 64 * 128, &tmWrs, &ctl->full[s], I.kc * KC, 128 * I.m + 64 * (int)rank,
                                          cl_mask);
                } else if (rank == 0) {                  // 64-row tail block (RS = H = 64 or 192): one box, from rank 0
                    tma_load_2d_multicast(sb, &tmWrs, &ctl->full[s], I.kc * KC, 128 * I.m, cl_mask);
                }
            };
            const int n_pre = min(WN_STAGES, n_all);
            for (int it = 0; it < n_pre; ++it) load_w(it, it);
            pdl_wait();
            mbar_arrive_expect_tx(&ctl->a_full, (uint32_t)(n_kh * a_tile));
            for (int kc = 0; kc < n_kh; ++kc) tma_load_3d(a_res + (size_t)kc * a_tile, &tmH, &ctl->a_full, kc * KC, t0 - halo, b);
            for (int it = n_pre; it < n_all; ++it) {
                const int s = it % WN_STAGES, use = it / WN_STAGES;
                mbar_wait(&ctl->empty[s], (use - 1) & 1);           // BOTH CTAs have consumed the stage's previous tile
                load_w(it, s);
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            const uint32_t idesc128 = umma_idesc_bf16(BT, 128), idesc64 = umma_idesc_bf16(BT, 64);
            mbar_wait(&ctl->a_full, 0);
            tc_fence_after();
            int it = 0;
            // GEMM 1: chunk j accumulates in TMEM region (j & 1) * 256: two accumulators of 128 columns take alternate k16
            // steps (back-to-back MMAs into one accumulator serialise on its read-modify-write latency); the epilogue adds them.
            for (int j = 0; j < n_kh; ++j) {
                if (j >= 2) { mbar_wait(&ctl->reg_empty[j & 1], ((j >> 1) - 1) & 1); tc_fence_after(); }
                const uint32_t dreg = tmem + (uint32_t)(j & 1) * 256;
                const int per = p.taps * n_kh;
                for (int q = 0; q < per; ++q, ++it) {
                    const int s = it % WN_STAGES, use = it / WN_STAGES;
                    const int tap = q / n_kh, kc = q - tap * n_kh;
                    mbar_wait(&ctl->full[s], use & 1);
                    tc_fence_after();
                    const uint64_t adesc = umma_desc_kmajor_sw128(smem_u32(a_res + (size_t)kc * a_tile) + (uint32_t)(tap * p.dil) * 128u);
                    const uint64_t bdesc = umma_desc_kmajor_sw128(smem_u32(ring + (size_t)s * WN_STAGE_BYTES));
#pragma unroll
                    for (int k = 0; k < KC / 16; ++k)
                        umma_bf16(dreg + (k & 1) * 128, adesc + 2 * k, bdesc + 2 * k, idesc128, q != 0 || k >= 2);
                    umma_commit_multicast(&ctl->empty[s], cl_mask);
                }
                umma_commit(&ctl->g1_full[j]);
            }
            // GEMM 2 reuses TMEM columns [0, RS): every gate epilogue has drained its region and written its acts
            mbar_wait(&ctl->acts_ready, 0);
            tc_fence_after();
            if (p.acts_save) {                           // bf16 gate output for the res_skip weight gradient
                for (int kc = 0; kc < n_kh; ++kc) tma_store_3d(&tmActs, acts + (size_t)kc * A_BYTES, kc * KC, t0, b);
                tma_store_commit();
            }
            for (int kc = 0; kc < n_kh; ++kc) {
                // the m2 output blocks are independent accumulators: their k16 steps interleave
                for (int m = 0; m < m2; ++m) {
                    const int s = (it + m) % WN_STAGES, use = (it + m) / WN_STAGES;
                    mbar_wait(&ctl->full[s], use & 1);
                }
                tc_fence_after();
                const uint64_t adesc = umma_desc_kmajor_sw128(smem_u32(acts + (size_t)kc * A_BYTES));
#pragma unroll
                for (int k = 0; k < KC / 16; ++k)
                    for (int m = 0; m < m2; ++m) {
                        const int s = (it + m) % WN_STAGES;
                        const uint64_t bdesc = umma_desc_kmajor_sw128(smem_u32(ring + (size_t)s * WN_STAGE_BYTES));
                        umma_bf16(tmem + 128 * m, adesc + 2 * k, bdesc + 2 * k, (RS - 128 * m >= 128) ? idesc128 : idesc64,
                                  kc != 0 || k != 0);
                    }
                for (int m = 0; m < m2; ++m) umma_commit_multicast(&ctl->empty[(it + m) % WN_STAGES], cl_mask);
                it += m2;
            }
            umma_commit(&ctl->g2_full);
        }
    }
    __syncwarp();

    if (warp >= 4) {
        // ------------------------------------------------------------------------------------ epilogue warps
        // Warp w owns TMEM lane quarter w % 4 (frames) and column part (w - 4) / 4 of 4.
        const int q = warp & 3, part = (warp - 4) >> 2;
        const int r = q * 32 + lane;                     // row of the tile == TMEM lane
        const int t = t0 + r;
        const bool t_ok = t < p.T;
        const float mk = (p.mask && t_ok) ? p.mask[(size_t)b * p.T + t] : 1.f;
        const bool live = t_ok && mk != 0.f;             // masked rows are written as exact zeros (never value * 0)
        const size_t frame = (size_t)b * p.T + (t_ok ? t : 0);
        // dropout counter: the un-packed frame index, so that packed and [B, T] runs draw the same mask
        const unsigned long long dframe = p.row_fr ? (unsigned long long)p.row_fr[frame] : (unsigned long long)frame;
        // conditioning of this row: speaker embedding projection (per utterance) and / or the WNP pitch-energy track
        const float* cond_row = p.cond ? p.cond + (size_t)(p.row_b ? p.row_b[frame] : b) * p.cond_bstride : nullptr;
        const float* trk_row = p.trk ? p.trk + frame * p.n_sqz : nullptr;
        const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
        RowIo io;
        io.tp = p.tile_layout; io.tile_id = b * ((p.T + BT - 1) / BT) + blockIdx.x; io.r = r; io.frame = frame;
        const unsigned long long dseed = p.drop_p > 0.f ? drop_resolve(p.drop_seed) : 0ull;

        // ---- epilogue 1, chunk by chunk: gate -> A operand of GEMM 2.  Chunk j = gate channels [64 j, 64 j + 64); this
        // thread takes 16 of them: tanh columns lc.., sigmoid columns 64 + lc.. of both accumulators of the chunk's region.
        for (int j = 0; j < n_kh; ++j) {
            mbar_wait(&ctl->g1_full[j], 0);
            tc_fence_after();
            if (j == 0) WN_TRACE(2);
            const uint32_t reg = trow + (uint32_t)(j & 1) * 256 + 16 * part;
            const int c0 = 64 * j + 16 * part;
            float a[16], g[16];
            tmem_ld16_sum(reg, reg + 128, a);
            tmem_ld16_sum(reg + 64, reg + 192, g);
            tc_fence_before();
            mbar_arrive(&ctl->reg_empty[j & 1]);          // the region may be refilled (chunk j + 2)
#pragma unroll
            for (int i = 0; i < 16; i += 4) {
                const float4 ba = *reinterpret_cast<const float4*>(&ctl->bias_in[c0 + i]);
                const float4 bg = *reinterpret_cast<const float4*>(&ctl->bias_in[H + c0 + i]);
                a[i] += ba.x; a[i + 1] += ba.y; a[i + 2] += ba.z; a[i + 3] += ba.w;
                g[i] += bg.x; g[i + 1] += bg.y; g[i + 2] += bg.z; g[i + 3] += bg.w;
            }
            if (p.drop_p > 0.f) {
                float sc[16];
                dropout_scale16(dseed, dframe * H2 + c0, p.drop_p, sc);
#pragma unroll
                for (int i = 0; i < 16; ++i) a[i] *= sc[i];
                dropout_scale16(dseed, dframe * H2 + H + c0, p.drop_p, sc);
#pragma unroll
                for (int i = 0; i < 16; ++i) g[i] *= sc[i];
            }
            if (cond_row) {                              // cond_layer(g) slice of this utterance (modules.py:148-157)
#pragma unroll
                for (int i = 0; i < 16; i += 4) {
                    const float4 ca = __ldg(reinterpret_cast<const float4*>(cond_row + c0 + i));
                    const float4 cg = __ldg(reinterpret_cast<const float4*>(cond_row + H + c0 + i));
                    a[i] += ca.x; a[i + 1] += ca.y; a[i + 2] += ca.z; a[i + 3] += ca.w;
                    g[i] += cg.x; g[i + 1] += cg.y; g[i + 2] += cg.z; g[i + 3] += cg.w;
                }
            }
            if (trk_row) {                               // WNP: cond_layer1(track) squeezed, rank 1 in time (modules.py:320-336)
                const float pa = trk_row[(p.cs_base + c0) / p.cs_div], pg = trk_row[(p.cs_base + H + c0) / p.cs_div];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    a[i] += fmaf(ctl->cw[c0 + i], pa, ctl->cb[c0 + i]);
                    g[i] += fmaf(ctl->cw[H + c0 + i], pg, ctl->cb[H + c0 + i]);
                }
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                a[i] = live ? tanh_fast(a[i]) : 0.f;
                g[i] = live ? sigmoid_fast(g[i]) : 0.f;
            }
            if (p.ts_save && t_ok) {
                row_store_bf16x16(p.ts_save, H2, c0, io, a);
                row_store_bf16x16(p.ts_save, H2, H + c0, io, g);
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] *= g[i];
            // gate output -> A operand of GEMM 2 (canonical K-major SWIZZLE_128B); the same tiles are TMA-stored to acts_save
            stage_bf16x16_sw128(acts, A_BYTES, r, c0, a);
        }
        fence_proxy_async_smem();
        tc_fence_before();
        mbar_arrive(&ctl->acts_ready);
        WN_TRACE(3);

        // ---- epilogue 2: residual / skip.  The bf16 operand copy of the result (next layer's input, or the masked WN
        // output for `end`) is staged over the activation tile (GEMM 1 is long done) and leaves as TMA stores.
        const int n_chunks = H / 16;
        const int ch_lo = n_chunks * part / 4, ch_hi = n_chunks * (part + 1) / 4;
        mbar_wait(&ctl->g2_full, 0);
        tc_fence_after();
        WN_TRACE(4);
        for (int cc = ch_lo; cc < ch_hi; ++cc) {
            const int c0 = cc * 16;
            float v[16], sk[16];
            if (RS > H) {
                tmem_ld16(trow + c0, v);
                tmem_ld16(trow + H + c0, sk);
            } else {
                tmem_ld16(trow + c0, sk);
            }
            if (!t_ok) continue;
            if (RS > H) {
                float x[16];
                row_load_f32x16(p.h_f32, H, c0, io, x);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float4 bb = *reinterpret_cast<const float4*>(&ctl->bias_rs[c0 + 4 * i]);
                    v[4 * i] = live ? v[4 * i] + bb.x + x[4 * i] : 0.f;
                    v[4 * i + 1] = live ? v[4 * i + 1] + bb.y + x[4 * i + 1] : 0.f;
                    v[4 * i + 2] = live ? v[4 * i + 2] + bb.z + x[4 * i + 2] : 0.f;
                    v[4 * i + 3] = live ? v[4 * i + 3] + bb.w + x[4 * i + 3] : 0.f;
                }
                row_store_f32x16(p.h_out_f32, H, c0, io, v);
                stage_bf16x16_sw128(a_res, (size_t)a_tile, r, c0, v);
            }
            const int bo = (RS > H) ? H : 0;
#pragma unroll
            for (int i = 0; i < 16; ++i) sk[i] += ctl->bias_rs[bo + c0 + i];
            if (!p.first) {
                float x[16];
                row_load_f32x16(p.skip_f32, H, c0, io, x);
#pragma unroll
                for (int i = 0; i < 16; ++i) sk[i] += x[i];
            }
            if (p.last) {                                 // WN output = skip * mask (modules.py:171), operand of `end`
#pragma unroll
                for (int i = 0; i < 16; ++i) sk[i] = live ? sk[i] : 0.f;
                stage_bf16x16_sw128(a_res, (size_t)a_tile, r, c0, sk);
                if (p.h_out_f32_last) row_store_f32x16(p.h_out_f32_last, H, c0, io, sk);   // fp32 master for a chained WN
            } else {
                row_store_f32x16(p.skip_f32, H, c0, io, sk);
            }
        }
        fence_proxy_async_smem();
        tc_fence_before();
        WN_TRACE(5);
        asm volatile("bar.sync 1, %0;" ::"n"(WN_EPI_NT) : "memory");      // the epilogue warps only
        if (warp == 4 && lane == 0) {
            for (int kc = 0; kc < n_kh; ++kc) tma_store_3d(&tmOut, a_res + (size_t)kc * a_tile, kc * KC, t0, b);
            tma_store_commit();
            tma_store_wait_read();                        // shared memory must outlive the bulk reads
        }
    } else if (warp == 1 && lane == 0) {
        if (p.acts_save) tma_store_wait_read();           // the acts tiles this thread handed to the TMA
    }
    tc_fence_before();
    cluster_sync_all();                                   // the peer's multicasts / commits may still be arriving here
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc<512>(tmem);
    }
}

#ifdef GTK_INSTRUMENT
int wn_trace_read(unsigned long long* out, int n_ctas) {
    GTK_REQUIRE(out && n_ctas >= 0 && n_ctas <= WN_TRACE_CTAS, "n_ctas");
    GTK_CUDA(cudaMemcpyFromSymbol(out, g_wn_trace, size_t(n_ctas) * 6 * sizeof(unsigned long long)));
    return GTK_OK;
}
#endif

static size_t wn_layer_smem(int H, int taps, int dil) {
    const int halo = (taps / 2) * dil, R = (BT + 2 * halo + 7) & ~7;
    return (size_t)(H / KC) * R * 128 + (size_t)(H / KC) * A_BYTES + (size_t)WN_STAGES * WN_STAGE_BYTES + sizeof(WnCtl) + 1024;
}

int wn_layer_launch(const __nv_bfloat16* h_bf16, const __nv_bfloat16* win_p, const __nv_bfloat16* wrs_p, WnLayerParams p,
                    cudaStream_t st) {
    GTK_REQUIRE(h_bf16 && win_p && wrs_p && p.bias_in && p.bias_rs, "null");
    GTK_REQUIRE(p.H % KC == 0 && p.H / KC <= 3 && p.H <= 256, "fused WN layer needs hidden in {64,128,192}");
    GTK_REQUIRE(p.RS == p.H || p.RS == 2 * p.H, "res_skip channels");
    GTK_REQUIRE(p.taps >= 1 && (p.taps & 1) && p.dil >= 1, "bad conv shape");
    GTK_REQUIRE(p.last ? (p.RS == p.H && p.skip_out_bf16) : (p.RS == 2 * p.H && p.h_f32 && p.h_out_f32 && p.h_out_bf16), "layer outputs");
    GTK_REQUIRE(p.first || p.skip_f32, "skip accumulator");
    GTK_REQUIRE(p.last || p.skip_f32, "skip accumulator");
    GTK_REQUIRE(!p.trk || (p.cw && p.cb && p.n_sqz >= 1 && p.cs_div >= 1), "time-varying conditioner arguments");
    if (p.B == 0 || p.T == 0) return GTK_OK;
    const int H = p.H, halo = (p.taps / 2) * p.dil, R = (BT + 2 * halo + 7) & ~7;
    GTK_REQUIRE(R <= 256, "dilation too large for the resident activation tile");
    const size_t smem = wn_layer_smem(H, p.taps, p.dil);
    GTK_REQUIRE(smem <= 227 * 1024, "tile does not fit in shared memory");
    CUtensorMap tmH, tmWin, tmWrs;
    {
        const uint64_t dims[3] = {(uint64_t)H, (uint64_t)p.T, (uint64_t)p.B};
        const uint64_t strides[2] = {(uint64_t)H * 2, (uint64_t)H * 2 * p.T};
        const uint32_t box[3] = {KC, (uint32_t)R, 1};
        int rc = make_tmap_bf16(&tmH, h_bf16, 3, dims, strides, box, true);
        if (rc) return rc;
    }
    {
        const uint64_t dims[3] = {(uint64_t)H, (uint64_t)2 * H, (uint64_t)p.taps};
        const uint64_t strides[2] = {(uint64_t)H * 2, (uint64_t)H * 2 * 2 * H};
        const uint32_t box[3] = {KC, 64, 1};
        int rc = make_tmap_bf16(&tmWin, win_p, 3, dims, strides, box, true);
        if (rc) return rc;
    }
    {
        const uint64_t dims[2] = {(uint64_t)H, (uint64_t)p.RS};
        const uint64_t strides[1] = {(uint64_t)H * 2};
        const uint32_t box[2] = {KC, 64};
        int rc = make_tmap_bf16(&tmWrs, wrs_p, 2, dims, strides, box, true);
        if (rc) return rc;
    }
    CUtensorMap tmActs, tmOut;
    {
        const uint64_t dims[3] = {(uint64_t)H, (uint64_t)p.T, (uint64_t)p.B};
        const uint64_t strides[2] = {(uint64_t)H * 2, (uint64_t)H * 2 * p.T};
        const uint32_t box[3] = {KC, BT, 1};
        __nv_bfloat16* out = p.last ? p.skip_out_bf16 : p.h_out_bf16;
        int rc = make_tmap_bf16(&tmOut, out, 3, dims, strides, box, true);
        if (rc) return rc;
        rc = make_tmap_bf16(&tmActs, p.acts_save ? p.acts_save : out, 3, dims, strides, box, true);
        if (rc) return rc;
    }
    GTK_CUDA(cudaFuncSetAttribute(wn_layer_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(ceil_div(ceil_div(p.T, BT), WN_CL) * WN_CL, p.B);           // an out-of-range tile only feeds its peer
    ProfScope prof("wn_layer_tc", st);
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid; cfg.blockDim = dim3(WN_NT); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = WN_CL; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = attr; cfg.numAttrs = 2;
    GTK_CUDA(cudaLaunchKernelEx(&cfg, wn_layer_kernel, tmH, tmWin, tmWrs, tmActs, tmOut, p));
    return check_launch("wn_layer_tc");
}

}  // namespace tc
}  // namespace gtk
