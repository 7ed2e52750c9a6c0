// glowtts_b200/csrc/wn_tc.cu -- one WaveNet layer (modules.py:151-170) as ONE tcgen05 kernel per 128-frame tile:
//
//   GEMM 1   xin[t, 0:2H]  = sum_{j,c} h[t + (j-K/2) dil, c] * Win[j][n][c]          (dilated conv, TMA-shifted taps)
//   epilogue acts[t, 0:H]  = tanh(drop(xin[:, :H] + b) + g) * sigmoid(drop(xin[:, H:] + b) + g)   (commons.py:61-68)
//            -> written as bf16 straight into the shared-memory A-operand slots (canonical SWIZZLE_128B layout)
//   GEMM 2   rs[t, 0:RS]   = acts[t, :] * Wrs[n][c]                                   (1x1 res/skip conv)
//   epilogue h' = (h + rs[:, :H] + b) * mask ;  skip (+)= rs[:, H:] + b               (modules.py:166-170)
//
// The 2H-channel pre-activation and the gate output never touch HBM (except the bf16 copies the backward needs,
// written only when training).  TMEM: the same 2H fp32 columns hold xin, then rs.  Thread <-> frame ownership
// (TMEM lane = frame) makes every epilogue step thread-local.
#include <string.h>

#include "tc.cuh"
#include "ops.h"

namespace gtk {
namespace tc {

struct WnCtl {
    uint64_t full[3], empty[3], acc_full, acts_ready;
    uint32_t tmem_base;
    alignas(16) float bias_in[512];     // [2H]
    alignas(16) float cw[512];          // [2H]  WNP conditioner weight / bias per channel (zeros when there is none)
    alignas(16) float cb[512];
    alignas(16) float bias_rs[512];     // [RS]
};

constexpr int WN_STAGES = 3;

// Instrumented builds only (-DGTK_INSTRUMENT): phase timestamps (globaltimer, ns) of every CTA: g_wn_trace[cta][start,
// setup done, GEMM1 done, epilogue1 done, GEMM2 done, end], cta = blockIdx.y * gridDim.x + blockIdx.x (first
// WN_TRACE_CTAS only); read with gtk_debug_wn_trace for the per-phase breakdown in profiles/.
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
// SM cycles per CTA: [MMA thread waiting for operand stages in GEMM 1, MMA thread GEMM-1 total, producer waiting for free
// stages, producer total]
__device__ long long g_wn_wait[WN_TRACE_CTAS * 4];
#define WN_CLK() clock64()
#define WN_WAIT_STORE(i, v) do { const int cta_ = blockIdx.y * gridDim.x + blockIdx.x; if (cta_ < WN_TRACE_CTAS) g_wn_wait[cta_ * 4 + (i)] = (v); } while (0)
#else
#define WN_TRACE(i) do { } while (0)
#define WN_CLK() 0ll
#define WN_WAIT_STORE(i, v) do { } while (0)
#endif

__global__ void __launch_bounds__(EPI_NT, 1)
wn_layer_kernel(const __grid_constant__ CUtensorMap tmH, const __grid_constant__ CUtensorMap tmWin,
                const __grid_constant__ CUtensorMap tmWrs, const __grid_constant__ CUtensorMap tmActs,
                const __grid_constant__ CUtensorMap tmOut, const WnLayerParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int H = p.H, H2 = 2 * p.H, RS = p.RS;
    const int n_kh = H / KC;                            // k-chunks of the hidden dimension (<= 3)
    const int stage_bytes = A_BYTES + H2 * KC * 2;      // A tile + up to 2H weight rows of one k-chunk
    WnCtl* ctl = reinterpret_cast<WnCtl*>(smem + (size_t)WN_STAGES * stage_bytes);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t0 = blockIdx.x * BT, b = blockIdx.y;
    pdl_trigger();
    if (p.n_rows && t0 >= *p.n_rows) return;            // varlen packing: nothing lives in this tile
    WN_TRACE(0);
    const int n1 = p.taps * n_kh;                       // GEMM-1 iterations
    const int n_all = n1 + n_kh;                        // + GEMM-2 iterations

    if (threadIdx.x == 0) {
        for (int s = 0; s < WN_STAGES; ++s) { mbar_init(&ctl->full[s], 1); mbar_init(&ctl->empty[s], 1); }
        mbar_init(&ctl->acc_full, 1);
        mbar_init(&ctl->acts_ready, EPI_NT);
        mbar_fence_init();
    }
    {
        for (int i = threadIdx.x; i < H2; i += EPI_NT) {
            ctl->bias_in[i] = p.bias_in[i];
            float cwv = 0.f, cbv = 0.f;
            if (p.trk) {                                 // WNP conditioner of channel i (modules.py:320-322, 353-362)
                const int c = (p.cs_base + i) % p.cs_div;
                const float v = p.cw_v[c];
                cwv = p.cw_g ? p.cw_g[c] * (v < 0.f ? -1.f : 1.f) : v;      // weight norm of a 1-element row: g v / |v|
                cbv = p.cb[c];
            }
            ctl->cw[i] = cwv;
            ctl->cb[i] = cbv;
        }
        for (int i = threadIdx.x; i < RS; i += EPI_NT) ctl->bias_rs[i] = p.bias_rs[i];
    }
    if (warp == 2) tmem_alloc<512>(&ctl->tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    // PDL: everything the previous kernel of the stream wrote (h, skip) may be touched only after pdl_wait().  The TMA
    // producer thread first requests the WEIGHT tiles of its first stages (packed long before the previous kernel
    // started), so they are in flight while the previous kernel drains, and waits afterwards.
    const bool is_producer = (warp == 0);
    if (!is_producer) pdl_wait();
    const uint32_t tmem = ctl->tmem_base;
    const uint32_t idesc = umma_idesc_bf16(BT, H);
    const int n_lo = min(H2, 256), n_hi = H2 - n_lo;          // GEMM-1 column split of a k16 step (see the MMA warp)
    const uint32_t idesc_lo = umma_idesc_bf16(BT, n_lo), idesc_hi = umma_idesc_bf16(BT, n_hi > 0 ? n_hi : 16);
    WN_TRACE(1);

    // ---------------------------------------------------------------- roles during GEMM 1.  The whole warp walks the
    // (warp-uniform) loops and ONE ELECTED lane issues the TMA / MMA instructions: their descriptors then live in uniform
    // registers.  Under `if (lane == 0)` the compiler cannot prove uniformity and wraps every tcgen05.mma / TMA in an R2UR
    // waterfall loop, ~15 instructions and > 100 cycles each -- more than the 96-cycle tensor-pipe time of one
    // 128 x 192 x 16 MMA, i.e. GEMM 1 was bound by instruction issue (profiles/mma_issue_r02.txt).
    if (warp == 0) {
        const int n_pre = min(WN_STAGES, n1);
        if (elect_one()) {
            tma_prefetch_desc(&tmH);
            tma_prefetch_desc(&tmWin);
            tma_prefetch_desc(&tmWrs);
            for (int it = 0; it < n_pre; ++it) {         // weights of the first stages: independent of the previous kernel
                unsigned char* sb = smem + (size_t)it * stage_bytes + A_BYTES;
                const int j = it / n_kh, kc = it % n_kh;
                mbar_arrive_expect_tx(&ctl->full[it], (uint32_t)(A_BYTES + H2 * KC * 2));
                tma_load_3d(sb, &tmWin, &ctl->full[it], kc * KC, 0, j);
                tma_load_3d(sb + (size_t)H * KC * 2, &tmWin, &ctl->full[it], kc * KC, H, j);
            }
        }
        __syncwarp();
        pdl_wait();
        long long w_prod = 0;
        const long long p0 = WN_CLK();
        for (int it = 0; it < n_all; ++it) {
            const int s = it % WN_STAGES, use = it / WN_STAGES;
            if (it == n1) { WN_WAIT_STORE(2, w_prod); WN_WAIT_STORE(3, WN_CLK() - p0); }
            const long long c0 = WN_CLK();
            if (use > 0) mbar_wait(&ctl->empty[s], (use - 1) & 1);
            w_prod += WN_CLK() - c0;
            unsigned char* sa = smem + (size_t)s * stage_bytes;
            unsigned char* sb = sa + A_BYTES;
            if (elect_one()) {
                if (it < n1) {
                    const int j = it / n_kh, kc = it % n_kh;
                    if (it >= n_pre) {
                        mbar_arrive_expect_tx(&ctl->full[s], (uint32_t)(A_BYTES + H2 * KC * 2));
                        tma_load_3d(sb, &tmWin, &ctl->full[s], kc * KC, 0, j);
                        tma_load_3d(sb + (size_t)H * KC * 2, &tmWin, &ctl->full[s], kc * KC, H, j);
                    }
                    tma_load_3d(sa, &tmH, &ctl->full[s], kc * KC, t0 + (j - p.taps / 2) * p.dil, b);
                } else {
                    const int kc = it - n1;             // the A slot of this stage is written by epilogue 1
                    mbar_arrive_expect_tx(&ctl->full[s], (uint32_t)(RS * KC * 2));
                    tma_load_2d(sb, &tmWrs, &ctl->full[s], kc * KC, 0);
                    if (RS > H) tma_load_2d(sb + (size_t)H * KC * 2, &tmWrs, &ctl->full[s], kc * KC, H);
                }
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
        long long w_mma = 0;
        const long long m0 = WN_CLK();
        for (int it = 0; it < n1; ++it) {
            const int s = it % WN_STAGES, use = it / WN_STAGES;
            const long long c0 = WN_CLK();
            mbar_wait(&ctl->full[s], use & 1);
            w_mma += WN_CLK() - c0;
            tc_fence_after();
            const uint32_t sa = smem_u32(smem + (size_t)s * stage_bytes);
            const uint32_t sb = sa + A_BYTES;
            if (elect_one()) {
                // The 2H gate columns of a k16 step go out as one MMA of N = min(2H, 256) and, beyond 256, one of
                // N = 2H - 256 (H = 192: 256 + 128), NOT as a tanh and a sigmoid MMA of N = H each.  Measured
                // (profiles/mma_issue_r02.txt): an accumulator accepts its next MMA ~193 cycles after the previous one, so
                // two N = 192 chains run at 127 cycles per MMA (254 per k16 step), while the 256 | 128 pair keeps the tensor
                // pipe busy for 128 + 64 = 192 cycles per step -- exactly the accumulators' turn-around.
                const uint64_t adesc = umma_desc_kmajor_sw128(sa), bdesc = umma_desc_kmajor_sw128(sb);
#pragma unroll
                for (int k = 0; k < KC / 16; ++k) {
                    umma_bf16(tm, adesc + 2 * k, bdesc + 2 * k, idesc_lo, (it | k) != 0);
                    if (n_hi) umma_bf16(tm + 256, adesc + 2 * k, bdesc + 2 * k + (256 * KC * 2 >> 4), idesc_hi, (it | k) != 0);
                }
                umma_commit(&ctl->empty[s]);
            }
            __syncwarp();
        }
        if (elect_one()) umma_commit(&ctl->acc_full);    // GEMM 1 done -> epilogue 1 may read TMEM
        __syncwarp();
        WN_WAIT_STORE(0, w_mma);
        WN_WAIT_STORE(1, WN_CLK() - m0);
    }
    __syncwarp();

    // ---------------------------------------------------------------- epilogues: ALL 16 warps.  Warp w owns TMEM lane
    // quarter w % 4 (frames) and column part w / 4 of EPI_PARTS.
    const int q = warp & 3, part = warp >> 2;
    const int r = q * 32 + lane;                         // row of the tile == TMEM lane
    const int t = t0 + r;
    const bool t_ok = t < p.T;
    const float mk = (p.mask && t_ok) ? p.mask[(size_t)b * p.T + t] : 1.f;
    const bool live = t_ok && mk != 0.f;                 // masked rows are written as exact zeros (never value * 0)
    const size_t frame = (size_t)b * p.T + (t_ok ? t : 0);
    // dropout counter: the un-packed frame index, so that packed and [B, T] runs draw the same mask
    const unsigned long long dframe = p.row_fr ? (unsigned long long)p.row_fr[frame] : (unsigned long long)frame;
    // conditioning of this row: speaker embedding projection (per utterance) and / or the WNP pitch-energy track
    const float* cond_row = p.cond ? p.cond + (size_t)(p.row_b ? p.row_b[frame] : b) * p.cond_bstride : nullptr;
    const float* trk_row = p.trk ? p.trk + frame * p.n_sqz : nullptr;
    const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
    RowIo io;
    io.tp = p.tile_layout; io.tile_id = b * ((p.T + BT - 1) / BT) + blockIdx.x; io.r = r; io.frame = frame;
    const int n_chunks = H / 16;
    const int ch_lo = n_chunks * part / EPI_PARTS, ch_hi = n_chunks * (part + 1) / EPI_PARTS;

    // ---- epilogue 1: gate -> A operand of GEMM 2
    const unsigned long long dseed = p.drop_p > 0.f ? drop_resolve(p.drop_seed) : 0ull;
    mbar_wait(&ctl->acc_full, 0);
    tc_fence_after();
    WN_TRACE(2);
    for (int cc = ch_lo; cc < ch_hi; ++cc) {
        const int c0 = cc * 16;
        float a[16], g[16];
        tmem_ld16(trow + c0, a);
        tmem_ld16(trow + H + c0, g);
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
        if (cond_row) {                                  // cond_layer(g) slice of this utterance (modules.py:148-157)
#pragma unroll
            for (int i = 0; i < 16; i += 4) {
                const float4 ca = __ldg(reinterpret_cast<const float4*>(cond_row + c0 + i));
                const float4 cg = __ldg(reinterpret_cast<const float4*>(cond_row + H + c0 + i));
                a[i] += ca.x; a[i + 1] += ca.y; a[i + 2] += ca.z; a[i + 3] += ca.w;
                g[i] += cg.x; g[i + 1] += cg.y; g[i + 2] += cg.z; g[i + 3] += cg.w;
            }
        }
        if (trk_row) {                                   // WNP: cond_layer1(track) squeezed, rank 1 in time (modules.py:320-336)
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
        // gate output -> A slot of GEMM 2 (canonical K-major SWIZZLE_128B); the same tile is TMA-stored to acts_save
        stage_bf16x16_sw128(smem + (size_t)((n1 + c0 / KC) % WN_STAGES) * stage_bytes, 0, r, c0 % KC, a);
    }
    fence_proxy_async_smem();
    tc_fence_before();
    mbar_arrive(&ctl->acts_ready);
    WN_TRACE(3);

    // ---- GEMM 2 (res_skip 1x1): issued by warp 1 once every thread has delivered its gate outputs
    if (warp == 1) {
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
        mbar_wait(&ctl->acts_ready, 0);                  // acts are in the A slots, the xin columns are free again
        tc_fence_after();
        for (int it = n1; it < n_all; ++it) {
            const int s = it % WN_STAGES, use = it / WN_STAGES;
            mbar_wait(&ctl->full[s], use & 1);
            tc_fence_after();
            const uint32_t sa = smem_u32(smem + (size_t)s * stage_bytes);
            const uint32_t sb = sa + A_BYTES;
            if (elect_one()) {
                const uint64_t adesc = umma_desc_kmajor_sw128(sa), bdesc = umma_desc_kmajor_sw128(sb);
                const int r_lo = min(RS, 256), r_hi = RS - r_lo;       // same column split as GEMM 1
                const uint32_t id_lo = umma_idesc_bf16(BT, r_lo), id_hi = umma_idesc_bf16(BT, r_hi > 0 ? r_hi : 16);
#pragma unroll
                for (int k = 0; k < KC / 16; ++k) {
                    umma_bf16(tm, adesc + 2 * k, bdesc + 2 * k, id_lo, (it != n1) || (k != 0));
                    if (r_hi) umma_bf16(tm + 256, adesc + 2 * k, bdesc + 2 * k + (256 * KC * 2 >> 4), id_hi, (it != n1) || (k != 0));
                }
            }
            __syncwarp();
        }
        if (elect_one()) umma_commit(&ctl->acc_full);
        __syncwarp();
        if (lane == 0 && p.acts_save) {                  // bf16 gate output for the res_skip weight gradient
            for (int kc = 0; kc < n_kh; ++kc)
                tma_store_3d(&tmActs, smem + (size_t)((n1 + kc) % WN_STAGES) * stage_bytes, kc * KC, t0, b);
            tma_store_commit();
        }
    }
    __syncwarp();

    // ---- epilogue 2: residual / skip.  The bf16 operand copy of the result (next layer's input, or the masked WN
    // output for `end`) is staged in the now idle weight slots and leaves as TMA stores.
    unsigned char* stage_out = smem + A_BYTES;
    mbar_wait(&ctl->acc_full, 1);
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
            stage_bf16x16_sw128(stage_out, stage_bytes, r, c0, v);
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
        if (p.last) {                                     // WN output = skip * mask (modules.py:171), operand of `end`
#pragma unroll
            for (int i = 0; i < 16; ++i) sk[i] = live ? sk[i] : 0.f;
            stage_bf16x16_sw128(stage_out, stage_bytes, r, c0, sk);
            if (p.h_out_f32_last) row_store_f32x16(p.h_out_f32_last, H, c0, io, sk);   // fp32 master for a chained WN
        } else {
            row_store_f32x16(p.skip_f32, H, c0, io, sk);
        }
    }
    fence_proxy_async_smem();
    tc_fence_before();
    WN_TRACE(5);
    __syncthreads();
    if (warp == 1 && lane == 0) {                         // the thread that also issued the acts stores
        for (int kc = 0; kc < n_kh; ++kc) tma_store_3d(&tmOut, stage_out + (size_t)kc * stage_bytes, kc * KC, t0, b);
        tma_store_commit();
        tma_store_wait_read();                            // shared memory must outlive the bulk reads
    }
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc<512>(tmem);
    }
}

#ifdef GTK_INSTRUMENT
int wn_wait_read(long long* out, int n_ctas) {
    GTK_REQUIRE(out && n_ctas >= 0 && n_ctas <= WN_TRACE_CTAS, "n_ctas");
    GTK_CUDA(cudaMemcpyFromSymbol(out, g_wn_wait, size_t(n_ctas) * 4 * sizeof(long long)));
    return GTK_OK;
}
int wn_trace_read(unsigned long long* out, int n_ctas) {
    GTK_REQUIRE(out && n_ctas >= 0 && n_ctas <= WN_TRACE_CTAS, "n_ctas");
    GTK_CUDA(cudaMemcpyFromSymbol(out, g_wn_trace, size_t(n_ctas) * 6 * sizeof(unsigned long long)));
    return GTK_OK;
}
#endif

int wn_layer_launch(const __nv_bfloat16* h_bf16, const __nv_bfloat16* win_p, const __nv_bfloat16* wrs_p, WnLayerParams p,
                    cudaStream_t st) {
    GTK_REQUIRE(h_bf16 && win_p && wrs_p && p.bias_in && p.bias_rs, "null");
    GTK_REQUIRE(p.H % KC == 0 && p.H / KC <= WN_STAGES && p.H <= 256, "fused WN layer needs hidden in {64,128,192}");
    GTK_REQUIRE(p.RS == p.H || p.RS == 2 * p.H, "res_skip channels");
    GTK_REQUIRE(p.taps >= 1 && (p.taps & 1) && p.dil >= 1, "bad conv shape");
    GTK_REQUIRE(p.last ? (p.RS == p.H && p.skip_out_bf16) : (p.RS == 2 * p.H && p.h_f32 && p.h_out_f32 && p.h_out_bf16), "layer outputs");
    GTK_REQUIRE(p.first || p.skip_f32, "skip accumulator");
    GTK_REQUIRE(!p.trk || (p.cw_v && p.cb && p.n_sqz >= 1 && p.cs_div >= 16 && p.cs_div % 16 == 0),
                "time-varying conditioner arguments (2 H L / n_sqz must be a multiple of 16)");
    GTK_REQUIRE(p.last || p.skip_f32, "skip accumulator");
    if (p.B == 0 || p.T == 0) return GTK_OK;
    const int H = p.H;
    CUtensorMap tmH, tmWin, tmWrs;
    {
        const uint64_t dims[3] = {(uint64_t)H, (uint64_t)p.T, (uint64_t)p.B};
        const uint64_t strides[2] = {(uint64_t)H * 2, (uint64_t)H * 2 * p.T};
        const uint32_t box[3] = {KC, BT, 1};
        int rc = make_tmap_bf16(&tmH, h_bf16, 3, dims, strides, box, true);
        if (rc) return rc;
    }
    {
        const uint64_t dims[3] = {(uint64_t)H, (uint64_t)2 * H, (uint64_t)p.taps};
        const uint64_t strides[2] = {(uint64_t)H * 2, (uint64_t)H * 2 * 2 * H};
        const uint32_t box[3] = {KC, (uint32_t)H, 1};
        int rc = make_tmap_bf16(&tmWin, win_p, 3, dims, strides, box, true);
        if (rc) return rc;
    }
    {
        const uint64_t dims[2] = {(uint64_t)H, (uint64_t)p.RS};
        const uint64_t strides[1] = {(uint64_t)H * 2};
        const uint32_t box[2] = {KC, (uint32_t)H};
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
    const size_t smem = (size_t)WN_STAGES * (A_BYTES + (size_t)2 * H * KC * 2) + sizeof(WnCtl) + 1024;
    GTK_REQUIRE(smem <= 227 * 1024, "tile does not fit in shared memory");
    GTK_CUDA(cudaFuncSetAttribute(wn_layer_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(ceil_div(p.T, BT), p.B);
    ProfScope prof("wn_layer_tc", st);
    GTK_CUDA(launch_pdl(wn_layer_kernel, grid, dim3(EPI_NT), smem, st, tmH, tmWin, tmWrs, tmActs, tmOut, p));
    return check_launch("wn_layer_tc");
}

}  // namespace tc
}  // namespace gtk
