// glowtts_b200/csrc/wn_bwd_tc.cu -- backward of one WaveNet layer (modules.py:151-170) as ONE tcgen05 kernel per tile:
//
//   GEMM a   d(acts)[t, 0:H] = drs[t, 0:RS] * Wrs                       (data gradient of the 1x1 res/skip conv)
//   epilogue d(xin)[t, 0:2H] = gate backward of d(acts) with the saved tanh | sigmoid, dropout mask re-applied
//            -> written as bf16 straight into a shared-memory A operand [128 frames x 2H] (SWIZZLE_128B), and from there
//               TMA-stored to HBM for the weight-gradient launch
//   GEMM b   d(h)[t, 0:H]    = sum_j d(xin)[t + j dil - halo, :] * Win_j^T   (data gradient of the dilated conv)
//            -- every tap reads the SAME shared-memory tile through a row-shifted UMMA descriptor (the swizzle is a
//               function of the address only, profiles/shift_probe.py), so d(xin) never comes back from L2
//   epilogue d(h) (+ fp32 carry of the residual path) * mask -> fp32 carry + bf16 operand of the next layer down
//
// Tiles overlap: a CTA computes d(xin) for 128 frames [t0 - halo, t0 + 128 - halo) but owns only the 128 - 2 halo output
// frames [t0, t0 + 128 - 2 halo), halo = (K/2) dil (base config: 124 of 128 frames, no inter-CTA dependency).
#include <stdlib.h>
#include <string.h>

#include "tc.cuh"
#include "ops.h"

namespace gtk {
namespace tc {

constexpr int WB_RING = 120 * 1024;          // operand ring behind the d(xin) tile: 3 x (16 + 24) KB or 5 x 24 KB
constexpr int WB_MAX_STAGES = 8;

struct WbCtl {
    uint64_t full[WB_MAX_STAGES], empty[WB_MAX_STAGES];        // GEMM a ring
    uint64_t full_b[WB_MAX_STAGES], empty_b[WB_MAX_STAGES];    // GEMM b ring (same memory, re-cut; cluster-wide `empty`)
    uint64_t acc_full, acts_ready;
    uint64_t full_c;                                           // weights of the fused `start` data gradient (GEMM c)
    uint32_t tmem_base;
};
constexpr int WB_CL = 2;                     // CTAs per cluster: every in_layer weight tile is fetched once per pair

// Phase timestamps (globaltimer, ns) per CTA of the last launch: [start, setup, GEMM a done, epilogue a done, GEMM b done,
// epilogue b done, end]; read with gtk_debug_wb_trace.
#ifdef GTK_INSTRUMENT
constexpr int WB_TRACE_CTAS = 512;
__device__ unsigned long long g_wb_trace[WB_TRACE_CTAS * 7];
__device__ __forceinline__ unsigned long long wb_timer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#define WB_TRACE(i)                                                                         \
    do {                                                                                    \
        const int cta_ = blockIdx.y * gridDim.x + blockIdx.x;                               \
        if (threadIdx.x == 128 && cta_ < WB_TRACE_CTAS) g_wb_trace[cta_ * 7 + (i)] = wb_timer(); \
    } while (0)
// SM cycles per CTA: [MMA thread waiting for weight stages in GEMM b, MMA thread GEMM-b total, producer waiting for free
// stages in GEMM b, producer GEMM-b total]
__device__ long long g_wb_wait[WB_TRACE_CTAS * 4];
#define WB_CLK() clock64()
#define WB_WAIT_STORE(i, v) do { const int cta_ = blockIdx.y * gridDim.x + blockIdx.x; if (cta_ < WB_TRACE_CTAS) g_wb_wait[cta_ * 4 + (i)] = (v); } while (0)
#else
#define WB_TRACE(i) do { } while (0)
#define WB_CLK() 0ll
#define WB_WAIT_STORE(i, v) do { } while (0)
#endif


__global__ void __launch_bounds__(EPI_NT, 1)
wn_bwd_layer_kernel(const __grid_constant__ CUtensorMap tmDrs, const __grid_constant__ CUtensorMap tmWrs,
                    const __grid_constant__ CUtensorMap tmWin, const __grid_constant__ CUtensorMap tmDxin,
                    const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmWpost,
                    const WnBwdParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int H = p.H, H2 = 2 * p.H;
    const int nka = (p.RS + KC - 1) / KC, nkb = H2 / KC;       // k-chunks of GEMM a / of one tap of GEMM b
    const int x_bytes = nkb * A_BYTES;                         // d(xin) tile: nkb chunks of [128 x 64] bf16
    unsigned char* ring = smem + x_bytes;
    const int sa_bytes = A_BYTES + H * KC * 2, sb_bytes = H * KC * 2;
    const int n_sa = min(WB_MAX_STAGES, WB_RING / sa_bytes), n_sb = min(WB_MAX_STAGES, WB_RING / sb_bytes);
    WbCtl* ctl = reinterpret_cast<WbCtl*>(ring + WB_RING);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int halo = (p.taps / 2) * p.dil, tile_out = BT - 2 * halo;
    const int t0 = blockIdx.x * tile_out, b = blockIdx.y;
    const int n_b = p.taps * nkb;
    pdl_trigger();
    // varlen packing: a pair of tiles with nothing alive leaves at once (pair-uniform: the peer's barriers are involved).
    // Every kernel uses the same limit, n_rows rounded up to whole 128-row tiles: all rows below it are always written.
    if (p.n_rows && (int)(blockIdx.x & ~(unsigned)(WB_CL - 1)) * tile_out >= ((*p.n_rows + BT - 1) & ~(BT - 1))) return;
    WB_TRACE(0);

    const uint32_t cta_rank = cluster_ctarank();
    const uint16_t cl_mask = (uint16_t)((1u << WB_CL) - 1);
    if (threadIdx.x == 0) {
        for (int s = 0; s < WB_MAX_STAGES; ++s) {
            mbar_init(&ctl->full[s], 1); mbar_init(&ctl->empty[s], 1);
            mbar_init(&ctl->full_b[s], 1); mbar_init(&ctl->empty_b[s], WB_CL);
        }
        mbar_init(&ctl->acc_full, 1);
        mbar_init(&ctl->acts_ready, EPI_NT);
        mbar_init(&ctl->full_c, 1);
        mbar_fence_init();
    }
    if (warp == 2) tmem_alloc<512>(&ctl->tmem_base);
    tc_fence_before();
    cluster_sync_all();                                        // the peer's barriers exist before anything is multicast
    tc_fence_after();
    // PDL: the producer thread requests the res_skip weight tiles of its first stages before pdl_wait() (see wn_tc.cu)
    const bool is_producer = (warp == 0);
    if (!is_producer) pdl_wait();
    const uint32_t tmem = ctl->tmem_base;
    const uint32_t idesc = umma_idesc_bf16(BT, H);
    WB_TRACE(1);

    // ---------------------------------------------------------------- TMA producer.  Whole warp in the uniform loops, ONE
    // ELECTED lane issues (TMA and MMA descriptors stay in uniform registers; see wn_tc.cu / profiles/mma_issue_r02.txt)
    if (warp == 0) {
        const int n_pre = min(n_sa, nka);
        if (elect_one()) {
            tma_prefetch_desc(&tmDrs);
            tma_prefetch_desc(&tmWrs);
            tma_prefetch_desc(&tmWin);
            for (int kc = 0; kc < n_pre; ++kc) {
                mbar_arrive_expect_tx(&ctl->full[kc], (uint32_t)sa_bytes);
                tma_load_2d(ring + (size_t)kc * sa_bytes + A_BYTES, &tmWrs, &ctl->full[kc], kc * KC, 0);
            }
        }
        __syncwarp();
        pdl_wait();
        for (int kc = 0; kc < nka; ++kc) {
            const int s = kc % n_sa, use = kc / n_sa;
            if (use > 0) mbar_wait(&ctl->empty[s], (use - 1) & 1);
            unsigned char* sa = ring + (size_t)s * sa_bytes;
            if (elect_one()) {
                if (kc >= n_pre) {
                    mbar_arrive_expect_tx(&ctl->full[s], (uint32_t)sa_bytes);
                    tma_load_2d(sa + A_BYTES, &tmWrs, &ctl->full[s], kc * KC, 0);
                }
                tma_load_3d(sa, &tmDrs, &ctl->full[s], kc * KC, t0 - halo, b);
            }
            __syncwarp();
        }
        // The ring is re-cut into weight-only stages for GEMM b.  Each CTA of the pair fetches half of the rows of every
        // weight tile and the TMA multicasts them into both shared memories; a stage may be written once BOTH CTAs have
        // released it (empty_b counts the pair: first release = each CTA's GEMM a has read all its operands).  The first
        // n_sb tiles go out now and land during the gate-backward epilogue; the rest follows below, after this thread
        // has delivered its share of that epilogue (GEMM b, which frees the stages, cannot start before).
        for (int it = 0; it < min(n_sb, n_b); ++it) {
            mbar_wait(&ctl->empty_b[it], 0);
            const int j = it / nkb, kc = it % nkb;
            if (elect_one()) {
                mbar_arrive_expect_tx(&ctl->full_b[it], (uint32_t)sb_bytes);
                tma_load_3d_multicast(ring + (size_t)it * sb_bytes + (size_t)cta_rank * (H / WB_CL) * 128, &tmWin,
                                      &ctl->full_b[it], kc * KC, (int)cta_rank * (H / WB_CL), j, cl_mask);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ---------------------------------------------------------------- GEMM a
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
        for (int kc = 0; kc < nka; ++kc) {
            const int s = kc % n_sa, use = kc / n_sa;
            mbar_wait(&ctl->full[s], use & 1);
            tc_fence_after();
            const uint32_t sa = smem_u32(ring + (size_t)s * sa_bytes);
            if (elect_one()) {
                const uint64_t adesc = umma_desc_kmajor_sw128(sa), bdesc = umma_desc_kmajor_sw128(sa + A_BYTES);
#pragma unroll
                for (int k = 0; k < KC / 16; ++k)              // two accumulators take alternate k16 steps (conv_tc.cu)
                    umma_bf16(tm + (k & 1) * 256, adesc + 2 * k, bdesc + 2 * k, idesc, kc != 0 || k >= 2);
                umma_commit(&ctl->empty[s]);
            }
            __syncwarp();
        }
        if (elect_one()) {
            umma_commit(&ctl->acc_full);
            for (int s = 0; s < n_sb; ++s) umma_commit_multicast(&ctl->empty_b[s], cl_mask);   // ring free on this CTA
        }
    }
    __syncwarp();

    // ---------------------------------------------------------------- epilogue a: gate backward -> d(xin) tile in smem
    const int q = warp & 3, part = warp >> 2;
    const int r = q * 32 + lane;
    const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
    const int nck = H / 16;
    {
        const int f = t0 - halo + r;
        const bool in_t = f >= 0 && f < p.T;
        // masked frames carry no gradient (d(skip) and d(h) are masked there): exact zeros, whatever the saved
        // activations of such a row hold
        const bool ok = in_t && (!p.mask || p.mask[(size_t)b * p.T + f] != 0.f);
        const RowIo io = frame_io(b, ok ? f : 0, p.T);
        const unsigned long long dframe = p.row_fr ? (unsigned long long)p.row_fr[io.frame] : (unsigned long long)io.frame;
        const bool own = in_t && r >= halo && r < halo + tile_out;      // tiles overlap by the halo: owned frames only
        __nv_bfloat16* dpre_row = (p.dpre && own) ? p.dpre + ((size_t)b * p.T + f) * H2 : nullptr;
        const unsigned long long dseed = p.drop_p > 0.f ? drop_resolve(p.drop_seed) : 0ull;
        // The saved tanh | sigmoid of this thread's columns (written by the forward pass long ago: they come from HBM) are
        // requested BEFORE the wait for GEMM a, so their latency hides under it (H <= 192: at most 3 chunks per thread).
        constexpr int WB_MAXC = 3;
        const int cc_lo = nck * part / EPI_PARTS, cc_hi = nck * (part + 1) / EPI_PARTS;
        uint4 pre[WB_MAXC][4];
#pragma unroll
        for (int i = 0; i < WB_MAXC; ++i) {
            if (ok && cc_lo + i < cc_hi) {
                const int c0 = (cc_lo + i) * 16;
                const uint4* st = reinterpret_cast<const uint4*>(p.ts) + ((size_t)io.tile_id * (H2 >> 3) + (c0 >> 3)) * 128 + io.r;
                const uint4* ss4 = reinterpret_cast<const uint4*>(p.ts) + ((size_t)io.tile_id * (H2 >> 3) + ((H + c0) >> 3)) * 128 + io.r;
                pre[i][0] = __ldg(st); pre[i][1] = __ldg(st + 128); pre[i][2] = __ldg(ss4); pre[i][3] = __ldg(ss4 + 128);
            }
        }
        mbar_wait(&ctl->acc_full, 0);
        tc_fence_after();
        WB_TRACE(2);
#pragma unroll
        for (int ci = 0; ci < WB_MAXC; ++ci) {
            const int cc = cc_lo + ci;
            if (cc >= cc_hi) break;
            const int c0 = cc * 16;
            float d[16], dt[16], ds[16];
            tmem_ld16_dual(trow + c0, true, d);
            if (ok) {
                float tt[16], ss[16];
                unpack_bf16x16(pre[ci][0], pre[ci][1], tt);
                unpack_bf16x16(pre[ci][2], pre[ci][3], ss);
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    dt[i] = d[i] * ss[i] * (1.f - tt[i] * tt[i]);
                    ds[i] = d[i] * tt[i] * ss[i] * (1.f - ss[i]);
                }
                if (dpre_row) { store_bf16x16(dpre_row + c0, dt); store_bf16x16(dpre_row + H + c0, ds); }
                if (p.drop_p > 0.f) {
                    float sc[16];
                    dropout_scale16(dseed, dframe * H2 + c0, p.drop_p, sc);
#pragma unroll
                    for (int i = 0; i < 16; ++i) dt[i] *= sc[i];
                    dropout_scale16(dseed, dframe * H2 + H + c0, p.drop_p, sc);
#pragma unroll
                    for (int i = 0; i < 16; ++i) ds[i] *= sc[i];
                }
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) { dt[i] = 0.f; ds[i] = 0.f; }   // outside the utterance: the taps read zeros
                if (dpre_row) { store_bf16x16(dpre_row + c0, dt); store_bf16x16(dpre_row + H + c0, ds); }
            }
            stage_bf16x16_sw128(smem, A_BYTES, r, c0, dt);
            stage_bf16x16_sw128(smem, A_BYTES, r, H + c0, ds);
        }
        fence_proxy_async_smem();
        tc_fence_before();
        mbar_arrive(&ctl->acts_ready);
        WB_TRACE(3);
    }

    // ---------------------------------------------------------------- GEMM b: taps = row-shifted views of the smem tile
    if (warp == 0) {
        long long w_prod = 0;
        const long long p0 = WB_CLK();
        for (int it = n_sb; it < n_b; ++it) {
            const int s = it % n_sb, use = it / n_sb;
            const long long c0 = WB_CLK();
            mbar_wait(&ctl->empty_b[s], use & 1);
            w_prod += WB_CLK() - c0;
            const int j = it / nkb, kc = it % nkb;
            if (elect_one()) {
                mbar_arrive_expect_tx(&ctl->full_b[s], (uint32_t)sb_bytes);
                tma_load_3d_multicast(ring + (size_t)s * sb_bytes + (size_t)cta_rank * (H / WB_CL) * 128, &tmWin,
                                      &ctl->full_b[s], kc * KC, (int)cta_rank * (H / WB_CL), j, cl_mask);
            }
            __syncwarp();
        }
        WB_WAIT_STORE(2, w_prod);
        WB_WAIT_STORE(3, WB_CLK() - p0);
    } else if (warp == 1) {
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
        mbar_wait(&ctl->acts_ready, 0);
        tc_fence_after();
        // d(xin) of the OWNED frames [t0, t0 + tile_out) for the weight-gradient launch: rows halo.. of the tile (a TMA
        // store must not start at a negative coordinate; the swizzle is address-based, so a row-offset source is fine)
        if (lane == 0 && t0 < p.T) {
            for (int kc = 0; kc < nkb; ++kc) tma_store_3d(&tmDxin, smem + (size_t)kc * A_BYTES + halo * 128, kc * KC, t0, b);
            tma_store_commit();
        }
        __syncwarp();
        long long w_mma = 0;
        const long long m0 = WB_CLK();
        for (int it = 0; it < n_b; ++it) {
            const int s = it % n_sb, use = it / n_sb;
            const long long c0 = WB_CLK();
            mbar_wait(&ctl->full_b[s], use & 1);
            w_mma += WB_CLK() - c0;
            tc_fence_after();
            const int j = it / nkb, kc = it % nkb;
            const uint32_t a_addr = smem_u32(smem + (size_t)kc * A_BYTES) + (uint32_t)(j * p.dil) * 128u;
            const uint32_t b_addr = smem_u32(ring + (size_t)s * sb_bytes);
            if (elect_one()) {
                const uint64_t adesc = umma_desc_kmajor_sw128(a_addr), bdesc = umma_desc_kmajor_sw128(b_addr);
#pragma unroll
                for (int k = 0; k < KC / 16; ++k)
                    umma_bf16(tm + (k & 1) * 256, adesc + 2 * k, bdesc + 2 * k, idesc, it != 0 || k >= 2);
                umma_commit_multicast(&ctl->empty_b[s], cl_mask);
            }
            __syncwarp();
        }
        if (p.post_w && lane == 0) tma_store_wait_read();       // the d(xin) store has read the tile GEMM c's weights will overwrite
        __syncwarp();
        if (elect_one()) umma_commit(&ctl->acc_full);
        __syncwarp();
        WB_WAIT_STORE(0, w_mma);
        WB_WAIT_STORE(1, WB_CLK() - m0);
    }
    __syncwarp();

    // ---------------------------------------------------------------- epilogue b: residual carry, mask, outputs
    {
        const int f = t0 + r;
        const bool ok = r < tile_out && f < p.T;
        const RowIo io = frame_io(b, ok ? f : 0, p.T);
        const float mk = (p.mask && ok) ? p.mask[(size_t)b * p.T + f] : 1.f;
        const bool live = mk != 0.f;
        // the fp32 carry of the layer above is requested before the wait for GEMM b (it was complete at kernel start)
        constexpr int WB_MAXC = 3;
        const int cc_lo = nck * part / EPI_PARTS, cc_hi = nck * (part + 1) / EPI_PARTS;
        float xpre[WB_MAXC][16];
        if (p.resid_f32 && ok) {
#pragma unroll
            for (int i = 0; i < WB_MAXC; ++i)
                if (cc_lo + i < cc_hi) row_load_f32x16(p.resid_f32, H, (cc_lo + i) * 16, io, xpre[i]);
        }
        mbar_wait(&ctl->acc_full, 1);
        tc_fence_after();
        WB_TRACE(4);
        if (p.post_w && warp == 0) {
            // fused `start` data gradient (GEMM c): its weight chunks go where the d(xin) tile was, now that GEMM b has read
            // it (and its TMA store too: the MMA warp waits for that store before it commits GEMM b); they land under this
            // epilogue
            if (elect_one()) {
                const int chp = (p.post_N * 128 + 1023) & ~1023;
                mbar_arrive_expect_tx(&ctl->full_c, (uint32_t)((H / KC) * p.post_N * 128));
                for (int kc = 0; kc < H / KC; ++kc) tma_load_2d(smem + (size_t)kc * chp, &tmWpost, &ctl->full_c, kc * KC, 0);
            }
            __syncwarp();
        }
#pragma unroll
        for (int ci = 0; ci < WB_MAXC; ++ci) {
            const int cc = cc_lo + ci;
            if (cc >= cc_hi) break;
            const int c0 = cc * 16;
            float v[16];
            tmem_ld16_dual(trow + c0, true, v);
            if (!ok) continue;
            if (p.resid_f32) {
                float (&x)[16] = xpre[ci];
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = live ? v[i] + x[i] : 0.f;     // resid and output both masked (modules.py:166)
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = live ? v[i] : 0.f;
            }
            if (p.out_f32) row_store_f32x16(p.out_f32, H, c0, io, v);
            stage_bf16x16_sw128(ring, A_BYTES, r, c0, v);
        }
        fence_proxy_async_smem();
        tc_fence_before();
        WB_TRACE(5);
    }
    __syncthreads();
    if (warp == 1 && lane == 0) {
        if (t0 < p.T) {
            for (int kc = 0; kc < H / KC; ++kc) tma_store_3d(&tmOut, ring + (size_t)kc * A_BYTES, kc * KC, t0, b);
            tma_store_commit();
        }
    }
    if (p.post_w) {
        // ---------------------------------------------------------------- GEMM c + epilogue c: data gradient of the `start`
        // 1x1 conv on the d(h_0) tile that sits in shared memory as an A operand (conv_tc.cu EPI_PLAIN with resid_tm / out_tm,
        // same sums: two accumulators take alternate k16 steps): dM[c, t] = dZ[c, t] + sum_n d(h_0)[t, n] Wstart[n, c]
        if (warp == 1) {
            const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
            mbar_wait(&ctl->full_c, 0);
            tc_fence_after();
            const int chp = (p.post_N * 128 + 1023) & ~1023;
            const uint32_t idc = umma_idesc_bf16(BT, p.post_N);
            if (elect_one()) {
                for (int kc = 0; kc < H / KC; ++kc) {
                    const uint64_t adesc = umma_desc_kmajor_sw128(smem_u32(ring + (size_t)kc * A_BYTES));
                    const uint64_t bdesc = umma_desc_kmajor_sw128(smem_u32(smem + (size_t)kc * chp));
#pragma unroll
                    for (int k = 0; k < KC / 16; ++k)
                        umma_bf16(tm + (k & 1) * 256, adesc + 2 * k, bdesc + 2 * k, idc, kc != 0 || k >= 2);
                }
                umma_commit(&ctl->acc_full);
            }
            __syncwarp();
        }
        const int f = t0 + r;
        const bool ok = r < tile_out && f < p.T;
        const int nck_c = p.post_N / 16;
        // the residual (the other half's gradient, complete long before this kernel) is requested before the wait
        const int cq_lo = nck_c * part / EPI_PARTS, cq_hi = nck_c * (part + 1) / EPI_PARTS;
        constexpr int WB_PREC = 2;
        float xc[WB_PREC][16];
        if (p.post_resid_tm && ok) {
#pragma unroll
            for (int ci = 0; ci < WB_PREC; ++ci)
                if (cq_lo + ci < cq_hi) {
                    const size_t o0 = ((size_t)b * p.post_C + (cq_lo + ci) * 16) * p.T + f;
#pragma unroll
                    for (int i = 0; i < 16; ++i) xc[ci][i] = __ldg(p.post_resid_tm + o0 + (size_t)i * p.T);
                }
        }
        mbar_wait(&ctl->acc_full, 0);                      // third completion of acc_full
        tc_fence_after();
        for (int cc = cq_lo; cc < cq_hi; ++cc) {
            const int c0 = cc * 16;
            float v[16];
            tmem_ld16_dual(trow + c0, true, v);
            if (!ok) continue;
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] += 0.f;      // conv_tc adds its (absent) bias here
            const size_t o0 = ((size_t)b * p.post_C + c0) * p.T + f;
            if (p.post_resid_tm) {
                if (cc - cq_lo < WB_PREC) {
#pragma unroll
                    for (int ci = 0; ci < WB_PREC; ++ci)
                        if (ci == cc - cq_lo) {
#pragma unroll
                            for (int i = 0; i < 16; ++i) v[i] += xc[ci][i];
                        }
                } else {
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] += p.post_resid_tm[o0 + (size_t)i * p.T];
                }
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) p.post_out_tm[o0 + (size_t)i * p.T] = v[i];
        }
        tc_fence_before();
    }
    if (warp == 1 && lane == 0) tma_store_wait_read();
    WB_TRACE(6);
    cluster_sync_all();                                        // the peer's commits may still be arriving on our barriers
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc<512>(tmem);
    }
}

#ifdef GTK_INSTRUMENT
int wb_wait_read(long long* out, int n_ctas) {
    GTK_REQUIRE(out && n_ctas >= 0 && n_ctas <= WB_TRACE_CTAS, "n_ctas");
    GTK_CUDA(cudaMemcpyFromSymbol(out, g_wb_wait, size_t(n_ctas) * 4 * sizeof(long long)));
    return GTK_OK;
}
int wb_trace_read(unsigned long long* out, int n_ctas) {
    GTK_REQUIRE(out && n_ctas >= 0 && n_ctas <= WB_TRACE_CTAS, "n_ctas");
    GTK_CUDA(cudaMemcpyFromSymbol(out, g_wb_trace, size_t(n_ctas) * 7 * sizeof(unsigned long long)));
    return GTK_OK;
}
#endif

bool wn_bwd_post_ok(int H, int N) { return N % 16 == 0 && N >= 16 && N <= 256 && (size_t)(H / KC) * ((N * 128 + 1023) & ~1023) <= (size_t)(2 * H / KC) * A_BYTES; }

int wn_bwd_layer_launch(const __nv_bfloat16* drs, int ld_drs, const __nv_bfloat16* wrs_t, const __nv_bfloat16* win_t,
                        __nv_bfloat16* dxin, __nv_bfloat16* out_bf16, int ld_out, WnBwdParams p, cudaStream_t st) {
    GTK_REQUIRE(drs && wrs_t && win_t && dxin && out_bf16 && p.ts && p.out_f32, "null");
    GTK_REQUIRE(p.H % KC == 0 && p.H <= 256 && (p.RS == p.H || p.RS == 2 * p.H), "hidden / res_skip channels");
    GTK_REQUIRE(p.taps >= 1 && (p.taps & 1) && p.dil >= 1 && wn_bwd_supported(p.H, p.taps, p.dil), "conv shape not supported by the fused backward");
    GTK_REQUIRE(ld_drs % 8 == 0 && ld_out % 8 == 0 && ld_drs >= p.RS && ld_out >= p.H, "frame pitch");
    if (p.B == 0 || p.T == 0) return GTK_OK;
    const int H = p.H, halo = (p.taps / 2) * p.dil, tile_out = BT - 2 * halo;
    const int rs_pad = ceil_div(p.RS, KC) * KC;
    CUtensorMap tmDrs, tmWrs, tmWin, tmDxin, tmOut, tmWpost;
    memset(&tmWpost, 0, sizeof(tmWpost));
    if (p.post_w) {     // data-gradient pack of the `start` weight: [post_N rows (N)][H columns (K)]
        GTK_REQUIRE(wn_bwd_post_ok(p.H, p.post_N) && p.post_out_tm && p.post_C >= p.post_N, "fused start data gradient arguments");
        const uint64_t dims[2] = {(uint64_t)p.H, (uint64_t)p.post_N};
        const uint64_t strides[1] = {(uint64_t)p.H * 2};
        const uint32_t box[2] = {KC, (uint32_t)p.post_N};
        int rc = make_tmap_bf16(&tmWpost, p.post_w, 2, dims, strides, box, true);
        if (rc) return rc;
    }
    {
        const uint64_t dims[3] = {(uint64_t)rs_pad, (uint64_t)p.T, (uint64_t)p.B};
        const uint64_t strides[2] = {(uint64_t)ld_drs * 2, (uint64_t)ld_drs * 2 * p.T};
        const uint32_t box[3] = {KC, BT, 1};
        int rc = make_tmap_bf16(&tmDrs, drs, 3, dims, strides, box, true);
        if (rc) return rc;
    }
    {   // data-gradient pack of the res/skip weight: [H rows (N)][rs_pad columns (K)]
        const uint64_t dims[2] = {(uint64_t)rs_pad, (uint64_t)H};
        const uint64_t strides[1] = {(uint64_t)rs_pad * 2};
        const uint32_t box[2] = {KC, (uint32_t)H};
        int rc = make_tmap_bf16(&tmWrs, wrs_t, 2, dims, strides, box, true);
        if (rc) return rc;
    }
    {   // data-gradient pack of the in_layer weight: [taps][H rows (N)][2H columns (K)]
        const uint64_t dims[3] = {(uint64_t)2 * H, (uint64_t)H, (uint64_t)p.taps};
        const uint64_t strides[2] = {(uint64_t)2 * H * 2, (uint64_t)2 * H * 2 * H};
        const uint32_t box[3] = {KC, (uint32_t)(H / WB_CL), 1};
        int rc = make_tmap_bf16(&tmWin, win_t, 3, dims, strides, box, true);
        if (rc) return rc;
    }
    {
        const uint64_t dims[3] = {(uint64_t)2 * H, (uint64_t)p.T, (uint64_t)p.B};
        const uint64_t strides[2] = {(uint64_t)2 * H * 2, (uint64_t)2 * H * 2 * p.T};
        const uint32_t box[3] = {KC, (uint32_t)tile_out, 1};
        int rc = make_tmap_bf16(&tmDxin, dxin, 3, dims, strides, box, true);
        if (rc) return rc;
    }
    {
        const uint64_t dims[3] = {(uint64_t)H, (uint64_t)p.T, (uint64_t)p.B};
        const uint64_t strides[2] = {(uint64_t)ld_out * 2, (uint64_t)ld_out * 2 * p.T};
        const uint32_t box[3] = {KC, (uint32_t)tile_out, 1};
        int rc = make_tmap_bf16(&tmOut, out_bf16, 3, dims, strides, box, true);
        if (rc) return rc;
    }
    const size_t smem = (size_t)(2 * H / KC) * A_BYTES + WB_RING + sizeof(WbCtl) + 1024;
    GTK_REQUIRE(smem <= 227 * 1024, "tile does not fit in shared memory");
    GTK_CUDA(cudaFuncSetAttribute(wn_bwd_layer_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(ceil_div(ceil_div(p.T, tile_out), WB_CL) * WB_CL, p.B);     // an out-of-range tile only feeds its peer
    ProfScope prof("wn_bwd_layer_tc", st);
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid; cfg.blockDim = dim3(EPI_NT); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = WB_CL; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = attr; cfg.numAttrs = 2;
    GTK_CUDA(cudaLaunchKernelEx(&cfg, wn_bwd_layer_kernel, tmDrs, tmWrs, tmWin, tmDxin, tmOut, tmWpost, p));
    return check_launch("wn_bwd_layer_tc");
}

// The fused backward needs the whole d(xin) tile (2H columns) plus a useful operand ring in shared memory and a halo
// that leaves most of the 128-frame tile as output.
bool wn_bwd_supported(int H, int taps, int dil) {
    const int halo = (taps / 2) * dil;
    return H % KC == 0 && (H / WB_CL) % 8 == 0 && (size_t)(2 * H / KC) * A_BYTES + WB_RING + sizeof(WbCtl) + 1024 <= 227 * 1024 && 2 * halo <= 32 &&
           WB_RING / (A_BYTES + H * KC * 2) >= 2;
}

}  // namespace tc
}  // namespace gtk
