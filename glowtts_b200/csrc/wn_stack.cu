// glowtts_b200/csrc/wn_stack.cu -- a whole WaveNet (attentions.py:151 `start` + modules.py:151-171, every layer) as ONE
// tcgen05 kernel: a CTA keeps a 128-row window of the residual stream resident in shared memory and walks all L layers
// on it, so the layers cost no launch, no grid drain and no operand re-fetch.
//
//   window     rows [t_own - HT, t_own - HT + 128) of the packed sequence, HT = sum over layers of (K/2) dil_l.  A layer's
//              dilated conv reads (K/2) dil rows to either side, so after layer l the outer sum_{i<=l} (K/2) dil_i rows of
//              the window are stale; the CTA OWNS the inner tile_out = 128 - 2 HT rows (base config: 112) and recomputes
//              the halo its neighbours own (a ragged B = 32 batch is 92 windows on 148 SMs: the redundant rows run on
//              SMs that would idle).  No inter-CTA dependency, any grid size.
//   tile       bf16 [lead + 128 + lead rows][H] in the canonical K-major SWIZZLE_128B layout (lead = max halo); tap j of
//              GEMM 1 is the same tile through a row-shifted UMMA descriptor (the swizzle is a function of the address).
//              The tile holds h_l during GEMM 1, the gate output during GEMM 2, h_{l+1} after epilogue 2.
//   ring       3 weight stages streamed by TMA; the producer runs ahead across epilogues and layers (weights depend on
//              nothing), so a GEMM never waits for its first operands.
//   TMEM       2H fp32 columns: x_in of GEMM 1, then res|skip of GEMM 2 (as wn_tc.cu).
//   masters    fp32 residual / skip sums live in window-private global scratch (written and re-read by the same thread).
//
// Epilogue arithmetic is wn_tc.cu's, statement for statement: results are bit-identical to the layer-by-layer path.
#include <stdlib.h>
#include <string.h>

#include "tc.cuh"
#include "ops.h"

#define GTK_TRY(expr)            \
    do {                         \
        int rc__ = (expr);       \
        if (rc__ != GTK_OK) return rc__; \
    } while (0)

namespace gtk {
namespace tc {

struct WsCtl {
    uint64_t full[WS_STAGES], empty[WS_STAGES], acc_full, acts_ready, tile_ready, tile_tma;
    uint32_t tmem_base;
};

// Instrumented builds only (-DGTK_INSTRUMENT): globaltimer stamps of every CTA of the last launch:
// g_ws_trace[cta][0 start, 1 prologue done, 2 start-conv epilogue done, then per layer l: 3+4l GEMM 1 complete,
// 4+4l epilogue 1 done, 5+4l GEMM 2 complete, 6+4l epilogue 2 done; last: end]; g_ws_wait[cta][l] = SM cycles the MMA
// warp waited for weight stages during GEMM 1 of layer l, [4 + l] = the cycles of that GEMM 1.
#ifdef GTK_INSTRUMENT
constexpr int WS_TRACE_CTAS = 512, WS_TRACE_N = 24;
__device__ unsigned long long g_ws_trace[WS_TRACE_CTAS * WS_TRACE_N];
__device__ long long g_ws_wait[WS_TRACE_CTAS * 8];
__device__ __forceinline__ unsigned long long ws_timer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#define WS_TRACE(i)                                                                     \
    do {                                                                                \
        const int cta_ = blockIdx.y * gridDim.x + blockIdx.x;                           \
        if (threadIdx.x == 128 && cta_ < WS_TRACE_CTAS && (i) < WS_TRACE_N) g_ws_trace[cta_ * WS_TRACE_N + (i)] = ws_timer(); \
    } while (0)
#define WS_CLK() clock64()
#define WS_WAIT_STORE(i, v) do { const int cta_ = blockIdx.y * gridDim.x + blockIdx.x; if (cta_ < WS_TRACE_CTAS && (i) < 8) g_ws_wait[cta_ * 8 + (i)] = (v); } while (0)
int ws_trace_read(unsigned long long* out, int n_ctas) {
    GTK_REQUIRE(out && n_ctas >= 0 && n_ctas <= WS_TRACE_CTAS, "n_ctas");
    GTK_CUDA(cudaMemcpyFromSymbol(out, g_ws_trace, size_t(n_ctas) * WS_TRACE_N * sizeof(unsigned long long)));
    return GTK_OK;
}
int ws_wait_read(long long* out, int n_ctas) {
    GTK_REQUIRE(out && n_ctas >= 0 && n_ctas <= WS_TRACE_CTAS, "n_ctas");
    GTK_CUDA(cudaMemcpyFromSymbol(out, g_ws_wait, size_t(n_ctas) * 8 * sizeof(long long)));
    return GTK_OK;
}
#else
#define WS_TRACE(i) do { } while (0)
#define WS_CLK() 0ll
#define WS_WAIT_STORE(i, v) do { } while (0)
#endif

__device__ __forceinline__ unsigned char* ws_align1024(unsigned char* p) {
    return reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(p) + 1023) & ~uintptr_t(1023));
}

__global__ void __launch_bounds__(EPI_NT, 1)
wn_stack_kernel(const __grid_constant__ WnStackMaps tm, const __grid_constant__ WnStackParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = ws_align1024(smem_raw);
    const int H = p.H, H2 = 2 * p.H, L = p.L;
    const int n_kh = H / KC;
    const int rows_buf = BT + 2 * p.lead;
    const int pitch = (rows_buf * 128 + 1023) & ~1023;                  // one k-chunk of the tile
    const int stage_bytes = max(H2 * KC * 2, A_BYTES + H * KC * 2);
    unsigned char* tile = smem;
    unsigned char* ring = tile + (size_t)n_kh * pitch;
    float* cst = reinterpret_cast<float*>(ring + (size_t)WS_STAGES * stage_bytes);   // [L][bias_in | bias_rs | cw | cb][2H], bias_start[H]
    float* c_start = cst + (size_t)L * 4 * H2;
    float* c_end = c_start + H;                                          // bias of `end` [256]
    WsCtl* ctl = reinterpret_cast<WsCtl*>(c_end + 256);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wi = blockIdx.x, b = blockIdx.y;
    const int t_own = wi * p.tile_out;                                  // first owned row
    const int t0w = t_own - p.halo_tot;                                 // row of window row 0
    pdl_trigger();
    if (p.n_rows && t_own >= ((*p.n_rows + BT - 1) & ~(BT - 1))) return;   // varlen packing: nothing below the common limit
    WS_TRACE(0);

    const int n_s = p.has_start ? p.n_ks : 0;                           // operand stages of the start conv
    const int n1 = p.taps * n_kh, n_l = n1 + n_kh;                      // ... of GEMM 1, of one layer
    const int n_lay = n_s + L * n_l;                                    // ... up to and including the last layer
    const int total = n_lay + (p.has_end ? n_kh : 0);                   // + the `end` conv

    if (threadIdx.x == 0) {
        for (int s = 0; s < WS_STAGES; ++s) { mbar_init(&ctl->full[s], 1); mbar_init(&ctl->empty[s], 1); }
        mbar_init(&ctl->acc_full, 1);
        mbar_init(&ctl->acts_ready, EPI_NT);
        mbar_init(&ctl->tile_ready, EPI_NT);
        mbar_init(&ctl->tile_tma, 1);
        mbar_fence_init();
    }
    for (int idx = threadIdx.x; idx < L * H2; idx += EPI_NT) {
        const int l = idx / H2, i = idx - l * H2;
        const WnStackLayer& ly = p.layer[l];
        float* c = cst + (size_t)l * 4 * H2;
        c[i] = ly.bias_in[i];
        c[H2 + i] = i < ly.RS ? ly.bias_rs[i] : 0.f;
        float cwv = 0.f, cbv = 0.f;
        if (p.trk) {                                     // WNP conditioner of gate channel i (modules.py:320-322, 353-362)
            const int ch = (ly.cs_base + i) % p.cs_div;
            const float v = p.cw_v[ch];
            cwv = p.cw_g ? p.cw_g[ch] * (v < 0.f ? -1.f : 1.f) : v;      // weight norm of a 1-element row: g v / |v|
            cbv = p.cb[ch];
        }
        c[2 * H2 + i] = cwv;
        c[3 * H2 + i] = cbv;
    }
    if (p.has_end)
        for (int i = threadIdx.x; i < p.C_end; i += EPI_NT) c_end[i] = p.bias_end[i];
    if (p.has_start) {
        for (int i = threadIdx.x; i < H; i += EPI_NT) c_start[i] = p.bias_start[i];
        // the lead-in / lead-out rows are only ever read (by the outermost taps): give them a finite value once
        uint4* z = reinterpret_cast<uint4*>(tile);
        for (int i = threadIdx.x; i < n_kh * pitch / 16; i += EPI_NT) z[i] = make_uint4(0u, 0u, 0u, 0u);
        fence_proxy_async_smem();
    }
    if (warp == 2) tmem_alloc<512>(&ctl->tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctl->tmem_base;
    const bool is_prod = warp == 0, is_mma = warp == 1;

    // ------------------------------------------------------------------ operand-stage sequence, walked by the producer
    // (it_p) and by the MMA warp (it_m): [start conv: n_s] then per layer [GEMM 1: taps x n_kh] [GEMM 2: n_kh]
    int it_p = 0, it_m = 0;
    auto issue_weights = [&](int it) {                   // one elected lane: the weight part of stage `it`
        const int s = it % WS_STAGES;
        unsigned char* sb = ring + (size_t)s * stage_bytes;
        if (it < n_s) {
            mbar_arrive_expect_tx(&ctl->full[s], (uint32_t)(A_BYTES + H * KC * 2));
            tma_load_2d(sb + A_BYTES, &tm.wstart, &ctl->full[s], it * KC, 0);
            return;
        }
        if (it >= n_lay) {                              // `end` weights: [C_end rows][64 channels] per k-chunk
            mbar_arrive_expect_tx(&ctl->full[s], (uint32_t)(p.C_end * KC * 2));
            tma_load_2d(sb, &tm.wend, &ctl->full[s], (it - n_lay) * KC, 0);
            return;
        }
        const int x = it - n_s, l = x / n_l, y = x - l * n_l;
        if (y < n1) {
            const int j = y / n_kh, kc = y - j * n_kh;
            mbar_arrive_expect_tx(&ctl->full[s], (uint32_t)(H2 * KC * 2));
            tma_load_3d(sb, &tm.win[l], &ctl->full[s], kc * KC, 0, j);
            tma_load_3d(sb + (size_t)H * KC * 2, &tm.win[l], &ctl->full[s], kc * KC, H, j);
        } else {
            const int kc = y - n1, RS = p.layer[l].RS;
            mbar_arrive_expect_tx(&ctl->full[s], (uint32_t)(RS * KC * 2));
            tma_load_2d(sb, &tm.wrs[l], &ctl->full[s], kc * KC, 0);
            if (RS > H) tma_load_2d(sb + (size_t)H * KC * 2, &tm.wrs[l], &ctl->full[s], kc * KC, H);
        }
    };
    auto issue_upto = [&](int end) {                     // whole producer warp; blocks on stages the MMA warp still reads
        for (; it_p < end; ++it_p) {
            const int s = it_p % WS_STAGES, use = it_p / WS_STAGES;
            if (use > 0) mbar_wait(&ctl->empty[s], (use - 1) & 1);
            if (elect_one()) issue_weights(it_p);
            __syncwarp();
        }
    };

    // ------------------------------------------------------------------ prologue: weights first (they depend on no earlier
    // kernel), then wait for the predecessor and fetch the window
    if (is_prod) {
        if (elect_one()) {
            tma_prefetch_desc(&tm.win[0]);
            tma_prefetch_desc(&tm.wrs[0]);
            tma_prefetch_desc(p.has_start ? &tm.a0 : &tm.hin);
        }
        __syncwarp();
        issue_upto(min(WS_STAGES, total));
        pdl_wait();
        if (elect_one()) {
            if (p.has_start) {
                for (int kc = 0; kc < n_s; ++kc)
                    tma_load_3d(ring + (size_t)kc * stage_bytes, &tm.a0, &ctl->full[kc], kc * KC, t0w, b);
            } else {
                mbar_arrive_expect_tx(&ctl->tile_tma, (uint32_t)(n_kh * rows_buf * 128));
                for (int kc = 0; kc < n_kh; ++kc)
                    tma_load_3d(tile + (size_t)kc * pitch, &tm.hin, &ctl->tile_tma, kc * KC, t0w - p.lead, b);
            }
        }
        __syncwarp();
    } else {
        pdl_wait();
    }
    WS_TRACE(1);

    // ------------------------------------------------------------------ epilogue geometry: warp w owns TMEM lane quarter
    // w % 4 (window rows) and column part w / 4
    const int q = warp & 3, part = warp >> 2;
    const int r = q * 32 + lane;                         // window row == TMEM lane
    const int t = t0w + r;
    const bool t_ok = t >= 0 && t < p.T;
    const float mk = (p.mask && t_ok) ? p.mask[(size_t)b * p.T + t] : 1.f;
    const bool live = t_ok && mk != 0.f;                 // dead rows are written as exact zeros (never value * 0)
    const bool own = t_ok && r >= p.halo_tot && r < p.halo_tot + p.tile_out;
    const size_t frame = (size_t)b * p.T + (t_ok ? t : 0);
    const unsigned long long dframe = p.row_fr ? (unsigned long long)p.row_fr[frame] : (unsigned long long)frame;
    const size_t cond_off = (size_t)(p.row_b ? p.row_b[frame] : b) * p.cond_bstride;
    const float* trk_row = p.trk ? p.trk + frame * p.n_sqz : nullptr;
    const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
    RowIo iow;                                           // window-private masters
    iow.tp = 1; iow.tile_id = b * gridDim.x + wi; iow.r = r; iow.frame = frame;
    const RowIo iof = frame_io(b, t_ok ? t : 0, p.T);    // tensors addressed by frame (saved activations, chained masters)
    const int n_chunks = H / 16;
    const int ch_lo = n_chunks * part / EPI_PARTS, ch_hi = n_chunks * (part + 1) / EPI_PARTS;
    const int rb = r + p.lead;                           // row of the tile buffer
    uint32_t acc_n = 0, tr_n = 0;
    const uint32_t own_src = (uint32_t)(p.lead + p.halo_tot) * 128u;   // byte offset of the first owned row inside a chunk

    // ------------------------------------------------------------------ start conv (attentions.py:151): h_0 = (W x0 + b) mask
    if (p.has_start) {
        if (is_mma) {
            const uint32_t tmw = __shfl_sync(0xffffffffu, tmem, 0);
            const uint32_t idesc = umma_idesc_bf16(BT, H);
            for (; it_m < n_s; ++it_m) {
                const int s = it_m % WS_STAGES;
                mbar_wait(&ctl->full[s], 0);
                tc_fence_after();
                const uint32_t sa = smem_u32(ring + (size_t)s * stage_bytes);
                if (elect_one()) {
                    const uint64_t adesc = umma_desc_kmajor_sw128(sa), bdesc = umma_desc_kmajor_sw128(sa + A_BYTES);
#pragma unroll
                    for (int k = 0; k < KC / 16; ++k)          // two accumulators take alternate k16 steps (as conv_tc.cu: same sums)
                        umma_bf16(tmw + (k & 1) * 256, adesc + 2 * k, bdesc + 2 * k, idesc, it_m != 0 || k >= 2);
                    umma_commit(&ctl->empty[s]);
                }
                __syncwarp();
            }
            if (elect_one()) umma_commit(&ctl->acc_full);
            __syncwarp();
        }
        mbar_wait(&ctl->acc_full, acc_n & 1); ++acc_n;
        tc_fence_after();
        if (is_prod) issue_upto(min(total, n_s + WS_STAGES));           // the start stages are free: does not block
        for (int cc = ch_lo; cc < ch_hi; ++cc) {
            const int c0 = cc * 16;
            float v[16];
            tmem_ld16_dual(trow + c0, true, v);
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = live ? v[i] + c_start[c0 + i] : 0.f;
            row_store_f32x16(p.hf, H, c0, iow, v);
            stage_bf16x16_sw128(tile, pitch, rb, c0, v);
        }
        fence_proxy_async_smem();
        tc_fence_before();
        mbar_arrive(&ctl->tile_ready);
    } else if (is_mma) {
        mbar_wait(&ctl->tile_tma, 0);
    }
    WS_TRACE(2);

    // ------------------------------------------------------------------ layers
    for (int l = 0; l < L; ++l) {
        const WnStackLayer& ly = p.layer[l];
        const int RS = ly.RS;
        const bool first = l == 0, last = l == L - 1;
        const float* c_bin = cst + (size_t)l * 4 * H2;
        const float* c_brs = c_bin + H2;
        const float* c_cw = c_bin + 2 * H2;
        const float* c_cb = c_bin + 3 * H2;
        const int end_l = n_s + (l + 1) * n_l;

        if (is_prod) {
            issue_upto(end_l);                           // the rest of this layer's weights (paced by the MMA warp)
        } else if (is_mma) {
            const uint32_t tmw = __shfl_sync(0xffffffffu, tmem, 0);
            if (p.has_start || l > 0) {
                mbar_wait(&ctl->tile_ready, tr_n & 1); ++tr_n;
                tc_fence_after();
                if (lane == 0 && p.save) {               // h_l of the owned rows: operand of the in_layer weight gradient
                    for (int kc = 0; kc < n_kh; ++kc) tma_store_3d(&tm.hsave[l], tile + (size_t)kc * pitch + own_src, kc * KC, t_own, b);
                    tma_store_commit();
                }
                __syncwarp();
            }
            // GEMM 1: taps are row-shifted views of the resident tile; 2H gate columns as N = 256 | 2H - 256 (wn_tc.cu)
            const int n_lo = min(H2, 256), n_hi = H2 - n_lo;
            const uint32_t idesc_lo = umma_idesc_bf16(BT, n_lo), idesc_hi = umma_idesc_bf16(BT, n_hi > 0 ? n_hi : 16);
            long long w_mma = 0;
            const long long m0 = WS_CLK();
            for (int y = 0; y < n1; ++y, ++it_m) {
                const int s = it_m % WS_STAGES, use = it_m / WS_STAGES;
                const int j = y / n_kh, kc = y - j * n_kh;
                const long long c0 = WS_CLK();
                mbar_wait(&ctl->full[s], use & 1);
                w_mma += WS_CLK() - c0;
                tc_fence_after();
                const uint32_t sa = smem_u32(tile + (size_t)kc * pitch) + (uint32_t)(p.lead + (j - p.taps / 2) * ly.dil) * 128u;
                const uint32_t sb = smem_u32(ring + (size_t)s * stage_bytes);
                if (elect_one()) {
                    const uint64_t adesc = umma_desc_kmajor_sw128(sa), bdesc = umma_desc_kmajor_sw128(sb);
#pragma unroll
                    for (int k = 0; k < KC / 16; ++k) {
                        umma_bf16(tmw, adesc + 2 * k, bdesc + 2 * k, idesc_lo, (y | k) != 0);
                        if (n_hi) umma_bf16(tmw + 256, adesc + 2 * k, bdesc + 2 * k + (256 * KC * 2 >> 4), idesc_hi, (y | k) != 0);
                    }
                    umma_commit(&ctl->empty[s]);
                }
                __syncwarp();
            }
            WS_WAIT_STORE(l, w_mma);
            WS_WAIT_STORE(4 + l, WS_CLK() - m0);
            if (lane == 0) tma_store_wait_read();        // epilogue 1 overwrites the tile: the h_l store has read it
            __syncwarp();
            if (elect_one()) umma_commit(&ctl->acc_full);
            __syncwarp();
        }
        __syncwarp();

        // ---- epilogue 1: gate (commons.py:61-68) -> bf16 A operand of GEMM 2, written over the tile
        const unsigned long long dseed = p.drop_p > 0.f ? drop_resolve(ly.drop_seed) : 0ull;
        const float* cond_row = ly.cond ? ly.cond + cond_off : nullptr;
        mbar_wait(&ctl->acc_full, acc_n & 1); ++acc_n;
        tc_fence_after();
        WS_TRACE(3 + 4 * l);
        for (int cc = ch_lo; cc < ch_hi; ++cc) {
            const int c0 = cc * 16;
            float a[16], g[16];
            tmem_ld16(trow + c0, a);
            tmem_ld16(trow + H + c0, g);
#pragma unroll
            for (int i = 0; i < 16; i += 4) {
                const float4 ba = *reinterpret_cast<const float4*>(&c_bin[c0 + i]);
                const float4 bg = *reinterpret_cast<const float4*>(&c_bin[H + c0 + i]);
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
                const float pa = trk_row[(ly.cs_base + c0) / p.cs_div], pg = trk_row[(ly.cs_base + H + c0) / p.cs_div];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    a[i] += fmaf(c_cw[c0 + i], pa, c_cb[c0 + i]);
                    g[i] += fmaf(c_cw[H + c0 + i], pg, c_cb[H + c0 + i]);
                }
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                a[i] = live ? tanh_fast(a[i]) : 0.f;
                g[i] = live ? sigmoid_fast(g[i]) : 0.f;
            }
            if (ly.ts_save && own) {
                row_store_bf16x16(ly.ts_save, H2, c0, iof, a);
                row_store_bf16x16(ly.ts_save, H2, H + c0, iof, g);
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] *= g[i];
            stage_bf16x16_sw128(tile, pitch, rb, c0, a);
        }
        fence_proxy_async_smem();
        tc_fence_before();
        mbar_arrive(&ctl->acts_ready);
        WS_TRACE(4 + 4 * l);

        // ---- GEMM 2 (res_skip 1x1)
        if (is_mma) {
            const uint32_t tmw = __shfl_sync(0xffffffffu, tmem, 0);
            mbar_wait(&ctl->acts_ready, l & 1);          // the gate outputs are in the tile, the x_in columns are free
            tc_fence_after();
            const int r_lo = min(RS, 256), r_hi = RS - r_lo;
            const uint32_t id_lo = umma_idesc_bf16(BT, r_lo), id_hi = umma_idesc_bf16(BT, r_hi > 0 ? r_hi : 16);
            for (int kc = 0; kc < n_kh; ++kc, ++it_m) {
                const int s = it_m % WS_STAGES, use = it_m / WS_STAGES;
                mbar_wait(&ctl->full[s], use & 1);
                tc_fence_after();
                const uint32_t sa = smem_u32(tile + (size_t)kc * pitch) + (uint32_t)p.lead * 128u;
                const uint32_t sb = smem_u32(ring + (size_t)s * stage_bytes);
                if (elect_one()) {
                    const uint64_t adesc = umma_desc_kmajor_sw128(sa), bdesc = umma_desc_kmajor_sw128(sb);
#pragma unroll
                    for (int k = 0; k < KC / 16; ++k) {
                        umma_bf16(tmw, adesc + 2 * k, bdesc + 2 * k, id_lo, (kc | k) != 0);
                        if (r_hi) umma_bf16(tmw + 256, adesc + 2 * k, bdesc + 2 * k + (256 * KC * 2 >> 4), id_hi, (kc | k) != 0);
                    }
                    umma_commit(&ctl->empty[s]);
                }
                __syncwarp();
            }
            if (lane == 0 && p.save) {                   // bf16 gate output of the owned rows for the res_skip weight gradient
                for (int kc = 0; kc < n_kh; ++kc) tma_store_3d(&tm.acts[l], tile + (size_t)kc * pitch + own_src, kc * KC, t_own, b);
                tma_store_commit();
                tma_store_wait_read();                   // epilogue 2 overwrites the tile
            }
            __syncwarp();
            if (elect_one()) umma_commit(&ctl->acc_full);
            __syncwarp();
        }
        __syncwarp();

        // ---- epilogue 2: residual / skip (modules.py:166-171); h_{l+1} (or the masked WN output) goes back into the tile
        mbar_wait(&ctl->acc_full, acc_n & 1); ++acc_n;
        tc_fence_after();
        WS_TRACE(5 + 4 * l);
        if (is_prod) issue_upto(min(total, end_l + WS_STAGES));             // GEMM 2 has released the ring: does not block
        for (int cc = ch_lo; cc < ch_hi; ++cc) {
            const int c0 = cc * 16;
            float v[16], sk[16];
            if (RS > H) {
                tmem_ld16(trow + c0, v);
                tmem_ld16(trow + H + c0, sk);
                float x[16];
                if (first && !p.has_start) row_load_f32x16(p.h_in_f32, H, c0, iof, x);
                else row_load_f32x16(p.hf, H, c0, iow, x);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float4 bb = *reinterpret_cast<const float4*>(&c_brs[c0 + 4 * i]);
                    v[4 * i] = live ? v[4 * i] + bb.x + x[4 * i] : 0.f;
                    v[4 * i + 1] = live ? v[4 * i + 1] + bb.y + x[4 * i + 1] : 0.f;
                    v[4 * i + 2] = live ? v[4 * i + 2] + bb.z + x[4 * i + 2] : 0.f;
                    v[4 * i + 3] = live ? v[4 * i + 3] + bb.w + x[4 * i + 3] : 0.f;
                }
                row_store_f32x16(p.hf, H, c0, iow, v);
                stage_bf16x16_sw128(tile, pitch, rb, c0, v);
            } else {
                tmem_ld16(trow + c0, sk);
            }
            const int bo = (RS > H) ? H : 0;
#pragma unroll
            for (int i = 0; i < 16; ++i) sk[i] += c_brs[bo + c0 + i];
            if (!first) {
                float x[16];
                row_load_f32x16(p.skipf, H, c0, iow, x);
#pragma unroll
                for (int i = 0; i < 16; ++i) sk[i] += x[i];
            }
            if (last) {                                   // WN output = skip * mask (modules.py:171), operand of `end`
#pragma unroll
                for (int i = 0; i < 16; ++i) sk[i] = live ? sk[i] : 0.f;
                stage_bf16x16_sw128(tile, pitch, rb, c0, sk);
                if (p.h_out_f32_last && own) row_store_f32x16(p.h_out_f32_last, H, c0, iof, sk);   // fp32 master for a chained WN
            } else {
                row_store_f32x16(p.skipf, H, c0, iow, sk);
            }
        }
        fence_proxy_async_smem();
        tc_fence_before();
        mbar_arrive(&ctl->tile_ready);
        WS_TRACE(6 + 4 * l);
    }

    // ------------------------------------------------------------------ masked WN output of the owned rows; `end` conv +
    // affine coupling on the resident tile (conv_tc.cu EPI_AFFINE, same sums: two accumulators take alternate k16 steps)
    if (is_mma) {
        const uint32_t tmw = __shfl_sync(0xffffffffu, tmem, 0);
        mbar_wait(&ctl->tile_ready, tr_n & 1);
        tc_fence_after();
        if (lane == 0 && (!p.has_end || p.save)) {        // with `end` in the kernel only the weight gradient reads it
            for (int kc = 0; kc < n_kh; ++kc) tma_store_3d(&tm.skip_out, tile + (size_t)kc * pitch + own_src, kc * KC, t_own, b);
            tma_store_commit();
        }
        __syncwarp();
        if (p.has_end) {
            const uint32_t idesc = umma_idesc_bf16(BT, p.C_end);
            for (int kc = 0; kc < n_kh; ++kc, ++it_m) {
                const int s = it_m % WS_STAGES, use = it_m / WS_STAGES;
                mbar_wait(&ctl->full[s], use & 1);
                tc_fence_after();
                const uint32_t sa = smem_u32(tile + (size_t)kc * pitch) + (uint32_t)p.lead * 128u;
                const uint32_t sb = smem_u32(ring + (size_t)s * stage_bytes);
                if (elect_one()) {
                    const uint64_t adesc = umma_desc_kmajor_sw128(sa), bdesc = umma_desc_kmajor_sw128(sb);
#pragma unroll
                    for (int k = 0; k < KC / 16; ++k)
                        umma_bf16(tmw + (k & 1) * 256, adesc + 2 * k, bdesc + 2 * k, idesc, kc != 0 || k >= 2);
                }
                __syncwarp();
            }
            if (elect_one()) umma_commit(&ctl->acc_full);
            __syncwarp();
        }
        if (lane == 0) tma_store_wait_read();             // shared memory must outlive the bulk reads
        __syncwarp();
    }
    __syncwarp();
    if (p.has_end) {
        const int Ch = p.C_end / 2, C = p.C_end;
        float ld_sum = 0.f;
        const int nck = Ch / 16;
        mbar_wait(&ctl->acc_full, acc_n & 1); ++acc_n;
        tc_fence_after();
        for (int cc = nck * part / EPI_PARTS; cc < nck * (part + 1) / EPI_PARTS; ++cc) {
            const int c0 = cc * 16;
            float m[16], lg[16];
            tmem_ld16_dual(trow + c0, true, m);
            tmem_ld16_dual(trow + Ch + c0, true, lg);
            if (!own) continue;
            const float* __restrict__ ax = p.aff_x;
            float* __restrict__ az = p.aff_z;
            const size_t o0 = ((size_t)b * C + c0) * p.T + t, o1 = o0 + (size_t)Ch * p.T;
            float x0[16], x1[16], lsave[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) { x0[i] = ax[o0 + (size_t)i * p.T]; x1[i] = ax[o1 + (size_t)i * p.T]; }
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const int c = c0 + i;
                float mm = m[i] + c_end[c];
                float ll = lg[i] + c_end[Ch + c];
                lsave[i] = ll;                             // raw end-conv output (pre sigmoid_scale), as the fp32 path keeps it
                if (p.aff_sigmoid_scale) ll = __logf(1e-6f + sigmoid_f(ll + 2.f));
                az[o0 + (size_t)i * p.T] = x0[i];
                if (p.aff_reverse) {
                    az[o1 + (size_t)i * p.T] = live ? (x1[i] - mm) * __expf(-ll) : 0.f;
                } else {
                    az[o1 + (size_t)i * p.T] = live ? mm + __expf(ll) * x1[i] : 0.f;
                    ld_sum += live ? ll : 0.f;
                }
            }
            if (p.aff_logs_save) {
                float4* o4 = reinterpret_cast<float4*>(p.aff_logs_save + frame * Ch + c0);
#pragma unroll
                for (int i = 0; i < 4; ++i) o4[i] = make_float4(lsave[4 * i], lsave[4 * i + 1], lsave[4 * i + 2], lsave[4 * i + 3]);
            }
        }
        if (!p.aff_reverse && p.aff_logdet) {
            if (!own) ld_sum = 0.f;
            if (p.row_b) {
                // packed rows: the owned live rows of a warp usually belong to one utterance; at a boundary each lane adds
                // its own
                const int ub = (own && live) ? p.row_b[frame] : -1;
                const unsigned act = __ballot_sync(0xffffffffu, ub >= 0);
                if (act) {
                    const int ub0 = __shfl_sync(0xffffffffu, ub, __ffs(act) - 1);
                    if (__all_sync(0xffffffffu, ub < 0 || ub == ub0)) {
                        ld_sum = warp_sum(ld_sum);
                        if (lane == 0 && ld_sum != 0.f) atomicAdd(p.aff_logdet + ub0, ld_sum);
                    } else if (ub >= 0 && ld_sum != 0.f) {
                        atomicAdd(p.aff_logdet + ub, ld_sum);
                    }
                }
            } else {
                ld_sum = warp_sum(ld_sum);
                if (lane == 0) atomicAdd(p.aff_logdet + b, ld_sum);
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    WS_TRACE(3 + 4 * L);
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc<512>(tmem);
    }
}

static size_t ws_smem_bytes(int H, int L, int lead) {
    const int n_kh = H / KC;
    const size_t pitch = ((size_t)(BT + 2 * lead) * 128 + 1023) & ~(size_t)1023;
    const size_t stage = (size_t)((2 * H * KC * 2) > (A_BYTES + H * KC * 2) ? (2 * H * KC * 2) : (A_BYTES + H * KC * 2));
    return (size_t)n_kh * pitch + WS_STAGES * stage + ((size_t)L * 4 * 2 * H + H + 256) * sizeof(float) + sizeof(WsCtl) + 1024;
}

// Window geometry of a WaveNet with L layers, kernel K, dilation rate dr; false if the stack kernel does not cover it.
bool wn_stack_geometry(int H, int K, int dr, int L, int n_ks, int* lead, int* halo_tot, int* tile_out) {
    if (H % KC != 0 || H > 192 || L < 1 || L > WS_MAX_LAYERS || !(K & 1) || n_ks > WS_STAGES) return false;
    long long tot = 0, d = 1, mx = 0;
    for (int l = 0; l < L; ++l) { const long long h = (long long)(K / 2) * d; tot += h; mx = h; d *= dr; if (tot > 64) return false; }
    if (BT - 2 * tot < 96 || mx > 16) return false;      // the owned share of a window must stay large
    if (ws_smem_bytes(H, L, (int)mx) > 227 * 1024) return false;
    if (lead) *lead = (int)mx;
    if (halo_tot) *halo_tot = (int)tot;
    if (tile_out) *tile_out = BT - 2 * (int)tot;
    return true;
}

// `end` + affine inside the kernel: the C_end weight rows of a k-chunk must fit a ring stage
bool wn_stack_end_ok(int H, int C_end) {
    const int stage = (2 * H * KC * 2) > (A_BYTES + H * KC * 2) ? (2 * H * KC * 2) : (A_BYTES + H * KC * 2);
    return C_end % 32 == 0 && C_end >= 32 && C_end <= 256 && C_end * KC * 2 <= stage;
}

static bool stack_env_on() {
    const char* e = getenv("GTK_WN_STACK");              // read per call: tests compare both paths in one process
    return !(e && e[0] == '0');
}
bool wn_stack_enabled() { return stack_env_on(); }

int wn_stack_launch(const WnStackIo& io, WnStackParams p, cudaStream_t st) {
    GTK_REQUIRE(p.L >= 1 && p.L <= WS_MAX_LAYERS, "layers");
    GTK_REQUIRE(p.hf && p.skipf && io.skip_out, "null");
    GTK_REQUIRE(p.has_start ? (io.a0 && io.wstart && p.bias_start && p.n_ks >= 1 && p.n_ks <= WS_STAGES) : (io.hin && p.h_in_f32),
                "window input");
    GTK_REQUIRE(!p.trk || (p.cw_v && p.cb && p.n_sqz >= 1 && p.cs_div >= 16 && p.cs_div % 16 == 0),
                "time-varying conditioner arguments (2 H L / n_sqz must be a multiple of 16)");
    if (p.B == 0 || p.T == 0) return GTK_OK;
    const int H = p.H;
    // geometry from the layers' dilations
    int tot = 0, mx = 0;
    for (int l = 0; l < p.L; ++l) {
        const int h = (p.taps / 2) * p.layer[l].dil;
        tot += h; mx = h > mx ? h : mx;
        GTK_REQUIRE(p.layer[l].bias_in && p.layer[l].bias_rs && io.win[l] && io.wrs[l], "null layer parameter");
        GTK_REQUIRE(p.layer[l].RS == ((l < p.L - 1) ? 2 * H : H), "res_skip channels");
        GTK_REQUIRE(!p.save || (io.acts[l] && (io.hsave[l] || (l == 0 && !p.has_start))), "training saves");
    }
    p.lead = mx; p.halo_tot = tot; p.tile_out = BT - 2 * tot;
    GTK_REQUIRE(H % KC == 0 && H <= 192 && p.tile_out >= 64, "shape not covered by the stack kernel");
    const size_t smem = ws_smem_bytes(H, p.L, p.lead);
    GTK_REQUIRE(smem <= 227 * 1024, "window does not fit in shared memory");

    static thread_local WnStackMaps tm;
    memset(&tm, 0, sizeof(tm));
    const int rows_buf = BT + 2 * p.lead;
    auto frames_map = [&](CUtensorMap* m, const void* base, int cols, int ld, int box_rows) -> int {
        const uint64_t dims[3] = {(uint64_t)cols, (uint64_t)p.T, (uint64_t)p.B};
        const uint64_t strides[2] = {(uint64_t)ld * 2, (uint64_t)ld * 2 * p.T};
        const uint32_t box[3] = {KC, (uint32_t)box_rows, 1};
        return make_tmap_bf16(m, base, 3, dims, strides, box, true);
    };
    if (p.has_start) {
        GTK_TRY(frames_map(&tm.a0, io.a0, p.n_ks * KC, io.ld_a0, BT));
        const uint64_t dims[2] = {(uint64_t)p.n_ks * KC, (uint64_t)H};
        const uint64_t strides[1] = {(uint64_t)p.n_ks * KC * 2};
        const uint32_t box[2] = {KC, (uint32_t)H};
        GTK_TRY(make_tmap_bf16(&tm.wstart, io.wstart, 2, dims, strides, box, true));
    } else {
        GTK_TRY(frames_map(&tm.hin, io.hin, H, H, rows_buf));
    }
    for (int l = 0; l < p.L; ++l) {
        {
            const uint64_t dims[3] = {(uint64_t)H, (uint64_t)2 * H, (uint64_t)p.taps};
            const uint64_t strides[2] = {(uint64_t)H * 2, (uint64_t)H * 2 * 2 * H};
            const uint32_t box[3] = {KC, (uint32_t)H, 1};
            GTK_TRY(make_tmap_bf16(&tm.win[l], io.win[l], 3, dims, strides, box, true));
        }
        {
            const uint64_t dims[2] = {(uint64_t)H, (uint64_t)p.layer[l].RS};
            const uint64_t strides[1] = {(uint64_t)H * 2};
            const uint32_t box[2] = {KC, (uint32_t)H};
            GTK_TRY(make_tmap_bf16(&tm.wrs[l], io.wrs[l], 2, dims, strides, box, true));
        }
        if (p.save) {
            if (io.hsave[l]) GTK_TRY(frames_map(&tm.hsave[l], io.hsave[l], H, H, p.tile_out));
            GTK_TRY(frames_map(&tm.acts[l], io.acts[l], H, H, p.tile_out));
        }
    }
    GTK_TRY(frames_map(&tm.skip_out, io.skip_out, H, H, p.tile_out));
    if (p.has_end) {
        GTK_REQUIRE(io.wend && p.bias_end && p.aff_x && p.aff_z && wn_stack_end_ok(H, p.C_end),
                    "`end` conv arguments (in_channels * n_sqz must be a multiple of 32 and fit a ring stage)");
        const uint64_t dims[2] = {(uint64_t)H, (uint64_t)p.C_end};
        const uint64_t strides[1] = {(uint64_t)H * 2};
        const uint32_t box[2] = {KC, (uint32_t)p.C_end};
        GTK_TRY(make_tmap_bf16(&tm.wend, io.wend, 2, dims, strides, box, true));
    }
    GTK_CUDA(cudaFuncSetAttribute(wn_stack_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(ceil_div(p.T, p.tile_out), p.B);
    ProfScope prof("wn_stack_tc", st);
    GTK_CUDA(launch_pdl(wn_stack_kernel, grid, dim3(EPI_NT), smem, st, tm, p));
    return check_launch("wn_stack_tc");
}

// fp32 rows of the window-private masters the kernel needs for T rows per utterance
size_t wn_stack_master_rows(int T, int tile_out) { return (size_t)ceil_div(T, tile_out) * BT; }

}  // namespace tc
}  // namespace gtk
