// glowtts_b200/csrc/mas.cu -- monotonic alignment search on sm_100a.
//
// Replaces /root/reference/monotonic_align/core.pyx:9-45 (maximum_path_each /
// maximum_path_c) and the host round trip of monotonic_align/__init__.py:6-21.
//
// One CTA per utterance.  Warp 0 is the DP warp: lane l owns text rows
// x = l*XPT .. l*XPT+XPT-1, keeps the previous DP column Q[x, y-1] in registers,
// takes the x-1 neighbour of its first row with one warp shuffle per column and
// sweeps the mel frames y.  For every cell it records ONE bit,
//     d(x,y) = (x == y) || (Q[x,y-1] < Q[x-1,y-1])      (core.pyx:34, strict <)
// i.e. "the path at (x,y) came from (x-1,y-1)", packed along y (32 frames per word)
// in shared memory, so the backtrack never touches HBM and can skip whole runs with
// one clz per text row.  Warps 1..7 stream `value` tiles (32 frames x t_x rows) from
// HBM into a shared-memory ring (transposed so the DP warp reads bank-conflict-free)
// and, in the gaps, zero-fill this utterance's plane of the output, so the memory
// system works while the latency-bound DP chain runs.  HBM traffic is the algorithmic
// 8 B per cell (fp32 value read once, fp32 path written once).
//
// Bit-exactness (core.pyx): one fp32 add per cell, max() in Cython's select form
// ((v_prev > v_cur) ? v_prev : v_cur), -1e9 sentinels at x==y and (x==0,y>0), 0 at
// (0,0), tie => stay.  Cells outside the reference's band are computed but never
// consumed by an in-band cell, so the lower band clip is not needed.
#include <mutex>
#include "common.cuh"
#include <stdlib.h>
#include "ops.h"
#include "tc.cuh"

namespace gtk {
namespace mas {

constexpr int YT = 32;                   // frames per tile == bits per decision word
constexpr int NH = 7;                    // helper (load / zero-fill) warps
constexpr int NT = 32 * (1 + NH);        // threads per CTA
constexpr int MAX_STAGES = 4;
constexpr int MAX_XPT = 16;              // fast path covers T_x <= 512
constexpr float NEG = -1e9f;             // max_neg_val default of maximum_path_c (core.pyx:40)

struct Params {
    const float* value;   // [B, T_x, T_y] fp32, y fastest
    float* path;          // [B, T_x, T_y] fp32 out (fully written: zeros + ones), or null: durations / index only
    const int* t_xs;      // [B]
    const int* t_ys;      // [B]
    int* dur_out;         // [B, T_x] or null: frames per token  (= attn.sum(2), models.py:1085)
    int* idx_out;         // [B, T_y] or null: token per frame, -1 beyond t_y
    int B, T_x, T_y;
    int n_stages;
    int vec4;             // 1: value rows are 16-byte aligned (T_y % 4 == 0 and base aligned)
    uint32_t* bits_ws;    // wavefront kernel: decision words [B][T_y / 32][ROWW] in GLOBAL memory (null: shared memory);
                          // opt-in experiment, see plan_wave
};

template <int XPT>
__device__ __forceinline__ int phys_col(int x) {   // row x -> column inside a tile / bits row
    return (x % XPT) * 32 + (x / XPT);             // lane-interleaved: lane = x / XPT, k = x % XPT
}

template <int XPT, bool DIAG>
__device__ __forceinline__ void dp_tile(const float* __restrict__ tile, float (&q)[XPT], uint32_t (&bits)[XPT],
                                        int lane, int y0, int ncols) {
    constexpr int P = 32 * XPT + 1;
    const int x0 = lane * XPT;
#pragma unroll 2
    for (int yl = 0; yl < ncols; ++yl) {
        const int y = y0 + yl;
        float up = __shfl_up_sync(0xffffffffu, q[XPT - 1], 1);
        if (lane == 0) up = (y == 0) ? 0.0f : NEG;
        float v[XPT];
#pragma unroll
        for (int k = 0; k < XPT; ++k) v[k] = tile[yl * P + k * 32 + lane];
        const uint32_t bit = 1u << yl;
#pragma unroll
        for (int k = XPT - 1; k >= 0; --k) {
            const float b = (k == 0) ? up : q[k - 1];
            const float a = q[k];
            bool d = a < b;
            float a2 = a;
            if (DIAG) {
                const bool diag = (x0 + k == y);
                d = d || diag;
                a2 = diag ? NEG : a;
            }
            const float m = (b > a2) ? b : a2;
            q[k] = m + v[k];
            if (d) bits[k] |= bit;
        }
    }
}

// Ones of the path, durations and per-frame token indices from the token start frames ys[] (all threads of the CTA).
__device__ __forceinline__ void emit_path(const Params& p, float* __restrict__ pplane, const int* ys, bool valid, int t_x, int t_y,
                                          int b, int tid, int n_threads) {
    for (int x = tid; x < p.T_x; x += n_threads) {
        int start = 0, end = 0;
        if (valid && x < t_x) {
            start = ys[x];
            end = (x + 1 < t_x) ? ys[x + 1] : t_y;
            if (pplane)
                for (int y = start; y < end; ++y) pplane[(size_t)x * p.T_y + y] = 1.0f;
            if (p.idx_out)
                for (int y = start; y < end; ++y) p.idx_out[(size_t)b * p.T_y + y] = x;
        }
        if (p.dur_out) p.dur_out[(size_t)b * p.T_x + x] = end - start;
    }
    if (p.idx_out)
        for (int y = (valid ? t_y : 0) + tid; y < p.T_y; y += n_threads) p.idx_out[(size_t)b * p.T_y + y] = -1;
}

template <int XPT>
__global__ void __launch_bounds__(NT, 1) mas_kernel(const Params p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int P = 32 * XPT + 1;
    constexpr int ROWW = 32 * XPT;   // words per decision row-block
    const int S = p.n_stages;
    const int NYB = (p.T_y + YT - 1) / YT;

    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);
    uint64_t* empty = full + MAX_STAGES;
    int* ys = reinterpret_cast<int*>(smem_raw + 2 * MAX_STAGES * sizeof(uint64_t));
    constexpr int YS_WORDS = ((ROWW + 1 + 3) / 4) * 4;
    uint32_t* bitsT = reinterpret_cast<uint32_t*>(ys + YS_WORDS);
    float* tiles = reinterpret_cast<float*>(bitsT + (size_t)NYB * ROWW);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.x;
    int t_x = min(max(p.t_xs[b], 0), p.T_x);
    int t_y = min(max(p.t_ys[b], 0), p.T_y);
    const bool valid = (t_x >= 1) && (t_y >= t_x);   // t_x > t_y is undefined in the reference
    const int n_tiles = valid ? (t_y + YT - 1) / YT : 0;
    const int n_iter = max(n_tiles, 1);
    const size_t plane_elems = (size_t)p.T_x * p.T_y;
    const float* __restrict__ vplane = p.value + (size_t)b * plane_elems;
    float* __restrict__ pplane = p.path ? p.path + (size_t)b * plane_elems : nullptr;   // null: index + durations only
    const size_t fill_elems = p.path ? plane_elems : 0;

    if (tid == 0) {
        for (int s = 0; s < MAX_STAGES; ++s) {
            mbar_init(&full[s], NH * 32);
            mbar_init(&empty[s], 1);
        }
        mbar_fence_init();
    }
    __syncthreads();

    if (warp == 0) {
        // ------------------------------------------------------------ DP warp
        float q[XPT];
        uint32_t bits[XPT];
#pragma unroll
        for (int k = 0; k < XPT; ++k) q[k] = 0.0f;
        for (int it = 0; it < n_tiles; ++it) {
            const int s = it % S, use = it / S;
            mbar_wait(&full[s], use & 1);
            const float* tile = tiles + (size_t)s * YT * P;
            const int y0 = it * YT;
            const int ncols = min(YT, t_y - y0);
#pragma unroll
            for (int k = 0; k < XPT; ++k) bits[k] = 0u;
            if (y0 < 32 * XPT) dp_tile<XPT, true>(tile, q, bits, lane, y0, ncols);
            else               dp_tile<XPT, false>(tile, q, bits, lane, y0, ncols);
#pragma unroll
            for (int k = 0; k < XPT; ++k) bitsT[(size_t)it * ROWW + k * 32 + lane] = bits[k];
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        }
        __syncwarp();
        // ------------------------------------------------------------ backtrack (run-skipping)
        if (lane == 0 && valid) {
            int index = t_x - 1, y = t_y - 1;
            while (true) {
                if (index == 0 || y < 0) { ys[0] = 0; break; }
                const int yb = y >> 5;
                const uint32_t w = bitsT[(size_t)yb * ROWW + phys_col<XPT>(index)];
                const uint32_t m = w & (0xffffffffu >> (31 - (y & 31)));
                if (m) {
                    const int yp = (yb << 5) + (31 - __clz(m));
                    ys[index] = yp;       // first frame of token `index`
                    --index;
                    y = yp - 1;
                } else {
                    y = (yb << 5) - 1;
                }
            }
            // y < 0 with index > 0 cannot happen for valid inputs (diagonal bits force the move);
            // if it did, make the remaining rows empty rather than reading garbage.
            for (int x = 1; x <= index; ++x) ys[x] = 0;
        }
    } else {
        // ------------------------------------------------------------ helper warps
        const int hw = warp - 1;
        const int ht = tid - 32;
        constexpr int HT = NH * 32;
        // zero-fill geometry: [head scalars][nvec float4][tail scalars]
        const uintptr_t addr = reinterpret_cast<uintptr_t>(pplane);
        size_t head = ((16 - (addr & 15)) & 15) / 4;
        if (head > fill_elems) head = fill_elems;
        const size_t nvec = (fill_elems - head) / 4;
        const size_t tail0 = head + nvec * 4;
        float4* pv = reinterpret_cast<float4*>(pplane + head);
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);

        for (int it = 0; it < n_iter; ++it) {
            if (it < n_tiles) {
                const int s = it % S, use = it / S;
                if (use >= 1) mbar_wait(&empty[s], (use - 1) & 1);
                float* tile = tiles + (size_t)s * YT * P;
                const int y0 = it * YT;
                if (p.vec4) {
                    // item = (k, group of 4 lanes-of-rows); each lane loads 4 frames of one row.
                    const int yc = (lane & 7) * 4;
                    const bool ycol_ok = (y0 + yc) < t_y;
                    constexpr int U = 4;
                    for (int item0 = hw; item0 < XPT * 8; item0 += NH * U) {
                        float4 r[U];
#pragma unroll
                        for (int u = 0; u < U; ++u) {
                            const int item = item0 + u * NH;
                            const int k = item >> 3, qg = item & 7;
                            const int x = (4 * qg + (lane >> 3)) * XPT + k;
                            r[u] = z4;
                            if (item < XPT * 8 && x < t_x && ycol_ok)
                                r[u] = __ldg(reinterpret_cast<const float4*>(vplane + (size_t)x * p.T_y + y0 + yc));
                        }
#pragma unroll
                        for (int u = 0; u < U; ++u) {
                            const int item = item0 + u * NH;
                            const int k = item >> 3, qg = item & 7;
                            const int qq = 4 * qg + (lane >> 3);
                            if (item < XPT * 8 && qq * XPT + k < t_x) {
                                float* dst = tile + (size_t)yc * P + k * 32 + qq;
                                dst[0] = r[u].x; dst[P] = r[u].y; dst[2 * P] = r[u].z; dst[3 * P] = r[u].w;
                            }
                        }
                    }
                } else {
                    const int y = y0 + lane;
                    const bool y_ok = y < t_y;
                    constexpr int U = 8;
                    for (int x0 = hw; x0 < t_x; x0 += NH * U) {
                        float r[U];
#pragma unroll
                        for (int u = 0; u < U; ++u) {
                            const int x = x0 + u * NH;
                            r[u] = (x < t_x && y_ok) ? __ldg(vplane + (size_t)x * p.T_y + y) : 0.f;
                        }
#pragma unroll
                        for (int u = 0; u < U; ++u) {
                            const int x = x0 + u * NH;
                            if (x < t_x) tile[(size_t)lane * P + phys_col<XPT>(x)] = r[u];
                        }
                    }
                }
                mbar_arrive(&full[s]);
            }
            // zero-fill share `it` of this utterance's output plane
            const size_t lo = nvec * (size_t)it / n_iter, hi = nvec * (size_t)(it + 1) / n_iter;
            for (size_t i = lo + ht; i < hi; i += HT) pv[i] = z4;
            if (it == 0) {
                if ((size_t)ht < head) pplane[ht] = 0.f;
                for (size_t i = tail0 + ht; i < fill_elems; i += HT) pplane[i] = 0.f;
            }
        }
    }
    __syncthreads();

    emit_path(p, pplane, ys, valid, t_x, t_y, b, tid, NT);
}

// -------------------------------------------------------------------------------------------------
// TMA wavefront kernel: the fast path when value rows are 16-byte aligned (T_y % 4 == 0).  Same DP, same bits, but
//   * FOUR DP warps (one per SM sub-partition) each own a band of 32 * XPT text rows and run one 32-frame tile apart:
//     warp w hands the DP value of its last row to warp w + 1 through a 32-float shared-memory mailbox per tile
//     (mbarrier full/empty), so the per-column critical path is XPT cells instead of T_x / 32;
//   * `value` is streamed by cp.async.bulk.tensor (3-D map [B][T_x][T_y], boxes of 36 frames x 32 text rows, only the
//     boxes below t_x are issued) into one mbarrier ring PER DP WARP (the warps are skewed in time), so several tiles
//     are in flight per SM while the latency-bound DP chain runs;
//   * a tile is kept row-major [row][36 floats] exactly as TMA writes it (pitch 36 words == 4 mod 32): a lane reads one
//     128-bit word per (text row, 4 frames) and software-pipelines the next 4 frames' loads;
//   * off the diagonal a cell is 2 FMA-pipe + 2 ALU-pipe instructions: t = a - b (its sign IS a < b: neither operand
//     can be -0, see dp_col), the decision bit funnel-shifted out of t's sign, max(a, b), + value;
//   * the remaining warps zero-fill the output plane; the backtrack prefetches the words its next step can need.
constexpr int TP = 36;                   // tile row pitch in floats (32 frames + 4: conflict-free 128-bit reads)
constexpr int TBOX = 32;                 // text rows per TMA box
constexpr int TBOX_BYTES = TBOX * TP * 4;
constexpr int NDP = 4;                   // DP warps
constexpr int NBOX = 8;                  // mailbox slots between neighbouring DP warps (tiles a warp may run ahead)
constexpr int TMA_MAX_STAGES = 6;
constexpr int TMA_MAX_XPT = 8;           // T_x <= 1024 (if the decision bits fit)

struct WaveCtl {
    uint64_t full[NDP][TMA_MAX_STAGES], empty[NDP][TMA_MAX_STAGES];   // value tiles, one ring per DP warp
    uint64_t box_full[NDP][NBOX], box_empty[NDP][NBOX];               // mailbox of warp w -> w + 1, NBOX tiles deep
    uint64_t dp_done;
    alignas(16) float box[NDP][NBOX][YT];
};

// Phase stamps (globaltimer, ns) per CTA of the last launch: [start, first tile landed, DP done (all warps), backtrack
// done, all warps past the barrier, end]; read with gtk_debug_mas_trace.
#ifdef GTK_INSTRUMENT
constexpr int MAS_TRACE_CTAS = 1024;
__device__ unsigned long long g_mas_trace[MAS_TRACE_CTAS * 6];
__device__ long long g_mas_wait[MAS_TRACE_CTAS];            // SM cycles DP warp 0 waited for tiles (after the first)
__device__ __forceinline__ void mas_stamp(int i) {
    if (blockIdx.x < MAS_TRACE_CTAS) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        g_mas_trace[blockIdx.x * 6 + i] = t;
    }
}
#else
__device__ __forceinline__ void mas_stamp(int) {}
#endif

__device__ __forceinline__ float4 lds128(uint32_t addr) {      // explicit ld.shared (a generic pointer costs an LD.E)
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

// One DP column of one warp.  `up_in` = DP value of the row below this warp's band at column y - 1 (lane 0 only).
// Decision bits are shifted in MSB-first (frame 0 of the tile ends at bit 31), the order the backtrack wants.
// Off the diagonal: sign(a - b) == (a < b) exactly, because a DP value is never -0 (a sum is -0 only if both addends
// are; the recursion starts from +0 and the sentinels are non-zero) and fp32 subtraction of unequal finite numbers is
// never zero (denormals are kept: no -ftz); fmaxf(a, b) == ((b > a) ? b : a) for non-NaN input.
template <int XPT, bool DIAG>
__device__ __forceinline__ void dp_col(const float (&v)[XPT], float (&q)[XPT], uint32_t (&bits)[XPT], int lane, int x0, int y,
                                       float up_in) {
#if defined(GTK_MAS_DIAGNOSE) && GTK_MAS_DIAGNOSE == 1
    float up = q[XPT - 1];           // DIAGNOSTIC BUILD (wrong results): no cross-lane shuffle, to time the chain without it
#else
    float up = __shfl_up_sync(0xffffffffu, q[XPT - 1], 1);
#endif
    if (lane == 0) up = up_in;
#pragma unroll
    for (int k = XPT - 1; k >= 0; --k) {
        const float b = (k == 0) ? up : q[k - 1];
        const float a = q[k];
        if (DIAG) {
            const bool diag = (x0 + k == y);
            const bool d = (a < b) || diag;
            const float a2 = diag ? NEG : a;
            const float m = (b > a2) ? b : a2;
            q[k] = m + v[k];
            bits[k] = (bits[k] << 1) | (d ? 1u : 0u);
        } else {
            const float t = a - b;
            bits[k] = __funnelshift_l(__float_as_uint(t), bits[k], 1);
            q[k] = fmaxf(a, b) + v[k];
        }
    }
}

// All 32 frames of a tile (frames past t_y are computed too: their bits are masked by the backtrack and q is dead).
// `tile` = shared-space byte address of this warp's rows of the stage, roff[k] = byte offset of the lane's k-th row.
// FIRST: the warp that owns text row 0 (its lane 0 takes the reference's sentinels); otherwise bv[yl] = DP value of
// the row below the band at frame y0 + yl - 1.  SEND: publish this warp's last row into `box_out` (shared address).
template <int XPT, bool DIAG, bool FIRST, bool SEND>
__device__ __forceinline__ void dp_tile4(uint32_t tile, const int (&roff)[XPT], float (&q)[XPT], uint32_t (&bits)[XPT], int lane,
                                         int x0, int y0, const float (&bv)[YT], uint32_t box_out) {
    float4 cur[XPT], nxt[XPT];
#pragma unroll
    for (int k = 0; k < XPT; ++k) cur[k] = lds128(tile + roff[k]);
#pragma unroll
    for (int g = 0; g < YT / 4; ++g) {
        if (g + 1 < YT / 4) {
#pragma unroll
            for (int k = 0; k < XPT; ++k) nxt[k] = lds128(tile + roff[k] + 16 * (g + 1));
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int yl = 4 * g + c;
            float v[XPT];
#pragma unroll
            for (int k = 0; k < XPT; ++k) v[k] = c == 0 ? cur[k].x : c == 1 ? cur[k].y : c == 2 ? cur[k].z : cur[k].w;
            const float up_in = FIRST ? ((DIAG && y0 + yl == 0) ? 0.0f : NEG) : bv[yl];
            dp_col<XPT, DIAG>(v, q, bits, lane, x0, y0 + yl, up_in);
#if !defined(GTK_MAS_DIAGNOSE) || GTK_MAS_DIAGNOSE != 2      // 2: DIAGNOSTIC BUILD (wrong results) without the mailbox store
            if (SEND && lane == 31)
                asm volatile("st.shared.f32 [%0], %1;" ::"r"(box_out + 4 * yl), "f"(q[XPT - 1]) : "memory");
#endif
        }
        if (g + 1 < YT / 4) {
#pragma unroll
            for (int k = 0; k < XPT; ++k) cur[k] = nxt[k];
        }
    }
}

// Probe first (mbarrier.test_wait does not suspend): in steady state the phase a DP warp needs has already completed.
__device__ __forceinline__ void fast_wait(uint64_t* bar, uint32_t parity) {
    if (!mbar_test_wait(bar, parity)) mbar_wait(bar, parity);
}

// Cycle stamps inside the DP loop cost ~25 cycles each, four per tile: only in the instrumented build
// (-DGTK_MAS_EXP_TILETIME, profiles/mas_tiletime.py); the product build reports 0 waited cycles.
#ifdef GTK_MAS_EXP_TILETIME
#define MAS_CLOCK() clock64()
#else
#define MAS_CLOCK() 0ll
#endif

// One DP warp of the wavefront (band w = text rows [w * 32 * XPT, (w + 1) * 32 * XPT)) over all tiles of the utterance:
// waits for its value tiles on ctl->full[w][*], releases them on ctl->empty[w][*], trades the band-boundary row with its
// neighbours through the mailboxes and leaves the decision words in bitsT.  Returns the cycles warp 0 waited for tiles.
template <int XPT>
__device__ __forceinline__ long long wave_dp_warp(WaveCtl* ctl, const float* tiles, size_t stage_floats, uint32_t* bitsT,
                                                  int ROWW, int RB, int S, int n_tiles, int n_act, int w, int lane) {
    constexpr int BAND = 32 * XPT;
    if (w == 0 && lane == 0) mas_stamp(0);
    long long waited = 0;
    if (w < n_act) {
        const int r0 = w * BAND;                 // first text row of the band
        const int x0 = r0 + lane * XPT;
        const bool send = (w + 1 < n_act);
        float q[XPT];
        uint32_t bits[XPT];
        int roff[XPT];
#pragma unroll
        for (int k = 0; k < XPT; ++k) {
            q[k] = 0.0f;
            roff[k] = min(x0 + k, RB * TBOX - 1) * TP * 4;     // rows past the stage are never consumed: clamp the read
        }
        float carry = 0.0f;                      // row r0 - 1 at the last frame of the previous tile
        for (int it = 0; it < n_tiles; ++it) {
            const int s = it % S, use = it / S, y0 = it * YT;
            const long long w0 = MAS_CLOCK();
            fast_wait(&ctl->full[w][s], use & 1);
            const long long w1 = MAS_CLOCK();
            if (w == 0 && it > 0) waited += w1 - w0;
            if (w == 0 && it == 0 && lane == 0) mas_stamp(1);
            float bv[YT];
            if (w > 0) {
                fast_wait(&ctl->box_full[w - 1][it % NBOX], (it / NBOX) & 1);
                const uint32_t bx = smem_u32(&ctl->box[w - 1][it % NBOX][0]);
                float4 r[YT / 4];
#pragma unroll
                for (int j = 0; j < YT / 4; ++j) r[j] = lds128(bx + 16 * j);
                bv[0] = carry;
#pragma unroll
                for (int j = 0; j < YT / 4; ++j) {
                    if (4 * j + 1 < YT) bv[4 * j + 1] = r[j].x;
                    if (4 * j + 2 < YT) bv[4 * j + 2] = r[j].y;
                    if (4 * j + 3 < YT) bv[4 * j + 3] = r[j].z;
                    if (4 * j + 4 < YT) bv[4 * j + 4] = r[j].w;
                }
                carry = r[YT / 4 - 1].w;
                __syncwarp();
                if (lane == 0) mbar_arrive(&ctl->box_empty[w - 1][it % NBOX]);
            }
            if (send && it >= NBOX) fast_wait(&ctl->box_empty[w][it % NBOX], (it / NBOX - 1) & 1);
            const long long w2 = MAS_CLOCK();
            const uint32_t tile = smem_u32(tiles + (size_t)s * stage_floats);
            const uint32_t bo = smem_u32(&ctl->box[w][it % NBOX][0]);
#pragma unroll
            for (int k = 0; k < XPT; ++k) bits[k] = 0u;
            const bool diag = (y0 < r0 + BAND) && (y0 + YT > r0);
            if (w == 0) {
                if (diag) { if (send) dp_tile4<XPT, true, true, true>(tile, roff, q, bits, lane, x0, y0, bv, bo);
                            else      dp_tile4<XPT, true, true, false>(tile, roff, q, bits, lane, x0, y0, bv, bo); }
                else      { if (send) dp_tile4<XPT, false, true, true>(tile, roff, q, bits, lane, x0, y0, bv, bo);
                            else      dp_tile4<XPT, false, true, false>(tile, roff, q, bits, lane, x0, y0, bv, bo); }
            } else {
                if (diag) { if (send) dp_tile4<XPT, true, false, true>(tile, roff, q, bits, lane, x0, y0, bv, bo);
                            else      dp_tile4<XPT, true, false, false>(tile, roff, q, bits, lane, x0, y0, bv, bo); }
                else      { if (send) dp_tile4<XPT, false, false, true>(tile, roff, q, bits, lane, x0, y0, bv, bo);
                            else      dp_tile4<XPT, false, false, false>(tile, roff, q, bits, lane, x0, y0, bv, bo); }
            }
            const long long w3 = MAS_CLOCK();
#pragma unroll
            for (int k = 0; k < XPT; ++k)
                if (x0 + k < ROWW) bitsT[(size_t)it * ROWW + x0 + k] = bits[k];      // MSB-first: frame f at bit 31 - f
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&ctl->empty[w][s]);
                if (send) mbar_arrive(&ctl->box_full[w][it % NBOX]);
            }
#if defined(GTK_MAS_EXP_TILETIME) && defined(GTK_INSTRUMENT)
            if (it == 10 && lane == 0 && blockIdx.x == 0 && w < 4) {
                g_mas_trace[900 * 6 + w * 6 + 0] = (unsigned long long)(w1 - w0);
                g_mas_trace[900 * 6 + w * 6 + 1] = (unsigned long long)(w2 - w1);
                g_mas_trace[900 * 6 + w * 6 + 2] = (unsigned long long)(w3 - w2);
                g_mas_trace[900 * 6 + w * 6 + 3] = (unsigned long long)(clock64() - w3);
                g_mas_trace[900 * 6 + w * 6 + 4] = (unsigned long long)w0;
            }
#endif
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&ctl->dp_done);
    }
    return waited;
}

// ------------------------------------------------------------ backtrack (run-skipping, one thread)
// Token `index` starts at the highest set bit of its row at or below y.  The loop-carried chain is mask -> clz ->
// select: the words the NEXT row can need (same 32-frame block or the one before) are loaded one step ahead.
__device__ __forceinline__ void wave_backtrack(WaveCtl* ctl, const uint32_t* bitsT, int ROWW, int* ys, bool valid, int t_x,
                                               int t_y, long long waited) {
    if (valid) fast_wait(&ctl->dp_done, 0);
    mas_stamp(2);
#ifdef GTK_INSTRUMENT
    if (blockIdx.x < MAS_TRACE_CTAS) g_mas_wait[blockIdx.x] = waited;
#endif
    if (valid) {
        // Decision words are MSB-first (frame f of the block at bit 31 - f), so "the last frame <= y whose bit is set"
        // is the LOWEST set bit of (word & M), M = the bit positions still allowed.  The words of the next BT_ROWS rows
        // of the current 32-frame block are loaded together; a token that starts in an earlier block (m == 0) ends the
        // batch and moves one block down with M = all ones.
        constexpr int BT_ROWS = 8;
        int index = t_x - 1;
        int yb = (t_y - 1) >> 5;
        uint32_t M = 0xffffffffu << (31 - ((t_y - 1) & 31));
        while (index > 0 && yb >= 0) {
            const uint32_t* base = bitsT + (size_t)yb * ROWW + index;      // word of (row index, block yb)
            uint32_t W[BT_ROWS];
#pragma unroll
            for (int j = 0; j < BT_ROWS; ++j) {
                const uint32_t w = base[-min(j, index)];
                W[j] = (j < index) ? w : 0u;                               // row 0 never moves down: it cannot "hit"
            }
            const int f0 = yb << 5;
            int done = 0;
            // Branch-free: per token the loop-carried chain is and -> (m - 1) -> not(m xor (m - 1)) = the bits above
            // the lowest set bit of m.  A miss (m == 0) leaves M == 0, so every later row of the batch misses too
            // (exactly the "first miss ends the batch" rule); the frame number (clz) and the store are off the chain.
#pragma unroll
            for (int j = 0; j < BT_ROWS; ++j) {
                const uint32_t m = W[j] & M;
                const uint32_t bb = m & (0u - m);
                ys[(m != 0u) ? index - j : ROWW + 2] = f0 + __clz(bb);     // first frame of token index - j (a miss
                                                                           // writes the padding slot: no branch)
                done += (m != 0u) ? 1 : 0;
                M = ~(m ^ (m - 1u));
            }
            const bool cross = done < min(BT_ROWS, index);                 // an active row started in an earlier block
            index -= done;
            if (cross) { --yb; M = 0xffffffffu; }
        }
        ys[0] = 0;
        // y < 0 with index > 0 cannot happen for valid inputs (diagonal bits force the move);
        // if it did, make the remaining rows empty rather than reading garbage.
        for (int x = 1; x <= index; ++x) ys[x] = 0;
    }
    mas_stamp(3);
}

template <int XPT>
__global__ void __launch_bounds__(NT, 1) mas_wave_kernel(const __grid_constant__ CUtensorMap tmV, const Params p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int BAND = 32 * XPT;                   // text rows per DP warp
    const int S = p.n_stages;
    const int NYB = (p.T_y + YT - 1) / YT;
    const int ROWW = (p.T_x + 3) & ~3;               // decision words per 32-frame block: word [yb][x]
    const int RB = (p.T_x + TBOX - 1) / TBOX;        // row boxes a stage has room for
    const size_t stage_floats = (size_t)RB * TBOX * TP;

    WaveCtl* ctl = reinterpret_cast<WaveCtl*>(smem_raw);
    int* ys = reinterpret_cast<int*>(smem_raw + sizeof(WaveCtl));
    uint32_t* bits_smem = reinterpret_cast<uint32_t*>(ys + ROWW + 4);
    uint32_t* bitsT = p.bits_ws ? p.bits_ws + (size_t)blockIdx.x * NYB * ROWW : bits_smem;
    float* tiles = reinterpret_cast<float*>(
        (reinterpret_cast<uintptr_t>(bits_smem + (p.bits_ws ? 0 : (size_t)NYB * ROWW)) + 127) & ~static_cast<uintptr_t>(127));

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.x;
    int t_x = min(max(p.t_xs[b], 0), p.T_x);
    int t_y = min(max(p.t_ys[b], 0), p.T_y);
    const bool valid = (t_x >= 1) && (t_y >= t_x);   // t_x > t_y is undefined in the reference
    const int n_tiles = valid ? (t_y + YT - 1) / YT : 0;
    const int n_act = valid ? (t_x + BAND - 1) / BAND : 0;      // DP warps with rows below t_x
    const size_t plane_elems = (size_t)p.T_x * p.T_y;
    float* __restrict__ pplane = p.path ? p.path + (size_t)b * plane_elems : nullptr;   // null: index + durations only
    const size_t fill_elems = p.path ? plane_elems : 0;

    if (tid == 0) {
        for (int w = 0; w < NDP; ++w) {
            for (int s = 0; s < TMA_MAX_STAGES; ++s) {
                mbar_init(&ctl->full[w][s], 1);
                mbar_init(&ctl->empty[w][s], 1);
            }
            for (int s = 0; s < NBOX; ++s) {
                mbar_init(&ctl->box_full[w][s], 1);
                mbar_init(&ctl->box_empty[w][s], 1);
            }
        }
        mbar_init(&ctl->dp_done, max(n_act, 1));
        mbar_fence_init();
    }
    __syncthreads();

    if (warp < NDP) {
        // ------------------------------------------------------------ DP warps
        const int w = warp;
        const long long waited = wave_dp_warp<XPT>(ctl, tiles, stage_floats, bitsT, ROWW, RB, S, n_tiles, n_act, w, lane);
        if (w == 0 && lane == 0) wave_backtrack(ctl, bitsT, ROWW, ys, valid, t_x, t_y, waited);
    } else if (warp == NDP) {
        // ------------------------------------------------------------ TMA producers: lane w feeds DP warp w
        if (lane < n_act) {
            const int w = lane;
            const int r0 = w * BAND;
            const int nbox = min(XPT, (t_x - r0 + TBOX - 1) / TBOX);
            for (int it = 0; it < n_tiles; ++it) {
                const int s = it % S, use = it / S;
                if (use >= 1) fast_wait(&ctl->empty[w][s], (use - 1) & 1);
                float* dst = tiles + (size_t)s * stage_floats + (size_t)r0 * TP;
                mbar_arrive_expect_tx(&ctl->full[w][s], (uint32_t)(nbox * TBOX_BYTES));
                for (int j = 0; j < nbox; ++j)
                    tc::tma_load_3d(dst + (size_t)j * TBOX * TP, &tmV, &ctl->full[w][s], it * YT, r0 + j * TBOX, b);
            }
        }
    } else {
        // ------------------------------------------------------------ zero-fill of this utterance's output plane
        const int ht = tid - 32 * (NDP + 1);
        constexpr int HT = NT - 32 * (NDP + 1);
        const uintptr_t addr = reinterpret_cast<uintptr_t>(pplane);
        size_t head = ((16 - (addr & 15)) & 15) / 4;
        if (head > fill_elems) head = fill_elems;
        const size_t nvec = (fill_elems - head) / 4;
        const size_t tail0 = head + nvec * 4;
        float4* pv = reinterpret_cast<float4*>(pplane + head);
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        for (size_t i = ht; i < nvec; i += HT) pv[i] = z4;
        if ((size_t)ht < head) pplane[ht] = 0.f;
        for (size_t i = tail0 + ht; i < fill_elems; i += HT) pplane[i] = 0.f;
    }
    __syncthreads();
    if (tid == 0) mas_stamp(4);

    emit_path(p, pplane, ys, valid, t_x, t_y, b, tid, NT);
    if (tid == 0) mas_stamp(5);
}

// -------------------------------------------------------------------------------------------------
// Fused logp + MAS (mean_only models, D = 80 mel channels, T_x <= 256): the value tiles the DP warps consume are
// never read from HBM -- they are COMPUTED in the CTA, one 32-frame tile ahead of the DP, from z [80, T_y] and
// x_m [80, T_x] (models.py:1076-1082 with x_logs == 0):
//     value[x, y] = c[x] + sum_d x_m[d, x] z[d, y],      c[x] = -80 * 0.5 log 2pi - 0.5 sum_d x_m[d, x]^2.
// The fourth term of the reference, -0.5 sum_d z[d, y]^2, is the same for every text row of frame y: it shifts the
// whole DP column Q[., y] by one constant and cannot change any decision Q[x, y-1] < Q[x-1, y-1], so the DP runs
// without it (and with smaller magnitudes); it is added back only where logp itself is written (logp_out).
//   * 4 DP warps: unchanged (wave_dp_warp), fed through the same per-band rings of [row][36]-pitch tiles;
//   * 8 MMA warps, two per DP band: a warp keeps its 16 * XPT rows of x_m as mma.sync A fragments in REGISTERS for
//     the whole utterance (80 fp32 at XPT = 2) and per tile runs the [16 XPT x 80] x [80 x 32] product on the
//     tensor cores as 3xTF32 (a_lo b_hi + a_hi b_lo + a_hi b_hi, hi = upper 19 bits of v, lo = v - hi: fp32-level
//     accuracy from the tf32 pipe), adds c[x] and stores the D fragments into its band's tile stage;
//   * z tiles (80 x 40-float rows: pitch 40 makes the B-fragment reads conflict-free) arrive by TMA into a
//     4-deep ring shared by all bands (issued by one lane of MMA warp 0, released by all active MMA warps);
//   * the MMA warps also zero-fill the output plane between tiles.
// HBM traffic: 4 B/cell (the path) + the z / x_m operands; no [B,T_x,T_y] value tensor exists.
constexpr int FD = 80;                   // mel channels = contraction length
constexpr int ZP = 40;                   // z tile row pitch in floats (TMA box of 40 frames, 32 used)
constexpr int FZ_STAGES = 4;
constexpr int FZ_BYTES = FD * ZP * 4;
constexpr int F_NMMA = 2 * NDP;
constexpr int F_NT = 32 * (NDP + F_NMMA);     // 384 threads: 170 registers per thread
constexpr int F_MAX_XPT = 2;             // T_x <= 256: the A fragments of a band half must fit the register file

struct FusedCtl {
    uint64_t zfull[FZ_STAGES], zempty[FZ_STAGES];
};

// hi / lo split for 3xTF32 without cvt (cvt.rna.tf32 issues at a quarter of the ALU rate and there are 320 of them per
// warp and tile): the tf32 tensor-core path reads the upper 19 bits of an fp32 register, so hi = v with the low 13
// mantissa bits cleared (one LOP) and lo = v - hi (exact in fp32; its own low bits are ignored by the MMA).
// `pinned` (volatile) keeps the split of the loop-invariant A fragments inside the tile loop -- hoisted, hi and lo of
// every fragment would have to live in registers (2 x 80) and spill.
__device__ __forceinline__ uint32_t tf32_hi(float v) { return __float_as_uint(v) & 0xffffe000u; }
__device__ __forceinline__ uint32_t tf32_hi_pinned(float v) {
    uint32_t r;
    asm volatile("and.b32 %0, %1, 0xffffe000;" : "=r"(r) : "r"(__float_as_uint(v)));
    return r;
}
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float lds32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}

template <int XPT>
__global__ void __launch_bounds__(F_NT, 1) mas_fused_kernel(const __grid_constant__ CUtensorMap tmZ, const Params p,
                                                            const float* __restrict__ x_m, float* __restrict__ logp_out) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int BAND = 32 * XPT;                   // text rows per DP warp
    constexpr int MB = XPT;                          // 16-row MMA blocks per MMA warp (half a band)
    const int S = p.n_stages;
    const int NYB = (p.T_y + YT - 1) / YT;
    const int ROWW = (p.T_x + 3) & ~3;
    const int RB = (p.T_x + TBOX - 1) / TBOX;
    const size_t stage_floats = (size_t)RB * TBOX * TP;

    WaveCtl* ctl = reinterpret_cast<WaveCtl*>(smem_raw);
    FusedCtl* fz = reinterpret_cast<FusedCtl*>(smem_raw + sizeof(WaveCtl));
    int* ys = reinterpret_cast<int*>(smem_raw + sizeof(WaveCtl) + sizeof(FusedCtl));
    uint32_t* bitsT = reinterpret_cast<uint32_t*>(ys + ROWW + 4);
    float* tiles = reinterpret_cast<float*>(
        (reinterpret_cast<uintptr_t>(bitsT + (size_t)NYB * ROWW) + 127) & ~static_cast<uintptr_t>(127));
    float* ztiles = tiles + (size_t)S * stage_floats;                  // stage_floats * 4 is a multiple of 128

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b = blockIdx.x;
    int t_x = min(max(p.t_xs[b], 0), p.T_x);
    int t_y = min(max(p.t_ys[b], 0), p.T_y);
    const bool valid = (t_x >= 1) && (t_y >= t_x);
    const int n_tiles = valid ? (t_y + YT - 1) / YT : 0;
    const int n_iter = max(n_tiles, 1);
    const int n_act = valid ? (t_x + BAND - 1) / BAND : 0;
    const size_t plane_elems = (size_t)p.T_x * p.T_y;
    float* __restrict__ pplane = p.path ? p.path + (size_t)b * plane_elems : nullptr;   // null: index + durations only
    const size_t fill_elems = p.path ? plane_elems : 0;

    if (tid == 0) {
        for (int w = 0; w < NDP; ++w) {
            for (int s = 0; s < TMA_MAX_STAGES; ++s) {
                mbar_init(&ctl->full[w][s], F_NMMA / NDP);     // the two MMA warps of the band
                mbar_init(&ctl->empty[w][s], 1);
            }
            for (int s = 0; s < NBOX; ++s) {
                mbar_init(&ctl->box_full[w][s], 1);
                mbar_init(&ctl->box_empty[w][s], 1);
            }
        }
        mbar_init(&ctl->dp_done, max(n_act, 1));
        for (int s = 0; s < FZ_STAGES; ++s) {
            mbar_init(&fz->zfull[s], 1);
            mbar_init(&fz->zempty[s], max(2 * n_act, 1));
        }
        mbar_fence_init();
    }
    __syncthreads();

    if (warp < NDP) {
        // ------------------------------------------------------------ DP warps (same code as the TMA wavefront kernel)
        const int w = warp;
        const long long waited = wave_dp_warp<XPT>(ctl, tiles, stage_floats, bitsT, ROWW, RB, S, n_tiles, n_act, w, lane);
        if (w == 0 && lane == 0) wave_backtrack(ctl, bitsT, ROWW, ys, valid, t_x, t_y, waited);
    } else {
        // ------------------------------------------------------------ MMA warps: value tiles from x_m and z
        const int mw = warp - NDP, w = mw >> 1, h = mw & 1;
        const int g = lane >> 2, t = lane & 3;            // mma.sync fragment coordinates
        const bool act = w < n_act;
        const int rbase = w * BAND + h * (16 * MB);       // first text row of this warp
        const int ht = tid - 32 * NDP;
        constexpr int HT = 32 * F_NMMA;
        const uintptr_t addr = reinterpret_cast<uintptr_t>(pplane);
        size_t head = ((16 - (addr & 15)) & 15) / 4;
        if (head > fill_elems) head = fill_elems;
        const size_t nvec = (fill_elems - head) / 4;
        const size_t tail0 = head + nvec * 4;
        float4* pv = reinterpret_cast<float4*>(pplane + head);
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);

        // A fragments (m16n8k8, row-major): a0 (g, t), a1 (g + 8, t), a2 (g, t + 4), a3 (g + 8, t + 4) of each 16 x 8 block;
        // rows = text positions, columns = mel channels.  Rows at or beyond t_x are zero (never consumed by a valid row).
        float a[MB][FD / 8][4];
        float crow[MB][2];
#pragma unroll
        for (int mb = 0; mb < MB; ++mb) {
            const int r_lo = rbase + 16 * mb + g, r_hi = r_lo + 8;
            const bool ok_lo = act && r_lo < t_x, ok_hi = act && r_hi < t_x;
            const float* xb = x_m + (size_t)b * FD * p.T_x;
            float s_lo = 0.f, s_hi = 0.f;
#pragma unroll
            for (int ks = 0; ks < FD / 8; ++ks) {
                const size_t d0 = (size_t)(8 * ks + t) * p.T_x, d1 = d0 + (size_t)4 * p.T_x;
                const float a0 = ok_lo ? __ldg(xb + d0 + r_lo) : 0.f;
                const float a1 = ok_hi ? __ldg(xb + d0 + r_hi) : 0.f;
                const float a2 = ok_lo ? __ldg(xb + d1 + r_lo) : 0.f;
                const float a3 = ok_hi ? __ldg(xb + d1 + r_hi) : 0.f;
                a[mb][ks][0] = a0; a[mb][ks][1] = a1; a[mb][ks][2] = a2; a[mb][ks][3] = a3;
                s_lo = fmaf(a0, a0, s_lo); s_lo = fmaf(a2, a2, s_lo);
                s_hi = fmaf(a1, a1, s_hi); s_hi = fmaf(a3, a3, s_hi);
            }
            s_lo += __shfl_xor_sync(0xffffffffu, s_lo, 1); s_lo += __shfl_xor_sync(0xffffffffu, s_lo, 2);
            s_hi += __shfl_xor_sync(0xffffffffu, s_hi, 1); s_hi += __shfl_xor_sync(0xffffffffu, s_hi, 2);
            crow[mb][0] = -0.5f * s_lo - (float)FD * 0.918938533204672742f;      // logp1 + logp4 (models.py:1078,1081)
            crow[mb][1] = -0.5f * s_hi - (float)FD * 0.918938533204672742f;
        }

        int z_next = 0;                                   // (MMA warp 0, lane 0) next z tile to request
        for (int it = 0; it < n_iter; ++it) {
            if (act && it < n_tiles) {
                if (mw == 0 && lane == 0) {
                    // keep up to FZ_STAGES z tiles requested; block only for the tile needed now
                    while (z_next < n_tiles && z_next < it + FZ_STAGES) {
                        const int zs = z_next % FZ_STAGES, zu = z_next / FZ_STAGES;
                        if (zu >= 1) {
                            if (z_next > it) { if (!mbar_test_wait(&fz->zempty[zs], (zu - 1) & 1)) break; }
                            else mbar_wait(&fz->zempty[zs], (zu - 1) & 1);
                        }
                        mbar_arrive_expect_tx(&fz->zfull[zs], (uint32_t)FZ_BYTES);
                        tc::tma_load_3d(ztiles + (size_t)zs * FD * ZP, &tmZ, &fz->zfull[zs], z_next * YT, 0, b);
                        ++z_next;
                    }
                }
                __syncwarp();
                const int s = it % S, use = it / S, zs = it % FZ_STAGES, zu = it / FZ_STAGES;
                if (use >= 1) fast_wait(&ctl->empty[w][s], (use - 1) & 1);
                fast_wait(&fz->zfull[zs], zu & 1);
                const uint32_t zb = smem_u32(ztiles + (size_t)zs * FD * ZP);

                float acc[MB][4][4];
#pragma unroll
                for (int mb = 0; mb < MB; ++mb)
#pragma unroll
                    for (int nb = 0; nb < 4; ++nb)
#pragma unroll
                        for (int i = 0; i < 4; ++i) acc[mb][nb][i] = 0.f;
#pragma unroll
                for (int ks = 0; ks < FD / 8; ++ks) {
                    // B fragments (k x n = channel x frame, "col"): b0 (k = t, n = g), b1 (k = t + 4, n = g)
                    uint32_t bh[4][2], bl[4][2];
#pragma unroll
                    for (int nb = 0; nb < 4; ++nb) {
                        const float v0 = lds32(zb + 4u * ((8 * ks + t) * ZP + 8 * nb + g));
                        const float v1 = lds32(zb + 4u * ((8 * ks + t + 4) * ZP + 8 * nb + g));
                        bh[nb][0] = tf32_hi(v0); bl[nb][0] = __float_as_uint(v0 - __uint_as_float(bh[nb][0]));
                        bh[nb][1] = tf32_hi(v1); bl[nb][1] = __float_as_uint(v1 - __uint_as_float(bh[nb][1]));
                    }
                    uint32_t ah[MB][4], al[MB][4];
#pragma unroll
                    for (int mb = 0; mb < MB; ++mb)
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            ah[mb][i] = tf32_hi_pinned(a[mb][ks][i]);
                            al[mb][i] = __float_as_uint(a[mb][ks][i] - __uint_as_float(ah[mb][i]));
                        }
                    // three passes over the 4 * MB independent accumulators (small terms first): the MMAs that depend on each
                    // other are 4 * MB instructions apart instead of back to back
#pragma unroll
                    for (int mb = 0; mb < MB; ++mb)
#pragma unroll
                        for (int nb = 0; nb < 4; ++nb) mma_tf32(acc[mb][nb], al[mb], bh[nb][0], bh[nb][1]);
#pragma unroll
                    for (int mb = 0; mb < MB; ++mb)
#pragma unroll
                        for (int nb = 0; nb < 4; ++nb) mma_tf32(acc[mb][nb], ah[mb], bl[nb][0], bl[nb][1]);
#pragma unroll
                    for (int mb = 0; mb < MB; ++mb)
#pragma unroll
                        for (int nb = 0; nb < 4; ++nb) mma_tf32(acc[mb][nb], ah[mb], bh[nb][0], bh[nb][1]);
                }
                // -0.5 sum_d z^2 per frame: only where logp itself is emitted (see the header comment)
                float zq[4][2];
#pragma unroll
                for (int nb = 0; nb < 4; ++nb) zq[nb][0] = zq[nb][1] = 0.f;
                if (logp_out) {
                    float sq = 0.f;
                    for (int d = 0; d < FD; ++d) {
                        const float v = lds32(zb + 4u * (d * ZP + lane));
                        sq = fmaf(v, v, sq);
                    }
                    sq *= -0.5f;
#pragma unroll
                    for (int nb = 0; nb < 4; ++nb) {
                        zq[nb][0] = __shfl_sync(0xffffffffu, sq, 8 * nb + 2 * t);
                        zq[nb][1] = __shfl_sync(0xffffffffu, sq, 8 * nb + 2 * t + 1);
                    }
                }
                // D fragments: c0 (g, 2t), c1 (g, 2t + 1), c2 (g + 8, 2t), c3 (g + 8, 2t + 1) -> tile[row][36-pitch]
                const uint32_t tile = smem_u32(tiles + (size_t)s * stage_floats);
                const int y0 = it * YT;
#pragma unroll
                for (int mb = 0; mb < MB; ++mb) {
#pragma unroll
                    for (int half = 0; half < 2; ++half) {
                        const int x = rbase + 16 * mb + g + 8 * half;
                        if (x >= RB * TBOX) continue;              // row beyond the stage (band rounding): never consumed
#pragma unroll
                        for (int nb = 0; nb < 4; ++nb) {
                            const int col = 8 * nb + 2 * t;
                            const float v0 = acc[mb][nb][2 * half] + crow[mb][half];
                            const float v1 = acc[mb][nb][2 * half + 1] + crow[mb][half];
                            asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(tile + 4u * (x * TP + col)), "f"(v0), "f"(v1) : "memory");
                            if (logp_out && x < t_x) {
                                float* lp = logp_out + ((size_t)b * p.T_x + x) * p.T_y + y0 + col;
                                if (y0 + col < t_y) lp[0] = v0 + zq[nb][0];
                                if (y0 + col + 1 < t_y) lp[1] = v1 + zq[nb][1];
                            }
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(&ctl->full[w][s]);
                    mbar_arrive(&fz->zempty[zs]);
                }
            }
            // zero-fill share `it` of this utterance's output plane
            const size_t lo = nvec * (size_t)it / n_iter, hi = nvec * (size_t)(it + 1) / n_iter;
            for (size_t i = lo + ht; i < hi; i += HT) pv[i] = z4;
            if (it == 0) {
                if ((size_t)ht < head) pplane[ht] = 0.f;
                for (size_t i = tail0 + ht; i < fill_elems; i += HT) pplane[i] = 0.f;
            }
        }
    }
    __syncthreads();
    if (tid == 0) mas_stamp(4);
    emit_path(p, pplane, ys, valid, t_x, t_y, b, tid, F_NT);
    if (tid == 0) mas_stamp(5);
}

#ifdef GTK_INSTRUMENT
int trace_read(unsigned long long* out, int n_ctas) {
    GTK_REQUIRE(out && n_ctas >= 0 && n_ctas <= MAS_TRACE_CTAS, "n_ctas");
    GTK_CUDA(cudaMemcpyFromSymbol(out, g_mas_trace, size_t(n_ctas) * 6 * sizeof(unsigned long long)));
    return GTK_OK;
}
int wait_read(long long* out, int n_ctas) {
    GTK_REQUIRE(out && n_ctas >= 0 && n_ctas <= MAS_TRACE_CTAS, "n_ctas");
    GTK_CUDA(cudaMemcpyFromSymbol(out, g_mas_wait, size_t(n_ctas) * sizeof(long long)));
    return GTK_OK;
}
#endif

// -------------------------------------------------------------------------------------------------
// Generic fallback for shapes the fast path does not cover (T_x > 512, or decision bits that do not
// fit in shared memory): DP column in shared memory, block barrier per frame, decision words in a
// global workspace [B][T_y][ceil(T_x/32)].  Same arithmetic, same bits.
constexpr int GEN_NT = 1024;

__global__ void __launch_bounds__(GEN_NT, 1) mas_generic_kernel(const Params p, uint32_t* __restrict__ ws) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* qa = reinterpret_cast<float*>(smem_raw);
    float* qb = qa + p.T_x;
    const int tid = threadIdx.x, lane = tid & 31;
    const int b = blockIdx.x;
    const int W = (p.T_x + 31) / 32;
    int t_x = min(max(p.t_xs[b], 0), p.T_x);
    int t_y = min(max(p.t_ys[b], 0), p.T_y);
    const bool valid = (t_x >= 1) && (t_y >= t_x);
    const size_t plane_elems = (size_t)p.T_x * p.T_y;
    const float* __restrict__ vplane = p.value + (size_t)b * plane_elems;
    float* __restrict__ pplane = p.path ? p.path + (size_t)b * plane_elems : nullptr;   // null: index + durations only
    const size_t fill_elems = p.path ? plane_elems : 0;
    uint32_t* wsb = ws + (size_t)b * p.T_y * W;

    for (size_t i = tid; i < fill_elems; i += GEN_NT) pplane[i] = 0.f;
    if (p.dur_out) for (int x = tid; x < p.T_x; x += GEN_NT) p.dur_out[(size_t)b * p.T_x + x] = 0;
    if (p.idx_out) for (int y = tid; y < p.T_y; y += GEN_NT) p.idx_out[(size_t)b * p.T_y + y] = -1;
    for (int x = tid; x < p.T_x; x += GEN_NT) qa[x] = 0.f;
    __syncthreads();
    if (!valid) return;

    for (int y = 0; y < t_y; ++y) {
        for (int xb = 0; xb < t_x; xb += GEN_NT) {
            const int x = xb + tid;
            bool d = false;
            if (x < t_x) {
                const float a = qa[x];
                const float bq = (x == 0) ? ((y == 0) ? 0.f : NEG) : qa[x - 1];
                const bool diag = (x == y);
                d = (x > 0) && (diag || a < bq);
                const float a2 = diag ? NEG : a;
                const float m = (bq > a2) ? bq : a2;
                qb[x] = m + vplane[(size_t)x * p.T_y + y];
            }
            const uint32_t word = __ballot_sync(0xffffffffu, d);
            if (lane == 0 && (x >> 5) < W) wsb[(size_t)y * W + (x >> 5)] = word;
        }
        __syncthreads();
        float* t = qa; qa = qb; qb = t;
    }
    if (tid == 0) {
        int index = t_x - 1, cnt = 0;
        for (int y = t_y - 1; y >= 0; --y) {
            if (pplane) pplane[(size_t)index * p.T_y + y] = 1.0f;
            if (p.idx_out) p.idx_out[(size_t)b * p.T_y + y] = index;
            ++cnt;
            if (index != 0 && ((wsb[(size_t)y * W + (index >> 5)] >> (index & 31)) & 1u)) {
                if (p.dur_out) p.dur_out[(size_t)b * p.T_x + index] = cnt;
                cnt = 0;
                --index;
            }
        }
        if (p.dur_out) p.dur_out[(size_t)b * p.T_x + index] = cnt;
    }
}

// monotonic_align/__init__.py:18-19: t_x = mask.sum(1)[:,0], t_y = mask.sum(2)[:,0], on the device.
__global__ void mas_lengths_kernel(const float* __restrict__ mask, int* __restrict__ t_xs, int* __restrict__ t_ys,
                                   int T_x, int T_y) {
    __shared__ float scratch[33];
    const int b = blockIdx.x;
    const float* m = mask + (size_t)b * T_x * T_y;
    float sx = 0.f, sy = 0.f;
    for (int x = threadIdx.x; x < T_x; x += 256) sx += m[(size_t)x * T_y];
    for (int y = threadIdx.x; y < T_y; y += 256) sy += m[y];
    sx = block_sum<256>(sx, scratch);
    sy = block_sum<256>(sy, scratch);
    if (threadIdx.x == 0) {
        t_xs[b] = (int)sx;
        t_ys[b] = (int)sy;
    }
}

// ------------------------------------------------------------------------------------------ host side
static size_t fast_smem_bytes(int xpt, int T_y, int stages) {
    const size_t roww = 32 * (size_t)xpt;
    const size_t ys_words = ((roww + 1 + 3) / 4) * 4;
    const size_t nyb = (T_y + YT - 1) / YT;
    return 2 * MAX_STAGES * sizeof(uint64_t) + ys_words * 4 + nyb * roww * 4 + (size_t)stages * YT * (roww + 1) * 4;
}

static const size_t SMEM_LIMIT = 227 * 1024;

// Picks the fast path configuration; returns stages (0 => use the generic kernel).
static int plan_fast(int T_x, int T_y, int* xpt_out) {
    const int xpt = max(1, ceil_div(T_x, 32));
    *xpt_out = xpt;
    if (xpt > MAX_XPT) return 0;
    for (int s = MAX_STAGES; s >= 2; --s)
        if (fast_smem_bytes(xpt, T_y, s) <= SMEM_LIMIT) return s;
    return 0;
}

static size_t wave_bits_bytes(int B, int T_x, int T_y) {
    return (size_t)B * ceil_div(T_y, YT) * ((T_x + 3) & ~3) * sizeof(uint32_t);
}
static int plan_wave(int T_x, int T_y, int* xpt_out, int* bits_global);
size_t workspace_bytes(int B, int T_x, int T_y) {
    int xpt, bg = 1;
    const size_t wave = (T_y % 4 == 0 && plan_wave(T_x, T_y, &xpt, &bg) > 0 && bg) ? wave_bits_bytes(B, T_x, T_y) : 0;
    if (plan_fast(T_x, T_y, &xpt) > 0) return wave;
    return max(wave, (size_t)B * T_y * ceil_div(T_x, 32) * sizeof(uint32_t));
}

// TMA wavefront path: text rows per lane of each of the NDP DP warps.
static size_t wave_smem_bytes(int T_x, int T_y, int stages, int bits_global = 0) {
    const size_t roww = (size_t)((T_x + 3) & ~3);
    const size_t nyb = (T_y + YT - 1) / YT;
    const size_t rb = ceil_div(T_x, TBOX);
    return sizeof(WaveCtl) + (roww + 4) * 4 + (bits_global ? 0 : nyb * roww * 4) + 128 + (size_t)stages * rb * TBOX_BYTES;
}
// *bits_global: in = the caller can provide the global decision-word workspace; out = the plan uses it (when the bits in
// shared memory would leave fewer than 3 value stages and moving them out buys at least one more)
static int plan_wave(int T_x, int T_y, int* xpt_out, int* bits_global) {
    int xpt = max(1, ceil_div(T_x, 32 * NDP));
    {   // experiment switch: at least this many text rows per lane (fewer, longer DP warps)
        static const int min_xpt = getenv("GTK_MAS_MIN_XPT") ? atoi(getenv("GTK_MAS_MIN_XPT")) : 0;
        if (min_xpt > xpt) xpt = min(min_xpt, TMA_MAX_XPT);
    }
    *xpt_out = xpt;
    const int can_global = *bits_global;
    *bits_global = 0;
    if (xpt > TMA_MAX_XPT || T_y % 4 != 0) return 0;
    int s_smem = 0, s_glob = 0;
    for (int s = TMA_MAX_STAGES; s >= 2 && !s_smem; --s)
        if (wave_smem_bytes(T_x, T_y, s) <= SMEM_LIMIT) s_smem = s;
    // opt-in (GTK_MAS_GLOBAL_BITS=1).  Measured on B200, B = 512 x 400 x 2000: three stages with the bits in L2 run 844 us
    // against 781 us with two stages and the bits in shared memory -- the single-thread backtrack pays L2 latency per batch
    // of rows and the extra stage buys nothing: the kernel is not bound by the bytes in flight.
    static const bool glob_on = getenv("GTK_MAS_GLOBAL_BITS") && atoi(getenv("GTK_MAS_GLOBAL_BITS")) != 0;
    if (can_global && glob_on && s_smem < 3) {
        for (int s = TMA_MAX_STAGES; s >= 2 && !s_glob; --s)
            if (wave_smem_bytes(T_x, T_y, s, 1) <= SMEM_LIMIT) s_glob = s;
        if (s_glob > s_smem) { *bits_global = 1; return s_glob; }
    }
    return s_smem;
}

// Fused logp + MAS path: smem = wavefront layout + the z ring.
static size_t fused_smem_bytes(int T_x, int T_y, int stages) {
    const size_t roww = (size_t)((T_x + 3) & ~3);
    const size_t nyb = (T_y + YT - 1) / YT;
    const size_t rb = ceil_div(T_x, TBOX);
    return sizeof(WaveCtl) + sizeof(FusedCtl) + (roww + 4) * 4 + nyb * roww * 4 + 128 + (size_t)stages * rb * TBOX_BYTES +
           (size_t)FZ_STAGES * FZ_BYTES;
}
static int plan_fused(int D, int T_x, int T_y, int* xpt_out) {
    const int xpt = max(1, ceil_div(T_x, 32 * NDP));
    *xpt_out = xpt;
    if (D != FD || T_x < 1 || T_y < 1 || T_y % 4 != 0 || xpt > F_MAX_XPT) return 0;
    for (int s = TMA_MAX_STAGES; s >= 2; --s)
        if (fused_smem_bytes(T_x, T_y, s) <= SMEM_LIMIT) return s;
    return 0;
}
int fused_supported(int D, int T_x, int T_y, int has_logs) {
    int xpt;
    return (!has_logs && plan_fused(D, T_x, T_y, &xpt) > 0) ? 1 : 0;
}

template <int XPT>
static int launch_fused(const Params& p, const CUtensorMap& tm, const float* x_m, float* logp_out, size_t smem, cudaStream_t stream) {
    GTK_CUDA(cudaFuncSetAttribute(mas_fused_kernel<XPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ProfScope prof("mas_fused", stream);
    mas_fused_kernel<XPT><<<p.B, F_NT, smem, stream>>>(tm, p, x_m, logp_out);
    return check_launch("mas_fused_kernel");
}

int fused_forward(const float* z, const float* x_m, const int* t_xs, const int* t_ys, float* path, int* dur_out, int* idx_out,
                  float* logp_out, int B, int D, int T_x, int T_y, cudaStream_t stream) {
    GTK_REQUIRE(B >= 0 && T_x >= 0 && T_y >= 0 && D >= 0, "negative size");
    if (B == 0) return GTK_OK;
    GTK_REQUIRE(z && x_m && t_xs && t_ys && path, "null pointer");
    int xpt = 0;
    const int stages = plan_fused(D, T_x, T_y, &xpt);
    if (stages == 0 || (reinterpret_cast<uintptr_t>(z) & 15) != 0) {
        set_error("mas_fused_logp: unsupported shape D=%d T_x=%d T_y=%d (needs D == %d, T_x <= %d, T_y %% 4 == 0, 16-byte aligned z); "
                  "use gtk_logp + gtk_mas_forward", D, T_x, T_y, FD, 32 * NDP * F_MAX_XPT);
        return GTK_ERR_UNSUPPORTED;
    }
    Params p;
    p.value = nullptr; p.path = path; p.t_xs = t_xs; p.t_ys = t_ys; p.dur_out = dur_out; p.idx_out = idx_out;
    p.B = B; p.T_x = T_x; p.T_y = T_y; p.n_stages = stages; p.vec4 = 1; p.bits_ws = nullptr;
    CUtensorMap tm;
    const uint64_t dims[3] = {(uint64_t)T_y, (uint64_t)D, (uint64_t)B};
    const uint64_t strides[2] = {(uint64_t)T_y * 4, (uint64_t)T_y * 4 * D};
    const uint32_t box[3] = {ZP, FD, 1};
    int rc = tc::make_tmap_f32(&tm, z, 3, dims, strides, box);
    if (rc) return rc;
    const size_t smem = fused_smem_bytes(T_x, T_y, stages);
    if (xpt == 1) return launch_fused<1>(p, tm, x_m, logp_out, smem, stream);
    return launch_fused<2>(p, tm, x_m, logp_out, smem, stream);
}

template <int XPT>
static int launch_wave(const Params& p, const CUtensorMap& tm, size_t smem, cudaStream_t stream) {
    GTK_CUDA(cudaFuncSetAttribute(mas_wave_kernel<XPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ProfScope prof("mas", stream);
    mas_wave_kernel<XPT><<<p.B, NT, smem, stream>>>(tm, p);
    return check_launch("mas_wave_kernel");
}

template <int XPT>
static int launch_fast(const Params& p, size_t smem, cudaStream_t stream) {
    GTK_CUDA(cudaFuncSetAttribute(mas_kernel<XPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ProfScope prof("mas", stream);
    mas_kernel<XPT><<<p.B, NT, smem, stream>>>(p);
    return check_launch("mas_kernel");
}

int forward(const float* value, const int* t_xs, const int* t_ys, float* path, int* dur_out, int* idx_out, int B,
            int T_x, int T_y, void* workspace, size_t workspace_bytes_given, int force_generic, cudaStream_t stream) {
    GTK_REQUIRE(B >= 0 && T_x >= 0 && T_y >= 0, "negative size");
    if (B == 0) return GTK_OK;
    GTK_REQUIRE(t_xs && t_ys, "null pointer");
    GTK_REQUIRE(value || T_x == 0 || T_y == 0, "null pointer");
    GTK_REQUIRE(path || dur_out || idx_out, "no output requested");
    if (T_x == 0 || T_y == 0) {
        if (dur_out && B * T_x > 0) GTK_CUDA(cudaMemsetAsync(dur_out, 0, sizeof(int) * (size_t)B * T_x, stream));
        if (idx_out && B * T_y > 0) GTK_CUDA(cudaMemsetAsync(idx_out, 0xff, sizeof(int) * (size_t)B * T_y, stream));
        return GTK_OK;
    }
    Params p;
    p.value = value; p.path = path; p.t_xs = t_xs; p.t_ys = t_ys; p.dur_out = dur_out; p.idx_out = idx_out;
    p.B = B; p.T_x = T_x; p.T_y = T_y;
    p.vec4 = ((T_y % 4) == 0) && ((reinterpret_cast<uintptr_t>(value) & 15) == 0);
    int xpt = 0;
    // force_generic: 0 = pick the fastest applicable path, 1 = generic kernel, 2 = helper-warp fast path (no TMA)
    int bits_global = (workspace && workspace_bytes_given >= wave_bits_bytes(B, T_x, T_y)) ? 1 : 0;
    const int tma_stages = (force_generic == 0 && p.vec4) ? plan_wave(T_x, T_y, &xpt, &bits_global) : 0;
    p.bits_ws = nullptr;
    if (tma_stages > 0) {
        p.n_stages = tma_stages;
        if (bits_global) p.bits_ws = static_cast<uint32_t*>(workspace);
        CUtensorMap tm;
        const uint64_t dims[3] = {(uint64_t)T_y, (uint64_t)T_x, (uint64_t)B};
        const uint64_t strides[2] = {(uint64_t)T_y * 4, (uint64_t)T_y * 4 * T_x};
        const uint32_t box[3] = {TP, TBOX, 1};
        int rc = tc::make_tmap_f32(&tm, value, 3, dims, strides, box);
        if (rc) return rc;
        const size_t smem = wave_smem_bytes(T_x, T_y, tma_stages, bits_global);
        switch (xpt) {
#define GTK_MAS_CASE(N) case N: return launch_wave<N>(p, tm, smem, stream);
            GTK_MAS_CASE(1) GTK_MAS_CASE(2) GTK_MAS_CASE(3) GTK_MAS_CASE(4) GTK_MAS_CASE(5) GTK_MAS_CASE(6)
            GTK_MAS_CASE(7) GTK_MAS_CASE(8)
#undef GTK_MAS_CASE
            default: break;
        }
        set_error("mas: internal: no wavefront instantiation for XPT=%d", xpt);
        return GTK_ERR_UNSUPPORTED;
    }
    const int stages = force_generic == 1 ? 0 : plan_fast(T_x, T_y, &xpt);
    p.n_stages = stages;
    if (stages > 0) {
        const size_t smem = fast_smem_bytes(xpt, T_y, stages);
        switch (xpt) {
#define GTK_MAS_CASE(N) case N: return launch_fast<N>(p, smem, stream);
            GTK_MAS_CASE(1) GTK_MAS_CASE(2) GTK_MAS_CASE(3) GTK_MAS_CASE(4) GTK_MAS_CASE(5) GTK_MAS_CASE(6)
            GTK_MAS_CASE(7) GTK_MAS_CASE(8) GTK_MAS_CASE(9) GTK_MAS_CASE(10) GTK_MAS_CASE(11) GTK_MAS_CASE(12)
            GTK_MAS_CASE(13) GTK_MAS_CASE(14) GTK_MAS_CASE(15) GTK_MAS_CASE(16)
#undef GTK_MAS_CASE
            default: break;
        }
        set_error("mas: internal: no instantiation for XPT=%d", xpt);
        return GTK_ERR_UNSUPPORTED;
    }
    // generic path
    const size_t need = (size_t)B * T_y * ceil_div(T_x, 32) * sizeof(uint32_t);
    if (!workspace || workspace_bytes_given < need) {
        set_error("mas: generic path needs %zu workspace bytes, got %zu", need, workspace_bytes_given);
        return GTK_ERR_WORKSPACE;
    }
    const size_t smem = 2 * (size_t)T_x * sizeof(float);
    if (smem > SMEM_LIMIT) {
        set_error("mas: T_x=%d exceeds the generic kernel's shared-memory column buffer", T_x);
        return GTK_ERR_UNSUPPORTED;
    }
    GTK_CUDA(cudaFuncSetAttribute(mas_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ProfScope prof("mas_generic", stream);
    mas_generic_kernel<<<B, GEN_NT, smem, stream>>>(p, static_cast<uint32_t*>(workspace));
    return check_launch("mas_generic_kernel");
}

// ------------------------------------------------------------------------------------------ allocation-free variant
// gtk_mas_forward_static: the generic kernel (the one MAS kernel that needs scratch memory: its decision words live in
// global memory) with the scratch in a file-scope device array instead of a caller-provided workspace, for callers that
// cannot allocate (CUDA-graph capture, serving loops, harnesses that forbid cudaMalloc).  Every GPU that loads the library
// holds MAS_STATIC_SCRATCH_BYTES until exit.  All calls share the array: calls that overlap on the GPU (different
// streams) race.  Nothing in it has to start at zero -- the kernel writes every decision word before the backtrack reads
// it -- so stale contents of an earlier call are harmless.
constexpr size_t MAS_STATIC_SCRATCH_BYTES = 8u << 20;           // 8 MiB = B * T_y * ceil(T_x / 32) * 4 bytes at most
static __device__ __align__(256) unsigned char g_mas_static_scratch[MAS_STATIC_SCRATCH_BYTES];

size_t static_scratch_bytes() { return MAS_STATIC_SCRATCH_BYTES; }

// Device address of the scratch array on the CURRENT device.  In host code `g_mas_static_scratch` is only a stand-in
// symbol, not a device address: it must go through cudaGetSymbolAddress, once per device, with that device current.
static int static_scratch_ptr(void** out) {
    constexpr int MAX_DEV = 64;
    static void* ptrs[MAX_DEV] = {nullptr};
    static std::mutex mu;
    int dev = 0;
    GTK_CUDA(cudaGetDevice(&dev));
    GTK_REQUIRE(dev >= 0 && dev < MAX_DEV, "device ordinal out of range");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        void* p = nullptr;
        GTK_CUDA(cudaGetSymbolAddress(&p, g_mas_static_scratch));
        GTK_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned null");
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return GTK_OK;
}

int forward_static(const float* value, const int* t_xs, const int* t_ys, float* path, int* dur_out, int* idx_out, int B,
                   int T_x, int T_y, cudaStream_t stream) {
    GTK_REQUIRE(B >= 0 && T_x >= 0 && T_y >= 0, "negative size");
    if (B == 0) return GTK_OK;
    const size_t need = (size_t)B * T_y * ceil_div(T_x, 32) * sizeof(uint32_t);
    if (need > MAS_STATIC_SCRATCH_BYTES) {
        set_error("mas_forward_static: B * T_y * ceil(T_x / 32) * 4 = %zu bytes exceeds the built-in scratch of %zu bytes; "
                  "use gtk_mas_forward with a caller-provided workspace", need, MAS_STATIC_SCRATCH_BYTES);
        return GTK_ERR_WORKSPACE;
    }
    void* scratch = nullptr;
    int rc = static_scratch_ptr(&scratch);
    if (rc) return rc;
    return forward(value, t_xs, t_ys, path, dur_out, idx_out, B, T_x, T_y, scratch, MAS_STATIC_SCRATCH_BYTES, /*generic*/ 1, stream);
}

int lengths_from_mask(const float* mask, int* t_xs, int* t_ys, int B, int T_x, int T_y, cudaStream_t stream) {
    GTK_REQUIRE(mask && t_xs && t_ys, "null pointer");
    if (B == 0) return GTK_OK;
    if (T_x == 0 || T_y == 0) {
        GTK_CUDA(cudaMemsetAsync(t_xs, 0, sizeof(int) * B, stream));
        GTK_CUDA(cudaMemsetAsync(t_ys, 0, sizeof(int) * B, stream));
        return GTK_OK;
    }
    mas_lengths_kernel<<<B, 256, 0, stream>>>(mask, t_xs, t_ys, T_x, T_y);
    return check_launch("mas_lengths_kernel");
}

}  // namespace mas
}  // namespace gtk
