// glowtts_b200/csrc/tc.cuh -- sm_100a tensor-core plumbing: TMA (cp.async.bulk.tensor), tcgen05.mma with
// TMEM accumulators, UMMA shared-memory / instruction descriptors.  Raw PTX, no CUTLASS dependency.
#pragma once
#include <cuda.h>   // CUtensorMap (types only; the encode entry point is fetched at run time)
#include "common.cuh"

namespace gtk {
namespace tc {

// ---------------------------------------------------------------------------------- host: tensor maps
// bf16 tensor viewed as up to 3 dims (innermost first).  box = tile written by one TMA load; swizzle 128B
// requires box[0] * 2 bytes == 128.
int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                   const uint32_t* box, bool swizzle128);
// fp32 tensor, no swizzle (box[0] * 4 bytes must be a multiple of 16)
int make_tmap_f32(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                  const uint32_t* box);

constexpr int BT = 128;            // frames per CTA tile (UMMA M)
constexpr int KC = 64;             // channels per k-chunk (128 B of bf16 = one swizzle row)
constexpr int A_BYTES = BT * KC * 2;
constexpr int NTHREADS = 256;          // weight-gradient kernels: 2 issue warps + 4 drain warps
constexpr int EPI_NT = 512;            // conv / fused-layer kernels: every warp drains TMEM (16 warps = 4 lane quarters x 4 column parts)
constexpr int EPI_PARTS = EPI_NT / 128;

struct ConvTcParams {
    int B, T, taps, dil, n_k;      // n_k = padded input channels / 64
    int tile_layout;               // 1: out_f32 / resid_f32 / ts use the tile-private layout (RowIo); ld = column count
    int N, NW, NCH;                // N = NW * NCH output columns, NW % 16 == 0, NW <= 256
    int stages;
    int cluster;                   // 1 or 2 CTAs per cluster (2: the weight tile is fetched once per pair, TMA multicast)
    int trace;                     // 1: record per-CTA phase timestamps (conv_trace_select)
    int ctl_off;                   // byte offset of the control block behind the operand stages / output staging area
    // ---- varlen packing (pack.cu; all null for plain [B, T] tensors)
    const int* n_rows;             // device: rows in use; tiles at or behind it exit at once
    const int* row_b;              // device [T]: utterance of a packed row (per-utterance logdet / dlogdet)
    const int* row_fr;             // device [T]: un-packed frame index b * T' + t of a packed row (dropout counter)
    // ---- plain epilogue: v = acc + bias[n] (+ resid) ; optional mask ; stores
    const float* bias;             // [N] or null
    const float* mask;             // [B, T] or null
    int mask_out, mask_resid;
    const float* resid_f32;        // frame-major [B, T, ld_resid] or null
    int ld_resid;
    float* out_f32;                // frame-major [B, T, ld_f32] or null
    int ld_f32;
    __nv_bfloat16* out_bf16;       // frame-major [B, T, ld_bf16] or null (columns N..ld_bf16 are left untouched)
    int ld_bf16;
    int rep_bf16;                  // > 1: the bf16 output is also written to rep_bf16 - 1 more buffers,
    size_t rep_stride;             //      rep_stride elements apart (plain epilogue only)
    float* out_tm;                 // t-major [B, tm_C, T]: channel tm_c0 + n, or null
    const float* resid_tm;         // t-major residual added before the store (same indexing), or null
    int tm_C, tm_c0;
    // ---- EPI_GATE_BWD (N = H): acc = d(acts); ts = saved [tanh | sigmoid] frames [B,T,2H] ->
    //      out_bf16[.., 0:2H] = d(conv output) (dropout-scaled), out2_bf16 = d(gate input) (pre-dropout; may be null)
    const __nv_bfloat16* ts;
    __nv_bfloat16* out2_bf16;
    float drop_p;                  // dropout probability on the conv output (modules.py:153); 0 = off
    DropSeed drop_seed;            // layer seed (or its salt + the device address of the call seed), common.cuh
    // ---- EPI_AFFINE (N = C, attentions.py:172-183): acc = [m | logs]; x / z t-major [B, C, T]
    const float* aff_x;
    float* aff_z;
    float* aff_logdet;             // [B], accumulated with atomicAdd (forward only)
    float* aff_logs_save;          // frame-major [B, T, C/2] or null: logs kept for the backward
    int aff_reverse, aff_sigmoid_scale;
};

enum { EPI_PLAIN = 0, EPI_GATE_BWD = 1, EPI_AFFINE = 2 };

// One fused WaveNet layer (wn_tc.cu).  All activations frame-major [B, T, H].
struct WnLayerParams {
    int B, T, H, RS, taps, dil;
    int tile_layout;               // 1: h_f32 / h_out_f32 / skip_f32 / ts_save use the tile-private layout (RowIo)
    int first, last;               // first: skip is written, not accumulated; last: RS == H, skip*mask -> skip_out_bf16
    const float* bias_in;          // [2H]
    const float* bias_rs;          // [RS]
    const float* cond;             // this layer's slice of cond_layer(g): cond[b * cond_bstride + n], n < 2H; or null
    size_t cond_bstride;
    const int* n_rows;             // varlen packing (pack.cu): rows in use / utterance of a row / un-packed frame of a row
    const int* row_b;
    const int* row_fr;
    // time-varying rank-1 conditioner of WNP (modules.py:320-322, 353-362): pre += cw[n] * trk[row * n_sqz + phase(n)] + cb[n]
    const float* trk;              // [T, n_sqz] packed pitch / energy samples of each row, or null
    const float* cw_g;             // cond_layer1.weight_g [C1] (null: weight norm removed), C1 = 2HL / n_sqz
    const float* cw_v;             // cond_layer1.weight_v (or .weight) [C1]: effective weight g v / |v|
    const float* cb;               // cond_layer1.bias [C1]
    int n_sqz, cs_div;             // squeezed channel cs_base + n reads sample (cs_base + n) / cs_div of the row's n_sqz
    int cs_base;                   // samples through conv channel (cs_base + n) % cs_div;  cs_div = C1, cs_base = layer * 2H
    float* h_out_f32_last;         // last layer: masked WN output also as fp32 tile-private master (input of a chained WN)
    int dbg_noshift;               // instrumented builds: 1 = every tap reads the un-shifted tile (timing experiments only)
    const float* mask;             // [B, T]
    float drop_p;
    DropSeed drop_seed;
    const float* h_f32;            // layer input (fp32 master copy) -- the bf16 copy is the TMA operand
    float* h_out_f32;              // next layer's input (not last)
    __nv_bfloat16* h_out_bf16;
    float* skip_f32;               // running skip sum (fp32)
    __nv_bfloat16* skip_out_bf16;  // last layer: masked WN output, operand of the `end` conv
    __nv_bfloat16* ts_save;        // [B,T,2H] tanh | sigmoid outputs for the backward, or null
    __nv_bfloat16* acts_save;      // [B,T,H] gate output for the res_skip weight gradient, or null
};
int wn_layer_launch(const __nv_bfloat16* h_bf16, const __nv_bfloat16* win_p, const __nv_bfloat16* wrs_p, WnLayerParams p,
                    cudaStream_t st);

// A whole WaveNet (start conv + every layer) in one launch (wn_stack.cu): windows of 128 rows with a recomputed halo.
constexpr int WS_MAX_LAYERS = 4;
constexpr int WS_STAGES = 3;
struct WnStackLayer {
    const float* bias_in;          // [2H]
    const float* bias_rs;          // [RS]
    const float* cond;             // this layer's slice of cond_layer(g): cond[b * cond_bstride + n], or null
    DropSeed drop_seed;
    __nv_bfloat16* ts_save;        // [rows, 2H] tanh | sigmoid, tile-private layout addressed by frame, or null
    int RS, dil, cs_base;          // cs_base = layer * 2H (WNP conditioner channel offset)
};
struct WnStackParams {
    int B, T, H, taps, L;
    int lead, halo_tot, tile_out;  // filled by wn_stack_launch
    int has_start, n_ks;           // 1: the window is start(x0) (n_ks k-chunks of padded input channels); 0: it is loaded from hin
    int save;                      // training: h_l and the gate outputs of the owned rows are stored
    const float* bias_start;       // [H]
    size_t cond_bstride;
    const int* n_rows;             // varlen packing (pack.cu)
    const int* row_b;
    const int* row_fr;
    const float* mask;             // [B, T]
    float drop_p;
    const float* trk;              // WNP (see WnLayerParams)
    const float* cw_g;
    const float* cw_v;
    const float* cb;
    int n_sqz, cs_div;
    float* hf;                     // window-private fp32 masters: wn_stack_master_rows(T, tile_out) * B rows of H floats
    float* skipf;
    const float* h_in_f32;         // !has_start: fp32 master of the input, tile-private layout addressed by frame
    float* h_out_f32_last;         // masked WN output as fp32 (same layout) for a chained WaveNet, or null
    // `end` 1x1 conv + affine coupling (attentions.py:170-183) on the resident WN output (last WaveNet of a block); see
    // ConvTcParams EPI_AFFINE for the meaning of the aff_* fields.  has_end: the bf16 WN output is stored only when save
    int has_end, C_end;            // C_end = in_channels * n_sqz (output channels of `end`, a multiple of 32, <= 256)
    const float* bias_end;         // [C_end]
    const float* aff_x;
    float* aff_z;
    float* aff_logdet;
    float* aff_logs_save;
    int aff_reverse, aff_sigmoid_scale;
    WnStackLayer layer[WS_MAX_LAYERS];
};
struct WnStackMaps {
    CUtensorMap a0, wstart, hin, skip_out, wend, win[WS_MAX_LAYERS], wrs[WS_MAX_LAYERS], hsave[WS_MAX_LAYERS], acts[WS_MAX_LAYERS];
};
struct WnStackIo {                 // bf16 operands / outputs, frames [B, T, H] unless noted
    const __nv_bfloat16* a0; int ld_a0;          // x0 frames [B, T, ld_a0] (has_start)
    const __nv_bfloat16* wstart;                 // [H][n_ks * 64]
    const __nv_bfloat16* hin;                    // window input (!has_start)
    const __nv_bfloat16* win[WS_MAX_LAYERS];     // [taps][2H][H]
    const __nv_bfloat16* wrs[WS_MAX_LAYERS];     // [RS][H]
    __nv_bfloat16* hsave[WS_MAX_LAYERS];         // h_l (save)
    __nv_bfloat16* acts[WS_MAX_LAYERS];          // gate output of layer l (save)
    __nv_bfloat16* skip_out;                     // masked WN output
    const __nv_bfloat16* wend;                   // [C_end][H] (has_end)
};
bool wn_stack_geometry(int H, int K, int dr, int L, int n_ks, int* lead, int* halo_tot, int* tile_out);
bool wn_stack_enabled();                         // GTK_WN_STACK != 0
bool wn_stack_end_ok(int H, int C_end);
size_t wn_stack_master_rows(int T, int tile_out);
int wn_stack_launch(const WnStackIo& io, WnStackParams p, cudaStream_t st);

// Backward of one WaveNet layer fused into one kernel (wn_bwd_tc.cu).
struct WnBwdParams {
    int B, T, H, RS, taps, dil;
    const float* mask;             // [B, T]
    const __nv_bfloat16* ts;       // saved tanh | sigmoid of the layer, tile-private layout (ld = 2H)
    float drop_p;
    DropSeed drop_seed;
    const float* resid_f32;        // fp32 carry of d(h) from the layer above (tile-private layout, ld = H) or null
    float* out_f32;                // fp32 carry written for the layer below (tile-private layout, ld = H)
    const int* n_rows;             // varlen packing (pack.cu), or null
    const int* row_fr;
    // layer 0 of the first WaveNet: the data gradient of the `start` 1x1 conv rides in the same kernel (GEMM c on the d(h_0)
    // tile): post_out_tm[b, c, t] = post_resid_tm[b, c, t] + sum_n d(h_0)[t, n] W[n, c], c < post_N, t-major with post_C
    // channels per utterance.  post_w = the data-gradient pack [post_N][H]; null: no fused conv.
    const __nv_bfloat16* post_w;
    int post_N, post_C;
    const float* post_resid_tm;
    float* post_out_tm;
    __nv_bfloat16* dpre;           // [B,T,2H] frames or null: gate-input gradient BEFORE the dropout scale (what the
                                   // conditioners' gradients are reduced from when p_dropout > 0; == d(xin) when it is 0)
};
bool wn_bwd_supported(int H, int taps, int dil);
bool wn_bwd_post_ok(int H, int N);       // the fused `start` data gradient fits (N = in_channels * n_sqz / 2)
// drs: [B,T,ld_drs] frames (columns [0,RS) = [d(h_next) mask | d(skip)]); wrs_t / win_t: data-gradient weight packs;
// dxin: [B,T,2H] frames out; out_bf16: [B,T,ld_out] frames, columns [0,H) = d(h) mask out.
int wn_bwd_layer_launch(const __nv_bfloat16* drs, int ld_drs, const __nv_bfloat16* wrs_t, const __nv_bfloat16* win_t,
                        __nv_bfloat16* dxin, __nv_bfloat16* out_bf16, int ld_out, WnBwdParams p, cudaStream_t st);

int wn_trace_read(unsigned long long* out, int n_ctas);
int wn_wait_read(long long* out, int n_ctas);
int wb_trace_read(unsigned long long* out, int n_ctas);
int ws_trace_read(unsigned long long* out, int n_ctas);
int ws_wait_read(long long* out, int n_ctas);
int gg_wait_read(long long* out, int n_ctas);
int wb_wait_read(long long* out, int n_ctas);
void conv_trace_select(int epi, int taps);
int conv_trace_read(unsigned long long* out, int n_ctas);
int conv_wait_read(long long* out, int n_ctas);
int shift_probe(const __nv_bfloat16* a, const __nv_bfloat16* b, float* out, cudaStream_t st);

int wgrad_tc_launch(const __nv_bfloat16* dy, int ld_y, int cols_y, const __nv_bfloat16* x, int ld_x, int cols_x, float* dw, int B,
                    int T, int M, int Cin, int K, int dil, cudaStream_t st);
size_t wgrad_tc_workspace_bytes();
int wgrad_tc_param(const __nv_bfloat16* dy, int ld_y, int cols_y, const __nv_bfloat16* x, int ld_x, int cols_x, const float* g,
                   const float* v, float* dg, float* dv, float* part_ws, int B, int T, int M, int Cin, int K, int dil,
                   cudaStream_t st);
// All weight / bias gradients of one coupling block in one stream-K launch + one reduce launch (wgrad_group.cu).
constexpr int WGG_MAX_JOBS = 16, WGG_MAX_TILES = 160, WGG_MAX_SEGS = 320, WGG_MAX_CTAS = 160;
struct WgradGroupJob {
    const __nv_bfloat16* dy;   // output-gradient frames [B, T, ld_y], columns [0, cols_y) exist, [0, M) are used
    int ld_y, cols_y;
    const __nv_bfloat16* x;    // input frames [B, T, ld_x], columns [0, cols_x) exist, [0, Cin) are used
    int ld_x, cols_x;
    const float* g;            // weight_g [M] or null (plain weight)
    const float* v;            // weight_v / weight [M][Cin][K]
    float* dg;                 // += (weight-normed only)
    float* dv;                 // +=
    float* db;                 // += bias gradient [M], or null
    int M, Cin, K, dil;
};
bool wgrad_red_enabled();                           // shares ADD into one slot per tile (GTK_WGRAD_RED=0: one partial tile per share)
size_t wgrad_group_workspace_bytes();               // partial tiles of ONE launch (worst case)
// Deferred reduce: the stream-K launches of several coupling blocks write their partial tiles side by side into one
// workspace of `capacity` launches x wgrad_group_workspace_bytes(), and ONE reduce launch finishes them all
// (wgrad_group_reduce_flush) -- the per-block reduce is a latency-bound kernel of 2.7 waves, twelve of them cost 0.36 ms
// per step.  The batch is host state owned by the caller for the duration of one backward call.
struct WgradReduceBatch;
WgradReduceBatch* wgrad_reduce_batch(float* part_ws, int capacity);     // thread-local object, reset
int wgrad_group_reduce_flush(WgradReduceBatch* rb, cudaStream_t st);
// n_rows: null, or (varlen packing, B == 1) device pointer to hdr[1] of the packing plan {0, n_rows}: rows in use
// max_ctas > 0 caps the persistent grid (a launch that overlaps other kernels must not claim every SM)
// defer: null = reduce right after the stream-K launch (part_ws = one launch's workspace); else append to the batch
int wgrad_group_launch(const WgradGroupJob* jobs, int n_jobs, int B, int T, float* part_ws, const int* n_rows, cudaStream_t st,
                       int max_ctas = 0, WgradReduceBatch* defer = nullptr);
constexpr int WGG_MAX_DEFER = 14;                   // launches per reduce batch (kernel parameter space)
// out[g * out_gstride + n] += sum of rows [seg[g], seg[g + 1]) (device-resident bounds) of frames in[.., n]
// n_buf > 1: the same reduction over n_buf buffers buf_stride ELEMENTS apart, outputs out_bstride floats apart
int colsum_frames_seg(const __nv_bfloat16* in, int ld, int n_cols, int groups, const int* seg, float* out, size_t out_gstride,
                      cudaStream_t st, int n_buf = 1, size_t buf_stride = 0, size_t out_bstride = 0);
// WNP conditioner backward (wgrad_tc.cu): dpre = L bf16 frame buffers [T, 2H] `lstride` elements apart
int wnp_cond_bwd(const __nv_bfloat16* dpre, size_t lstride, int L, int H2, const float* trk, int n_sqz, int C1, const float* c1_g,
                 const float* c1_v, float* d_g, float* d_v, float* d_b, float* dtrk, const int* n_rows, int T, cudaStream_t st);
int frames_fanout(const __nv_bfloat16* src, int ld_src, __nv_bfloat16* dst, int ld_dst, size_t dst_stride, int n_dst, int cols,
                  const int* n_rows, int T, cudaStream_t st);
int zero_guard_rows(__nv_bfloat16* base, size_t stride_bytes, int n_buf, size_t stride2_bytes, int n_buf2, int ld,
                    const int* n_rows, int n_guard, int T, cudaStream_t st);
int colsum_frames(const __nv_bfloat16* in, int ld, int n_cols, int groups, long long rows_per_group, float* out, size_t out_gstride,
                  cudaStream_t st);
int affine_bwd_frames(const float* x, const float* logs_raw, const float* mask, const float* dz, const float* dlogdet,
                      __nv_bfloat16* dend, int ld, float* dx, int B, int C, int T, int sigmoid_scale, const int* row_b,
                      const int* n_rows, cudaStream_t st);
// Fold (weight norm) + bf16 pack of up to FOLD_PACK_MAX_JOBS conv weights in one launch (conv_tc.cu).
constexpr int FOLD_PACK_MAX_JOBS = 320;
struct FoldPackJob {
    const float* g;            // weight_g [M] or null (plain weight)
    const float* v;            // weight_v / weight [M][Cin][K]
    __nv_bfloat16* fwd;        // [K][rows_f][cols_f]
    __nv_bfloat16* tr;         // [K][rows_t][cols_t] or null
    int M, Cin, K, rows_f, cols_f, rows_t, cols_t;
    int groups;                // filled by fold_pack: CTAs (32-row groups) of this job
};
struct FoldPackJobs { FoldPackJob job[FOLD_PACK_MAX_JOBS]; int n; };
int fold_pack(FoldPackJobs& jobs, cudaStream_t st);
int conv_tc_launch(const __nv_bfloat16* in, int ld_in, const __nv_bfloat16* wp, ConvTcParams p, int epi, cudaStream_t st);
int pack_weight_bf16(const float* w, __nv_bfloat16* out, int M, int Cin, int K, int transpose, int rows, int cols, cudaStream_t st);
int frames_from_tmajor(const float* x, __nv_bfloat16* out, int B, int C, int T, int c0, int nc, int ld, cudaStream_t st,
                       const int* n_rows = nullptr);

#ifdef __CUDACC__
// ---------------------------------------------------------------------------------- device: TMA
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
// generic-proxy smem writes -> visible to the async proxy (TMA / tcgen05.mma operand reads)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---------------------------------------------------------------------------------- device: TMEM
template <int COLS>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result) {   // one full warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)), "n"(COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {        // the same warp that allocated
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// 32 lanes x 16 consecutive fp32 columns: thread l of warp w gets row 32*(w%4)+l (taddr lane field must match)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// v = columns [ta, ta + 16) + columns [tb, tb + 16): an accumulator kept as two halves that took alternate k16 steps.
// Both loads are in flight before the one wait.
__device__ __forceinline__ void tmem_ld16_sum(uint32_t ta, uint32_t tb, float (&v)[16]) {
    uint32_t r[16], s[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(ta));
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(s[0]), "=r"(s[1]), "=r"(s[2]), "=r"(s[3]), "=r"(s[4]), "=r"(s[5]), "=r"(s[6]), "=r"(s[7]), "=r"(s[8]),
          "=r"(s[9]), "=r"(s[10]), "=r"(s[11]), "=r"(s[12]), "=r"(s[13]), "=r"(s[14]), "=r"(s[15])
        : "r"(tb));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]) + __uint_as_float(s[i]);
}

// ---------------------------------------------------------------------------------- device: UMMA
// K-major operand tile in the canonical SWIZZLE_128B layout (what a TMA box {64 bf16, rows} with
// CU_TENSOR_MAP_SWIZZLE_128B writes): row r at r*128 B, 8-row groups 1024 B apart (SBO), 16-byte chunks
// XOR-ed with (r % 8).  The tile base must be 1024-byte aligned; advancing along K by 16 elements = +32 B.
__device__ __forceinline__ uint64_t umma_desc_kmajor_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);          // start address, bits [0,14)
    d |= (uint64_t)1 << 16;                               // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;                     // stride byte offset = 1024 B between 8-row groups
    d |= (uint64_t)1 << 46;                               // descriptor version (Blackwell)
    d |= (uint64_t)2 << 61;                               // SWIZZLE_128B
    return d;
}
// MN-major operand tile, SWIZZLE_128B: 64 MN-elements (128 B) contiguous per K row, 8 K rows = one 1024-B atom.
// lbo_bytes = distance between successive 64-element MN blocks, sbo_bytes = distance between 8-row K groups.
__device__ __forceinline__ uint64_t umma_desc_mnmajor_sw128(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// kind::f16 instruction descriptor: bf16 x bf16 -> fp32, M x N tile; a_mn / b_mn = 1 for MN-major operands.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N, int a_mn = 0, int b_mn = 0) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread.
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, bool accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"((uint32_t)accumulate)
        : "memory");
}
// mbarrier arrive when all MMAs issued so far by this thread have completed (implies fence::before_thread_sync)
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// ... the same, arriving on the barrier at this offset in every CTA of `cta_mask` (cluster multicast)
__device__ __forceinline__ void umma_commit_multicast(uint64_t* bar, uint16_t cta_mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"(cta_mask) : "memory");
}
// TMA load delivered to the same shared-memory offset (and signalling the same mbarrier offset) in every CTA of cta_mask
__device__ __forceinline__ void tma_load_3d_multicast(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2,
                                                      uint16_t cta_mask) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
        " [%0], [%1, {%3, %4, %5}], [%2], %6;"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "h"(cta_mask)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d_multicast(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1,
                                                      uint16_t cta_mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
        " [%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(cta_mask)
        : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ float sigmoid_f(float x) { return 1.f / (1.f + __expf(-x)); }
// single-MUFU forms used by the bf16 path (results are rounded to bf16 anyway; abs error ~5e-4)
__device__ __forceinline__ float tanh_fast(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float sigmoid_fast(float x) { return fmaf(0.5f, tanh_fast(0.5f * x), 0.5f); }
__device__ __forceinline__ void store_bf16x16(__nv_bfloat16* dst, const float (&v)[16]) {
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    uint4* o = reinterpret_cast<uint4*>(dst);
    o[0] = make_uint4(w[0], w[1], w[2], w[3]);
    o[1] = make_uint4(w[4], w[5], w[6], w[7]);
}
__device__ __forceinline__ void load_bf16x16(const __nv_bfloat16* src, float (&v)[16]) {
    const uint4* s4 = reinterpret_cast<const uint4*>(src);
    const uint4 a = s4[0], b = s4[1];
    const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&w[i]);
        v[2 * i] = __low2float(h);
        v[2 * i + 1] = __high2float(h);
    }
}

__device__ __forceinline__ void unpack_bf16x16(const uint4& a, const uint4& b, float (&v)[16]) {
    const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&w[i]);
        v[2 * i] = __low2float(h);
        v[2 * i + 1] = __high2float(h);
    }
}

// accumulator split over two TMEM column ranges (base and base + 256): load both halves and add
__device__ __forceinline__ void tmem_ld16_dual(uint32_t taddr, bool dual, float (&v)[16]) {
    tmem_ld16(taddr, v);
    if (dual) {
        float w[16];
        tmem_ld16(taddr + 256, w);
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] += w[i];
    }
}

// ---------------------------------------------------------------------------------- device: epilogue row I/O
// Tensors that only ever travel between epilogues (fp32 residual / skip masters, the saved tanh|sigmoid pairs, the
// fp32 data-gradient carry) need no particular layout, only that the thread that owns (tile, row) finds its data.
// "Tile-private" layout: 16-byte vector v of row r of tile `tile_id` lives at ((tile_id * n_vec + v) * 128 + r) * 16
// bytes, so the 32 lanes of a warp (32 consecutive rows) move 512 contiguous bytes per instruction instead of 32
// row-strided 16-byte pieces.  tp == 0 keeps the plain frame-major addressing (standalone C-ABI entry points).
struct RowIo {
    int tp;            // 1: tile-private layout, 0: frame-major
    int tile_id;       // b * tiles_per_utterance + tile
    int r;             // row inside the tile (== TMEM lane)
    size_t frame;      // b * T + t (frame-major row)
};
// RowIo of frame f of utterance b for the tile-private tensors (addressed by frame, whatever the tiling of the kernel)
__device__ __forceinline__ RowIo frame_io(int b, int f, int T) {
    RowIo io;
    io.tp = 1;
    io.tile_id = b * ((T + BT - 1) / BT) + (f >> 7);
    io.r = f & 127;
    io.frame = (size_t)b * T + f;
    return io;
}
__device__ __forceinline__ void row_load_f32x16(const float* base, int ld, int c0, const RowIo& io, float (&v)[16]) {
    const float4* s;
    size_t step;
    if (io.tp) { s = reinterpret_cast<const float4*>(base) + ((size_t)io.tile_id * (ld >> 2) + (c0 >> 2)) * 128 + io.r; step = 128; }
    else       { s = reinterpret_cast<const float4*>(base + io.frame * ld + c0); step = 1; }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float4 x = s[i * step];
        v[4 * i] = x.x; v[4 * i + 1] = x.y; v[4 * i + 2] = x.z; v[4 * i + 3] = x.w;
    }
}
__device__ __forceinline__ void row_store_f32x16(float* base, int ld, int c0, const RowIo& io, const float (&v)[16]) {
    float4* d;
    size_t step;
    if (io.tp) { d = reinterpret_cast<float4*>(base) + ((size_t)io.tile_id * (ld >> 2) + (c0 >> 2)) * 128 + io.r; step = 128; }
    else       { d = reinterpret_cast<float4*>(base + io.frame * ld + c0); step = 1; }
#pragma unroll
    for (int i = 0; i < 4; ++i) d[i * step] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
}
__device__ __forceinline__ void row_load_bf16x16(const __nv_bfloat16* base, int ld, int c0, const RowIo& io, float (&v)[16]) {
    const uint4* s;
    size_t step;
    if (io.tp) { s = reinterpret_cast<const uint4*>(base) + ((size_t)io.tile_id * (ld >> 3) + (c0 >> 3)) * 128 + io.r; step = 128; }
    else       { s = reinterpret_cast<const uint4*>(base + io.frame * ld + c0); step = 1; }
    const uint4 a = s[0], b = s[step];
    const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&w[i]);
        v[2 * i] = __low2float(h);
        v[2 * i + 1] = __high2float(h);
    }
}
__device__ __forceinline__ void row_store_bf16x16(__nv_bfloat16* base, int ld, int c0, const RowIo& io, const float (&v)[16]) {
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    uint4* d;
    size_t step;
    if (io.tp) { d = reinterpret_cast<uint4*>(base) + ((size_t)io.tile_id * (ld >> 3) + (c0 >> 3)) * 128 + io.r; step = 128; }
    else       { d = reinterpret_cast<uint4*>(base + io.frame * ld + c0); step = 1; }
    d[0] = make_uint4(w[0], w[1], w[2], w[3]);
    d[step] = make_uint4(w[4], w[5], w[6], w[7]);
}
// 16 bf16 columns [c0, c0 + 16) of tile row r into a staging area laid out as consecutive [128 rows x 64 columns]
// SWIZZLE_128B tiles of `tile_stride` bytes (the canonical K-major operand layout == what a TMA store expects).
__device__ __forceinline__ void stage_bf16x16_sw128(unsigned char* area, size_t tile_stride, int r, int c0, const float (&v)[16]) {
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    unsigned char* row = area + (size_t)(c0 >> 6) * tile_stride + (size_t)r * 128;
    const int ci = (c0 & 63) >> 3;
    *reinterpret_cast<uint4*>(row + ((ci ^ (r & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
    *reinterpret_cast<uint4*>(row + (((ci + 1) ^ (r & 7)) << 4)) = make_uint4(w[4], w[5], w[6], w[7]);
}
// TMA store of one staged [128 x 64] bf16 tile to frames [.., t0 .. t0+128, c0 .. c0+64) of utterance b; rows past the
// tensor's T extent are clipped by the hardware.
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* m, const void* smem_src, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* m, const void* smem_src, int c0, int c1, int c2, int c3) {
    asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
                 : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

#endif  // __CUDACC__

}  // namespace tc
}  // namespace gtk
