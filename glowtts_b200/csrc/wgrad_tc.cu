// glowtts_b200/csrc/wgrad_tc.cu -- weight gradients of the frame-major convolutions on tcgen05.
//
//   dW[m][c][j] += sum_{b,t} dY[b, t, m] * X[b, t + (j - K/2) dil, c]
//
// Both operands are frame-major bf16 ([B, T, channels]), i.e. the reduction dimension (t) is the ROW index and the
// M / N dimension is contiguous: "MN-major" UMMA operands.  A TMA box {64 channels, 64 frames} lands as 64 rows of
// 128 B in the SWIZZLE_128B layout; 8 rows form one 1024-byte atom (SBO), successive 64-channel blocks are one box
// (8192 B) apart (LBO).  The conv tap is again a shifted TMA box of X (zero-filled outside the utterance).
// Grid = (M tiles of 128, taps, split-K); each CTA reduces its share of the (b, 64-frame chunk) steps into a
// [128 x N] fp32 TMEM accumulator and adds it to dW with red.global.add.f32.
#include <string.h>

#include "tc.cuh"
#include "ops.h"

namespace gtk {
namespace tc {

constexpr int WG_TK = 64;                       // frames per k-step
constexpr int WG_BOX = WG_TK * 128;             // bytes of one {64 ch, 64 frames} box
constexpr int WG_STAGES = 4;

struct WgCtl {
    uint64_t full[WG_STAGES], empty[WG_STAGES], acc_full;
    uint32_t tmem_base;
};

struct WgradTcParams {
    int B, T, M, Cin, taps, dil;
    int n_blk;                 // N = 64 * n_blk columns of X are reduced (>= Cin, zero padded)
    int steps_total, splits;
    float* dw;                 // [M][Cin][taps] fp32, accumulated with atomics (part == null)
    float* part;               // or: partial tiles [split][tap][m_tiles*128][N] fp32, plain stores (reduced by wgrad_reduce)
};

__global__ void __launch_bounds__(NTHREADS, 1)
wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmY, const __grid_constant__ CUtensorMap tmX, const WgradTcParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int a_bytes = 2 * WG_BOX, b_bytes = p.n_blk * WG_BOX;
    const int stage_bytes = a_bytes + b_bytes;
    WgCtl* ctl = reinterpret_cast<WgCtl*>(smem + (size_t)WG_STAGES * stage_bytes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * 128, j = blockIdx.y, split = blockIdx.z;
    const int chunks_t = (p.T + WG_TK - 1) / WG_TK;
    const int s_lo = (int)((long long)p.steps_total * split / p.splits);
    const int s_hi = (int)((long long)p.steps_total * (split + 1) / p.splits);
    const int n_iter = s_hi - s_lo;
    const int N = p.n_blk * 64;

    if (threadIdx.x == 0) {
        for (int s = 0; s < WG_STAGES; ++s) { mbar_init(&ctl->full[s], 1); mbar_init(&ctl->empty[s], 1); }
        mbar_init(&ctl->acc_full, 1);
        mbar_fence_init();
    }
    if (warp == 2) tmem_alloc<256>(&ctl->tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctl->tmem_base;

    if (warp == 0) {
        if (lane == 0) {
            tma_prefetch_desc(&tmY);
            tma_prefetch_desc(&tmX);
            for (int it = 0; it < n_iter; ++it) {
                const int s = it % WG_STAGES, use = it / WG_STAGES;
                if (use > 0) mbar_wait(&ctl->empty[s], (use - 1) & 1);
                const int step = s_lo + it;
                const int b = step / chunks_t, t0 = (step % chunks_t) * WG_TK;
                unsigned char* sa = smem + (size_t)s * stage_bytes;
                unsigned char* sb = sa + a_bytes;
                mbar_arrive_expect_tx(&ctl->full[s], (uint32_t)stage_bytes);
                tma_load_3d(sa, &tmY, &ctl->full[s], m0, t0, b);
                tma_load_3d(sa + WG_BOX, &tmY, &ctl->full[s], m0 + 64, t0, b);
                for (int nb = 0; nb < p.n_blk; ++nb)
                    tma_load_3d(sb + (size_t)nb * WG_BOX, &tmX, &ctl->full[s], nb * 64, t0 + (j - p.taps / 2) * p.dil, b);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_bf16(128, N, 1, 1);
            for (int it = 0; it < n_iter; ++it) {
                const int s = it % WG_STAGES, use = it / WG_STAGES;
                mbar_wait(&ctl->full[s], use & 1);
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + (size_t)s * stage_bytes);
                const uint32_t sb = sa + a_bytes;
#pragma unroll
                for (int k = 0; k < WG_TK / 16; ++k) {
                    const uint64_t adesc = umma_desc_mnmajor_sw128(sa + k * 2048, WG_BOX, 1024);
                    const uint64_t bdesc = umma_desc_mnmajor_sw128(sb + k * 2048, WG_BOX, 1024);
                    umma_bf16(tmem, adesc, bdesc, idesc, (it | k) != 0);
                }
                umma_commit(&ctl->empty[s]);
            }
            umma_commit(&ctl->acc_full);
        }
    } else if (warp >= 4 && n_iter > 0) {
        mbar_wait(&ctl->acc_full, 0);
        tc_fence_after();
        const int q = warp & 3;
        const int m = m0 + q * 32 + lane;
        const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
        float* prow = p.part ? p.part + (((size_t)split * p.taps + j) * (gridDim.x * 128) + (m - 0)) * N : nullptr;
        for (int c0 = 0; c0 < N; c0 += 16) {
            float v[16];
            tmem_ld16(trow + c0, v);
            if (prow) {
                float4* o4 = reinterpret_cast<float4*>(prow + c0);
#pragma unroll
                for (int i = 0; i < 4; ++i) o4[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                continue;
            }
            if (m >= p.M) continue;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const int c = c0 + i;
                if (c < p.Cin) atomicAdd(p.dw + ((size_t)m * p.Cin + c) * p.taps + j, v[i]);
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc<256>(tmem);
    }
}

// dy: frames [B, T, ld_y] (columns [0, M) used; M may exceed... columns >= M read as zero must exist or be OOB);
// x: frames [B, T, ld_x] with at least Cin valid columns (the rest zero or OOB).
// Sum the split-K partial tiles of one weight and finish the parameter gradient in the same pass: one block per output
// row m (R = Cin*taps elements).  Plain weight: dv[m][c][j] += sum.  Weight-normed (g != null; torch.nn.utils.weight_norm,
// modules.py:124,131,141):  dg[m] += <dW_m, v_m>/||v_m||,  dv_m += g_m/||v_m|| (dW_m - v_m <dW_m, v_m>/||v_m||^2).
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float* __restrict__ part, int splits, int m_pad, int N, int M, int Cin,
                                                           int taps, const float* __restrict__ g, const float* __restrict__ v,
                                                           float* __restrict__ dg, float* __restrict__ dv) {
    extern __shared__ float row[];                 // [Cin * taps] reduced dW row, torch order (c, j)
    __shared__ float scratch[33];
    const int m = blockIdx.x, R = Cin * taps;
    float ss = 0.f, dd = 0.f;
    for (int e = threadIdx.x; e < taps * Cin; e += 256) {
        const int j = e / Cin, c = e % Cin;        // c fastest: coalesced reads of the partial tiles
        float acc = 0.f;
        for (int s = 0; s < splits; ++s) acc += part[(((size_t)s * taps + j) * m_pad + m) * N + c];
        row[c * taps + j] = acc;
        if (g) {
            const float t = v[(size_t)m * R + c * taps + j];
            ss += t * t;
            dd += t * acc;
        }
    }
    __syncthreads();
    if (!g) {
        for (int e = threadIdx.x; e < R; e += 256) dv[(size_t)m * R + e] += row[e];
        return;
    }
    ss = block_sum<256>(ss, scratch);
    dd = block_sum<256>(dd, scratch);
    const float nrm = sqrtf(ss);
    const float sc = g[m] / nrm, proj = dd / ss;
    for (int e = threadIdx.x; e < R; e += 256) dv[(size_t)m * R + e] += sc * (row[e] - v[(size_t)m * R + e] * proj);
    if (threadIdx.x == 0) dg[m] += dd / nrm;
}

size_t wgrad_tc_workspace_bytes() { return (size_t)(sm_count() + 16) * 128 * 256 * sizeof(float); }

// Weight gradient straight into the parameter gradients (dg, dv) -- or dv alone for a plain weight (g == null) --
// through split-K partial tiles in `part_ws` (wgrad_tc_workspace_bytes()).
int wgrad_tc_param(const __nv_bfloat16* dy, int ld_y, int cols_y, const __nv_bfloat16* x, int ld_x, int cols_x, const float* g,
                   const float* v, float* dg, float* dv, float* part_ws, int B, int T, int M, int Cin, int K, int dil,
                   cudaStream_t st);

static int wgrad_tc_run(const __nv_bfloat16* dy, int ld_y, int cols_y, const __nv_bfloat16* x, int ld_x, int cols_x, float* dw,
                        float* part, int* splits_out, int B, int T, int M, int Cin, int K, int dil, cudaStream_t st);

int wgrad_tc_launch(const __nv_bfloat16* dy, int ld_y, int cols_y, const __nv_bfloat16* x, int ld_x, int cols_x, float* dw, int B,
                    int T, int M, int Cin, int K, int dil, cudaStream_t st) {
    GTK_REQUIRE(dw, "null");
    return wgrad_tc_run(dy, ld_y, cols_y, x, ld_x, cols_x, dw, nullptr, nullptr, B, T, M, Cin, K, dil, st);
}

int wgrad_tc_param(const __nv_bfloat16* dy, int ld_y, int cols_y, const __nv_bfloat16* x, int ld_x, int cols_x, const float* g,
                   const float* v, float* dg, float* dv, float* part_ws, int B, int T, int M, int Cin, int K, int dil,
                   cudaStream_t st) {
    GTK_REQUIRE(dv && part_ws && (!g || (v && dg)), "null");
    if ((long long)B * T == 0 || M == 0 || Cin == 0) return GTK_OK;
    int splits = 0;
    int rc = wgrad_tc_run(dy, ld_y, cols_y, x, ld_x, cols_x, nullptr, part_ws, &splits, B, T, M, Cin, K, dil, st);
    if (rc) return rc;
    const int m_pad = ceil_div(M, 128) * 128, N = ceil_div(Cin, 64) * 64;
    const size_t smem = (size_t)Cin * K * sizeof(float);
    GTK_REQUIRE(smem <= 48 * 1024, "weight row too long for the reduce kernel");
    wgrad_reduce_kernel<<<M, 256, smem, st>>>(part_ws, splits, m_pad, N, M, Cin, K, g, v, dg, dv);
    return check_launch("wgrad_reduce");
}

static int wgrad_tc_run(const __nv_bfloat16* dy, int ld_y, int cols_y, const __nv_bfloat16* x, int ld_x, int cols_x, float* dw,
                        float* part, int* splits_out, int B, int T, int M, int Cin, int K, int dil, cudaStream_t st) {
    GTK_REQUIRE(dy && x && (dw || part), "null");
    GTK_REQUIRE(K >= 1 && (K & 1) && dil >= 1, "bad conv shape");
    GTK_REQUIRE(ld_y % 8 == 0 && ld_x % 8 == 0 && cols_y >= M && cols_x >= Cin, "frame pitch");
    if ((long long)B * T == 0 || M == 0 || Cin == 0) return GTK_OK;
    WgradTcParams p;
    memset(&p, 0, sizeof(p));
    p.B = B; p.T = T; p.M = M; p.Cin = Cin; p.taps = K; p.dil = dil; p.dw = dw; p.part = part;
    p.n_blk = ceil_div(Cin, 64);
    GTK_REQUIRE(p.n_blk * 64 <= 256, "Cin <= 256");
    p.steps_total = B * ceil_div(T, WG_TK);
    const int m_tiles = ceil_div(M, 128);
    int splits = max(1, sm_count() / max(1, m_tiles * K));
    if (splits > p.steps_total) splits = p.steps_total;
    p.splits = splits;
    if (splits_out) *splits_out = splits;
    CUtensorMap tmY, tmX;
    {
        const uint64_t dims[3] = {(uint64_t)cols_y, (uint64_t)T, (uint64_t)B};
        const uint64_t strides[2] = {(uint64_t)ld_y * 2, (uint64_t)ld_y * 2 * T};
        const uint32_t box[3] = {64, WG_TK, 1};
        int rc = make_tmap_bf16(&tmY, dy, 3, dims, strides, box, true);
        if (rc) return rc;
    }
    {
        const uint64_t dims[3] = {(uint64_t)cols_x, (uint64_t)T, (uint64_t)B};
        const uint64_t strides[2] = {(uint64_t)ld_x * 2, (uint64_t)ld_x * 2 * T};
        const uint32_t box[3] = {64, WG_TK, 1};
        int rc = make_tmap_bf16(&tmX, x, 3, dims, strides, box, true);
        if (rc) return rc;
    }
    const size_t smem = (size_t)WG_STAGES * (2 + p.n_blk) * WG_BOX + sizeof(WgCtl) + 1024;
    GTK_REQUIRE(smem <= 227 * 1024, "tile does not fit in shared memory");
    GTK_CUDA(cudaFuncSetAttribute(wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(m_tiles, K, splits);
    ProfScope prof(K > 1 ? "wgrad_tc_kN" : "wgrad_tc_k1", st);
    wgrad_tc_kernel<<<grid, NTHREADS, smem, st>>>(tmY, tmX, p);
    return check_launch("wgrad_tc");
}

// out[g][n] (+)= sum over the rows of group g of in[row][n]   (bias / conditioning gradients from bf16 frames)
// rows_per_group = T for a per-utterance sum (groups = B), or B*T for a whole-batch sum (groups = 1).
__global__ void colsum_frames_kernel(const __nv_bfloat16* __restrict__ in, int ld, int n_cols, long long rows_per_group,
                                     float* __restrict__ out, size_t out_gstride, int row_splits) {
    const int g = blockIdx.y / row_splits, rs = blockIdx.y % row_splits;
    const int n = blockIdx.x * 64 + (threadIdx.x & 63);
    const int rl = threadIdx.x >> 6;                                // 4 row lanes
    const long long r_lo = rows_per_group * rs / row_splits, r_hi = rows_per_group * (rs + 1) / row_splits;
    float acc = 0.f;
    if (n < n_cols)
        for (long long r = r_lo + rl; r < r_hi; r += 4) acc += __bfloat162float(in[((size_t)g * rows_per_group + r) * ld + n]);
    __shared__ float red[256];
    red[threadIdx.x] = acc;
    __syncthreads();
    if (rl == 0 && n < n_cols) atomicAdd(out + (size_t)g * out_gstride + n, red[threadIdx.x] + red[threadIdx.x + 64] + red[threadIdx.x + 128] + red[threadIdx.x + 192]);
}
int colsum_frames(const __nv_bfloat16* in, int ld, int n_cols, int groups, long long rows_per_group, float* out, size_t out_gstride,
                  cudaStream_t st) {
    GTK_REQUIRE(in && out, "null");
    if (groups == 0 || rows_per_group == 0 || n_cols == 0) return GTK_OK;
    int row_splits = (int)max(1LL, min((long long)64, rows_per_group / 256));
    dim3 grid(ceil_div(n_cols, 64), groups * row_splits);
    colsum_frames_kernel<<<grid, 256, 0, st>>>(in, ld, n_cols, rows_per_group, out, out_gstride, row_splits);
    return check_launch("colsum_frames");
}

// The same with device-resident group bounds (varlen packing: utterance b = rows [off[b], off[b+1]); whole batch =
// {0, n_rows}), vectorised: a thread owns 8 consecutive columns (one 16-byte load per row) and every 8th row of its
// slice; grid (column blocks of 256, groups * SEG_SPLITS row slices, n_buf buffers `buf_stride` elements apart whose
// outputs are `out_bstride` floats apart -- the L per-layer gradients of a coupling block in one launch).
constexpr int SEG_SPLITS = 8, SEG_RL = 8;
__global__ void __launch_bounds__(256) colsum_frames_seg_kernel(const __nv_bfloat16* __restrict__ in, int ld, int n_cols,
                                                                const int* __restrict__ seg, float* __restrict__ out,
                                                                size_t out_gstride, size_t buf_stride, size_t out_bstride) {
    const int g = blockIdx.y / SEG_SPLITS, rs = blockIdx.y % SEG_SPLITS;
    const int cv = threadIdx.x & 31, rl = threadIdx.x >> 5;          // 32 column vectors x 8 row lanes
    const int n0 = (blockIdx.x * 32 + cv) * 8;
    in += (size_t)blockIdx.z * buf_stride;
    out += (size_t)blockIdx.z * out_bstride;
    const long long lo = seg[g], len = (long long)seg[g + 1] - lo;
    const long long r_lo = lo + len * rs / SEG_SPLITS, r_hi = lo + len * (rs + 1) / SEG_SPLITS;
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    const bool vec = (ld % 8 == 0) && n0 + 8 <= n_cols;
    if (n0 < n_cols) {
        for (long long r = r_lo + rl; r < r_hi; r += SEG_RL) {
            const __nv_bfloat16* src = in + (size_t)r * ld + n0;
            if (vec) {
                const uint4 q = *reinterpret_cast<const uint4*>(src);
                const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&w[i]);
                    acc[2 * i] += __low2float(h);
                    acc[2 * i + 1] += __high2float(h);
                }
            } else {
                for (int i = 0; i < 8 && n0 + i < n_cols; ++i) acc[i] += __bfloat162float(src[i]);
            }
        }
    }
    __shared__ float red[SEG_RL][32][9];
#pragma unroll
    for (int i = 0; i < 8; ++i) red[rl][cv][i] = acc[i];
    __syncthreads();
    for (int e = threadIdx.x; e < 256; e += 256) {
        const int c = e >> 3, i = e & 7, n = (blockIdx.x * 32 + c) * 8 + i;
        if (n >= n_cols) continue;
        float s = 0.f;
#pragma unroll
        for (int q = 0; q < SEG_RL; ++q) s += red[q][c][i];
        if (s != 0.f) atomicAdd(out + (size_t)g * out_gstride + n, s);
    }
}
int colsum_frames_seg(const __nv_bfloat16* in, int ld, int n_cols, int groups, const int* seg, float* out, size_t out_gstride,
                      cudaStream_t st, int n_buf, size_t buf_stride, size_t out_bstride) {
    GTK_REQUIRE(in && out && seg && n_buf >= 1, "null");
    if (groups == 0 || n_cols == 0) return GTK_OK;
    dim3 grid(ceil_div(n_cols, 256), groups * SEG_SPLITS, n_buf);
    ProfScope prof("colsum_seg", st);
    colsum_frames_seg_kernel<<<grid, 256, 0, st>>>(in, ld, n_cols, seg, out, out_gstride, buf_stride, out_bstride);
    return check_launch("colsum_frames_seg");
}

// Zero rows [round_up(n_rows, 128), + n_guard) of `n_buf` bf16 frame buffers (row pitch ld, `stride` bytes apart).
// Varlen packing: kernels never write behind the last 128-row tile in use, but the weight-gradient GEMM reads the conv
// input up to (K/2) dil rows past its last k-step; those rows only ever meet zero output-gradient rows, and 0 x garbage
// must not be 0 x NaN.
__global__ void zero_guard_rows_kernel(unsigned char* base, size_t stride, size_t stride2, int ld, const int* __restrict__ n_rows,
                                       int n_guard, int T) {
    const int r0 = (*n_rows + 127) & ~127;
    const int rows = min(n_guard, T - r0);
    if (rows <= 0) return;
    uint4* dst = reinterpret_cast<uint4*>(base + (size_t)blockIdx.x * stride + (size_t)blockIdx.y * stride2 + (size_t)r0 * ld * 2);
    const int n16 = rows * ld * 2 / 16;
    for (int i = threadIdx.x; i < n16; i += blockDim.x) dst[i] = make_uint4(0u, 0u, 0u, 0u);
}
int zero_guard_rows(__nv_bfloat16* base, size_t stride_bytes, int n_buf, size_t stride2_bytes, int n_buf2, int ld,
                    const int* n_rows, int n_guard, int T, cudaStream_t st) {
    GTK_REQUIRE(base && n_rows && ld % 8 == 0, "bad args");
    if (n_buf == 0 || n_buf2 == 0 || n_guard == 0) return GTK_OK;
    zero_guard_rows_kernel<<<dim3(n_buf, n_buf2), 256, 0, st>>>(reinterpret_cast<unsigned char*>(base), stride_bytes, stride2_bytes,
                                                               ld, n_rows, n_guard, T);
    return check_launch("zero_guard_rows");
}

// ---------------------------------------------------------------------------------- WNP conditioner gradients
// Backward of the time-varying conditioner of modules.WNP (modules.py:320-336, 353-362) from the gate-input gradients
// dpre [rows, 2H] (bf16 frames) of its L layers (buffers `lstride` elements apart):
//   cond[row, layer, n] = w[c] trk[row, s] + b[c],   cch = layer 2H + n,  s = cch / C1,  c = cch % C1,   w = g v / |v|
//   d b[c] += sum_rows dpre ;  d w[c] += sum_rows dpre trk[row, s]  (-> d weight_g = d w sign(v), d weight_v = 0; or d weight
//   when the norm was removed) ;  d trk[row, s] += sum_{layer, n} dpre w[c]   (accumulated over blocks, plain +=: one
//   thread owns a row's sample).  One CTA = 32 rows, a warp = 4 of them; a lane owns column vectors of 8.
constexpr int WC_NT = 256, WC_ROWS = 32;
__global__ void __launch_bounds__(WC_NT) wnp_cond_bwd_kernel(const __nv_bfloat16* __restrict__ dpre, size_t lstride, int L, int H2,
                                                             const float* __restrict__ trk, int n_sqz, int C1,
                                                             const float* __restrict__ c1_g, const float* __restrict__ c1_v,
                                                             float* __restrict__ d_g, float* __restrict__ d_v, float* __restrict__ d_b,
                                                             const int* __restrict__ n_rows) {
    __shared__ float red_b[WC_NT / 32][512], red_w[WC_NT / 32][512];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row0 = blockIdx.x * WC_ROWS;
    if (row0 >= *n_rows) return;
    const int nvec = H2 / 8;                               // 16-byte column vectors per row (<= 64)
    for (int l = 0; l < L; ++l) {
        const __nv_bfloat16* src = dpre + (size_t)l * lstride;
        for (int v0 = 0; v0 < nvec; v0 += 32) {            // column vectors lane + v0
            const int vec = v0 + lane;
            if (vec >= nvec) continue;
            const int n0 = vec * 8;
            const int s = (l * H2 + n0) / C1;              // 8 consecutive squeezed channels share s (C1 % 16 == 0)
            float ab[8], aw[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) ab[i] = aw[i] = 0.f;
#pragma unroll
            for (int q = 0; q < WC_ROWS / 8; ++q) {
                const int row = row0 + warp + 8 * q;
                if (row >= *n_rows) continue;
                const uint4 qv = *reinterpret_cast<const uint4*>(src + (size_t)row * H2 + n0);
                const uint32_t u[4] = {qv.x, qv.y, qv.z, qv.w};
                const float tv = trk[(size_t)row * n_sqz + s];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&u[i]);
                    const float a = __low2float(h), bq = __high2float(h);
                    ab[2 * i] += a; ab[2 * i + 1] += bq;
                    aw[2 * i] = fmaf(a, tv, aw[2 * i]); aw[2 * i + 1] = fmaf(bq, tv, aw[2 * i + 1]);
                }
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) { red_b[warp][n0 + i] = ab[i]; red_w[warp][n0 + i] = aw[i]; }
        }
        __syncthreads();
        for (int n = threadIdx.x; n < H2; n += WC_NT) {
            float sb = 0.f, sw = 0.f;
#pragma unroll
            for (int w8 = 0; w8 < WC_NT / 32; ++w8) { sb += red_b[w8][n]; sw += red_w[w8][n]; }
            const int cch = l * H2 + n, s = cch / C1, c = cch - s * C1;
            if (sb != 0.f) atomicAdd(d_b + c, sb);
            if (sw != 0.f) {
                if (c1_g) atomicAdd(d_g + c, sw * (c1_v[c] < 0.f ? -1.f : 1.f));
                else atomicAdd(d_v + c, sw);
            }
        }
        __syncthreads();
    }
}
// d trk[row, s] += sum_{layer, n in phase s} dpre[row, n] w[c]: one warp per row, lanes stride the 2H L columns.
__global__ void __launch_bounds__(256) wnp_row_bwd_kernel(const __nv_bfloat16* __restrict__ dpre, size_t lstride, int L, int H2,
                                                          int n_sqz, int C1, const float* __restrict__ c1_g,
                                                          const float* __restrict__ c1_v, float* __restrict__ dtrk,
                                                          const int* __restrict__ n_rows) {
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= *n_rows) return;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const int nvec = H2 / 8;
    for (int l = 0; l < L; ++l) {
        const __nv_bfloat16* src = dpre + (size_t)l * lstride + (size_t)row * H2;
        for (int vec = lane; vec < nvec; vec += 32) {
            const int n0 = vec * 8, cch = l * H2 + n0, s = cch / C1, c0 = cch - s * C1;
            const uint4 qv = *reinterpret_cast<const uint4*>(src + n0);
            const uint32_t u[4] = {qv.x, qv.y, qv.z, qv.w};
            float part = 0.f;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&u[i]);
                const float v0 = c1_v[c0 + 2 * i], v1 = c1_v[c0 + 2 * i + 1];
                const float w0 = c1_g ? c1_g[c0 + 2 * i] * (v0 < 0.f ? -1.f : 1.f) : v0;
                const float w1 = c1_g ? c1_g[c0 + 2 * i + 1] * (v1 < 0.f ? -1.f : 1.f) : v1;
                part = fmaf(__low2float(h), w0, part);
                part = fmaf(__high2float(h), w1, part);
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[q] += (q == s) ? part : 0.f;
        }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[q] = warp_sum(acc[q]);
    if (lane < n_sqz) dtrk[(size_t)row * n_sqz + lane] += acc[lane == 0 ? 0 : lane == 1 ? 1 : lane == 2 ? 2 : 3];
}
int wnp_cond_bwd(const __nv_bfloat16* dpre, size_t lstride, int L, int H2, const float* trk, int n_sqz, int C1, const float* c1_g,
                 const float* c1_v, float* d_g, float* d_v, float* d_b, float* dtrk, const int* n_rows, int T, cudaStream_t st) {
    GTK_REQUIRE(dpre && trk && c1_v && d_b && dtrk && n_rows, "null");
    GTK_REQUIRE(H2 % 8 == 0 && H2 <= 512 && n_sqz >= 1 && n_sqz <= 4 && C1 % 16 == 0, "shape");
    GTK_REQUIRE(c1_g ? d_g != nullptr : d_v != nullptr, "conditioner weight gradient pointer missing");
    if (T == 0) return GTK_OK;
    wnp_cond_bwd_kernel<<<ceil_div(T, WC_ROWS), WC_NT, 0, st>>>(dpre, lstride, L, H2, trk, n_sqz, C1, c1_g, c1_v, d_g, d_v, d_b, n_rows);
    int rc = check_launch("wnp_cond_bwd");
    if (rc) return rc;
    wnp_row_bwd_kernel<<<ceil_div(T, 8), 256, 0, st>>>(dpre, lstride, L, H2, n_sqz, C1, c1_g, c1_v, dtrk, n_rows);
    return check_launch("wnp_row_bwd");
}

// dst_k[row, c] = src[row, c] for k < n_dst destinations `dst_stride` elements apart (bf16 frames, 8 columns per thread):
// the data gradient of a chained WaveNet's input is the d(skip) operand of every layer of the WaveNet before it.
__global__ void __launch_bounds__(256) frames_fanout_kernel(const __nv_bfloat16* __restrict__ src, int ld_src, __nv_bfloat16* dst,
                                                            int ld_dst, size_t dst_stride, int n_dst, int cols,
                                                            const int* __restrict__ n_rows, int T) {
    const int nv = cols / 8;
    const long long lim = n_rows ? (long long)min(T, (*n_rows + 127) & ~127) : (long long)T;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < lim * nv; i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / nv;
        const int v = (int)(i - row * nv);
        const uint4 q = *reinterpret_cast<const uint4*>(src + (size_t)row * ld_src + v * 8);
        for (int k = 0; k < n_dst; ++k) *reinterpret_cast<uint4*>(dst + (size_t)k * dst_stride + (size_t)row * ld_dst + v * 8) = q;
    }
}
int frames_fanout(const __nv_bfloat16* src, int ld_src, __nv_bfloat16* dst, int ld_dst, size_t dst_stride, int n_dst, int cols,
                  const int* n_rows, int T, cudaStream_t st) {
    GTK_REQUIRE(src && dst && cols % 8 == 0 && ld_src % 8 == 0 && ld_dst % 8 == 0, "bad args");
    if (T == 0 || n_dst == 0) return GTK_OK;
    frames_fanout_kernel<<<min(1024, ceil_div(T * (cols / 8), 256)), 256, 0, st>>>(src, ld_src, dst, ld_dst, dst_stride, n_dst, cols, n_rows, T);
    return check_launch("frames_fanout");
}

}  // namespace tc
}  // namespace gtk
