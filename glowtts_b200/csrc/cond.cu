// glowtts_b200/csrc/cond.cu -- WN.cond_layer(g) (modules.py:122-124, 148-149) of EVERY coupling block in one launch.
//
//   cond[k][b][n] = bias_k[n] + sum_c W_k[n][c] g[b][c],   W_k = weight_g_k v_k / ||v_k||  (torch.nn.utils.weight_norm)
//
// k < NB blocks, n < N = 2 H L channels, c < G = gin_channels, b < B.  It is a [N x G] x [G x B] product per block
// (1536 x 256 x 32 at configs/base_blank_ms.json): 0.15 GFLOP for all 12 blocks, so CUDA cores and one launch -- what
// matters is that the ~100 small launches of twelve weight_norm + addmm + stack calls (and as many again in autograd)
// disappear from the multi-speaker step.  One warp owns a weight row: the row lives in registers (G / 32 values per lane),
// its norm is a warp reduction, g[b][:] is broadcast from shared memory.
//
// Backward (one launch): per row  dW[n][c] = sum_b dcond[b][n] g[b][c], dbias[n] = sum_b dcond[b][n], folded through the
// weight norm into (d weight_g, d weight_v); and  dg[b][c] += sum_n dcond[b][n] W[n][c]  with the CTA's 64 effective rows
// staged in shared memory (one global atomicAdd per (b, c) per CTA).
#include "common.cuh"
#include "ops.h"

namespace gtk {
namespace flow {

constexpr int CD_NT = 256, CD_ROWS = 64, CD_RPW = CD_ROWS / (CD_NT / 32), CD_KPL = 16;   // G <= 32 * CD_KPL = 512

// a row of v into registers (+ its effective-weight scale)
__device__ __forceinline__ float cond_load_row(const float* __restrict__ v, const float* __restrict__ wg, int n, int G, int lane,
                                               float (&w)[CD_KPL], float* norm_out) {
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < CD_KPL; ++i) {
        const int c = lane + 32 * i;
        w[i] = c < G ? __ldg(v + (size_t)n * G + c) : 0.f;
        ss = fmaf(w[i], w[i], ss);
    }
    ss = warp_sum(ss);
    const float nrm = sqrtf(ss);
    if (norm_out) *norm_out = nrm;
    return wg ? __ldg(wg + n) / nrm : 1.f;
}

__global__ void __launch_bounds__(CD_NT) cond_all_kernel(const CondJobs J, const float* __restrict__ g, float* __restrict__ cond,
                                                         int B, int G, int N) {
    extern __shared__ float cd_sm[];
    float* gs = cd_sm;                               // [B][G]
    float* outs = cd_sm + (size_t)B * G;             // [B][CD_ROWS]
    const int k = blockIdx.y, n0 = blockIdx.x * CD_ROWS;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < B * G; i += CD_NT) gs[i] = g[i];
    __syncthreads();
    const float* v = J.v[k];
    const float* wg = J.g[k];
    const float* bias = J.b[k];
    for (int rr = 0; rr < CD_RPW; ++rr) {
        const int rl = warp * CD_RPW + rr, n = n0 + rl;
        if (n >= N) break;
        float w[CD_KPL];
        const float sc = cond_load_row(v, wg, n, G, lane, w, nullptr);
        const float bn = bias ? __ldg(bias + n) : 0.f;
        for (int b = 0; b < B; ++b) {
            float d = 0.f;
#pragma unroll
            for (int i = 0; i < CD_KPL; ++i) {
                const int c = lane + 32 * i;
                if (c < G) d = fmaf(w[i], gs[(size_t)b * G + c], d);
            }
            d = warp_sum(d);
            if (lane == 0) outs[b * CD_ROWS + rl] = fmaf(d, sc, bn);
        }
    }
    __syncthreads();
    const int rows = min(CD_ROWS, N - n0);
    for (int i = threadIdx.x; i < B * CD_ROWS; i += CD_NT) {
        const int b = i / CD_ROWS, rl = i % CD_ROWS;
        if (rl < rows) cond[((size_t)k * B + b) * N + n0 + rl] = outs[i];
    }
}

int cond_all(const CondJobs& J, const float* g, float* cond, int B, int G, int N, cudaStream_t st) {
    GTK_REQUIRE(g && cond && J.n >= 1 && J.n <= COND_MAX_BLOCKS, "bad args");
    GTK_REQUIRE(G >= 1 && G <= 32 * CD_KPL, "gin_channels must be <= 512");
    if (B == 0 || N == 0) return GTK_OK;
    const size_t smem = ((size_t)B * G + (size_t)B * CD_ROWS) * sizeof(float);
    GTK_REQUIRE(smem <= 200 * 1024, "batch x gin_channels too large for the conditioning kernel");
    GTK_CUDA(cudaFuncSetAttribute(cond_all_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ProfScope prof("cond_all", st);
    cond_all_kernel<<<dim3(ceil_div(N, CD_ROWS), J.n), CD_NT, smem, st>>>(J, g, cond, B, G, N);
    return check_launch("cond_all");
}

// dcond [NB][B][N] -> dv / dgw / dbias of every block (WRITTEN, not accumulated) and dg [B][G] (ACCUMULATED with atomics:
// zero it first).
__global__ void __launch_bounds__(CD_NT) cond_all_bwd_kernel(const CondJobs J, const CondGrads D, const float* __restrict__ g,
                                                             const float* __restrict__ dcond, float* __restrict__ dg, int B, int G,
                                                             int N) {
    extern __shared__ float cd_sm[];
    float* gs = cd_sm;                               // [B][G]
    float* dcs = gs + (size_t)B * G;                 // [B][CD_ROWS]
    float* ws = dcs + (size_t)B * CD_ROWS;           // [CD_ROWS][G] effective weight rows of this CTA
    const int k = blockIdx.y, n0 = blockIdx.x * CD_ROWS;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rows = min(CD_ROWS, N - n0);
    for (int i = threadIdx.x; i < B * G; i += CD_NT) gs[i] = g[i];
    for (int i = threadIdx.x; i < B * CD_ROWS; i += CD_NT) {
        const int b = i / CD_ROWS, rl = i % CD_ROWS;
        dcs[i] = rl < rows ? dcond[((size_t)k * B + b) * N + n0 + rl] : 0.f;
    }
    __syncthreads();
    const float* v = J.v[k];
    const float* wg = J.g[k];
    for (int rr = 0; rr < CD_RPW; ++rr) {
        const int rl = warp * CD_RPW + rr, n = n0 + rl;
        if (n >= N) {                                // padding rows of the last CTA contribute nothing
#pragma unroll
            for (int i = 0; i < CD_KPL; ++i) { const int c = lane + 32 * i; if (c < G) ws[(size_t)rl * G + c] = 0.f; }
            continue;
        }
        float w[CD_KPL], dW[CD_KPL], nrm;
        const float sc = cond_load_row(v, wg, n, G, lane, w, &nrm);
#pragma unroll
        for (int i = 0; i < CD_KPL; ++i) dW[i] = 0.f;
        float db = 0.f;
        for (int b = 0; b < B; ++b) {
            const float dc = dcs[b * CD_ROWS + rl];
            db += dc;
#pragma unroll
            for (int i = 0; i < CD_KPL; ++i) {
                const int c = lane + 32 * i;
                if (c < G) dW[i] = fmaf(dc, gs[(size_t)b * G + c], dW[i]);
            }
        }
        if (lane == 0 && D.b[k]) D.b[k][n] = db;
        float dd = 0.f;
#pragma unroll
        for (int i = 0; i < CD_KPL; ++i) dd = fmaf(dW[i], w[i], dd);
        dd = warp_sum(dd);
        const float proj = dd / (nrm * nrm);
#pragma unroll
        for (int i = 0; i < CD_KPL; ++i) {
            const int c = lane + 32 * i;
            if (c < G) {
                ws[(size_t)rl * G + c] = w[i] * sc;
                if (D.v[k]) D.v[k][(size_t)n * G + c] = wg ? sc * (dW[i] - w[i] * proj) : dW[i];
            }
        }
        if (lane == 0 && wg && D.g[k]) D.g[k][n] = dd / nrm;
    }
    __syncthreads();
    if (dg) {
        for (int b = warp; b < B; b += CD_NT / 32) {
            float acc[CD_KPL];
#pragma unroll
            for (int i = 0; i < CD_KPL; ++i) acc[i] = 0.f;
            for (int rl = 0; rl < rows; ++rl) {
                const float dc = dcs[b * CD_ROWS + rl];
#pragma unroll
                for (int i = 0; i < CD_KPL; ++i) {
                    const int c = lane + 32 * i;
                    if (c < G) acc[i] = fmaf(dc, ws[(size_t)rl * G + c], acc[i]);
                }
            }
#pragma unroll
            for (int i = 0; i < CD_KPL; ++i) {
                const int c = lane + 32 * i;
                if (c < G) atomicAdd(dg + (size_t)b * G + c, acc[i]);
            }
        }
    }
}

int cond_all_bwd(const CondJobs& J, const CondGrads& D, const float* g, const float* dcond, float* dg, int B, int G, int N,
                 cudaStream_t st) {
    GTK_REQUIRE(g && dcond && J.n >= 1 && J.n <= COND_MAX_BLOCKS && D.n == J.n, "bad args");
    GTK_REQUIRE(G >= 1 && G <= 32 * CD_KPL, "gin_channels must be <= 512");
    if (B == 0 || N == 0) return GTK_OK;
    const size_t smem = ((size_t)B * G + (size_t)B * CD_ROWS + (size_t)CD_ROWS * G) * sizeof(float);
    GTK_REQUIRE(smem <= 220 * 1024, "batch x gin_channels too large for the conditioning kernel");
    GTK_CUDA(cudaFuncSetAttribute(cond_all_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ProfScope prof("cond_all_bwd", st);
    cond_all_bwd_kernel<<<dim3(ceil_div(N, CD_ROWS), J.n), CD_NT, smem, st>>>(J, D, g, dcond, dg, B, G, N);
    return check_launch("cond_all_bwd");
}

}  // namespace flow
}  // namespace gtk
