// glowtts_b200/csrc/logp.cu -- the Gaussian log-likelihood that feeds MAS (models.py:1076-1082) and
// commons.generate_path (commons.py:127-143), the inference-time neighbour of MAS.
//
// logp[b,x,y] = sum_d [ -0.5 log 2pi - logs - 0.5 m^2 s ](d,x)  +  sum_d (-0.5 s)(d,x) z^2(d,y)
//             + sum_d (m s)(d,x) z(d,y),   s = exp(-2 logs)
// i.e. one [T_x, 2D] x [2D, T_y] contraction per utterance plus a per-token constant; D = 80, so the
// kernel is bound by writing the [B,T_x,T_y] fp32 output (4 B/cell).  fp32 FFMA keeps it within fp32
// rounding of the reference's two matmuls.
#include "common.cuh"
#include "ops.h"

namespace gtk {
namespace logp {

constexpr int BX = 64, BY = 64, BD = 16, NT = 256;

// MEAN_ONLY (x_logs == NULL, every configs/base*.json): s == 1, so the z^2 contraction is the per-frame constant
// -0.5 sum_d z[d,y]^2 -- accumulated once per column instead of as a second full set of FMAs.
template <bool MEAN_ONLY>
__global__ void __launch_bounds__(NT) logp_kernel(const float* __restrict__ x_m, const float* __restrict__ x_logs,
                                                  const float* __restrict__ z, float* __restrict__ out, int D, int T_x,
                                                  int T_y) {
    __shared__ __align__(16) float P1[BD][BX];   // -0.5 s
    __shared__ __align__(16) float P2[BD][BX];   // m s
    __shared__ __align__(16) float P3[BD][BX];   // -0.5 log 2pi - logs - 0.5 m^2 s
    __shared__ float Z1[BD][BY];                 // z^2
    __shared__ float Z2[BD][BY];                 // z
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int y0 = blockIdx.x * BY, x0 = blockIdx.y * BX, b = blockIdx.z;
    const float* mb = x_m + (size_t)b * D * T_x;
    const float* lb = x_logs ? x_logs + (size_t)b * D * T_x : nullptr;
    const float* zb = z + (size_t)b * D * T_y;

    float acc[4][4], cst[4], zsq[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        cst[i] = 0.f;
        zsq[i] = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    }
    for (int d0 = 0; d0 < D; d0 += BD) {
        for (int e = tid; e < BD * BX; e += NT) {
            const int xx = e % BX, dd = e / BX;
            float p1 = 0.f, p2 = 0.f, p3 = 0.f;
            if (d0 + dd < D && x0 + xx < T_x) {
                const float m = mb[(size_t)(d0 + dd) * T_x + x0 + xx];
                const float l = lb ? lb[(size_t)(d0 + dd) * T_x + x0 + xx] : 0.f;
                const float s = expf(-2.f * l);
                p1 = -0.5f * s;
                p2 = m * s;
                p3 = -0.918938533204672742f - l - 0.5f * m * m * s;
            }
            P1[dd][xx] = p1; P2[dd][xx] = p2; P3[dd][xx] = p3;
        }
        for (int e = tid; e < BD * BY; e += NT) {
            const int yy = e % BY, dd = e / BY;
            float v = 0.f;
            if (d0 + dd < D && y0 + yy < T_y) v = zb[(size_t)(d0 + dd) * T_y + y0 + yy];
            Z1[dd][yy] = v * v;
            Z2[dd][yy] = v;
        }
        __syncthreads();
#pragma unroll
        for (int dd = 0; dd < BD; ++dd) {
            const float4 a1 = *reinterpret_cast<const float4*>(&P1[dd][ty * 4]);
            const float4 a2 = *reinterpret_cast<const float4*>(&P2[dd][ty * 4]);
            const float4 a3 = *reinterpret_cast<const float4*>(&P3[dd][ty * 4]);
            cst[0] += a3.x; cst[1] += a3.y; cst[2] += a3.z; cst[3] += a3.w;
            const float p1[4] = {a1.x, a1.y, a1.z, a1.w};
            const float p2[4] = {a2.x, a2.y, a2.z, a2.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float q2 = Z2[dd][tx + 16 * j];
                if (MEAN_ONLY) {
                    zsq[j] = fmaf(q2, q2, zsq[j]);
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[i][j] = fmaf(p2[i], q2, acc[i][j]);
                } else {
                    const float q1 = Z1[dd][tx + 16 * j];
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[i][j] += p1[i] * q1 + p2[i] * q2;
                }
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int x = x0 + ty * 4 + i;
        if (x >= T_x) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int y = y0 + tx + 16 * j;
            if (y < T_y) out[((size_t)b * T_x + x) * T_y + y] = acc[i][j] + (cst[i] - 0.5f * zsq[j]);
        }
    }
}

int forward(const float* x_m, const float* x_logs, const float* z, float* out, int B, int D, int T_x, int T_y,
            cudaStream_t st) {
    GTK_REQUIRE(x_m && z && out, "null");
    if ((long long)B * T_x * T_y == 0) return GTK_OK;
    dim3 grid(ceil_div(T_y, BY), ceil_div(T_x, BX), B);
    ProfScope prof("logp", st);
    if (x_logs) logp_kernel<false><<<grid, NT, 0, st>>>(x_m, x_logs, z, out, D, T_x, T_y);
    else        logp_kernel<true><<<grid, NT, 0, st>>>(x_m, x_logs, z, out, D, T_x, T_y);
    return check_launch("logp");
}

// commons.generate_path: cum = cumsum(duration); path[x,y] = (y < cum[x]) - (y < cum[x-1]), times mask.
// One block per utterance: serial-free inclusive scan over T_x in shared memory, then a coalesced fill.
__global__ void generate_path_kernel(const float* __restrict__ dur, const float* __restrict__ mask, float* __restrict__ path,
                                     int T_x, int T_y) {
    extern __shared__ float cum[];   // [T_x + 1]
    const int b = blockIdx.x;
    if (threadIdx.x == 0) {
        float s = 0.f;
        cum[0] = 0.f;
        for (int x = 0; x < T_x; ++x) { s += dur[(size_t)b * T_x + x]; cum[x + 1] = s; }   // same order as torch.cumsum
    }
    __syncthreads();
    const size_t plane = (size_t)T_x * T_y;
    for (size_t i = threadIdx.x; i < plane; i += blockDim.x) {
        const int y = (int)(i % T_y), x = (int)(i / T_y);
        const float fy = (float)y;
        const float v = ((fy < cum[x + 1]) ? 1.f : 0.f) - ((fy < cum[x]) ? 1.f : 0.f);
        path[(size_t)b * plane + i] = v * (mask ? mask[(size_t)b * plane + i] : 1.f);
    }
}
int generate_path(const float* dur, const float* mask, float* path, int B, int T_x, int T_y, cudaStream_t st) {
    GTK_REQUIRE(dur && path, "null");
    if ((long long)B * T_x * T_y == 0) return GTK_OK;
    const size_t smem = (size_t)(T_x + 1) * sizeof(float);
    GTK_REQUIRE(smem <= 48 * 1024, "T_x too large");
    generate_path_kernel<<<B, 512, smem, st>>>(dur, mask, path, T_x, T_y);
    return check_launch("generate_path");
}

}  // namespace logp
}  // namespace gtk
