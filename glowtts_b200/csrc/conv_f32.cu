// glowtts_b200/csrc/conv_f32.cu -- fp32 (CUDA-core) conv1d family: the exact-parity path.
//
// Covers every dense contraction of the coupling block in full fp32 so that flow outputs and
// gradients match the reference within 1e-4 (BASELINE.json): WN in_layers (k=5 dilated conv,
// modules.py:129-132,152), res_skip 1x1 (modules.py:140-142,165), start / end 1x1
// (attentions.py:102-110,151,170), their data gradients (same kernel on transposed+flipped
// weights) and their weight gradients.  The bf16 tcgen05 path (conv_tc.cu) is the fast path and
// is checked against this one on the GPU as well as against the oracle.
//
// Layout: activations [B, C, T] (t fastest, as the reference); weights are consumed PACKED as
// Wp[c][j][m] (m fastest) so that tile loads are coalesced and shared-memory stores conflict-free.
#include "common.cuh"
#include "ops.h"

namespace gtk {
namespace conv {

constexpr int BM = 64, BT = 64, BC = 16, NT = 256;

// W[M][Cin][K] (torch layout) -> Wp[Cin][K][M]                      (forward operand)
// W[M][Cin][K] -> Wp[M][K][Cin] with taps flipped                   (data-gradient operand: conv^T)
__global__ void pack_weight_kernel(const float* __restrict__ w, float* __restrict__ wp, int M, int Cin, int K, int transpose) {
    const long long total = (long long)M * Cin * K;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int j = (int)(i % K);
        long long r = i / K;
        const int c = (int)(r % Cin), m = (int)(r / Cin);
        if (!transpose) wp[((size_t)c * K + j) * M + m] = w[i];
        else            wp[((size_t)m * K + (K - 1 - j)) * Cin + c] = w[i];
    }
}
int pack_weight(const float* w, float* wp, int M, int Cin, int K, int transpose, cudaStream_t st) {
    GTK_REQUIRE(w && wp, "null");
    const long long total = (long long)M * Cin * K;
    if (total == 0) return GTK_OK;
    long long g = ceil_div_ll(total, 256);
    int grid = (int)(g > 1024 ? 1024 : g);
    pack_weight_kernel<<<grid, 256, 0, st>>>(w, wp, M, Cin, K, transpose);
    return check_launch("pack_weight");
}

__global__ void __launch_bounds__(NT) conv1d_f32_kernel(const ConvParams p) {
    extern __shared__ __align__(16) float smem[];
    const int K = p.K, dil = p.dil;
    const int halo = ((K - 1) / 2) * dil;
    const int BW = BT + (K - 1) * dil;
    float* As = smem;                       // [BC][K][BM]
    float* Bs = smem + BC * K * BM;         // [BC][BW]
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int t0 = blockIdx.x * BT, m0 = blockIdx.y * BM, b = blockIdx.z;
    const float* inb = p.in + (size_t)b * (p.in_bstride ? p.in_bstride : (size_t)p.Cin * p.T);
    const float* maskb = p.mask ? p.mask + (size_t)b * p.T : nullptr;

    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int c0 = 0; c0 < p.Cin; c0 += BC) {
        // A tile: rows (c, j), BM consecutive m each
        for (int e = tid; e < BC * K * BM; e += NT) {
            const int m = e % BM, cj = e / BM;
            const int c = c0 + cj / K;
            float v = 0.f;
            if (c < p.Cin && m0 + m < p.M) v = p.wp[((size_t)c * K + (cj % K)) * p.M + m0 + m];
            As[e] = v;
        }
        for (int e = tid; e < BC * BW; e += NT) {
            const int w = e % BW, c = c0 + e / BW;
            const int t = t0 - halo + w;
            float v = 0.f;
            if (c < p.Cin && t >= 0 && t < p.T) {
                v = inb[(size_t)c * p.T + t];
                if (p.mask_in && maskb) v *= maskb[t];
            }
            Bs[e] = v;
        }
        __syncthreads();
        for (int c = 0; c < BC; ++c) {
            for (int j = 0; j < K; ++j) {
                const float4 a = *reinterpret_cast<const float4*>(&As[(c * K + j) * BM + ty * 4]);
                const float* brow = &Bs[c * BW + j * dil + tx];
                const float b0 = brow[0], b1 = brow[16], b2 = brow[32], b3 = brow[48];
                acc[0][0] += a.x * b0; acc[0][1] += a.x * b1; acc[0][2] += a.x * b2; acc[0][3] += a.x * b3;
                acc[1][0] += a.y * b0; acc[1][1] += a.y * b1; acc[1][2] += a.y * b2; acc[1][3] += a.y * b3;
                acc[2][0] += a.z * b0; acc[2][1] += a.z * b1; acc[2][2] += a.z * b2; acc[2][3] += a.z * b3;
                acc[3][0] += a.w * b0; acc[3][1] += a.w * b1; acc[3][2] += a.w * b2; acc[3][3] += a.w * b3;
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = m0 + ty * 4 + i;
        if (m >= p.M) continue;
        const float bs = p.bias ? p.bias[m] : 0.f;
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
            const int t = t0 + tx + 16 * jj;
            if (t >= p.T) continue;
            const size_t o = (size_t)b * (p.out_bstride ? p.out_bstride : (size_t)p.M * p.T) + (size_t)m * p.T + t;
            const float mk = maskb ? maskb[t] : 1.f;
            float v = acc[i][jj] + bs;
            if (p.resid)
                v += p.resid[(size_t)b * (p.resid_bstride ? p.resid_bstride : (size_t)p.M * p.T) + (size_t)m * p.T + t] *
                     (p.mask_resid ? mk : 1.f);
            if (p.mask_out) v *= mk;
            p.out[o] = v;
        }
    }
}

int conv1d_f32(const ConvParams& p, cudaStream_t st) {
    GTK_REQUIRE(p.in && p.wp && p.out, "null");
    GTK_REQUIRE(p.K >= 1 && (p.K % 2) == 1 && p.dil >= 1, "kernel size must be odd, dilation >= 1");
    if ((long long)p.B * p.M * p.T == 0) return GTK_OK;
    const int BW = BT + (p.K - 1) * p.dil;
    const size_t smem = ((size_t)BC * p.K * BM + (size_t)BC * BW) * sizeof(float);
    GTK_REQUIRE(smem <= 200 * 1024, "kernel_size * dilation too large for the fp32 conv tile");
    if (smem > 48 * 1024)
        GTK_CUDA(cudaFuncSetAttribute(conv1d_f32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(ceil_div(p.T, BT), ceil_div(p.M, BM), p.B);
    ProfScope prof(p.K > 1 ? "conv_f32_kN" : "conv_f32_k1", st);
    conv1d_f32_kernel<<<grid, NT, smem, st>>>(p);
    return check_launch("conv1d_f32");
}

// ------------------------------------------------------------------------------------------ weight gradient
// dW[m][c][j] += sum_{b,t} dout[b,m,t] * in[b,c,t + (j - (K-1)/2) * dil]     (torch layout [M][Cin][K])
// Block = 64 m x 16 c x K taps; the (b, t-chunk) reduction is split over blockIdx.z and finished with
// atomicAdd, so dW must be zeroed by the caller once per step.
template <int K>
__global__ void __launch_bounds__(NT) wgrad_f32_kernel(const WgradParams p) {
    extern __shared__ __align__(16) float smem[];
    constexpr int TT = 64, DP = TT + 1;
    const int dil = p.dil, halo = ((K - 1) / 2) * dil;
    const int IW = TT + (K - 1) * dil;
    float* Ds = smem;              // [BM][DP]
    float* Is = smem + BM * DP;    // [BC][IW]
    const int tid = threadIdx.x, ml = tid & 63, cq = tid >> 6;
    const int m0 = blockIdx.x * BM, c0 = blockIdx.y * BC;
    const int chunks_t = ceil_div(p.T, TT);
    const int n_chunks = p.B * chunks_t;

    float acc[4][K];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < K; ++j) acc[i][j] = 0.f;

    for (int ch = blockIdx.z; ch < n_chunks; ch += gridDim.z) {
        const int b = ch / chunks_t, t0 = (ch % chunks_t) * TT;
        const float* maskb = p.mask ? p.mask + (size_t)b * p.T : nullptr;
        for (int e = tid; e < BM * TT; e += NT) {
            const int t = e % TT, m = e / TT;
            float v = 0.f;
            if (m0 + m < p.M && t0 + t < p.T) {
                v = p.dout[(size_t)b * (p.dout_bstride ? p.dout_bstride : (size_t)p.M * p.T) + (size_t)(m0 + m) * p.T + t0 + t];
                if (p.mask_dout && maskb) v *= maskb[t0 + t];
            }
            Ds[m * DP + t] = v;
        }
        for (int e = tid; e < BC * IW; e += NT) {
            const int w = e % IW, c = e / IW;
            const int t = t0 - halo + w;
            float v = 0.f;
            if (c0 + c < p.Cin && t >= 0 && t < p.T) {
                v = p.in[(size_t)b * (p.in_bstride ? p.in_bstride : (size_t)p.Cin * p.T) + (size_t)(c0 + c) * p.T + t];
                if (p.mask_in && maskb) v *= maskb[t];
            }
            Is[c * IW + w] = v;
        }
        __syncthreads();
        for (int t = 0; t < TT; ++t) {
            const float d = Ds[ml * DP + t];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < K; ++j) acc[i][j] += d * Is[(cq * 4 + i) * IW + t + j * dil];
        }
        __syncthreads();
    }
    const int m = m0 + ml;
    if (m < p.M) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int c = c0 + cq * 4 + i;
            if (c >= p.Cin) continue;
#pragma unroll
            for (int j = 0; j < K; ++j) atomicAdd(&p.dw[((size_t)m * p.Cin + c) * K + j], acc[i][j]);
        }
    }
}

template <int K>
static int wgrad_launch(const WgradParams& p, cudaStream_t st) {
    const int IW = 64 + (K - 1) * p.dil;
    const size_t smem = ((size_t)BM * 65 + (size_t)BC * IW) * sizeof(float);
    GTK_REQUIRE(smem <= 200 * 1024, "dilation too large for the fp32 wgrad tile");
    if (smem > 48 * 1024)
        GTK_CUDA(cudaFuncSetAttribute(wgrad_f32_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int gx = ceil_div(p.M, BM), gy = ceil_div(p.Cin, BC);
    const int n_chunks = p.B * ceil_div(p.T, 64);
    int gz = max(1, min(n_chunks, (sm_count() * 4) / max(1, gx * gy)));
    ProfScope prof(K > 1 ? "wgrad_f32_kN" : "wgrad_f32_k1", st);
    wgrad_f32_kernel<K><<<dim3(gx, gy, gz), NT, smem, st>>>(p);
    return check_launch("wgrad_f32");
}

int wgrad_f32(const WgradParams& p, int K, cudaStream_t st) {
    GTK_REQUIRE(p.dout && p.in && p.dw, "null");
    GTK_REQUIRE(p.dil >= 1, "dilation >= 1");
    if ((long long)p.M * p.Cin * K == 0 || (long long)p.B * p.T == 0) return GTK_OK;
    switch (K) {
        case 1: return wgrad_launch<1>(p, st);
        case 3: return wgrad_launch<3>(p, st);
        case 5: return wgrad_launch<5>(p, st);
        case 7: return wgrad_launch<7>(p, st);
        case 9: return wgrad_launch<9>(p, st);
        default:
            set_error("wgrad_f32: kernel_size %d not instantiated (1,3,5,7,9)", K);
            return GTK_ERR_UNSUPPORTED;
    }
}

}  // namespace conv
}  // namespace gtk
