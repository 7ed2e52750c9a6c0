// glowtts_b200/csrc/flow_pointwise.cu -- memory-bound pieces of the flow decoder (fp32, [B,C,T], t fastest).
//
//   squeeze / unsqueeze        commons.py:339-364
//   mix (ActNorm + InvConvNear fused, fwd / reverse / bwd)   modules.py:584-599, 635-665
//   invconv_prep (W^-1, logdet W of the n_split x n_split matrix, on device)  modules.py:651,658,668
//   flow_logdet_const (+bwd)   logdet terms that only depend on parameters and lengths
//   actnorm_ddi_stats          modules.py:607-619
//   gate fwd / bwd             commons.py:61-68 (fused_add_tanh_sigmoid_multiply)
//   coupling affine fwd / rev / bwd   attentions.py:172-183
//   resskip, cat_mask, channel_sum    glue of WN (modules.py:165-171) and its backward
//   weightnorm fwd / bwd       torch.nn.utils.weight_norm as used at modules.py:124,131,141
//   mle_loss fwd / bwd         commons.py:28-33
//
// All kernels are grid-stride over (b, c, t) with t innermost so that a warp touches
// consecutive addresses; reductions go warp-shuffle -> shared -> one atomicAdd per block.
#include "common.cuh"
#include "ops.h"

namespace gtk {
namespace flow {

constexpr int PW_NT = 256;

static inline int pw_grid(long long work, int per_block = PW_NT) {
    long long g = ceil_div_ll(work, per_block);
    long long cap = (long long)sm_count() * 16;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

// ------------------------------------------------------------------------------------------ squeeze
// out[b, s*C + c, tq] = x[b, c, n*tq + s] * mask[b, n*tq + n-1];  mask_out[b, tq] = mask[b, n*tq + n-1]
__global__ void squeeze_kernel(const float* __restrict__ x, const float* __restrict__ mask, float* __restrict__ out,
                               float* __restrict__ mask_out, int B, int C, int T, int n, int msq) {
    const int Tq = T / n;
    const long long total = (long long)B * C * n * Tq;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int tq = (int)(i % Tq);
        long long r = i / Tq;
        const int cs = (int)(r % (C * n));
        const int b = (int)(r / (C * n));
        const int s = cs / C, c = cs % C;
        const float m = mask ? (msq ? mask[(size_t)b * Tq + tq] : mask[(size_t)b * T + n * tq + n - 1]) : 1.f;
        out[i] = x[((size_t)b * C + c) * T + n * tq + s] * m;
        if (mask_out && cs == 0) mask_out[(size_t)b * Tq + tq] = m;
    }
}

// out[b, c, n*t + s] = x[b, s*Cq + c, t] * mask[b, t];  mask_out[b, n*t + s] = mask[b, t]
__global__ void unsqueeze_kernel(const float* __restrict__ x, const float* __restrict__ mask, float* __restrict__ out,
                                 float* __restrict__ mask_out, int B, int C, int T, int n, int out_T) {
    const int Cq = C / n, To = T * n;
    const long long total = (long long)B * Cq * To;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int to = (int)(i % To);
        long long r = i / To;
        const int c = (int)(r % Cq);
        const int b = (int)(r / Cq);
        const int t = to / n, s = to % n;
        const float m = mask ? mask[(size_t)b * T + t] : 1.f;
        out[((size_t)b * Cq + c) * out_T + to] = x[((size_t)b * C + s * Cq + c) * T + t] * m;
        if (mask_out && c == 0) mask_out[(size_t)b * To + to] = m;
    }
}

// n_sqz == 2 (every config of the reference): a thread owns the frame pair (2 tq, 2 tq + 1) of one channel -- one 64-bit
// access on the un-squeezed side, two coalesced 32-bit accesses on the squeezed side -- instead of stride-2 scalars.
// SQUEEZE: out[b, s*C + c, tq] = x[b, c, 2 tq + s] * m;  otherwise out[b, c, 2 t + s] = x[b, s*Cq + c, t] * m.
template <bool SQUEEZE>
__global__ void __launch_bounds__(PW_NT) sqz2_kernel(const float* __restrict__ x, const float* __restrict__ mask,
                                                     float* __restrict__ out, float* __restrict__ mask_out, int Cq, int Tq,
                                                     int T_long, int msq) {
    // Cq = channels on the un-squeezed side, Tq = squeezed frames, T_long = row pitch of the un-squeezed tensor (even)
    const int c = blockIdx.y, b = blockIdx.z;
    const float* m_row = mask ? (SQUEEZE && !msq ? mask + (size_t)b * T_long : mask + (size_t)b * Tq) : nullptr;
    for (int tq = blockIdx.x * blockDim.x + threadIdx.x; tq < Tq; tq += gridDim.x * blockDim.x) {
        const float m = m_row ? ((SQUEEZE && !msq) ? m_row[2 * tq + 1] : m_row[tq]) : 1.f;
        if (SQUEEZE) {
            const float2 v = *reinterpret_cast<const float2*>(x + ((size_t)b * Cq + c) * T_long + 2 * tq);
            out[((size_t)b * 2 * Cq + c) * Tq + tq] = v.x * m;
            out[((size_t)b * 2 * Cq + Cq + c) * Tq + tq] = v.y * m;
            if (mask_out && c == 0) mask_out[(size_t)b * Tq + tq] = m;
        } else {
            const float a = x[((size_t)b * 2 * Cq + c) * Tq + tq], d = x[((size_t)b * 2 * Cq + Cq + c) * Tq + tq];
            *reinterpret_cast<float2*>(out + ((size_t)b * Cq + c) * T_long + 2 * tq) = make_float2(a * m, d * m);
            if (mask_out && c == 0) *reinterpret_cast<float2*>(mask_out + (size_t)b * 2 * Tq + 2 * tq) = make_float2(m, m);
        }
    }
}
static inline bool sqz2_ok(int n, int T_long, const void* long_side, int B, int Cq) {
    return n == 2 && T_long % 2 == 0 && (reinterpret_cast<uintptr_t>(long_side) & 7) == 0 && Cq <= 65535 && B <= 65535;
}

int squeeze(const float* x, const float* mask, float* out, float* mask_out, int B, int C, int T, int n, int mask_is_squeezed,
            cudaStream_t st) {
    GTK_REQUIRE(x && out && n >= 1, "bad args");
    const long long total = (long long)B * C * n * (T / n);
    if (total == 0) return GTK_OK;
    if (sqz2_ok(n, T, x, B, C)) {
        const int Tq = T / 2;
        sqz2_kernel<true><<<dim3(ceil_div(Tq, PW_NT), C, B), PW_NT, 0, st>>>(x, mask, out, mask_out, C, Tq, T, mask_is_squeezed);
        return check_launch("squeeze");
    }
    squeeze_kernel<<<pw_grid(total), PW_NT, 0, st>>>(x, mask, out, mask_out, B, C, T, n, mask_is_squeezed);
    return check_launch("squeeze");
}
int unsqueeze(const float* x, const float* mask, float* out, float* mask_out, int B, int C, int T, int n, int out_T,
              cudaStream_t st) {
    if (out_T <= 0) out_T = T * n;
    GTK_REQUIRE(x && out && n >= 1 && C % n == 0, "bad args");
    const long long total = (long long)B * C * T;
    if (total == 0) return GTK_OK;
    if (sqz2_ok(n, out_T, out, B, C / 2) && out_T == 2 * T && (!mask_out || (reinterpret_cast<uintptr_t>(mask_out) & 7) == 0)) {
        sqz2_kernel<false><<<dim3(ceil_div(T, PW_NT), C / 2, B), PW_NT, 0, st>>>(x, mask, out, mask_out, C / 2, T, out_T, 1);
        return check_launch("unsqueeze");
    }
    unsqueeze_kernel<<<pw_grid(total), PW_NT, 0, st>>>(x, mask, out, mask_out, B, C, T, n, out_T);
    return check_launch("unsqueeze");
}

// ------------------------------------------------------------------------------------------ lengths
// x_len[b] = sum_t mask[b, t]   (modules.py:587, 642)
__global__ void mask_len_kernel(const float* __restrict__ mask, float* __restrict__ x_len, int T) {
    __shared__ float scratch[33];
    const int b = blockIdx.x;
    float s = 0.f;
    for (int t = threadIdx.x; t < T; t += PW_NT) s += mask[(size_t)b * T + t];
    s = block_sum<PW_NT>(s, scratch);
    if (threadIdx.x == 0) x_len[b] = s;
}
int mask_len(const float* mask, float* x_len, int B, int T, cudaStream_t st) {
    GTK_REQUIRE(mask && x_len, "null");
    if (B == 0) return GTK_OK;
    mask_len_kernel<<<B, PW_NT, 0, st>>>(mask, x_len, T);
    return check_launch("mask_len");
}

// ------------------------------------------------------------------------------------------ invconv prep
// Single thread: LU with partial pivoting of the NS x NS weight -> inverse and log(det) (NaN for det<0,
// like torch.logdet).  winv and logdet may be null.
__device__ void invconv_lu(const float* __restrict__ w, int ns, float* __restrict__ winv, float* __restrict__ logdet) {
    float a[8][16];
    for (int i = 0; i < ns; ++i)
        for (int j = 0; j < ns; ++j) {
            a[i][j] = w[i * ns + j];
            a[i][ns + j] = (i == j) ? 1.f : 0.f;
        }
    float det = 1.f;
    for (int c = 0; c < ns; ++c) {
        int piv = c;
        float best = fabsf(a[c][c]);
        for (int r = c + 1; r < ns; ++r)
            if (fabsf(a[r][c]) > best) { best = fabsf(a[r][c]); piv = r; }
        if (piv != c) {
            for (int j = 0; j < 2 * ns; ++j) { float t = a[c][j]; a[c][j] = a[piv][j]; a[piv][j] = t; }
            det = -det;
        }
        const float d = a[c][c];
        det *= d;
        const float inv = 1.f / d;
        for (int j = 0; j < 2 * ns; ++j) a[c][j] *= inv;
        for (int r = 0; r < ns; ++r) {
            if (r == c) continue;
            const float f = a[r][c];
            for (int j = 0; j < 2 * ns; ++j) a[r][j] -= f * a[c][j];
        }
    }
    if (winv)
        for (int i = 0; i < ns; ++i)
            for (int j = 0; j < ns; ++j) winv[i * ns + j] = a[i][ns + j];
    if (logdet) *logdet = logf(det);
}
__global__ void invconv_prep_kernel(const float* __restrict__ w, int ns, float* __restrict__ winv,
                                    float* __restrict__ logdet) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    invconv_lu(w, ns, winv, logdet);
}

// All flow blocks at once (the decoder's per-call constants): warp k inverts block k's InvConvNear weight and sums its
// ActNorm logs; then logdet[b] (+)= x_len[b] * sum_k (sum_c logs_kc + logdetW_k * C/NS), blocks added in index order.
__global__ void __launch_bounds__(1024) flow_consts_kernel(const FlowConstJobs J, int ns, int C, float group_count,
                                                           const float* __restrict__ x_len, float* __restrict__ logdet, int B,
                                                           int accumulate) {
    __shared__ float s_blk[FLOW_CONST_MAX_BLOCKS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int k = warp; k < J.n; k += blockDim.x >> 5) {
        float ld = 0.f;
        if (lane == 0) invconv_lu(J.w[k], ns, J.winv ? J.winv + (size_t)k * 64 : nullptr, &ld);
        float s = 0.f;
        if (J.logs[k])
            for (int c = lane; c < C; c += 32) s += J.logs[k][c];
        s = warp_sum(s);
        if (lane == 0) {
            s_blk[k] = s + ld * group_count;
            if (J.icld) J.icld[k] = ld;
        }
    }
    __syncthreads();
    if (!logdet) return;
    float tot = 0.f;
    for (int k = 0; k < J.n; ++k) tot += s_blk[k];
    for (int b = threadIdx.x; b < B; b += blockDim.x) {
        const float v = tot * x_len[b];
        logdet[b] = accumulate ? logdet[b] + v : v;
    }
}
int flow_consts(const FlowConstJobs& J, int ns, int C, const float* x_len, float* logdet, int B, int accumulate, cudaStream_t st) {
    GTK_REQUIRE(ns >= 1 && ns <= 8 && J.n >= 0 && J.n <= FLOW_CONST_MAX_BLOCKS, "bad args");
    GTK_REQUIRE(!logdet || x_len, "null");
    if (J.n == 0) return GTK_OK;
    const int threads = min(1024, max(32, J.n * 32));
    flow_consts_kernel<<<1, threads, 0, st>>>(J, ns, C, (float)(C / ns), x_len, logdet, B, accumulate);
    return check_launch("flow_consts");
}
int invconv_prep(const float* w, int ns, float* winv, float* logdet, cudaStream_t st) {
    GTK_REQUIRE(w && ns >= 1 && ns <= 8, "n_split must be in [1, 8]");
    invconv_prep_kernel<<<1, 32, 0, st>>>(w, ns, winv, logdet);
    return check_launch("invconv_prep");
}

// ------------------------------------------------------------------------------------------ mix
// Channel c = h*(C/2) + g*(NS/2) + k is entry j = h*(NS/2) + k of group g's NS-vector.
// forward : y_j = bias_j + exp(logs_j) x_j ;  z_i = mask * sum_j M[i][j] y_j
// reverse : v_i = sum_j M[i][j] x_j        ;  z_i = mask * (v_i - bias_i) * exp(-logs_i)
// logs/bias null => no ActNorm part; M null => no InvConvNear part.
template <int NS>
__device__ __forceinline__ int mix_channel(int g, int j, int C) {
    if constexpr (NS == 1) {
        return g;   // plain per-channel affine (ActNorm on its own, any channel count)
    } else {
        constexpr int HALF = NS / 2;
        return (j / HALF) * (C / 2) + g * HALF + (j % HALF);
    }
}

// VEC consecutive frames per thread (128-bit accesses when T % 4 == 0, 64-bit when T is even -- the squeezed T' = 430
// of the BASELINE configs -- scalar otherwise); one block row per group g, so the per-channel constants (e^logs, bias)
// are set up once per thread, and no 64-bit division runs per element.
template <int VEC> struct VecF;
template <> struct VecF<1> { float v[1]; };
template <> struct alignas(8) VecF<2> { float v[2]; };
template <> struct alignas(16) VecF<4> { float v[4]; };

template <int NS, bool REVERSE, int VEC>
__global__ void __launch_bounds__(PW_NT) mix_kernel(const float* __restrict__ x, const float* __restrict__ mask,
                                                    const float* __restrict__ logs, const float* __restrict__ bias,
                                                    const float* __restrict__ M, float* __restrict__ z, int B, int C, int T,
                                                    const int* __restrict__ n_rows) {
    __shared__ float sM[NS * NS];
    pdl_trigger();
    pdl_wait();
    // varlen packing (B == 1): rows at or behind n_rows rounded up to a whole 128-row tile hold nothing
    const int t_lim = n_rows ? ((*n_rows + 127) & ~127) : T;
    if (threadIdx.x < NS * NS) sM[threadIdx.x] = M ? M[threadIdx.x] : ((threadIdx.x / NS == threadIdx.x % NS) ? 1.f : 0.f);
    __syncthreads();
    const int g = blockIdx.y;
    int ch[NS];
    float e[NS], bs[NS];
#pragma unroll
    for (int j = 0; j < NS; ++j) {
        ch[j] = mix_channel<NS>(g, j, C);
        e[j] = logs ? expf(REVERSE ? -logs[ch[j]] : logs[ch[j]]) : 1.f;
        bs[j] = bias ? bias[ch[j]] : 0.f;
    }
    const int TV = T / VEC;
    const int total = B * TV;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int b = i / TV, t = (i - b * TV) * VEC;
        if (t >= t_lim) continue;
        VecF<VEC> m, v[NS], o[NS];
        if (mask) m = *reinterpret_cast<const VecF<VEC>*>(mask + (size_t)b * T + t);
        else {
#pragma unroll
            for (int u = 0; u < VEC; ++u) m.v[u] = 1.f;
        }
#pragma unroll
        for (int j = 0; j < NS; ++j) v[j] = *reinterpret_cast<const VecF<VEC>*>(x + ((size_t)b * C + ch[j]) * T + t);
#pragma unroll
        for (int u = 0; u < VEC; ++u) {
            float y[NS];
#pragma unroll
            for (int j = 0; j < NS; ++j) y[j] = (!REVERSE && logs) ? bs[j] + e[j] * v[j].v[u] : v[j].v[u];
#pragma unroll
            for (int a = 0; a < NS; ++a) {
                float acc = 0.f;
#pragma unroll
                for (int j = 0; j < NS; ++j) acc += sM[a * NS + j] * y[j];
                if (REVERSE && logs) acc = (acc - bs[a]) * e[a];
                o[a].v[u] = acc * m.v[u];
            }
        }
#pragma unroll
        for (int a = 0; a < NS; ++a) *reinterpret_cast<VecF<VEC>*>(z + ((size_t)b * C + ch[a]) * T + t) = o[a];
    }
}

// widest vector the row pitch and the base pointers allow
static inline int pw_vec(int T, const void* a, const void* b, const void* c = nullptr, const void* d = nullptr) {
    const uintptr_t bits = reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(c) |
                           reinterpret_cast<uintptr_t>(d);
    if (T % 4 == 0 && (bits & 15) == 0) return 4;
    if (T % 2 == 0 && (bits & 7) == 0) return 2;
    return 1;
}

template <int NS>
static int mix_launch(const float* x, const float* mask, const float* logs, const float* bias, const float* M, float* z,
                      int B, int C, int T, int reverse, const int* n_rows, cudaStream_t st) {
    const int G = C / NS;
    if ((long long)B * G * T == 0) return GTK_OK;
    GTK_REQUIRE((long long)B * T < (1ll << 31), "B * T too large");
    const int vec = pw_vec(T, x, z, mask);
    int gx = (int)ceil_div_ll((long long)B * (T / vec), PW_NT * 2);
    const int cap = max(1, sm_count() * 8 / G);
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    const dim3 grid(gx, G), block(PW_NT);
    ProfScope prof("mix", st);
#define GTK_MIX(R, V) GTK_CUDA(launch_pdl(mix_kernel<NS, R, V>, grid, block, 0, st, x, mask, logs, bias, M, z, B, C, T, n_rows))
    if (reverse) { if (vec == 4) GTK_MIX(true, 4); else if (vec == 2) GTK_MIX(true, 2); else GTK_MIX(true, 1); }
    else         { if (vec == 4) GTK_MIX(false, 4); else if (vec == 2) GTK_MIX(false, 2); else GTK_MIX(false, 1); }
#undef GTK_MIX
    return check_launch("mix");
}

int mix(const float* x, const float* mask, const float* logs, const float* bias, const float* M, int ns, float* z, int B,
        int C, int T, int reverse, cudaStream_t st, const int* n_rows) {
    GTK_REQUIRE(!n_rows || B == 1, "packed rows are one long sequence");
    GTK_REQUIRE(x && z, "null");
    GTK_REQUIRE((logs == nullptr) == (bias == nullptr), "logs and bias go together");
    GTK_REQUIRE(ns == 1 || ns == 2 || ns == 4 || ns == 8, "n_split must be 1 (no mix), 2, 4 or 8");
    GTK_REQUIRE(C % ns == 0, "channels % n_split != 0");
    switch (ns) {
        case 1: return mix_launch<1>(x, mask, logs, bias, M, z, B, C, T, reverse, n_rows, st);
        case 2: return mix_launch<2>(x, mask, logs, bias, M, z, B, C, T, reverse, n_rows, st);
        case 4: return mix_launch<4>(x, mask, logs, bias, M, z, B, C, T, reverse, n_rows, st);
        default: return mix_launch<8>(x, mask, logs, bias, M, z, B, C, T, reverse, n_rows, st);
    }
}

// Backward of the forward mix.  One block row per group g (blockIdx.y) so that the per-channel
// reductions (dlogs, dbias) and the NS x NS dM stay block-local until one atomicAdd per value.
// dx may alias dz.  dlogs/dbias/dM are ACCUMULATED into (caller zeroes them once per step).
template <int NS, int VEC>
__global__ void __launch_bounds__(PW_NT) mix_bwd_kernel(const float* __restrict__ x, const float* __restrict__ mask, const float* __restrict__ logs,
                               const float* __restrict__ bias, const float* __restrict__ M, const float* dz, float* dx,
                               float* __restrict__ dlogs, float* __restrict__ dbias, float* __restrict__ dM, int B, int C,
                               int T, const int* __restrict__ n_rows) {
    __shared__ float sM[NS * NS];
    __shared__ float red[(2 * NS + NS * NS) * (PW_NT / 32)];
    pdl_trigger();
    pdl_wait();
    const int t_lim = n_rows ? ((*n_rows + 127) & ~127) : T;      // varlen packing (B == 1): see mix_kernel
    if (threadIdx.x < NS * NS) sM[threadIdx.x] = M ? M[threadIdx.x] : ((threadIdx.x / NS == threadIdx.x % NS) ? 1.f : 0.f);
    __syncthreads();
    const int g = blockIdx.y;
    int ch[NS];
    float e[NS], bs[NS];
#pragma unroll
    for (int j = 0; j < NS; ++j) {
        ch[j] = mix_channel<NS>(g, j, C);
        e[j] = logs ? expf(logs[ch[j]]) : 1.f;
        bs[j] = bias ? bias[ch[j]] : 0.f;
    }
    float a_logs[NS], a_bias[NS], a_M[NS * NS];
#pragma unroll
    for (int j = 0; j < NS; ++j) a_logs[j] = a_bias[j] = 0.f;
#pragma unroll
    for (int j = 0; j < NS * NS; ++j) a_M[j] = 0.f;
    const int TV = T / VEC;
    const int total = B * TV;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int b = i / TV, t = (i - b * TV) * VEC;
        if (t >= t_lim) continue;
        VecF<VEC> m, xq[NS], gq[NS], oq[NS];
        if (mask) m = *reinterpret_cast<const VecF<VEC>*>(mask + (size_t)b * T + t);
        else {
#pragma unroll
            for (int u = 0; u < VEC; ++u) m.v[u] = 1.f;
        }
#pragma unroll
        for (int j = 0; j < NS; ++j) {
            const size_t off = ((size_t)b * C + ch[j]) * T + t;
            xq[j] = *reinterpret_cast<const VecF<VEC>*>(x + off);
            gq[j] = *reinterpret_cast<const VecF<VEC>*>(dz + off);
        }
#pragma unroll
        for (int u = 0; u < VEC; ++u) {
            float xv[NS], y[NS], gz[NS];
#pragma unroll
            for (int j = 0; j < NS; ++j) {
                xv[j] = xq[j].v[u];
                y[j] = bs[j] + e[j] * xv[j];
                gz[j] = gq[j].v[u] * m.v[u];
            }
#pragma unroll
            for (int j = 0; j < NS; ++j) {
                float dy = 0.f;
#pragma unroll
                for (int a = 0; a < NS; ++a) {
                    dy += sM[a * NS + j] * gz[a];
                    a_M[a * NS + j] += gz[a] * y[j];
                }
                a_bias[j] += dy;
                a_logs[j] += dy * e[j] * xv[j];
                oq[j].v[u] = dy * e[j];
            }
        }
#pragma unroll
        for (int j = 0; j < NS; ++j) *reinterpret_cast<VecF<VEC>*>(dx + ((size_t)b * C + ch[j]) * T + t) = oq[j];
    }
    // block reduction of 2*NS + NS*NS values
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int NV = 2 * NS + NS * NS;
#pragma unroll
    for (int j = 0; j < NS; ++j) {
        float v = warp_sum(a_logs[j]);
        if (lane == 0) red[j * (PW_NT / 32) + warp] = v;
        v = warp_sum(a_bias[j]);
        if (lane == 0) red[(NS + j) * (PW_NT / 32) + warp] = v;
    }
#pragma unroll
    for (int j = 0; j < NS * NS; ++j) {
        float v = warp_sum(a_M[j]);
        if (lane == 0) red[(2 * NS + j) * (PW_NT / 32) + warp] = v;
    }
    __syncthreads();
    if (threadIdx.x < NV) {
        float s = 0.f;
        for (int w = 0; w < PW_NT / 32; ++w) s += red[threadIdx.x * (PW_NT / 32) + w];
        const int j = threadIdx.x;
        if (j < NS) { if (dlogs) atomicAdd(&dlogs[mix_channel<NS>(g, j, C)], s); }
        else if (j < 2 * NS) { if (dbias) atomicAdd(&dbias[mix_channel<NS>(g, j - NS, C)], s); }
        else if (dM) atomicAdd(&dM[j - 2 * NS], s);
    }
}

int mix_bwd(const float* x, const float* mask, const float* logs, const float* bias, const float* M, int ns,
            const float* dz, float* dx, float* dlogs, float* dbias, float* dM, int B, int C, int T, cudaStream_t st,
            const int* n_rows) {
    GTK_REQUIRE(x && dz && dx, "null");
    GTK_REQUIRE(!n_rows || B == 1, "packed rows are one long sequence");
    GTK_REQUIRE(ns == 1 || ns == 2 || ns == 4 || ns == 8, "n_split must be 1 (no mix), 2, 4 or 8");
    GTK_REQUIRE(C % ns == 0, "channels % n_split != 0");
    if ((long long)B * C * T == 0) return GTK_OK;
    const int G = C / ns;
    GTK_REQUIRE((long long)B * T < (1ll << 31), "B * T too large");
    const int vec = pw_vec(T, x, dz, dx, mask);
    int gx = (int)ceil_div_ll((long long)B * (T / vec), PW_NT * 2);
    const int cap = max(1, sm_count() * 8 / G);
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    dim3 grid(gx, G);
    ProfScope prof("mix_bwd", st);
#define GTK_MIXB(N, V) GTK_CUDA(launch_pdl(mix_bwd_kernel<N, V>, grid, dim3(PW_NT), 0, st, x, mask, logs, bias, M, dz, dx, dlogs, dbias, dM, B, C, T, n_rows))
#define GTK_MIXB_V(N) do { if (vec == 4) GTK_MIXB(N, 4); else if (vec == 2) GTK_MIXB(N, 2); else GTK_MIXB(N, 1); } while (0)
    switch (ns) {
        case 1: GTK_MIXB_V(1); break;
        case 2: GTK_MIXB_V(2); break;
        case 4: GTK_MIXB_V(4); break;
        default: GTK_MIXB_V(8); break;
    }
#undef GTK_MIXB_V
#undef GTK_MIXB
    return check_launch("mix_bwd");
}

// ------------------------------------------------------------------------------------------ logdet consts
// logdet[b] (+)= (sum_c logs_c + logdetW * C/NS) * x_len[b]     (modules.py:597, 658)
__global__ void logdet_const_kernel(const float* __restrict__ logs, int C, const float* __restrict__ logdetW,
                                    float group_count, const float* __restrict__ x_len, float* __restrict__ logdet,
                                    int B, int accumulate) {
    __shared__ float scratch[33];
    float s = 0.f;
    if (logs)
        for (int c = threadIdx.x; c < C; c += PW_NT) s += logs[c];
    s = block_sum<PW_NT>(s, scratch);
    if (logdetW) s += (*logdetW) * group_count;
    for (int b = threadIdx.x; b < B; b += PW_NT) {
        const float v = s * x_len[b];
        logdet[b] = accumulate ? logdet[b] + v : v;
    }
}
int logdet_const(const float* logs, int C, const float* logdetW, float group_count, const float* x_len, float* logdet,
                 int B, int accumulate, cudaStream_t st) {
    GTK_REQUIRE(x_len && logdet, "null");
    if (B == 0) return GTK_OK;
    logdet_const_kernel<<<1, PW_NT, 0, st>>>(logs, C, logdetW, group_count, x_len, logdet, B, accumulate);
    return check_launch("logdet_const");
}

// d/dlogs_c += S, d/dW += S * group_count * W^-T   with S = sum_b dlogdet[b] * x_len[b]
__global__ void logdet_const_bwd_kernel(const float* __restrict__ dlogdet, const float* __restrict__ x_len, int B,
                                        float* __restrict__ dlogs, int C, const float* __restrict__ winv, int ns,
                                        float group_count, float* __restrict__ dW) {
    __shared__ float scratch[33];
    float s = 0.f;
    for (int b = threadIdx.x; b < B; b += PW_NT) s += dlogdet[b] * x_len[b];
    s = block_sum<PW_NT>(s, scratch);
    if (dlogs)
        for (int c = threadIdx.x; c < C; c += PW_NT) dlogs[c] += s;
    if (dW && winv && threadIdx.x < ns * ns) {
        const int i = threadIdx.x / ns, j = threadIdx.x % ns;
        dW[i * ns + j] += s * group_count * winv[j * ns + i];
    }
}
// The same for every block in one launch: block k adds to dlogs[k] / dW[k] (winv [k][64]).
__global__ void logdet_const_bwd_all_kernel(const float* __restrict__ dlogdet, const float* __restrict__ x_len, int B,
                                            const FlowConstGrads G, int C, const float* __restrict__ winv, int ns,
                                            float group_count) {
    __shared__ float scratch[33];
    const int k = blockIdx.x;
    float s = 0.f;
    for (int b = threadIdx.x; b < B; b += PW_NT) s += dlogdet[b] * x_len[b];
    s = block_sum<PW_NT>(s, scratch);
    if (G.dlogs[k])
        for (int c = threadIdx.x; c < C; c += PW_NT) G.dlogs[k][c] += s;
    if (G.dW[k] && threadIdx.x < ns * ns) {
        const int i = threadIdx.x / ns, j = threadIdx.x % ns;
        G.dW[k][i * ns + j] += s * group_count * winv[(size_t)k * 64 + j * ns + i];
    }
}
int logdet_const_bwd_all(const float* dlogdet, const float* x_len, int B, const FlowConstGrads& G, int C, const float* winv, int ns,
                         cudaStream_t st) {
    GTK_REQUIRE(dlogdet && x_len && winv && G.n >= 0 && G.n <= FLOW_CONST_MAX_BLOCKS, "bad args");
    if (G.n == 0) return GTK_OK;
    logdet_const_bwd_all_kernel<<<G.n, PW_NT, 0, st>>>(dlogdet, x_len, B, G, C, winv, ns, (float)(C / ns));
    return check_launch("logdet_const_bwd_all");
}
int logdet_const_bwd(const float* dlogdet, const float* x_len, int B, float* dlogs, int C, const float* winv, int ns,
                     float group_count, float* dW, cudaStream_t st) {
    GTK_REQUIRE(dlogdet && x_len, "null");
    logdet_const_bwd_kernel<<<1, PW_NT, 0, st>>>(dlogdet, x_len, B, dlogs, C, winv, ns, group_count, dW);
    return check_launch("logdet_const_bwd");
}

}  // namespace flow
}  // namespace gtk
