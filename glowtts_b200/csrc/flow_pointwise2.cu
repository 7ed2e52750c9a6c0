// glowtts_b200/csrc/flow_pointwise2.cu -- second half of the memory-bound flow kernels (see flow_pointwise.cu).
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

__device__ __forceinline__ float sigmoidf_(float v) { return 1.f / (1.f + expf(-v)); }

// ------------------------------------------------------------------------------------------ ActNorm DDI
// stats[0][c] = sum mask, stats[1][c] = sum x*mask, stats[2][c] = sum x*x*mask   (modules.py:609-611)
__global__ void ddi_stats_kernel(const float* __restrict__ x, const float* __restrict__ mask, float* __restrict__ stats,
                                 int B, int C, int T) {
    __shared__ float scratch[33];
    const int c = blockIdx.x;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f;
    for (long long i = threadIdx.x; i < (long long)B * T; i += PW_NT) {
        const int b = (int)(i / T), t = (int)(i % T);
        const float m = mask ? mask[(size_t)b * T + t] : 1.f;
        const float v = x[((size_t)b * C + c) * T + t];
        s0 += m; s1 += v * m; s2 += v * v * m;
    }
    s0 = block_sum<PW_NT>(s0, scratch);
    s1 = block_sum<PW_NT>(s1, scratch);
    s2 = block_sum<PW_NT>(s2, scratch);
    if (threadIdx.x == 0) { stats[c] = s0; stats[C + c] = s1; stats[2 * C + c] = s2; }
}
// logs = -0.5*log(max(var,1e-6)), bias = -mean*exp(logs)   (modules.py:612-619)
__global__ void ddi_finalize_kernel(const float* __restrict__ stats, float* __restrict__ logs, float* __restrict__ bias, int C) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float denom = stats[c];
    const float mean = stats[C + c] / denom;
    const float var = stats[2 * C + c] / denom - mean * mean;
    const float hl = 0.5f * logf(fmaxf(var, 1e-6f));
    logs[c] = -hl;
    bias[c] = -mean * expf(-hl);
}
int actnorm_ddi_stats(const float* x, const float* mask, float* stats, int B, int C, int T, cudaStream_t st) {
    GTK_REQUIRE(x && stats && C > 0, "bad args");
    ddi_stats_kernel<<<C, PW_NT, 0, st>>>(x, mask, stats, B, C, T);
    return check_launch("ddi_stats");
}
int actnorm_ddi_finalize(const float* stats, float* logs, float* bias, int C, cudaStream_t st) {
    GTK_REQUIRE(stats && logs && bias && C > 0, "bad args");
    ddi_finalize_kernel<<<ceil_div(C, 128), 128, 0, st>>>(stats, logs, bias, C);
    return check_launch("ddi_finalize");
}

// ------------------------------------------------------------------------------------------ gate
// cond_mode: 0 none, 1 cond[B,2H] broadcast over t, 2 cond[B,2H,T]
__global__ void gate_fwd_kernel(const float* __restrict__ a, const float* __restrict__ cond, int cond_mode, size_t cbs,
                                float* __restrict__ out, int B, int H, int T, float drop_p, DropSeed dseed) {
    const unsigned long long seed = drop_p > 0.f ? drop_resolve(dseed) : 0ull;
    const long long total = (long long)B * H * T;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int t = (int)(i % T);
        long long r = i / T;
        const int h = (int)(r % H), b = (int)(r / H);
        const size_t o1 = ((size_t)b * 2 * H + h) * T + t, o2 = o1 + (size_t)H * T;
        float s1 = a[o1], s2 = a[o2];
        if (drop_p > 0.f) {                                  // nn.Dropout on the in_layer output (modules.py:153)
            const unsigned long long e = ((unsigned long long)b * T + t) * (2ull * H) + h;
            s1 *= dropout_scale1(seed, e, drop_p);
            s2 *= dropout_scale1(seed, e + H, drop_p);
        }
        if (cond_mode == 1) { s1 += cond[(size_t)b * cbs + h]; s2 += cond[(size_t)b * cbs + H + h]; }
        else if (cond_mode == 2) { s1 += cond[o1]; s2 += cond[o2]; }
        out[i] = tanhf(s1) * sigmoidf_(s2);
    }
}
// scale_out = 1: da = d(in_layer output) (dropout-scaled); 0: da = d(gate pre-activation) (what the conditioning sees).
__global__ void gate_bwd_kernel(const float* __restrict__ a, const float* __restrict__ cond, int cond_mode, size_t cbs,
                                const float* __restrict__ dout, float* __restrict__ da, int B, int H, int T, float drop_p,
                                DropSeed dseed, int scale_out) {
    const unsigned long long seed = drop_p > 0.f ? drop_resolve(dseed) : 0ull;
    const long long total = (long long)B * H * T;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int t = (int)(i % T);
        long long r = i / T;
        const int h = (int)(r % H), b = (int)(r / H);
        const size_t o1 = ((size_t)b * 2 * H + h) * T + t, o2 = o1 + (size_t)H * T;
        float s1 = a[o1], s2 = a[o2];
        float k1 = 1.f, k2 = 1.f;
        if (drop_p > 0.f) {
            const unsigned long long e = ((unsigned long long)b * T + t) * (2ull * H) + h;
            k1 = dropout_scale1(seed, e, drop_p);
            k2 = dropout_scale1(seed, e + H, drop_p);
            s1 *= k1; s2 *= k2;
            if (!scale_out) k1 = k2 = 1.f;
        }
        if (cond_mode == 1) { s1 += cond[(size_t)b * cbs + h]; s2 += cond[(size_t)b * cbs + H + h]; }
        else if (cond_mode == 2) { s1 += cond[o1]; s2 += cond[o2]; }
        const float th = tanhf(s1), sg = sigmoidf_(s2), g = dout[i];
        da[o1] = g * sg * (1.f - th * th) * k1;
        da[o2] = g * th * sg * (1.f - sg) * k2;
    }
}
int gate_fwd(const float* a, const float* cond, int cond_mode, size_t cond_bstride, float* out, int B, int H, int T,
             float drop_p, DropSeed seed, cudaStream_t st) {
    if (cond_bstride == 0) cond_bstride = 2 * (size_t)H;
    GTK_REQUIRE(a && out && (cond_mode == 0 || cond), "bad args");
    GTK_REQUIRE(drop_p >= 0.f && drop_p < 1.f, "dropout probability");
    const long long total = (long long)B * H * T;
    if (total == 0) return GTK_OK;
    gate_fwd_kernel<<<pw_grid(total), PW_NT, 0, st>>>(a, cond, cond_mode, cond_bstride, out, B, H, T, drop_p, seed);
    return check_launch("gate_fwd");
}
int gate_bwd(const float* a, const float* cond, int cond_mode, size_t cond_bstride, const float* dout, float* da, int B, int H,
             int T, float drop_p, DropSeed seed, int scale_out, cudaStream_t st) {
    if (cond_bstride == 0) cond_bstride = 2 * (size_t)H;
    GTK_REQUIRE(a && dout && da && (cond_mode == 0 || cond), "bad args");
    GTK_REQUIRE(drop_p >= 0.f && drop_p < 1.f, "dropout probability");
    const long long total = (long long)B * H * T;
    if (total == 0) return GTK_OK;
    gate_bwd_kernel<<<pw_grid(total), PW_NT, 0, st>>>(a, cond, cond_mode, cond_bstride, dout, da, B, H, T, drop_p, seed,
                                                       scale_out);
    return check_launch("gate_bwd");
}
// scale[i] = dropout keep-scale of element i (0 or 1/(1-p)) for i in [0, n): the mask the decoder kernels apply.
__global__ void dropout_scale_kernel(unsigned long long seed, float p, long long n, float* __restrict__ out) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = dropout_scale1(seed, (unsigned long long)i, p);
}
int dropout_scale(unsigned long long seed, float p, long long n, float* out, cudaStream_t st) {
    GTK_REQUIRE(n >= 0 && (n == 0 || out) && p >= 0.f && p < 1.f, "bad args");
    if (n == 0) return GTK_OK;
    dropout_scale_kernel<<<pw_grid(n), PW_NT, 0, st>>>(seed, p, n, out);
    return check_launch("dropout_scale");
}

// ------------------------------------------------------------------------------------------ reductions
// out[r] (+)= sum_t in[r, t]
__global__ void row_sum_kernel(const float* __restrict__ in, float* __restrict__ out, int T, int accumulate, int rpg,
                               size_t ogs) {
    __shared__ float scratch[33];
    const size_t r = blockIdx.x;
    float s = 0.f;
    for (int t = threadIdx.x; t < T; t += 128) s += in[r * T + t];
    s = block_sum<128>(s, scratch);
    if (threadIdx.x == 0) {
        const size_t o = (r / rpg) * ogs + (r % rpg);
        out[o] = accumulate ? out[o] + s : s;
    }
}
// out index of row r: (r / rows_per_group) * out_group_stride + r % rows_per_group  (0,0 => out[r])
int row_sum(const float* in, float* out, int R, int T, int accumulate, int rows_per_group, size_t out_group_stride,
            cudaStream_t st) {
    if (rows_per_group <= 0) { rows_per_group = R > 0 ? R : 1; out_group_stride = 0; }
    GTK_REQUIRE(in && out, "null");
    if (R == 0) return GTK_OK;
    row_sum_kernel<<<R, 128, 0, st>>>(in, out, T, accumulate, rows_per_group, out_group_stride);
    return check_launch("row_sum");
}
// out[m] (+)= sum_{b,t} in[b, m, t]      (bias gradients)
__global__ void channel_sum_kernel(const float* __restrict__ in, const float* __restrict__ mask, float* __restrict__ out, int B,
                                   int M, int T, int accumulate) {
    __shared__ float scratch[33];
    const int m = blockIdx.x;
    float s = 0.f;
    for (long long i = threadIdx.x; i < (long long)B * T; i += PW_NT) {
        const int b = (int)(i / T), t = (int)(i % T);
        s += in[((size_t)b * M + m) * T + t] * (mask ? mask[(size_t)b * T + t] : 1.f);
    }
    s = block_sum<PW_NT>(s, scratch);
    if (threadIdx.x == 0) out[m] = accumulate ? out[m] + s : s;
}
int channel_sum(const float* in, const float* mask, float* out, int B, int M, int T, int accumulate, cudaStream_t st) {
    GTK_REQUIRE(in && out, "null");
    if (M == 0) return GTK_OK;
    channel_sum_kernel<<<M, PW_NT, 0, st>>>(in, mask, out, B, M, T, accumulate);
    return check_launch("channel_sum");
}

// ------------------------------------------------------------------------------------------ coupling affine
// end_out = [m | logs] (C channels), x1 = x[:, C/2:].  mode 0 forward, 1 reverse.   attentions.py:172-183
__global__ void affine_kernel(const float* __restrict__ x, const float* __restrict__ end_out,
                              const float* __restrict__ mask, float* __restrict__ z, float* __restrict__ logdet, int B,
                              int C, int T, int reverse, int sigmoid_scale) {
    __shared__ float scratch[33];
    const int b = blockIdx.y;
    const int Ch = C / 2;
    float ld = 0.f;
    const long long total = (long long)Ch * T;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int t = (int)(i % T), c = (int)(i / T);
        const float mk = mask ? mask[(size_t)b * T + t] : 1.f;
        const size_t o0 = ((size_t)b * C + c) * T + t, o1 = o0 + (size_t)Ch * T;
        const float m = end_out[o0];
        float logs = end_out[o1];
        if (sigmoid_scale) logs = logf(1e-6f + sigmoidf_(logs + 2.f));
        const float x1 = x[o1];
        z[o0] = x[o0];
        if (reverse) {
            z[o1] = (x1 - m) * expf(-logs) * mk;
        } else {
            z[o1] = (m + expf(logs) * x1) * mk;
            ld += logs * mk;
        }
    }
    if (!reverse && logdet) {
        ld = block_sum<PW_NT>(ld, scratch);
        if (threadIdx.x == 0) atomicAdd(&logdet[b], ld);
    }
}
int affine(const float* x, const float* end_out, const float* mask, float* z, float* logdet, int B, int C, int T,
           int reverse, int sigmoid_scale, cudaStream_t st) {
    GTK_REQUIRE(x && end_out && z && C % 2 == 0, "bad args");
    if ((long long)B * C * T == 0) return GTK_OK;
    int gx = (int)ceil_div_ll((long long)(C / 2) * T, PW_NT * 4);
    gx = max(1, min(gx, max(1, sm_count() * 8 / max(B, 1))));
    affine_kernel<<<dim3(gx, B), PW_NT, 0, st>>>(x, end_out, mask, z, logdet, B, C, T, reverse, sigmoid_scale);
    return check_launch("affine");
}
// dz -> d(end_out) = [dm | dlogs] and dx1 (written into dx[:, C/2:]); dx[:, :C/2] is left to the
// start-conv dgrad, which adds dz0 as its residual.
__global__ void affine_bwd_kernel(const float* __restrict__ x, const float* __restrict__ end_out,
                                  const float* __restrict__ mask, const float* __restrict__ dz,
                                  const float* __restrict__ dlogdet, float* __restrict__ dend, float* __restrict__ dx,
                                  int B, int C, int T, int sigmoid_scale) {
    const int Ch = C / 2;
    const long long total = (long long)B * Ch * T;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int t = (int)(i % T);
        long long r = i / T;
        const int c = (int)(r % Ch), b = (int)(r / Ch);
        const float mk = mask ? mask[(size_t)b * T + t] : 1.f;
        const size_t o0 = ((size_t)b * C + c) * T + t, o1 = o0 + (size_t)Ch * T;
        const float raw = end_out[o1];
        float logs = raw, dscale = 1.f;
        if (sigmoid_scale) {
            const float sg = sigmoidf_(raw + 2.f);
            logs = logf(1e-6f + sg);
            dscale = sg * (1.f - sg) / (1e-6f + sg);
        }
        const float e = expf(logs);
        const float g = dz[o1] * mk;
        dend[o0] = g;
        dend[o1] = (g * e * x[o1] + (dlogdet ? dlogdet[b] * mk : 0.f)) * dscale;
        dx[o1] = g * e;
    }
}
int affine_bwd(const float* x, const float* end_out, const float* mask, const float* dz, const float* dlogdet,
               float* dend, float* dx, int B, int C, int T, int sigmoid_scale, cudaStream_t st) {
    GTK_REQUIRE(x && end_out && dz && dend && dx && C % 2 == 0, "bad args");
    const long long total = (long long)B * (C / 2) * T;
    if (total == 0) return GTK_OK;
    affine_bwd_kernel<<<pw_grid(total), PW_NT, 0, st>>>(x, end_out, mask, dz, dlogdet, dend, dx, B, C, T, sigmoid_scale);
    return check_launch("affine_bwd");
}

// ------------------------------------------------------------------------------------------ WN glue
// rs [B,RS,T]: RS == 2H: x_out = (x + rs[:, :H]) * mask ; skip (+)= rs[:, H:]
//              RS ==  H: skip (+)= rs                                   (modules.py:166-171)
// mask_skip: multiply the updated skip by the mask (the WN's final `output * x_mask`).
__global__ void resskip_kernel(const float* __restrict__ rs, const float* __restrict__ x, const float* __restrict__ mask,
                               float* __restrict__ x_out, float* __restrict__ skip, int B, int H, int RS, int T, int first,
                               int mask_skip) {
    const long long total = (long long)B * H * T;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int t = (int)(i % T);
        long long r = i / T;
        const int h = (int)(r % H), b = (int)(r / H);
        const float mk = mask ? mask[(size_t)b * T + t] : 1.f;
        const size_t ors = ((size_t)b * RS + h) * T + t;
        float sk;
        if (RS == 2 * H) {
            x_out[i] = (x[i] + rs[ors]) * mk;
            sk = rs[ors + (size_t)H * T];
        } else {
            sk = rs[ors];
        }
        if (!first) sk += skip[i];
        skip[i] = mask_skip ? sk * mk : sk;
    }
}
int resskip(const float* rs, const float* x, const float* mask, float* x_out, float* skip, int B, int H, int RS, int T,
            int first, int mask_skip, cudaStream_t st) {
    GTK_REQUIRE(rs && skip && (RS == H || (RS == 2 * H && x && x_out)), "bad args");
    const long long total = (long long)B * H * T;
    if (total == 0) return GTK_OK;
    resskip_kernel<<<pw_grid(total), PW_NT, 0, st>>>(rs, x, mask, x_out, skip, B, H, RS, T, first, mask_skip);
    return check_launch("resskip");
}
// out[B, Ca+Cb, T] = cat(a * mask, b * (mask_b ? mask : 1))   (backward glue: drs = cat(dh*mask, dskip))
__global__ void cat_mask_kernel(const float* __restrict__ a, const float* __restrict__ bsrc, const float* __restrict__ mask,
                                float* __restrict__ out, int B, int Ca, int Cb, int T, int mask_b) {
    const int Ct = Ca + Cb;
    const long long total = (long long)B * Ct * T;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int t = (int)(i % T);
        long long r = i / T;
        const int c = (int)(r % Ct), b = (int)(r / Ct);
        const float mk = mask ? mask[(size_t)b * T + t] : 1.f;
        float v;
        if (c < Ca) v = a[((size_t)b * Ca + c) * T + t] * mk;
        else        v = bsrc[((size_t)b * Cb + (c - Ca)) * T + t] * (mask_b ? mk : 1.f);
        out[i] = v;
    }
}
int cat_mask(const float* a, const float* b, const float* mask, float* out, int B, int Ca, int Cb, int T, int mask_b,
             cudaStream_t st) {
    GTK_REQUIRE(out && (Ca == 0 || a) && (Cb == 0 || b), "bad args");
    const long long total = (long long)B * (Ca + Cb) * T;
    if (total == 0) return GTK_OK;
    cat_mask_kernel<<<pw_grid(total), PW_NT, 0, st>>>(a, b, mask, out, B, Ca, Cb, T, mask_b);
    return check_launch("cat_mask");
}

// ------------------------------------------------------------------------------------------ weight norm
// w[m,:] = g[m] * v[m,:] / ||v[m,:]||  ; one warp per output channel.  norm_out (nullable) keeps ||v||.
__global__ void weightnorm_fwd_kernel(const float* __restrict__ g, const float* __restrict__ v, float* __restrict__ w,
                                      float* __restrict__ norm_out, int M, int R) {
    const int m = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (m >= M) return;
    float s = 0.f;
    for (int i = lane; i < R; i += 32) { const float t = v[(size_t)m * R + i]; s += t * t; }
    s = warp_sum(s);
    const float nrm = sqrtf(s);
    const float sc = g[m] / nrm;
    for (int i = lane; i < R; i += 32) w[(size_t)m * R + i] = v[(size_t)m * R + i] * sc;
    if (norm_out && lane == 0) norm_out[m] = nrm;
}
// dg[m] (+)= <dw, v>/||v|| ; dv (+)= g/||v|| * (dw - v <dw,v>/||v||^2)
__global__ void weightnorm_bwd_kernel(const float* __restrict__ g, const float* __restrict__ v, const float* __restrict__ dw,
                                      float* __restrict__ dg, float* __restrict__ dv, int M, int R, int accumulate) {
    const int m = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (m >= M) return;
    float s = 0.f, d = 0.f;
    for (int i = lane; i < R; i += 32) {
        const float t = v[(size_t)m * R + i];
        s += t * t;
        d += t * dw[(size_t)m * R + i];
    }
    s = warp_sum(s);
    d = warp_sum(d);
    const float nrm = sqrtf(s);
    const float sc = g[m] / nrm, proj = d / s;
    for (int i = lane; i < R; i += 32) {
        const float val = sc * (dw[(size_t)m * R + i] - v[(size_t)m * R + i] * proj);
        dv[(size_t)m * R + i] = accumulate ? dv[(size_t)m * R + i] + val : val;
    }
    if (lane == 0) dg[m] = accumulate ? dg[m] + d / nrm : d / nrm;
}
int weightnorm_fwd(const float* g, const float* v, float* w, float* norm_out, int M, int R, cudaStream_t st) {
    GTK_REQUIRE(g && v && w, "null");
    if (M == 0) return GTK_OK;
    weightnorm_fwd_kernel<<<ceil_div(M, 4), 128, 0, st>>>(g, v, w, norm_out, M, R);
    return check_launch("weightnorm_fwd");
}
int weightnorm_bwd(const float* g, const float* v, const float* dw, float* dg, float* dv, int M, int R, int accumulate,
                   cudaStream_t st) {
    GTK_REQUIRE(g && v && dw && dg && dv, "null");
    if (M == 0) return GTK_OK;
    weightnorm_bwd_kernel<<<ceil_div(M, 4), 128, 0, st>>>(g, v, dw, dg, dv, M, R, accumulate);
    return check_launch("weightnorm_bwd");
}

// ------------------------------------------------------------------------------------------ mle loss
// acc[0] += sum logs, acc[1] += sum exp(-2 logs)(z-m)^2, acc[2] += sum mask   (commons.py:29-31)
// m / logs may be null (standard-normal prior: m = 0, logs = 0).
__global__ void mle_partial_kernel(const float* __restrict__ z, const float* __restrict__ m, const float* __restrict__ logs,
                                   const float* __restrict__ mask, float* __restrict__ acc, long long n_z, long long n_mask) {
    __shared__ float scratch[33];
    float s0 = 0.f, s1 = 0.f, s2 = 0.f;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_z; i += (long long)gridDim.x * blockDim.x) {
        const float l = logs ? logs[i] : 0.f;
        const float d = z[i] - (m ? m[i] : 0.f);
        s0 += l;
        s1 += expf(-2.f * l) * d * d;
    }
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_mask; i += (long long)gridDim.x * blockDim.x)
        s2 += mask[i];
    s0 = block_sum<PW_NT>(s0, scratch);
    s1 = block_sum<PW_NT>(s1, scratch);
    s2 = block_sum<PW_NT>(s2, scratch);
    if (threadIdx.x == 0) { atomicAdd(&acc[0], s0); atomicAdd(&acc[1], s1); atomicAdd(&acc[2], s2); }
}
// loss = (S0 + 0.5 S1 - sum logdet) / (C * S2) + 0.5 log(2 pi);  acc[3] = denominator (kept for backward)
__global__ void mle_finalize_kernel(float* __restrict__ acc, const float* __restrict__ logdet, int B, int C,
                                    float* __restrict__ loss) {
    __shared__ float scratch[33];
    float s = 0.f;
    for (int b = threadIdx.x; b < B; b += PW_NT) s += logdet[b];
    s = block_sum<PW_NT>(s, scratch);
    if (threadIdx.x == 0) {
        const float den = (float)C * acc[2];
        acc[3] = den;
        *loss = (acc[0] + 0.5f * acc[1] - s) / den + 0.918938533204672742f;
    }
}
int mle_loss_fwd(const float* z, const float* m, const float* logs, const float* logdet, const float* mask, float* acc4,
                 float* loss, int B, int C, int T, cudaStream_t st) {
    GTK_REQUIRE(z && logdet && mask && acc4 && loss, "null");
    GTK_CUDA(cudaMemsetAsync(acc4, 0, 4 * sizeof(float), st));
    const long long n = (long long)B * C * T;
    mle_partial_kernel<<<pw_grid(n > 0 ? n : 1), PW_NT, 0, st>>>(z, m, logs, mask, acc4, n, (long long)B * T);
    int rc = check_launch("mle_partial");
    if (rc) return rc;
    mle_finalize_kernel<<<1, PW_NT, 0, st>>>(acc4, logdet, B, C, loss);
    return check_launch("mle_finalize");
}
// ------------------------------------------------------------------------------------------ clip_grad_value_
// commons.clip_grad_value_ (commons.py:320-336) over ONE contiguous range of gradients: acc[0] += sum |g|^p, and
// g = clamp(g, -clip, clip) in place when do_clip.  The reference loops over parameters with a .item() host sync each;
// the caller here passes the flat gradient buffer (or merges neighbouring tensors) and takes acc^(1/p) on the device.
template <int MODE>   // 0: p == 2, 1: p == 1, 2: general p
__global__ void clip_grad_kernel(float* __restrict__ g, long long n, float clip, int do_clip, float p, float* __restrict__ acc) {
    __shared__ float scratch[33];
    float s = 0.f;
    const long long stride = (long long)gridDim.x * blockDim.x, i0 = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const uintptr_t a = reinterpret_cast<uintptr_t>(g);
    long long head = (long long)(((16 - (a & 15)) & 15) / 4);
    if (head > n) head = n;
    const long long nvec = (n - head) / 4;
    float4* gv = reinterpret_cast<float4*>(g + head);
    auto term = [&](float v) { return MODE == 0 ? v * v : MODE == 1 ? fabsf(v) : powf(fabsf(v), p); };
    auto cl = [&](float v) { return fminf(fmaxf(v, -clip), clip); };
    for (long long i = i0; i < nvec; i += stride) {
        float4 v = gv[i];
        s += term(v.x) + term(v.y) + term(v.z) + term(v.w);
        if (do_clip) { v.x = cl(v.x); v.y = cl(v.y); v.z = cl(v.z); v.w = cl(v.w); gv[i] = v; }
    }
    for (long long i = i0; i < head + (n - head - nvec * 4); i += stride) {      // unaligned head and tail scalars
        const long long j = i < head ? i : head + nvec * 4 + (i - head);
        const float v = g[j];
        s += term(v);
        if (do_clip) g[j] = cl(v);
    }
    s = block_sum<PW_NT>(s, scratch);
    if (threadIdx.x == 0) atomicAdd(acc, s);
}
int clip_grad_value(float* grad, long long n, float clip, int do_clip, float norm_type, float* acc, cudaStream_t st) {
    GTK_REQUIRE(acc && (grad || n == 0) && n >= 0 && norm_type > 0.f, "clip_grad_value arguments");
    if (n == 0) return GTK_OK;
    const int grid = pw_grid(ceil_div_ll(n, 4));
    if (norm_type == 2.f)      clip_grad_kernel<0><<<grid, PW_NT, 0, st>>>(grad, n, clip, do_clip, norm_type, acc);
    else if (norm_type == 1.f) clip_grad_kernel<1><<<grid, PW_NT, 0, st>>>(grad, n, clip, do_clip, norm_type, acc);
    else                       clip_grad_kernel<2><<<grid, PW_NT, 0, st>>>(grad, n, clip, do_clip, norm_type, acc);
    return check_launch("clip_grad_value");
}

// gloss: device scalar upstream gradient.  dz, dm, dlogs, dlogdet nullable.
__global__ void mle_bwd_kernel(const float* __restrict__ z, const float* __restrict__ m, const float* __restrict__ logs,
                               const float* __restrict__ acc4, const float* __restrict__ gloss, float* __restrict__ dz,
                               float* __restrict__ dm, float* __restrict__ dlogs, float* __restrict__ dlogdet, long long n,
                               int B) {
    const float sc = (*gloss) / acc4[3];
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float l = logs ? logs[i] : 0.f;
        const float d = z[i] - (m ? m[i] : 0.f);
        const float e = expf(-2.f * l);
        if (dz) dz[i] = sc * e * d;
        if (dm) dm[i] = -sc * e * d;
        if (dlogs) dlogs[i] = sc * (1.f - e * d * d);
    }
    if (dlogdet)
        for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < B; i += (long long)gridDim.x * blockDim.x)
            dlogdet[i] = -sc;
}
int mle_loss_bwd(const float* z, const float* m, const float* logs, const float* acc4, const float* gloss, float* dz,
                 float* dm, float* dlogs, float* dlogdet, int B, int C, int T, cudaStream_t st) {
    GTK_REQUIRE(z && acc4 && gloss, "null");
    const long long n = (long long)B * C * T;
    mle_bwd_kernel<<<pw_grid(n > 0 ? n : 1), PW_NT, 0, st>>>(z, m, logs, acc4, gloss, dz, dm, dlogs, dlogdet, n, B);
    return check_launch("mle_bwd");
}

// ------------------------------------------------------------------------------------------ prior expansion
// out[b, c, y] = src[b, c, idx[b, y]] (0 where idx < 0): models.py:1118-1119 as the gather it is.
__global__ void gather_tokens_kernel(const float* __restrict__ src, const int* __restrict__ idx, float* __restrict__ out,
                                     int B, int C, int T_x, int T_y) {
    const long long total = (long long)B * C * T_y;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int y = (int)(i % T_y);
        long long r = i / T_y;
        const int c = (int)(r % C), b = (int)(r / C);
        const int x = idx[(size_t)b * T_y + y];
        out[i] = (x >= 0 && x < T_x) ? src[((size_t)b * C + c) * T_x + x] : 0.f;
    }
}
int gather_tokens(const float* src, const int* idx, float* out, int B, int C, int T_x, int T_y, cudaStream_t st) {
    GTK_REQUIRE(src && idx && out, "null");
    const long long total = (long long)B * C * T_y;
    if (total == 0) return GTK_OK;
    gather_tokens_kernel<<<pw_grid(total), PW_NT, 0, st>>>(src, idx, out, B, C, T_x, T_y);
    return check_launch("gather_tokens");
}

}  // namespace flow
}  // namespace gtk
