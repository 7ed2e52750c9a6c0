// glowtts_b200/csrc/logp.cu -- the Gaussian log-likelihood that feeds MAS (models.py:1076-1082) and
// commons.generate_path (commons.py:127-143), the inference-time neighbour of MAS.
//
// logp[b,x,y] = sum_d [ -0.5 log 2pi - logs - 0.5 m^2 s ](d,x)  +  sum_d (-0.5 s)(d,x) z^2(d,y)
//             + sum_d (m s)(d,x) z(d,y),   s = exp(-2 logs)
// i.e. one [T_x, 2D] x [2D, T_y] contraction per utterance plus a per-token constant; D = 80, so the
// kernel is bound by writing the [B,T_x,T_y] fp32 output (4 B/cell).  fp32 FFMA keeps it within fp32
// rounding of the reference's two matmuls.
#include <stdlib.h>

#include "tc.cuh"
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

// -------------------------------------------------------------------------------------------------
// Tensor-core logp (D % 8 == 0, D <= 80): the contraction runs on the legacy tensor path as 3xTF32
// (a_lo b_hi + a_hi b_lo + a_hi b_hi with hi = upper 19 bits, lo = v - hi: fp32-level accuracy), one 64 (text) x 128
// (frames) output tile per CTA, operands staged ONCE per CTA in shared memory as fp32 [k][x] / [k][y] rows (pitch = 8
// mod 32: conflict-free fragment reads).  General models contract over 2 D rows ([m s | -0.5 s] x [z ; z^2]);
// mean_only models (x_logs == NULL, s == 1) over D rows, the z^2 term being the per-frame constant -0.5 sum z^2.
// The FFMA kernel above is LDS-bound (76 us at B=32, 180 x 860); this one is bound by HMMA issue (8 clk per
// HMMA.1688.TF32 per SM sub-partition, profiles/micro/hmma_rate.cu) and by writing the 4 B/cell output.
constexpr int TX = 64, TY = 128, TNT = 256;
constexpr int AP = TX + 8, BP = TY + 8;         // row pitches (floats)
constexpr int TC_MAX_D = 80;

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t tf32_hi(float v) { return __float_as_uint(v) & 0xffffe000u; }
__device__ __forceinline__ uint32_t tf32_lo(float v, uint32_t hi) { return __float_as_uint(v - __uint_as_float(hi)); }

template <bool MEAN_ONLY>
__global__ void __launch_bounds__(TNT, 2) logp_tc_kernel(const float* __restrict__ x_m, const float* __restrict__ x_logs,
                                                      const float* __restrict__ z, float* __restrict__ out, int D, int T_x,
                                                      int T_y, int vec4) {
    extern __shared__ __align__(16) float sm[];
    const int KT = MEAN_ONLY ? D : 2 * D;
    float* As = sm;                              // [KT][AP]
    float* Bs = As + (size_t)KT * AP;            // [KT][BP]
    float* Cs = Bs + (size_t)KT * BP;            // general: [D][TX] per-(d, x) constants; then crow[TX], zq[TY]
    float* crow = Cs + (MEAN_ONLY ? 0 : (size_t)D * TX);
    float* zq = crow + TX;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int y0 = blockIdx.x * TY, x0 = blockIdx.y * TX, b = blockIdx.z;
    const float* mb = x_m + (size_t)b * D * T_x;
    const float* lb = MEAN_ONLY ? nullptr : x_logs + (size_t)b * D * T_x;
    const float* zb = z + (size_t)b * D * T_y;

    auto put_a = [&](int d, int xx, float m, float l, bool ok) {
        if (MEAN_ONLY) {
            As[d * AP + xx] = ok ? m : 0.f;
        } else {
            const float sc = expf(-2.f * l);
            As[d * AP + xx] = ok ? m * sc : 0.f;
            As[(D + d) * AP + xx] = ok ? -0.5f * sc : 0.f;
            Cs[d * TX + xx] = ok ? (-0.918938533204672742f - l - 0.5f * m * m * sc) : 0.f;
        }
    };
    auto put_b = [&](int d, int yy, float v) {
        Bs[d * BP + yy] = v;
        if (!MEAN_ONLY) Bs[(D + d) * BP + yy] = v * v;
    };
    if (vec4) {
        // every operand load of the CTA is in flight before the first shared-memory store: one memory latency per tile
        constexpr int NA = TC_MAX_D * (TX / 4) / TNT, NB = TC_MAX_D * (TY / 4) / TNT;      // 5, 10 float4 per thread
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        float4 ra[NA], rl[NA], rb[NB];
#pragma unroll
        for (int i = 0; i < NA; ++i) {
            const int e = tid + i * TNT, d = e / (TX / 4), xx = (e % (TX / 4)) * 4;
            ra[i] = z4; rl[i] = z4;
            if (d < D && x0 + xx < T_x) {
                ra[i] = __ldg(reinterpret_cast<const float4*>(mb + (size_t)d * T_x + x0 + xx));
                if (!MEAN_ONLY) rl[i] = __ldg(reinterpret_cast<const float4*>(lb + (size_t)d * T_x + x0 + xx));
            }
        }
#pragma unroll
        for (int i = 0; i < NB; ++i) {
            const int e = tid + i * TNT, d = e / (TY / 4), yy = (e % (TY / 4)) * 4;
            rb[i] = z4;
            if (d < D && y0 + yy < T_y) rb[i] = __ldg(reinterpret_cast<const float4*>(zb + (size_t)d * T_y + y0 + yy));
        }
#pragma unroll
        for (int i = 0; i < NA; ++i) {
            const int e = tid + i * TNT, d = e / (TX / 4), xx = (e % (TX / 4)) * 4;
            if (d < D) {
                const bool ok = x0 + xx < T_x;
                put_a(d, xx, ra[i].x, rl[i].x, ok); put_a(d, xx + 1, ra[i].y, rl[i].y, ok);
                put_a(d, xx + 2, ra[i].z, rl[i].z, ok); put_a(d, xx + 3, ra[i].w, rl[i].w, ok);
            }
        }
#pragma unroll
        for (int i = 0; i < NB; ++i) {
            const int e = tid + i * TNT, d = e / (TY / 4), yy = (e % (TY / 4)) * 4;
            if (d < D) { put_b(d, yy, rb[i].x); put_b(d, yy + 1, rb[i].y); put_b(d, yy + 2, rb[i].z); put_b(d, yy + 3, rb[i].w); }
        }
    } else {
        for (int e = tid; e < D * TX; e += TNT) {
            const int d = e / TX, xx = e - d * TX;
            const bool ok = x0 + xx < T_x;
            const float m = ok ? mb[(size_t)d * T_x + x0 + xx] : 0.f;
            const float l = (ok && !MEAN_ONLY) ? lb[(size_t)d * T_x + x0 + xx] : 0.f;
            put_a(d, xx, m, l, ok);
        }
        for (int e = tid; e < D * TY; e += TNT) {
            const int d = e / TY, yy = e - d * TY;
            put_b(d, yy, (y0 + yy < T_y) ? zb[(size_t)d * T_y + y0 + yy] : 0.f);
        }
    }
    __syncthreads();
    if (tid < TX) {                              // per-token constant, summed over d in a fixed order
        float c = 0.f;
        if (MEAN_ONLY) {
            for (int d = 0; d < D; ++d) { const float m = As[d * AP + tid]; c = fmaf(m, m, c); }
            c = -0.5f * c - (float)D * 0.918938533204672742f;
        } else {
            for (int d = 0; d < D; ++d) c += Cs[d * TX + tid];
        }
        crow[tid] = c;
    } else if (tid >= 128) {                     // per-frame constant (mean_only): -0.5 sum_d z^2
        const int yy = tid - 128;
        float q = 0.f;
        if (MEAN_ONLY)
            for (int d = 0; d < D; ++d) { const float v = Bs[d * BP + yy]; q = fmaf(v, v, q); }
        zq[yy] = -0.5f * q;
    }

    // warp tile: 32 text rows x 32 frames = 2 x 4 m16n8k8 blocks
    const int g = lane >> 2, t = lane & 3;
    const int wx = (warp & 1) * 32, wy = (warp >> 1) * 32;
    float acc[2][4][4];
#pragma unroll
    for (int mb2 = 0; mb2 < 2; ++mb2)
#pragma unroll
        for (int nb = 0; nb < 4; ++nb)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[mb2][nb][i] = 0.f;
    for (int k0 = 0; k0 < KT; k0 += 8) {
        uint32_t ah[2][4], al[2][4], bh[4][2], bl[4][2];
#pragma unroll
        for (int mb2 = 0; mb2 < 2; ++mb2) {
            const float* ap = As + (size_t)(k0 + t) * AP + wx + 16 * mb2 + g;     // a0 (g, t) a1 (g+8, t) a2 (g, t+4) a3 (g+8, t+4)
            const float a0 = ap[0], a1 = ap[8], a2 = ap[4 * AP], a3 = ap[4 * AP + 8];
            ah[mb2][0] = tf32_hi(a0); al[mb2][0] = tf32_lo(a0, ah[mb2][0]);
            ah[mb2][1] = tf32_hi(a1); al[mb2][1] = tf32_lo(a1, ah[mb2][1]);
            ah[mb2][2] = tf32_hi(a2); al[mb2][2] = tf32_lo(a2, ah[mb2][2]);
            ah[mb2][3] = tf32_hi(a3); al[mb2][3] = tf32_lo(a3, ah[mb2][3]);
        }
#pragma unroll
        for (int nb = 0; nb < 4; ++nb) {
            const float* bp = Bs + (size_t)(k0 + t) * BP + wy + 8 * nb + g;        // b0 (k = t, n = g) b1 (k = t+4, n = g)
            const float b0 = bp[0], b1 = bp[4 * BP];
            bh[nb][0] = tf32_hi(b0); bl[nb][0] = tf32_lo(b0, bh[nb][0]);
            bh[nb][1] = tf32_hi(b1); bl[nb][1] = tf32_lo(b1, bh[nb][1]);
        }
#pragma unroll
        for (int mb2 = 0; mb2 < 2; ++mb2)
#pragma unroll
            for (int nb = 0; nb < 4; ++nb) mma_tf32(acc[mb2][nb], al[mb2], bh[nb][0], bh[nb][1]);
#pragma unroll
        for (int mb2 = 0; mb2 < 2; ++mb2)
#pragma unroll
            for (int nb = 0; nb < 4; ++nb) mma_tf32(acc[mb2][nb], ah[mb2], bl[nb][0], bl[nb][1]);
#pragma unroll
        for (int mb2 = 0; mb2 < 2; ++mb2)
#pragma unroll
            for (int nb = 0; nb < 4; ++nb) mma_tf32(acc[mb2][nb], ah[mb2], bh[nb][0], bh[nb][1]);
    }
    __syncthreads();                             // crow / zq
    const bool pair_ok = (T_y & 1) == 0;         // (x T_y + y) even for even y: 8-byte aligned float2 stores
#pragma unroll
    for (int mb2 = 0; mb2 < 2; ++mb2)
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const int xx = wx + 16 * mb2 + g + 8 * half, x = x0 + xx;
            if (x >= T_x) continue;
            const float c = crow[xx];
            float* orow = out + ((size_t)b * T_x + x) * T_y;
#pragma unroll
            for (int nb = 0; nb < 4; ++nb) {
                const int yy = wy + 8 * nb + 2 * t, y = y0 + yy;
                const float v0 = acc[mb2][nb][2 * half] + (c + zq[yy]);
                const float v1 = acc[mb2][nb][2 * half + 1] + (c + zq[yy + 1]);
                if (pair_ok && y + 1 < T_y) *reinterpret_cast<float2*>(orow + y) = make_float2(v0, v1);
                else {
                    if (y < T_y) orow[y] = v0;
                    if (y + 1 < T_y) orow[y + 1] = v1;
                }
            }
        }
}


// -------------------------------------------------------------------------------------------------
// tcgen05 logp (round 2): the same 3xTF32 contraction on the 5th-generation tensor cores.  One 128 (text) x 128 (frames)
// output tile per CTA, fp32 accumulator in TMEM (128 columns: three CTAs share an SM).  The operands are COMPUTED by the
// CTA (scaled means, squares, hi / lo split), so the threads stage them, not TMA: thread = one row (threads 0..127 text
// rows of A, 128..255 frames of B), walking k in chunks of 32: 32 coalesced global loads (lanes = consecutive rows), the
// hi / lo split, sixteen 16-byte stores into K-major SWIZZLE_128B tiles (32 fp32 per 128-byte row, 16-byte chunks XOR-ed
// with row % 8 -- the layout conv_tc.cu's bf16 tiles have) for A_hi, A_lo, B_hi, B_lo.  One elected lane then issues, per
// k8 step,  D += A_lo B_hi;  D += A_hi B_lo;  D += A_hi B_hi   (tcgen05.mma.kind::tf32, M = N = 128, K = 8; hi = upper
// 19 bits = exactly a tf32, lo = v - hi), while the threads already load the next chunk into registers; the 64 KB staging
// area is rewritten once the MMAs have completed.  The row owners also sum the per-token / per-frame constants over d in
// a fixed order.  Drain: warp = TMEM lane quarter x frame half, thread = text row, 64 consecutive frames.
constexpr int UX = 128, UY = 128, UNT = 256;
constexpr int U_KC = 32;                          // fp32 elements per swizzle row = k per pass
constexpr int U_TILE = UX * 128;                  // bytes of one [128 rows x 32 fp32] tile

__host__ __device__ constexpr uint32_t umma_idesc_tf32(int M, int N) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, bool accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"((uint32_t)accumulate)
        : "memory");
}

struct LogpUmmaCtl {
    uint64_t mma_done;
    uint32_t tmem_base;
    float crow[UX], crow2[UX], zq[UY];      // per-token constant (two partial sums, added in a fixed order), per-frame constant
};

template <bool MEAN_ONLY>
__global__ void __launch_bounds__(UNT, 3) logp_umma_kernel(const float* __restrict__ x_m, const float* __restrict__ x_logs,
                                                        const float* __restrict__ z, float* __restrict__ out, int D, int T_x,
                                                        int T_y) {
    extern __shared__ __align__(1024) unsigned char usm_raw[];
    unsigned char* usm = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(usm_raw) + 1023) & ~uintptr_t(1023));
    unsigned char* a_hi = usm;
    unsigned char* a_lo = a_hi + U_TILE;
    unsigned char* b_hi = a_lo + U_TILE;
    unsigned char* b_lo = b_hi + U_TILE;
    LogpUmmaCtl* ctl = reinterpret_cast<LogpUmmaCtl*>(b_lo + U_TILE);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int y0 = blockIdx.x * UY, x0 = blockIdx.y * UX, b = blockIdx.z;
    // k layout.  mean_only: k = d, chunks of 32.  General models: chunk c holds d = 16 c .. 16 c + 15 twice -- slots [0, 16) =
    // (m s)(d) against z(d), slots [16, 32) = (-0.5 s)(d) against z^2(d) -- so that exp(-2 logs) is evaluated once per (row, d).
    const int n_chunks = MEAN_ONLY ? (D + U_KC - 1) / U_KC : (D + 15) / 16;

    if (tid == 0) { mbar_init(&ctl->mma_done, 1); mbar_fence_init(); }
    if (warp == 0) tc::tmem_alloc<128>(&ctl->tmem_base);

    // Thread <-> rows.  mean_only: thread = one row (threads 0..127 text rows of A, 128..255 frames of B), all 32 slots of
    // a chunk.  General: thread = text row AND frame `tid % 128`, the 8 d's of half `tid / 128` of the chunk for both
    // operands (the expf work is spread over all 256 threads).
    const bool is_b = MEAN_ONLY && tid >= UX;
    const int row = tid % UX, half = tid / UX;
    const int gx = x0 + row, gy = y0 + row;
    const bool okx = gx < T_x, oky = gy < T_y;
    const float* srcm = x_m + (size_t)b * D * T_x + (okx ? gx : 0);
    const float* srcl = MEAN_ONLY ? nullptr : x_logs + (size_t)b * D * T_x + (okx ? gx : 0);
    const float* srcz = z + (size_t)b * D * T_y + (oky ? gy : 0);

    float cst = 0.f;                                  // per-token constant (partial); mean_only B rows: sum_d z^2
    float v[U_KC];                                    // general: v[0..15] = A values (m s | -0.5 s), v[16..31] = B values (z | z^2)
    auto load_chunk = [&](int c) {
        if constexpr (MEAN_ONLY) {
#pragma unroll
            for (int i = 0; i < U_KC; ++i) {
                const int d = c * U_KC + i;
                float val = 0.f;
                if (d < D && (is_b ? oky : okx)) {
                    val = __ldg((is_b ? srcz : srcm) + (size_t)d * (is_b ? T_y : T_x));
                    cst = fmaf(val, val, cst);
                }
                v[i] = val;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int d = c * 16 + half * 8 + i;
                float ms = 0.f, hs = 0.f, zz = 0.f;
                if (d < D) {
                    if (okx) {
                        const float m = __ldg(srcm + (size_t)d * T_x), l = __ldg(srcl + (size_t)d * T_x);
                        const float sc = expf(-2.f * l);
                        ms = m * sc;
                        hs = -0.5f * sc;
                        cst += -0.918938533204672742f - l - 0.5f * m * m * sc;
                    }
                    if (oky) zz = __ldg(srcz + (size_t)d * T_y);
                }
                v[i] = ms; v[8 + i] = hs; v[16 + i] = zz; v[24 + i] = zz * zz;
            }
        }
    };
    auto put_quad = [&](unsigned char* t_hi, unsigned char* t_lo, int quad, const float* q) {
        uint32_t h[4], l[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { h[i] = tf32_hi(q[i]); l[i] = tf32_lo(q[i], h[i]); }
        const size_t off = (size_t)row * 128 + (size_t)((quad ^ (row & 7)) << 4);
        *reinterpret_cast<uint4*>(t_hi + off) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4*>(t_lo + off) = make_uint4(l[0], l[1], l[2], l[3]);
    };
    load_chunk(0);
    const uint32_t idesc = umma_idesc_tf32(UX, UY);
    for (int c = 0; c < n_chunks; ++c) {
        if (c > 0) mbar_wait(&ctl->mma_done, (c - 1) & 1);          // the MMAs of the previous chunk have read the staging area
        if constexpr (MEAN_ONLY) {
#pragma unroll
            for (int sl = 0; sl < U_KC / 4; ++sl) put_quad(is_b ? b_hi : a_hi, is_b ? b_lo : a_lo, sl, v + 4 * sl);
        } else {
            // slots [8 half, 8 half + 8) and [16 + 8 half, 16 + 8 half + 8) of both operands
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                put_quad(a_hi, a_lo, 2 * half + u, v + 4 * u);
                put_quad(a_hi, a_lo, 4 + 2 * half + u, v + 8 + 4 * u);
                put_quad(b_hi, b_lo, 2 * half + u, v + 16 + 4 * u);
                put_quad(b_hi, b_lo, 4 + 2 * half + u, v + 24 + 4 * u);
            }
        }
        if (c + 1 == n_chunks) {
            if constexpr (MEAN_ONLY) {
                if (is_b) { ctl->zq[row] = -0.5f * cst; }
                else { ctl->crow[row] = -0.5f * cst - (float)D * 0.918938533204672742f; ctl->crow2[row] = 0.f; }
            } else {
                if (half == 0) { ctl->crow[row] = cst; ctl->zq[row] = 0.f; }
                else ctl->crow2[row] = cst;
            }
        }
        tc::fence_proxy_async_smem();
        tc::tc_fence_before();
        __syncthreads();
        tc::tc_fence_after();
        if (warp == 1) {
            const uint32_t tmem = ctl->tmem_base;
            if (tc::elect_one()) {
                const uint64_t ah = tc::umma_desc_kmajor_sw128(smem_u32(a_hi)), al = tc::umma_desc_kmajor_sw128(smem_u32(a_lo));
                const uint64_t bh = tc::umma_desc_kmajor_sw128(smem_u32(b_hi)), bl = tc::umma_desc_kmajor_sw128(smem_u32(b_lo));
#pragma unroll
                for (int k = 0; k < U_KC / 8; ++k) {              // K = 8 per tf32 MMA = 32 bytes = 2 descriptor units
                    umma_tf32(tmem, al + 2 * k, bh + 2 * k, idesc, (c | k) != 0);
                    umma_tf32(tmem, ah + 2 * k, bl + 2 * k, idesc, true);
                    umma_tf32(tmem, ah + 2 * k, bh + 2 * k, idesc, true);
                }
                tc::umma_commit(&ctl->mma_done);
            }
            __syncwarp();
        }
        if (c + 1 < n_chunks) load_chunk(c + 1);      // in flight while the tensor core works
    }
    mbar_wait(&ctl->mma_done, (n_chunks - 1) & 1);
    tc::tc_fence_after();
    // ---- drain: warp w owns TMEM lanes 32 (w % 4) .. (text rows) and frames 64 (w / 4) ..
    {
        const uint32_t tmem = ctl->tmem_base;
        const int q = warp & 3, half = warp >> 2;
        const int r = q * 32 + lane, x = x0 + r;
        const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(half * 64);
        const float c = ctl->crow[r] + ctl->crow2[r];
        float* orow = out + ((size_t)b * T_x + (x < T_x ? x : 0)) * T_y + y0 + half * 64;
        const bool vec = ((T_y & 3) == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float w[16];
            tc::tmem_ld16(trow + 16 * j, w);
            if (x >= T_x) continue;
            const int yy = half * 64 + 16 * j;
#pragma unroll
            for (int i = 0; i < 16; ++i) w[i] += c + ctl->zq[yy + i];
            if (vec && y0 + yy + 16 <= T_y) {
                float4* o4 = reinterpret_cast<float4*>(orow + 16 * j);
#pragma unroll
                for (int i = 0; i < 4; ++i) o4[i] = make_float4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]);
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    if (y0 + yy + i < T_y) orow[16 * j + i] = w[i];
            }
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc::tc_fence_after();
        tc::tmem_dealloc<128>(ctl->tmem_base);
    }
}

static size_t logp_umma_smem() { return (size_t)4 * U_TILE + sizeof(LogpUmmaCtl) + 1024; }

static size_t logp_tc_smem(int D, bool mean_only) {
    const int KT = mean_only ? D : 2 * D;
    return sizeof(float) * ((size_t)KT * AP + (size_t)KT * BP + (mean_only ? 0 : (size_t)D * TX) + TX + TY);
}

int forward(const float* x_m, const float* x_logs, const float* z, float* out, int B, int D, int T_x, int T_y,
            cudaStream_t st) {
    GTK_REQUIRE(x_m && z && out, "null");
    if ((long long)B * T_x * T_y == 0) return GTK_OK;
    // tcgen05 kind::tf32 kernel: the default (B = 32, 180 x 860 on B200: mean_only 33 us, general models 46 us against 52 us
    // for the mma.sync kernel).  GTK_LOGP_LEGACY=1 selects the mma.sync kernel, GTK_LOGP_FFMA=1 the FFMA one.
    if (D >= 1 && D <= 256 && !getenv("GTK_LOGP_FFMA") && !getenv("GTK_LOGP_LEGACY")) {
        const bool mo = x_logs == nullptr;
        const size_t smem = logp_umma_smem();
        dim3 grid(ceil_div(T_y, UY), ceil_div(T_x, UX), B);
        ProfScope prof("logp", st);
        if (mo) {
            GTK_CUDA(cudaFuncSetAttribute(logp_umma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            logp_umma_kernel<true><<<grid, UNT, smem, st>>>(x_m, x_logs, z, out, D, T_x, T_y);
        } else {
            GTK_CUDA(cudaFuncSetAttribute(logp_umma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            logp_umma_kernel<false><<<grid, UNT, smem, st>>>(x_m, x_logs, z, out, D, T_x, T_y);
        }
        return check_launch("logp_umma");
    }
    if (D % 8 == 0 && D >= 8 && D <= TC_MAX_D && !getenv("GTK_LOGP_FFMA")) {
        const bool mo = x_logs == nullptr;
        const size_t smem = logp_tc_smem(D, mo);
        const int vec4 = ((T_x | T_y) & 3) == 0 &&
                         ((reinterpret_cast<uintptr_t>(x_m) | reinterpret_cast<uintptr_t>(z) | reinterpret_cast<uintptr_t>(x_logs)) & 15) == 0;
        dim3 grid(ceil_div(T_y, TY), ceil_div(T_x, TX), B);
        ProfScope prof("logp", st);
        if (mo) {
            GTK_CUDA(cudaFuncSetAttribute(logp_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            logp_tc_kernel<true><<<grid, TNT, smem, st>>>(x_m, x_logs, z, out, D, T_x, T_y, vec4);
        } else {
            GTK_CUDA(cudaFuncSetAttribute(logp_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            logp_tc_kernel<false><<<grid, TNT, smem, st>>>(x_m, x_logs, z, out, D, T_x, T_y, vec4);
        }
        return check_launch("logp_tc");
    }
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
