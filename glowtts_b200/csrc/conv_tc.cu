// glowtts_b200/csrc/conv_tc.cu -- bf16 tcgen05 implicit-GEMM conv1d on "frame-major" activations.
//
// The WaveNet / 1x1 contractions of the coupling block (modules.py:129-142,152,165; attentions.py:102-110,151,170)
// as D[t, n] = sum_{j, c} X[t + (j - K/2) dil, c] * W[j][n][c]:
//   * activations are kept frame-major ([B, T, C], channel fastest) in bf16, so a tile of 128 frames x 64 channels
//     is a K-major UMMA operand that TMA writes straight into the canonical SWIZZLE_128B layout; the conv taps are
//     just shifted TMA boxes (out-of-range frames are zero-filled by TMA = the conv's zero padding, and the batch
//     index is a separate tensor-map dimension so a halo never crosses utterances);
//   * weights are packed [tap][n][c] bf16 (K-major B operand);
//   * the accumulator D (128 frames x N <= 512 fp32 columns) lives in TMEM: lane = frame, column = out channel.
//     An epilogue thread therefore owns one frame and sees all of its channels, which is what lets the
//     gate / residual / affine-coupling epilogues (later in this file) run without any cross-thread exchange.
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane), warp 2 = TMEM allocator,
// warps 4..7 = epilogue (TMEM lane quarter = warp % 4).
#include <string.h>

#include "tc.cuh"
#include "ops.h"

namespace gtk {
namespace tc {

// ---------------------------------------------------------------------------------- host: tensor maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

static int make_tmap(CUtensorMap* out, CUtensorMapDataType dt, const void* base, int rank, const uint64_t* dims,
                     const uint64_t* strides_bytes, const uint32_t* box, bool swizzle128) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) { set_error("cuTensorMapEncodeTiled entry point not available"); return GTK_ERR_LAUNCH; }
    cuuint64_t gdim[5], gstr[5];
    cuuint32_t bx[5], estr[5];
    for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bx[i] = box[i]; estr[i] = 1; }
    for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];   // stride of dim i+1
    CUresult r = fn(out, dt, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bx, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (%d)", (int)r); return GTK_ERR_LAUNCH; }
    return GTK_OK;
}
int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                   const uint32_t* box, bool swizzle128) {
    return make_tmap(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, base, rank, dims, strides_bytes, box, swizzle128);
}
int make_tmap_f32(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                  const uint32_t* box) {
    return make_tmap(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, base, rank, dims, strides_bytes, box, false);
}

// ---------------------------------------------------------------------------------- kernel
struct SmemCtl {
    uint64_t full[4], empty[4], acc_full;
    uint32_t tmem_base;
    alignas(16) float bias[512];        // [N] (zeros when there is none)
};

// Phase timestamps (globaltimer, ns) per CTA of the last launch selected with conv_trace_select():
// [start, setup done, accumulator complete, epilogue math done, stores issued].  Read with gtk_debug_conv_trace.
#ifdef GTK_INSTRUMENT
constexpr int CONV_TRACE_CTAS = 512;
__device__ unsigned long long g_conv_trace[CONV_TRACE_CTAS * 5];
__device__ long long g_conv_wait[CONV_TRACE_CTAS * 2];      // SM cycles the producer waited on `empty` / the MMA thread on `full`
static int g_trace_epi = -1, g_trace_taps = -1;
__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#define CONV_TRACE(i)                                                                                       \
    do {                                                                                                    \
        const int cta_ = blockIdx.y * gridDim.x + blockIdx.x;                                               \
        if (p.trace && threadIdx.x == 128 && cta_ < CONV_TRACE_CTAS) g_conv_trace[cta_ * 5 + (i)] = gtimer(); \
    } while (0)
#define CONV_TRACING(p) ((p).trace)
#else
#define CONV_TRACE(i) do { } while (0)
#define CONV_TRACING(p) false
#endif

// CL = CTAs per cluster.  CL == 2: two neighbouring frame tiles share every weight tile -- each CTA fetches half of
// its rows and the TMA multicasts them into both shared memories, so the L2 -> SM traffic of the weights (60 % of the
// bytes of the k = 5 data gradient, which is L2-latency bound, not MMA bound) is halved.
template <int TMEM_COLS, int EPI, int CL>
__global__ void __launch_bounds__(EPI_NT, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmO, const ConvTcParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    // 1024-byte alignment of the operand tiles (SWIZZLE_128B atoms)
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int stage_bytes = A_BYTES + p.N * KC * 2;
    SmemCtl* ctl = reinterpret_cast<SmemCtl*>(smem + p.ctl_off);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t0 = blockIdx.x * BT, b = blockIdx.y;
    int out_chunks = 0;                                  // staged [128 x 64] bf16 output tiles to TMA-store at the end
    pdl_trigger();
    // varlen packing: tiles (pairs of tiles when clustered) with nothing alive leave at once
    if (p.n_rows && (int)(blockIdx.x & ~(unsigned)(CL - 1)) * BT >= *p.n_rows) return;
    CONV_TRACE(0);
    const int n_iter = p.taps * p.n_k;
    const bool dual = (p.NCH == 1);                      // two interleaved accumulators (see the MMA loop)

    const uint32_t cta_rank = CL > 1 ? cluster_ctarank() : 0u;
    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) { mbar_init(&ctl->full[s], 1); mbar_init(&ctl->empty[s], CL); }
        mbar_init(&ctl->acc_full, 1);
        mbar_fence_init();
    }
    for (int i = threadIdx.x; i < p.N; i += EPI_NT) ctl->bias[i] = p.bias ? p.bias[i] : 0.f;
    if (warp == 2) tmem_alloc<512>(&ctl->tmem_base);
    tc_fence_before();
    if constexpr (CL > 1) cluster_sync_all();             // the peer's barriers exist before anything is multicast
    else __syncthreads();
    tc_fence_after();
    // PDL: the producer thread requests the weight tiles of its first stages before pdl_wait() (see wn_tc.cu)
    const bool is_producer = (warp == 0);
    if (!is_producer) pdl_wait();
    const uint32_t tmem = ctl->tmem_base;
    CONV_TRACE(1);

    // Whole warp in the uniform loops, ONE ELECTED lane issues the TMA / MMA instructions: their descriptors stay in
    // uniform registers (under `if (lane == 0)` every one is wrapped in an R2UR waterfall loop, profiles/mma_issue_r02.txt)
    if (warp == 0) {
        {
            if (elect_one()) {
                tma_prefetch_desc(&tmA);
                tma_prefetch_desc(&tmB);
            }
            long long w_empty = 0;
            auto load_weights = [&](int s, int j, int kc) {
                unsigned char* sb = smem + (size_t)s * stage_bytes + A_BYTES;
                mbar_arrive_expect_tx(&ctl->full[s], (uint32_t)stage_bytes);
                if constexpr (CL > 1) {
                    const int rows = p.NW / CL;            // this CTA's share of the weight rows, delivered to the whole cluster
                    for (int ch = 0; ch < p.NCH; ++ch)
                        tma_load_3d_multicast(sb + ((size_t)ch * p.NW + cta_rank * rows) * KC * 2, &tmB, &ctl->full[s], kc * KC,
                                              ch * p.NW + cta_rank * rows, j, (uint16_t)((1u << CL) - 1));
                } else {
                    for (int ch = 0; ch < p.NCH; ++ch)
                        tma_load_3d(sb + (size_t)ch * p.NW * KC * 2, &tmB, &ctl->full[s], kc * KC, ch * p.NW, j);
                }
            };
            const int n_pre = min(p.stages, n_iter);
            if (elect_one())
                for (int it = 0; it < n_pre; ++it) load_weights(it, it / p.n_k, it % p.n_k);
            __syncwarp();
            pdl_wait();
            for (int it = 0; it < n_iter; ++it) {
                const int s = it % p.stages, use = it / p.stages;
                if (use > 0) {
                    const long long c0 = CONV_TRACING(p) ? clock64() : 0;
                    mbar_wait(&ctl->empty[s], (use - 1) & 1);
                    if (CONV_TRACING(p)) w_empty += clock64() - c0;
                }
                const int j = it / p.n_k, kc = it % p.n_k;
                unsigned char* sa = smem + (size_t)s * stage_bytes;
                if (elect_one()) {
                    if (it >= n_pre) load_weights(s, j, kc);
                    tma_load_3d(sa, &tmA, &ctl->full[s], kc * KC, t0 + (j - p.taps / 2) * p.dil, b);
                }
                __syncwarp();
            }
            #ifdef GTK_INSTRUMENT
            if (p.trace) { const int cta_ = blockIdx.y * gridDim.x + blockIdx.x; if (cta_ < CONV_TRACE_CTAS) g_conv_wait[cta_ * 2] = w_empty; }
#endif
        }
    } else if (warp == 1) {
        {
            const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
            const uint32_t idesc = umma_idesc_bf16(BT, p.NW);
            long long w_full = 0;
            for (int it = 0; it < n_iter; ++it) {
                const int s = it % p.stages, use = it / p.stages;
                const long long c0 = CONV_TRACING(p) ? clock64() : 0;
                mbar_wait(&ctl->full[s], use & 1);
                if (CONV_TRACING(p)) w_full += clock64() - c0;
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + (size_t)s * stage_bytes);
                const uint32_t sb = sa + A_BYTES;
                if (elect_one()) {
                    const uint64_t adesc = umma_desc_kmajor_sw128(sa);
                    // Back-to-back MMAs into ONE accumulator serialise on its read-modify-write latency.  With N <= 256 the
                    // k16 steps alternate between two accumulators (TMEM columns 0.. and 256..), which the epilogue adds.
                    if (dual) {
                        const uint64_t bdesc = umma_desc_kmajor_sw128(sb);
#pragma unroll
                        for (int k = 0; k < KC / 16; ++k)
                            umma_bf16(tm + (k & 1) * 256, adesc + 2 * k, bdesc + 2 * k, idesc, it != 0 || k >= 2);
                    } else {
#pragma unroll
                        for (int k = 0; k < KC / 16; ++k) {
                            for (int ch = 0; ch < p.NCH; ++ch) {
                                const uint64_t bdesc = umma_desc_kmajor_sw128(sb + ch * p.NW * KC * 2);
                                umma_bf16(tm + ch * p.NW, adesc + 2 * k, bdesc + 2 * k, idesc, (it | k) != 0);
                            }
                        }
                    }
                    if constexpr (CL > 1) umma_commit_multicast(&ctl->empty[s], (uint16_t)((1u << CL) - 1));
                    else umma_commit(&ctl->empty[s]);
                }
                __syncwarp();
            }
            if (elect_one()) umma_commit(&ctl->acc_full);
            __syncwarp();
            #ifdef GTK_INSTRUMENT
            if (p.trace) { const int cta_ = blockIdx.y * gridDim.x + blockIdx.x; if (cta_ < CONV_TRACE_CTAS) g_conv_wait[cta_ * 2 + 1] = w_full; }
#endif
        }
    }
    __syncwarp();
    {   // epilogue: ALL warps; warp w owns TMEM lane quarter w % 4 (frames) and column part w / 4 of EPI_PARTS
        mbar_wait(&ctl->acc_full, 0);
        tc_fence_after();
        CONV_TRACE(2);
        const int q = warp & 3, part = warp >> 2;
        const int t = t0 + q * 32 + lane;
        const bool t_ok = t < p.T;
        const float mk = (p.mask && t_ok) ? p.mask[(size_t)b * p.T + t] : 1.f;
        const bool live = mk != 0.f;                     // masked rows are written as exact zeros (never value * 0)
        const size_t frame = (size_t)b * p.T + (t_ok ? t : 0);
        const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
        RowIo io;
        io.tp = p.tile_layout; io.tile_id = b * ((p.T + BT - 1) / BT) + blockIdx.x; io.r = q * 32 + lane; io.frame = frame;
        // The operand stages are idle once the accumulator is complete: bf16 outputs that the next kernel reads as TMA
        // operands are staged there in the SWIZZLE_128B layout and leave as bulk tensor stores (no row-strided stores).
        if constexpr (EPI == EPI_PLAIN) {
            const int nck = p.N / 16;
            for (int cc = nck * part / EPI_PARTS; cc < nck * (part + 1) / EPI_PARTS; ++cc) {
                const int c0 = cc * 16;
                float v[16];
                tmem_ld16_dual(trow + c0, dual, v);
                if (!t_ok) continue;
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] += ctl->bias[c0 + i];
                if (p.resid_f32) {
                    float x[16];
                    row_load_f32x16(p.resid_f32, p.ld_resid, c0, io, x);
                    const bool rl = !p.mask_resid || live;
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] += rl ? x[i] : 0.f;
                }
                if (p.resid_tm) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] += p.resid_tm[((size_t)b * p.tm_C + p.tm_c0 + c0 + i) * p.T + t];
                }
                if (p.mask_out) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = live ? v[i] : 0.f;
                }
                if (p.out_f32) row_store_f32x16(p.out_f32, p.ld_f32, c0, io, v);
                if (p.out_bf16) stage_bf16x16_sw128(smem, A_BYTES, io.r, c0, v);
                if (p.out_tm) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) p.out_tm[((size_t)b * p.tm_C + p.tm_c0 + c0 + i) * p.T + t] = v[i];
                }
            }
            if (p.out_bf16) out_chunks = (p.N + KC - 1) / KC;
        } else if constexpr (EPI == EPI_GATE_BWD) {
            const int H = p.N, nck = H / 16;
            const unsigned long long dseed = p.drop_p > 0.f ? drop_resolve(p.drop_seed) : 0ull;
            for (int cc = nck * part / EPI_PARTS; cc < nck * (part + 1) / EPI_PARTS; ++cc) {
                const int c0 = cc * 16;
                float d[16];
                tmem_ld16_dual(trow + c0, dual, d);
                if (!t_ok) continue;
                float tt[16], ss[16], dt[16], ds[16];
                row_load_bf16x16(p.ts, 2 * H, c0, io, tt);
                row_load_bf16x16(p.ts, 2 * H, H + c0, io, ss);
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    dt[i] = live ? d[i] * ss[i] * (1.f - tt[i] * tt[i]) : 0.f;
                    ds[i] = live ? d[i] * tt[i] * ss[i] * (1.f - ss[i]) : 0.f;
                }
                if (p.out2_bf16) {
                    store_bf16x16(p.out2_bf16 + frame * (2 * H) + c0, dt);
                    store_bf16x16(p.out2_bf16 + frame * (2 * H) + H + c0, ds);
                }
                if (p.drop_p > 0.f) {
                    float sc[16];
                    const unsigned long long dframe = p.row_fr ? (unsigned long long)p.row_fr[frame] : (unsigned long long)frame;
                    dropout_scale16(dseed, dframe * (2 * H) + c0, p.drop_p, sc);
#pragma unroll
                    for (int i = 0; i < 16; ++i) dt[i] *= sc[i];
                    dropout_scale16(dseed, dframe * (2 * H) + H + c0, p.drop_p, sc);
#pragma unroll
                    for (int i = 0; i < 16; ++i) ds[i] *= sc[i];
                }
                stage_bf16x16_sw128(smem, A_BYTES, io.r, c0, dt);
                stage_bf16x16_sw128(smem, A_BYTES, io.r, H + c0, ds);
            }
            out_chunks = 2 * H / KC;
        } else {   // EPI_AFFINE
            const int Ch = p.N / 2, C = p.N;
            float ld_sum = 0.f;
            const int nck = Ch / 16;
            for (int cc = nck * part / EPI_PARTS; cc < nck * (part + 1) / EPI_PARTS; ++cc) {
                const int c0 = cc * 16;
                float m[16], lg[16];
                tmem_ld16_dual(trow + c0, dual, m);
                tmem_ld16_dual(trow + Ch + c0, dual, lg);
                if (!t_ok) continue;
                // all loads of the chunk first (x and z may alias as far as the compiler knows: loads behind stores
                // would serialise into one L2 round trip per channel)
                const float* __restrict__ ax = p.aff_x;
                float* __restrict__ az = p.aff_z;
                const size_t o0 = ((size_t)b * C + c0) * p.T + t, o1 = o0 + (size_t)Ch * p.T;
                float x0[16], x1[16], lsave[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) { x0[i] = ax[o0 + (size_t)i * p.T]; x1[i] = ax[o1 + (size_t)i * p.T]; }
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int c = c0 + i;
                    float mm = m[i] + ctl->bias[c];
                    float ll = lg[i] + ctl->bias[Ch + c];
                    lsave[i] = ll;     // raw end-conv output (pre sigmoid_scale), as the fp32 path keeps it
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
                if (!t_ok) ld_sum = 0.f;
                if (p.row_b) {
                    // packed rows: a warp's 32 rows usually belong to one utterance; at a boundary each lane adds its own
                    const int ub = p.row_b[frame];
                    const int ub0 = __shfl_sync(0xffffffffu, ub, 0);
                    if (__all_sync(0xffffffffu, ub == ub0)) {
                        ld_sum = warp_sum(ld_sum);
                        if (lane == 0 && ld_sum != 0.f) atomicAdd(p.aff_logdet + ub0, ld_sum);
                    } else if (ld_sum != 0.f) {
                        atomicAdd(p.aff_logdet + ub, ld_sum);
                    }
                } else {
                    ld_sum = warp_sum(ld_sum);
                    if (lane == 0) atomicAdd(p.aff_logdet + b, ld_sum);
                }
            }
        }
        if (out_chunks) fence_proxy_async_smem();
        tc_fence_before();
    }
    CONV_TRACE(3);
    __syncthreads();
    if (out_chunks && warp == 1 && lane == 0) {
        const int reps = p.rep_bf16 > 1 ? p.rep_bf16 : 1;
        for (int rp = 0; rp < reps; ++rp)
            for (int ch = 0; ch < out_chunks; ++ch) tma_store_4d(&tmO, smem + (size_t)ch * A_BYTES, ch * KC, t0, b, rp);
        tma_store_commit();
        tma_store_wait_read();                           // shared memory must outlive the bulk reads
    }
    CONV_TRACE(4);
    if constexpr (CL > 1) cluster_sync_all();             // no CTA leaves while the peer's commits may still arrive here
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc<512>(tmem);
    }
}

#ifdef GTK_INSTRUMENT
void conv_trace_select(int epi, int taps) { g_trace_epi = epi; g_trace_taps = taps; }
int conv_trace_read(unsigned long long* out, int n_ctas) {
    GTK_REQUIRE(out && n_ctas >= 0 && n_ctas <= CONV_TRACE_CTAS, "n_ctas");
    GTK_CUDA(cudaMemcpyFromSymbol(out, g_conv_trace, size_t(n_ctas) * 5 * sizeof(unsigned long long)));
    return GTK_OK;
}
int conv_wait_read(long long* out, int n_ctas) {
    GTK_REQUIRE(out && n_ctas >= 0 && n_ctas <= CONV_TRACE_CTAS, "n_ctas");
    GTK_CUDA(cudaMemcpyFromSymbol(out, g_conv_wait, size_t(n_ctas) * 2 * sizeof(long long)));
    return GTK_OK;
}
#endif

// ---------------------------------------------------------------------------------- helpers: packing / layout
// W fp32 [M][Cin][K] (torch) -> bf16 [K][rows][cols] zero padded.
//   transpose = 0 (forward operand):        out[j][m][c] = W[m][c][j]            rows >= M,   cols >= Cin
//   transpose = 1 (data-gradient operand):  out[j][c][m] = W[m][c][K-1-j]        rows >= Cin, cols >= M
__global__ void pack_weight_bf16_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out, int M, int Cin, int K,
                                        int transpose, int rows, int cols) {
    const long long total = (long long)K * rows * cols;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % cols);
        long long r = i / cols;
        const int row = (int)(r % rows), j = (int)(r / rows);
        float v = 0.f;
        if (!transpose) { if (row < M && c < Cin) v = w[((size_t)row * Cin + c) * K + j]; }
        else            { if (row < Cin && c < M) v = w[((size_t)c * Cin + row) * K + (K - 1 - j)]; }
        out[i] = __float2bfloat16_rn(v);
    }
}
int pack_weight_bf16(const float* w, __nv_bfloat16* out, int M, int Cin, int K, int transpose, int rows, int cols, cudaStream_t st) {
    GTK_REQUIRE(w && out, "null");
    const long long total = (long long)K * rows * cols;
    if (total == 0) return GTK_OK;
    long long g = ceil_div_ll(total, 256);
    pack_weight_bf16_kernel<<<(int)(g > 2048 ? 2048 : g), 256, 0, st>>>(w, out, M, Cin, K, transpose, rows, cols);
    return check_launch("pack_weight_bf16");
}

// x fp32 t-major [B, C, T] channels [c0, c0+nc) -> bf16 frame-major out[b][t][0..ld): columns >= nc are zero.
__global__ void frames_from_tmajor_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int C, int T, int c0,
                                          int nc, int ld, const int* __restrict__ n_rows) {
    __shared__ float tile[32][33];
    pdl_trigger();
    pdl_wait();
    const int b = blockIdx.z, tt = blockIdx.x * 32, cc = blockIdx.y * 32;
    if (n_rows && tt >= ((*n_rows + 127) & ~127)) return;       // varlen packing: nothing lives behind the last 128-row tile
    const int tx = threadIdx.x, ty = threadIdx.y;   // 32 x 8
    for (int i = ty; i < 32; i += 8) {
        const int c = cc + i, t = tt + tx;
        tile[i][tx] = (c < nc && t < T) ? x[((size_t)b * C + c0 + c) * T + t] : 0.f;
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        const int t = tt + i, c = cc + tx;
        if (t < T && c < ld) out[((size_t)b * T + t) * ld + c] = __float2bfloat16_rn(tile[tx][i]);
    }
}
int frames_from_tmajor(const float* x, __nv_bfloat16* out, int B, int C, int T, int c0, int nc, int ld, cudaStream_t st,
                       const int* n_rows) {
    GTK_REQUIRE(x && out && c0 >= 0 && c0 + nc <= C && ld >= nc, "bad args");
    if ((long long)B * T * ld == 0) return GTK_OK;
    dim3 grid(ceil_div(T, 32), ceil_div(ld, 32), B);
    GTK_CUDA(launch_pdl(frames_from_tmajor_kernel, grid, dim3(32, 8), 0, st, x, out, C, T, c0, nc, ld, n_rows));
    return check_launch("frames_from_tmajor");
}

// ---------------------------------------------------------------------------------- fold + pack, one launch per block
// Every conv weight of the decoder: weight-norm fold (w = g v / ||v||, modules.py:124,131,141) and bf16 packing into the
// forward ([j][m][c]) and data-gradient ([j][c][m], taps flipped) operand layouts, zero padding included, in ONE launch.
// One CTA per (weight, group of FP_ROWS output rows m).  The CTA's rows are read from HBM ONCE into shared memory (each
// warp streams whole rows, 8 independent loads in flight per lane, and reduces the row's sum of squares on the way), then
// both operand layouts are written from shared memory as bf16 pairs with the row scale applied at the store: contiguous
// 128-byte runs for fwd (c fastest), 32-byte runs for tr (m fastest).  No per-element integer division.
constexpr int FP_ROWS = 16, FP_NT = 256;
__global__ void __launch_bounds__(FP_NT) fold_pack_kernel(const __grid_constant__ FoldPackJobs jobs) {
    extern __shared__ float fp_tile[];               // [FP_ROWS][Cin * K + 1]
    __shared__ float sc_sm[FP_ROWS];
    int grp = blockIdx.x, w = 0;
    while (w < jobs.n && grp >= jobs.job[w].groups) { grp -= jobs.job[w].groups; ++w; }
    if (w >= jobs.n) return;
    const FoldPackJob& J = jobs.job[w];
    const int K = J.K, Cin = J.Cin, R = Cin * K, P = R + 1, m0 = grp * FP_ROWS;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // ---- load + row scales (rows >= M are padding rows of the transposed pack: zeros, scale 0).  A warp owns rows
    // `warp` and `warp + 8` and keeps 2 x 16 loads in flight per lane: a 960-float row arrives in two round trips.
    {
        constexpr int NW = FP_NT / 32, U = 16;
        static_assert(FP_ROWS == 2 * NW, "two rows per warp");
        const float* vrow[2];
        float* trow[2];
        bool live[2];
        float ss[2] = {0.f, 0.f};
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int rl = warp + q * NW, m = m0 + rl;
            live[q] = m < J.M;
            vrow[q] = J.v + (size_t)(live[q] ? m : 0) * R;
            trow[q] = fp_tile + (size_t)rl * P;
        }
        for (int e0 = lane; e0 < R; e0 += 32 * U) {
            float t[2][U];
#pragma unroll
            for (int q = 0; q < 2; ++q)
#pragma unroll
                for (int u = 0; u < U; ++u) t[q][u] = (live[q] && e0 + 32 * u < R) ? __ldg(vrow[q] + e0 + 32 * u) : 0.f;
#pragma unroll
            for (int q = 0; q < 2; ++q)
#pragma unroll
                for (int u = 0; u < U; ++u)
                    if (e0 + 32 * u < R) { trow[q][e0 + 32 * u] = t[q][u]; ss[q] = fmaf(t[q][u], t[q][u], ss[q]); }
        }
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int rl = warp + q * NW, m = m0 + rl;
            const float s2 = warp_sum(ss[q]);
            if (lane == 0) sc_sm[rl] = live[q] ? (J.g ? J.g[m] / sqrtf(s2) : 1.f) : 0.f;
        }
    }
    __syncthreads();
    // ---- forward operand: out[j][m][c], c fastest (columns Cin..cols_f are zero padding)
    for (int idx = warp; idx < K * FP_ROWS; idx += FP_NT / 32) {
        const int j = idx / FP_ROWS, rl = idx % FP_ROWS, m = m0 + rl;
        if (m >= J.rows_f) continue;
        const float sc = sc_sm[rl];
        const float* trow = fp_tile + (size_t)rl * P + j;
        __nv_bfloat16* dst = J.fwd + ((size_t)j * J.rows_f + m) * J.cols_f;
        for (int c = 2 * lane; c < J.cols_f; c += 64) {
            const float a = c < Cin ? trow[c * K] * sc : 0.f;
            const float b = c + 1 < Cin ? trow[(c + 1) * K] * sc : 0.f;
            if (c + 1 < J.cols_f && (J.cols_f & 1) == 0) *reinterpret_cast<__nv_bfloat162*>(dst + c) = __floats2bfloat162_rn(a, b);
            else {
                dst[c] = __float2bfloat16_rn(a);
                if (c + 1 < J.cols_f) dst[c + 1] = __float2bfloat16_rn(b);
            }
        }
    }
    // ---- data-gradient operand: out[j][c][m] = W[m][c][K-1-j], m fastest (columns M..cols_t are zero padding)
    if (J.tr) {
        const int sub = lane >> 3, rp = (lane & 7) * 2, m = m0 + rp;
        const float s0 = sc_sm[rp], s1 = sc_sm[rp + 1];
        const bool pair = (J.cols_t & 1) == 0;
        for (int j = 0; j < K; ++j) {
            const float* tj = fp_tile + (K - 1 - j);
            for (int c = warp * 4 + sub; c < Cin; c += (FP_NT / 32) * 4) {
                if (m >= J.cols_t) continue;
                const float a = tj[(size_t)rp * P + c * K] * s0;
                const float b = tj[(size_t)(rp + 1) * P + c * K] * s1;
                __nv_bfloat16* dst = J.tr + ((size_t)j * J.rows_t + c) * J.cols_t + m;
                if (pair) *reinterpret_cast<__nv_bfloat162*>(dst) = __floats2bfloat162_rn(a, b);
                else {
                    dst[0] = __float2bfloat16_rn(a);
                    if (m + 1 < J.cols_t) dst[1] = __float2bfloat16_rn(b);
                }
            }
        }
    }
}
int fold_pack(FoldPackJobs& jobs, cudaStream_t st) {
    int groups = 0, kmax = 1, rmax = 1;
    for (int i = 0; i < jobs.n; ++i) {
        FoldPackJob& J = jobs.job[i];
        GTK_REQUIRE(J.rows_f == J.M && J.rows_t == J.Cin, "pack shapes");
        const int rows = J.tr ? (J.cols_t > J.M ? J.cols_t : J.M) : J.M;
        J.groups = ceil_div(rows, FP_ROWS);
        groups += J.groups;
        kmax = max(kmax, J.K);
        rmax = max(rmax, J.Cin * J.K);
    }
    (void)kmax;
    if (groups == 0) return GTK_OK;
    const size_t smem = (size_t)FP_ROWS * ((size_t)rmax + 1) * sizeof(float);
    GTK_REQUIRE(smem <= 200 * 1024, "weight rows too long for the pack kernel");
    GTK_CUDA(cudaFuncSetAttribute(fold_pack_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ProfScope prof("fold_pack", st);
    fold_pack_kernel<<<groups, FP_NT, smem, st>>>(jobs);
    return check_launch("fold_pack");
}

// ---------------------------------------------------------------------------------- affine coupling backward
// attentions.py:172-183 backward, one thread per frame: g = dz1 * mask; d(end_out) = [g | (g e^logs x1 + dlogdet mask) ds]
// as bf16 frames (operand of the `end` data/weight gradients), dx1 = g e^logs (t-major fp32).
__global__ void affine_bwd_frames_kernel(const float* __restrict__ x, const float* __restrict__ logs_raw,
                                         const float* __restrict__ mask, const float* __restrict__ dz,
                                         const float* __restrict__ dlogdet, __nv_bfloat16* __restrict__ dend, int ld,
                                         float* __restrict__ dx, int B, int C, int T, int sigmoid_scale,
                                         const int* __restrict__ row_b, const int* __restrict__ n_rows) {
    const int Ch = C / 2;
    pdl_trigger();
    pdl_wait();
    const int n_chunk = (Ch + 7) / 8, n_pad = (ld - C + 7) / 8;      // channel chunks + zero-padding chunks
    const long long frames = (long long)B * T, total = frames * (n_chunk + n_pad);
    const long long f_lim = n_rows ? (long long)((*n_rows + 127) & ~127) : frames;      // varlen packing (B == 1)
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long f = i % frames;
        if (f >= f_lim) continue;
        const int chunk = (int)(i / frames);
        const int b = (int)(f / T), t = (int)(f % T);
        __nv_bfloat16* row = dend + (size_t)f * ld;
        if (chunk >= n_chunk) {
            const int c0 = C + (chunk - n_chunk) * 8;
            if (c0 + 8 <= ld && (C % 8 == 0) && (ld % 8 == 0)) *reinterpret_cast<uint4*>(row + c0) = make_uint4(0u, 0u, 0u, 0u);
            else for (int c = c0; c < min(c0 + 8, ld); ++c) row[c] = __float2bfloat16_rn(0.f);
            continue;
        }
        const float mk = mask ? mask[f] : 1.f;
        const float dl = dlogdet ? dlogdet[row_b ? row_b[f] : b] * mk : 0.f;
        const int c0 = chunk * 8;
        const bool vec = (Ch % 8 == 0) && (ld % 8 == 0);          // 16-byte aligned 8-channel groups
        float gm[8], gl[8], raw8[8], dz8[8], x8[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) { raw8[k] = 0.f; dz8[k] = 0.f; x8[k] = 0.f; }
        if (vec) {
            const float4* r4 = reinterpret_cast<const float4*>(logs_raw + (size_t)f * Ch + c0);
            const float4 a = r4[0], bq = r4[1];
            raw8[0] = a.x; raw8[1] = a.y; raw8[2] = a.z; raw8[3] = a.w;
            raw8[4] = bq.x; raw8[5] = bq.y; raw8[6] = bq.z; raw8[7] = bq.w;
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {                              // all loads before any store
            const int c = c0 + k;
            if (c < Ch) {
                const size_t o1 = ((size_t)b * C + Ch + c) * T + t;
                if (!vec) raw8[k] = logs_raw[(size_t)f * Ch + c];
                dz8[k] = dz[o1];
                x8[k] = x[o1];
            }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int c = c0 + k;
            gm[k] = gl[k] = 0.f;
            if (c < Ch) {
                const size_t o1 = ((size_t)b * C + Ch + c) * T + t;
                float lg = raw8[k], dscale = 1.f;
                if (sigmoid_scale) {
                    const float sg = sigmoid_f(raw8[k] + 2.f);
                    lg = __logf(1e-6f + sg);
                    dscale = sg * (1.f - sg) / (1e-6f + sg);
                }
                const float e = __expf(lg);
                const float g = dz8[k] * mk;
                gm[k] = g;
                gl[k] = (g * e * x8[k] + dl) * dscale;
                dx[o1] = g * e;
            }
        }
        if (vec) {
            uint32_t wm[4], wl[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                __nv_bfloat162 hm = __floats2bfloat162_rn(gm[2 * k], gm[2 * k + 1]);
                __nv_bfloat162 hl = __floats2bfloat162_rn(gl[2 * k], gl[2 * k + 1]);
                wm[k] = *reinterpret_cast<uint32_t*>(&hm);
                wl[k] = *reinterpret_cast<uint32_t*>(&hl);
            }
            *reinterpret_cast<uint4*>(row + c0) = make_uint4(wm[0], wm[1], wm[2], wm[3]);
            *reinterpret_cast<uint4*>(row + Ch + c0) = make_uint4(wl[0], wl[1], wl[2], wl[3]);
        } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int c = c0 + k;
                if (c < Ch) { row[c] = __float2bfloat16_rn(gm[k]); row[Ch + c] = __float2bfloat16_rn(gl[k]); }
            }
        }
    }
}
int affine_bwd_frames(const float* x, const float* logs_raw, const float* mask, const float* dz, const float* dlogdet,
                      __nv_bfloat16* dend, int ld, float* dx, int B, int C, int T, int sigmoid_scale, const int* row_b,
                      const int* n_rows, cudaStream_t st) {
    GTK_REQUIRE(x && logs_raw && dz && dend && dx && C % 2 == 0 && ld >= C, "bad args");
    const long long total = (long long)B * T * ((C / 2 + 7) / 8 + (ld - C + 7) / 8);
    if (total == 0) return GTK_OK;
    ProfScope prof("affine_bwd_frames", st);
    GTK_CUDA(launch_pdl(affine_bwd_frames_kernel, dim3((unsigned)ceil_div_ll(total, 128)), dim3(128), 0, st, x, logs_raw, mask, dz,
                        dlogdet, dend, ld, dx, B, C, T, sigmoid_scale, row_b, n_rows));
    return check_launch("affine_bwd_frames");
}

// ---------------------------------------------------------------------------------- launcher
static size_t conv_tc_stage_area(int N, int stages, int out_cols) {
    const size_t ops = (size_t)stages * (A_BYTES + (size_t)N * KC * 2);
    const size_t staging = (size_t)ceil_div(out_cols, KC) * A_BYTES;      // bf16 output tiles staged for the TMA stores
    return ops > staging ? ops : staging;
}
static size_t conv_tc_smem(int N, int stages, int out_cols) { return conv_tc_stage_area(N, stages, out_cols) + sizeof(SmemCtl) + 1024; }

// in: frames with row pitch ld_in elements, of which the first p.n_k * 64 columns are the conv's input channels
// (the weights are packed with exactly that many columns).
int conv_tc_launch(const __nv_bfloat16* in, int ld_in, const __nv_bfloat16* wp, ConvTcParams p, int epi, cudaStream_t st) {
    GTK_REQUIRE(in && wp, "null");
    GTK_REQUIRE(p.taps >= 1 && (p.taps & 1) && p.dil >= 1 && p.n_k >= 1, "bad conv shape");
    GTK_REQUIRE(ld_in >= p.n_k * KC && ld_in % 8 == 0, "input frames: row pitch must cover n_k * 64 channels, 16-byte aligned");
    GTK_REQUIRE(p.N >= 16 && p.N % 16 == 0 && p.N <= 512, "N must be a multiple of 16, <= 512");
    if (p.B == 0 || p.T == 0) return GTK_OK;
    p.NCH = (p.N > 256) ? 2 : 1;
    p.NW = p.N / p.NCH;
    GTK_REQUIRE(p.NW % 16 == 0 && p.NW <= 256, "N chunk must be a multiple of 16");
    // weight-tile sharing between two frame tiles pays when the mainloop is long (k > 1 taps)
    p.cluster = (epi == EPI_PLAIN && p.taps > 1 && p.N <= 256 && (p.NW / 2) % 8 == 0 && ceil_div(p.T, BT) > 1) ? 2 : 1;
    const int out_cols = (epi == EPI_GATE_BWD) ? 2 * p.N : ((epi == EPI_PLAIN && p.out_bf16) ? p.N : 0);
    int stages = 4;
    while (stages > 2 && conv_tc_smem(p.N, stages, out_cols) > 200 * 1024) --stages;
    p.stages = stages;
    p.ctl_off = (int)conv_tc_stage_area(p.N, stages, out_cols);
    const size_t smem = conv_tc_smem(p.N, stages, out_cols);
    GTK_REQUIRE(smem <= 227 * 1024, "tile does not fit in shared memory");
    GTK_REQUIRE(!p.tile_layout || ((!p.out_f32 || p.ld_f32 == p.N) && (!p.resid_f32 || p.ld_resid == p.N)),
                "tile-private layout: row pitch must equal the column count");

    CUtensorMap tmA, tmB, tmO;
    if (out_cols) {
        GTK_REQUIRE(p.out_bf16 && p.ld_bf16 % 8 == 0 && p.ld_bf16 >= out_cols, "bf16 output frames: row pitch");
        const int reps = p.rep_bf16 > 1 ? p.rep_bf16 : 1;
        const uint64_t dims[4] = {(uint64_t)out_cols, (uint64_t)p.T, (uint64_t)p.B, (uint64_t)reps};
        const uint64_t strides[3] = {(uint64_t)p.ld_bf16 * 2, (uint64_t)p.ld_bf16 * 2 * p.T,
                                     reps > 1 ? (uint64_t)p.rep_stride * 2 : (uint64_t)p.ld_bf16 * 2 * p.T * p.B};
        const uint32_t box[4] = {KC, BT, 1, 1};
        int rc = make_tmap_bf16(&tmO, p.out_bf16, 4, dims, strides, box, true);
        if (rc) return rc;
    } else {
        memset(&tmO, 0, sizeof(tmO));
    }
    {
        const uint64_t dims[3] = {(uint64_t)p.n_k * KC, (uint64_t)p.T, (uint64_t)p.B};
        const uint64_t strides[2] = {(uint64_t)ld_in * 2, (uint64_t)ld_in * 2 * p.T};
        const uint32_t box[3] = {KC, BT, 1};
        int rc = make_tmap_bf16(&tmA, in, 3, dims, strides, box, true);
        if (rc) return rc;
    }
    {
        const uint64_t kcols = (uint64_t)p.n_k * KC;
        const uint64_t dims[3] = {kcols, (uint64_t)p.N, (uint64_t)p.taps};
        const uint64_t strides[2] = {kcols * 2, kcols * 2 * p.N};
        const uint32_t box[3] = {KC, (uint32_t)(p.NW / p.cluster), 1};
        int rc = make_tmap_bf16(&tmB, wp, 3, dims, strides, box, true);
        if (rc) return rc;
    }
    dim3 grid(ceil_div(ceil_div(p.T, BT), p.cluster) * p.cluster, p.B);   // an out-of-range tile only feeds its peer
#ifdef GTK_INSTRUMENT
    p.trace = (epi == g_trace_epi && p.taps == g_trace_taps) ? 1 : 0;
#endif
    ProfScope prof(epi == EPI_GATE_BWD ? "conv_tc_gate_bwd" : epi == EPI_AFFINE ? "conv_tc_affine" : p.taps > 1 ? "conv_tc_kN" : "conv_tc_k1", st);
#define GTK_TC_LAUNCH(COLS, E)                                                                                       \
    do {                                                                                                             \
        GTK_CUDA(cudaFuncSetAttribute(conv_tc_kernel<COLS, E, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        GTK_CUDA(launch_pdl(conv_tc_kernel<COLS, E, 1>, grid, dim3(EPI_NT), smem, st, tmA, tmB, tmO, p));          \
    } while (0)
    if (epi == EPI_PLAIN && p.cluster == 2) {
        GTK_CUDA(cudaFuncSetAttribute(conv_tc_kernel<256, EPI_PLAIN, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = grid; cfg.blockDim = dim3(EPI_NT); cfg.dynamicSmemBytes = smem; cfg.stream = st;
        cudaLaunchAttribute attr[2];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
        cfg.attrs = attr; cfg.numAttrs = 2;
        GTK_CUDA(cudaLaunchKernelEx(&cfg, conv_tc_kernel<256, EPI_PLAIN, 2>, tmA, tmB, tmO, p));
    }
    else if (epi == EPI_PLAIN) { if (p.N > 256) GTK_TC_LAUNCH(512, EPI_PLAIN); else GTK_TC_LAUNCH(256, EPI_PLAIN); }
    else if (epi == EPI_GATE_BWD) { GTK_REQUIRE(p.N <= 256 && p.ts && p.out_bf16, "gate_bwd epilogue arguments"); GTK_TC_LAUNCH(256, EPI_GATE_BWD); }
    else { GTK_REQUIRE(p.N <= 256 && p.N % 32 == 0 && p.aff_x && p.aff_z, "affine epilogue arguments"); GTK_TC_LAUNCH(256, EPI_AFFINE); }
#undef GTK_TC_LAUNCH
    return check_launch("conv_tc");
}

int conv1d_plain(const __nv_bfloat16* in, int ld_in, const __nv_bfloat16* wp, const float* bias, const float* mask, int B, int T,
                 int K, int dil, int N, float* out_f32, int ld_f32, __nv_bfloat16* out_bf16, int ld_bf16, int mask_out,
                 cudaStream_t st) {
    ConvTcParams p;
    memset(&p, 0, sizeof(p));
    p.B = B; p.T = T; p.taps = K; p.dil = dil; p.n_k = ld_in / KC; p.N = N;
    p.bias = bias; p.mask = mask; p.mask_out = mask_out;
    p.out_f32 = out_f32; p.ld_f32 = ld_f32; p.out_bf16 = out_bf16; p.ld_bf16 = ld_bf16;
    return conv_tc_launch(in, ld_in, wp, p, EPI_PLAIN, st);
}

}  // namespace tc
}  // namespace gtk
