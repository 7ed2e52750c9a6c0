// glowtts_b200/csrc/common.cuh -- shared device/host helpers for the sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#define GTK_OK 0
#define GTK_ERR_INVALID (-1)      // bad argument (null pointer, negative size, unsupported shape)
#define GTK_ERR_LAUNCH (-2)       // CUDA reported an error at launch / attribute set
#define GTK_ERR_WORKSPACE (-3)    // workspace too small
#define GTK_ERR_UNSUPPORTED (-4)  // shape outside what the kernels cover

namespace gtk {

void set_error(const char* fmt, ...);   // thread-local message for gtk_last_error()

void count_launch();                    // bumps the library's kernel-launch counter (gtk_launch_count)

// Optional per-kernel timing (gtk_prof_*): when enabled, a ProfScope around a launch records a CUDA event
// pair on the launching stream under `tag`; bench.py reads the totals for the roofline numbers.
void prof_begin(const char* tag, cudaStream_t st);
void prof_end(const char* tag, cudaStream_t st);
struct ProfScope {
    const char* tag; cudaStream_t st;
    ProfScope(const char* t, cudaStream_t s) : tag(t), st(s) { prof_begin(tag, st); }
    ~ProfScope() { prof_end(tag, st); }
};

inline int check_launch(const char* what) {
    count_launch();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return GTK_ERR_LAUNCH;
    }
    return GTK_OK;
}

#define GTK_CUDA(call)                                                        \
    do {                                                                      \
        cudaError_t e__ = (call);                                             \
        if (e__ != cudaSuccess) {                                             \
            gtk::set_error("%s: %s", #call, cudaGetErrorString(e__));         \
            return GTK_ERR_LAUNCH;                                            \
        }                                                                     \
    } while (0)

#define GTK_REQUIRE(cond, msg)                                                \
    do {                                                                      \
        if (!(cond)) {                                                        \
            gtk::set_error("%s: requirement failed: %s", __func__, msg);      \
            return GTK_ERR_INVALID;                                           \
        }                                                                     \
    } while (0)

int sm_count();   // cached multiprocessor count of the current device

// Programmatic dependent launch (PDL).  The decoder step is ~250 dependent launches of 5-30 us; launched with
// cudaLaunchAttributeProgrammaticStreamSerialization a kernel's CTAs may become resident -- and run their preamble
// (mbarrier init, TMEM allocation, constant staging) -- while the previous kernel drains, instead of after a full
// grid drain + launch latency.  Protocol: every PDL kernel executes pdl_trigger() on entry (lets ITS successor
// start early) and pdl_wait() before its first access to memory an earlier kernel of the stream may have written
// (griddepcontrol.wait returns once the predecessor grid has completed and flushed).  Nothing before pdl_wait() may
// write global memory.  GTK_PDL=0 disables the attribute (the instructions are then no-ops).
bool pdl_enabled();
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
__host__ __device__ inline long long ceil_div_ll(long long a, long long b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------- device helpers
#ifdef __CUDACC__

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// mbarrier (shared::cta) -- producer/consumer pipelines without block-wide barriers.
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ bool mbar_test_wait(uint64_t* bar, uint32_t parity) {    // non-suspending probe
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a pipeline bug traps (launch error) instead of hanging the GPU box.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > (1u << 24)) {
            printf("gtk: mbarrier wait timed out (block %d thread %d)\n", (int)blockIdx.x, (int)threadIdx.x);
            __trap();
        }
    }
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Block-wide sum; result valid in thread 0 (and broadcast through smem to all).
template <int NT>
__device__ __forceinline__ float block_sum(float v, float* scratch /* >= 33 floats */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = warp_sum(v);
    if (lane == 0) scratch[warp] = v;
    __syncthreads();
    if (warp == 0) {
        float t = (lane < NT / 32) ? scratch[lane] : 0.f;
        t = warp_sum(t);
        if (lane == 0) scratch[32] = t;
    }
    __syncthreads();
    float r = scratch[32];
    __syncthreads();
    return r;
}

#endif  // __CUDACC__
// ---------------------------------------------------------------------------------------------- dropout
// Counter-based dropout shared by every kernel that applies or back-propagates nn.Dropout on the WaveNet in_layer
// output (modules.py:153).  Element (row = b*T + t, col in [0, 2H)) has index idx = row * 2H + col (< 2^32); one 32-bit
// hash (two multiplies) of (layer seed, idx >> 1) yields two 16-bit uniforms, element idx keeps iff its field >= thr =
// round(p * 65536), and kept values are scaled by 1 / (1 - thr/65536) (exactly unbiased for the quantised p).  The
// epilogues that apply it are instruction-bound, hence the cheap 32-bit mixer instead of a 64-bit one.
__host__ __device__ __forceinline__ unsigned long long drop_hash(unsigned long long seed, unsigned long long ctr) {
    unsigned long long z = seed + (ctr + 1ull) * 0x9E3779B97F4A7C15ull;      // splitmix64: seed derivation only
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ unsigned drop_bits32(unsigned long long seed, unsigned pair) {
    unsigned x = pair * 0x9E3779B1u + (unsigned)seed;
    x ^= x >> 16; x *= 0x7FEB352Du;
    x ^= x >> 15; x ^= (unsigned)(seed >> 32); x *= 0x846CA68Bu;
    return x ^ (x >> 16);
}
__host__ __device__ __forceinline__ unsigned drop_threshold(float p) { return (unsigned)(p * 65536.f + 0.5f); }
__host__ __device__ __forceinline__ float drop_inv_keep(float p) { return 65536.f / (65536.f - (float)drop_threshold(p)); }
__host__ __device__ __forceinline__ float dropout_scale1(unsigned long long seed, unsigned long long idx, float p) {
    const unsigned h = drop_bits32(seed, (unsigned)(idx >> 1));
    const unsigned u = (h >> (16 * (unsigned)(idx & 1))) & 0xffffu;
    return u >= drop_threshold(p) ? drop_inv_keep(p) : 0.f;
}
// 16 consecutive elements starting at idx0 (a multiple of 16): eight hashes.
__device__ __forceinline__ void dropout_scale16(unsigned long long seed, unsigned long long idx0, float p, float (&s)[16]) {
    const unsigned thr = drop_threshold(p);
    const float keep = drop_inv_keep(p);
    const unsigned pair0 = (unsigned)(idx0 >> 1);
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const unsigned h = drop_bits32(seed, pair0 + q);
        s[2 * q] = ((h & 0xffffu) >= thr) ? keep : 0.f;
        s[2 * q + 1] = ((h >> 16) >= thr) ? keep : 0.f;
    }
}
// Seed of layer `layer` of block `blk` derived from the per-call seed in gtk_decoder_desc.
__host__ __device__ __forceinline__ unsigned long long drop_layer_seed(unsigned long long seed, int blk, int layer) {
    return drop_hash(seed ^ 0xD6E8FEB86659FD93ull, (unsigned long long)blk * 64ull + (unsigned long long)layer);
}
// The per-call seed may live in device memory (gtk_decoder_desc.seed_dev: CUDA-graph replays then see a fresh seed
// without re-capturing).  Kernels take (value, dev): dev == null -> value is the layer seed; otherwise value is the
// layer's salt blk * 64 + layer and the layer seed is derived on the device from *dev.
struct DropSeed {
    unsigned long long value;
    const unsigned long long* dev;
};
inline DropSeed make_drop_seed(unsigned long long call_seed, const unsigned long long* dev, int blk, int layer) {
    DropSeed s;
    s.dev = dev;
    s.value = dev ? (unsigned long long)blk * 64ull + (unsigned long long)layer : drop_layer_seed(call_seed, blk, layer);
    return s;
}
__device__ __forceinline__ unsigned long long drop_resolve(const DropSeed& s) {
    return s.dev ? drop_hash(*s.dev ^ 0xD6E8FEB86659FD93ull, s.value) : s.value;
}

}  // namespace gtk
