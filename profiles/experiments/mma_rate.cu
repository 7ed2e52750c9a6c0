// Microbenchmark (round 2): steady-state cycles per tcgen05.mma (cta_group::1, kind::f16, M = 128, K = 16) issued by one
// elected lane from operands resident in shared memory (no TMA in the loop), for the operand layouts / N / accumulator
// rotations the library's kernels use.  One CTA per SM on all SMs (so that clocks and power behave as in a real kernel).
// Build:  nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -I glowtts_b200/csrc -I include \
//              profiles/experiments/mma_rate.cu -o profiles/experiments/mma_rate
// Run  :  profiles/experiments/mma_rate        (prints one line per variant; results in profiles/mma_issue_r02.txt)
#include <cstdio>
#include <vector>
#include "tc.cuh"
using namespace gtk;
using namespace gtk::tc;

struct Ctl { uint64_t done; uint32_t tmem_base; };

// mn: 1 = MN-major A and B (weight-gradient kernels), 0 = K-major (conv kernels).  n_acc accumulators at column stride
// acc_stride take consecutive MMAs in rotation; b_shift_rows: B start address advances by this many 128-byte rows per
// accumulator (the all-taps weight gradient), a_same: every MMA of a rotation reads the same A.
__global__ void __launch_bounds__(128, 1) rate_kernel(int N, int mn, int n_acc, int acc_stride, int b_shift_rows, int n_mma,
                                                      long long* out) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    Ctl* ctl = reinterpret_cast<Ctl*>(smem + 4 * 49152);
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 4 * 49152 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { mbar_init(&ctl->done, 1); mbar_fence_init(); }
    if (warp == 2) tmem_alloc<512>(&ctl->tmem_base);
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctl->tmem_base;
    if (warp == 1) {
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
        const uint32_t idesc = umma_idesc_bf16(128, N > 0 ? N : 256, mn, mn), idesc_hi = umma_idesc_bf16(128, 128, mn, mn);
        long long t0 = 0, t1 = 0;
        for (int rep = 0; rep < 2; ++rep) {                    // rep 0 = warm-up
            t0 = clock64();
            for (int i = 0; i < n_mma; i += 4 * n_acc) {
                const int s = (i / (4 * n_acc)) & (N < 0 ? 1 : 3);      // the 256 | 128 variant reads 48 KB of B per stage
                const uint32_t sa = smem_u32(smem + (size_t)s * (N < 0 ? 98304 : 49152)), sb = sa + 16384;
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint64_t adesc = mn ? umma_desc_mnmajor_sw128(sa + k * 2048, 8192, 1024) : umma_desc_kmajor_sw128(sa) + 2 * k;
                        if (N < 0) {                          // the 256 | 128 column split of a 384-column k16 step
                            const uint64_t bdesc = umma_desc_kmajor_sw128(sb) + 2 * k;
                            umma_bf16(tm, adesc, bdesc, idesc, i > 0 || k > 0);
                            umma_bf16(tm + 256, adesc, bdesc + 2048, idesc_hi, i > 0 || k > 0);
                            continue;
                        }
                        for (int j = 0; j < n_acc; ++j) {
                            const uint32_t boff = (uint32_t)(j * b_shift_rows) * 128u;
                            const uint64_t bdesc = mn ? umma_desc_mnmajor_sw128(sb + boff + k * 2048, 8192, 1024)
                                                      : umma_desc_kmajor_sw128(sb + boff) + 2 * k;
                            umma_bf16(tm + j * acc_stride, adesc, bdesc, idesc, i > 0 || k > 0);
                        }
                    }
                }
                __syncwarp();
            }
            if (elect_one()) umma_commit(&ctl->done);
            __syncwarp();
            mbar_wait(&ctl->done, rep & 1);
            tc_fence_after();
            t1 = clock64();
        }
        if ((threadIdx.x & 31) == 0) out[blockIdx.x] = t1 - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    long long* d_out;
    cudaMalloc(&d_out, sms * sizeof(long long));
    const size_t smem = 4 * 49152 + sizeof(Ctl) + 1024;
    cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    struct V { const char* name; int N, mn, n_acc, stride, shift; };
    const V vs[] = {
        {"K-major  N=256 1 acc", 256, 0, 1, 0, 0},
        {"K-major  N=256 2 acc", 256, 0, 2, 256, 0},
        {"K-major  N=192 1 acc", 192, 0, 1, 0, 0},
        {"K-major  N=192 2 acc (fwd GEMM 1: tanh|sigmoid halves; bwd: alternate k16)", 192, 0, 2, 256, 0},
        {"K-major  N=256 | N=128 pair per k16 step (384 columns; cycles per MMA, x2 per step)", -1, 0, 2, 256, 0},
        {"K-major  N=128 3 acc", 128, 0, 3, 128, 0},
        {"K-major  N=96  4 acc", 96, 0, 4, 96, 0},
        {"K-major  N=128 2 acc", 128, 0, 2, 256, 0},
        {"K-major  N=128 4 acc", 128, 0, 4, 128, 0},
        {"K-major  N=64  5 acc", 64, 0, 5, 64, 0},
        {"MN-major N=256 2 acc (per-tap weight gradient with the ones block)", 256, 1, 2, 256, 0},
        {"MN-major N=192 2 acc (per-tap weight gradient)", 192, 1, 2, 256, 0},
        {"MN-major N=128 3 acc", 128, 1, 3, 128, 0},
        {"MN-major N=96  5 acc", 96, 1, 5, 96, 1},
        {"MN-major N=64  5 acc row-shifted B (all-taps weight gradient)", 64, 1, 5, 64, 1},
        {"MN-major N=64  5 acc same B", 64, 1, 5, 64, 0},
        {"MN-major N=64  1 acc", 64, 1, 1, 0, 0},
    };
    const int n_mma = 4000;
    for (const V& v : vs) {
        const int n = n_mma / (4 * v.n_acc) * (4 * v.n_acc);
        rate_kernel<<<sms, 128, smem>>>(v.N, v.mn, v.n_acc, v.stride, v.shift, n, d_out);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%s: %s\n", v.name, cudaGetErrorString(e)); return 1; }
        std::vector<long long> h(sms);
        cudaMemcpy(h.data(), d_out, sms * sizeof(long long), cudaMemcpyDeviceToHost);
        long long mx = 0, mnv = 1ll << 62;
        for (long long x : h) { mx = x > mx ? x : mx; mnv = x < mnv ? x : mnv; }
        const double floor_c = 128.0 * v.N / 256.0;
        printf("%-78s cycles/MMA min %.1f max %.1f  (tensor floor %.0f, smem A+B bytes/MMA %d)\n", v.name, (double)mnv / n,
               (double)mx / n, floor_c, 4096 + v.N * 32);
    }
    return 0;
}
