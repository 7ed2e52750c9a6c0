// Codegen experiment (round 2): how many SASS instructions does the MMA-issuing thread spend per tcgen05.mma?
//   variant A: `if (lane == 0) { loop { mma } }`         -- what the round-1 kernels did
//   variant B: whole warp runs the loop, `if (elect_one()) mma`  -- descriptors stay in uniform registers
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -cubin -I ../../glowtts_b200/csrc mma_issue_codegen.cu -o /tmp/x.cubin
//        cuobjdump -sass /tmp/x.cubin | grep -c ...    (see profiles/mma_issue_r02.txt)
#include "tc.cuh"
using namespace gtk;
using namespace gtk::tc;

struct Ctl { uint64_t full[4], empty[4], done; uint32_t tmem_base; };

template <int VARIANT>
__global__ void __launch_bounds__(192, 1) issue_kernel(const int* n_ptr, int taps, int dil) {
    int n_steps = *n_ptr;
    if (VARIANT == 2) n_steps = __shfl_sync(0xffffffffu, n_steps, 0);
    extern __shared__ __align__(1024) unsigned char smem[];
    Ctl* ctl = reinterpret_cast<Ctl*>(smem + 4 * 49152);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < 4; ++s) { mbar_init(&ctl->full[s], 1); mbar_init(&ctl->empty[s], 1); }
        mbar_fence_init();
    }
    if (warp == 2) tmem_alloc<512>(&ctl->tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctl->tmem_base;
    if (warp == 1) {
        if (VARIANT == 0) {
            if (lane == 0) {
                const uint32_t idesc = umma_idesc_bf16(128, 64, 1, 1);
                for (int it = 0; it < n_steps; ++it) {
                    const int s = it % 4, use = it / 4;
                    mbar_wait(&ctl->full[s], use & 1);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + (size_t)s * 49152), sb = sa + 16384;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint64_t adesc = umma_desc_mnmajor_sw128(sa + k * 2048, 8192, 1024);
                        for (int j = 0; j < taps; ++j) {
                            const uint64_t bdesc = umma_desc_mnmajor_sw128(sb + (uint32_t)(j * dil) * 128u + k * 2048, 8192, 1024);
                            umma_bf16(tmem + j * 64, adesc, bdesc, idesc, it > 0 || k > 0);
                        }
                    }
                    umma_commit(&ctl->empty[s]);
                }
                umma_commit(&ctl->done);
            }
        } else {
            const uint32_t idesc = umma_idesc_bf16(128, 64, 1, 1);
            const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
            for (int it = 0; it < n_steps; ++it) {
                const int s = it % 4, use = it / 4;
                mbar_wait(&ctl->full[s], use & 1);
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + (size_t)s * 49152), sb = sa + 16384;
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint64_t adesc = umma_desc_mnmajor_sw128(sa + k * 2048, 8192, 1024);
                        for (int j = 0; j < taps; ++j) {
                            const uint64_t bdesc = umma_desc_mnmajor_sw128(sb + (uint32_t)(j * dil) * 128u + k * 2048, 8192, 1024);
                            umma_bf16(tm + j * 64, adesc, bdesc, idesc, it > 0 || k > 0);
                        }
                    }
                    umma_commit(&ctl->empty[s]);
                }
                __syncwarp();
            }
            if (elect_one()) umma_commit(&ctl->done);
        }
    }
    __syncthreads();
    if (warp == 2) tmem_dealloc<512>(tmem);
}
template __global__ void issue_kernel<0>(const int*, int, int);
template __global__ void issue_kernel<1>(const int*, int, int);
template __global__ void issue_kernel<2>(const int*, int, int);
