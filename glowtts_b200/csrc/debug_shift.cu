// glowtts_b200/csrc/debug_shift.cu -- hardware probe: may a K-major SWIZZLE_128B UMMA operand descriptor start at a row
// that is not a multiple of 8 (i.e. not on a 1024-byte swizzle-atom boundary)?  If so, one TMA box with a halo serves
// every tap of a dilated convolution (rows shifted by the tap offset), and layer-to-layer fusion in shared memory becomes
// possible.  Computes D_s = A[s : s + 128, :] * B^T for s = 1..4 with the descriptor's "base offset" field (bits 49-51)
// either 0 or (start address >> 7) & 7.
#ifdef GTK_INSTRUMENT
#include <string.h>

#include "tc.cuh"
#include "ops.h"

namespace gtk {
namespace tc {

struct ShiftCtl {
    uint64_t full, done;
    uint32_t tmem_base;
};

__global__ void __launch_bounds__(128, 1) shift_probe_kernel(const __grid_constant__ CUtensorMap tmA,
                                                              const __grid_constant__ CUtensorMap tmB, float* out) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    unsigned char* sa = smem;                    // [136 rows][128 B] = 17 x 1024
    unsigned char* sb = smem + 17 * 1024 + 1024; // [64 rows][128 B], 1024-aligned
    ShiftCtl* ctl = reinterpret_cast<ShiftCtl*>(sb + 8 * 1024);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { mbar_init(&ctl->full, 1); mbar_init(&ctl->done, 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc<512>(&ctl->tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctl->tmem_base;
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&ctl->full, 136 * 128 + 64 * 128);
        tma_load_2d(sa, &tmA, &ctl->full, 0, 0);
        tma_load_2d(sb, &tmB, &ctl->full, 0, 0);
        mbar_wait(&ctl->full, 0);
        tc_fence_after();
        const uint32_t idesc = umma_idesc_bf16(128, 64);
        const uint64_t bdesc = umma_desc_kmajor_sw128(smem_u32(sb));
        for (int s = 1; s <= 4; ++s)
            for (int mode = 0; mode < 2; ++mode) {
                const uint32_t addr = smem_u32(sa) + s * 128;
                uint64_t adesc = umma_desc_kmajor_sw128(addr);
                if (mode) adesc |= (uint64_t)((addr >> 7) & 7) << 49;
                const uint32_t acc = tmem + ((s - 1) * 2 + mode) * 64;
                for (int k = 0; k < 4; ++k) umma_bf16(acc, adesc + 2 * k, bdesc + 2 * k, idesc, k != 0);
            }
        umma_commit(&ctl->done);
    }
    __syncwarp();
    mbar_wait(&ctl->done, 0);
    tc_fence_after();
    const int r = warp * 32 + lane;
    for (int v = 0; v < 8; ++v)
        for (int c0 = 0; c0 < 64; c0 += 16) {
            float x[16];
            tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + v * 64 + c0, x);
            for (int i = 0; i < 16; ++i) out[((size_t)v * 128 + r) * 64 + c0 + i] = x[i];
        }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

// a: [136][64] bf16 row-major, b: [64][64] bf16 row-major (N x K), out: [8][128][64] fp32 (index (s-1)*2 + mode)
int shift_probe(const __nv_bfloat16* a, const __nv_bfloat16* b, float* out, cudaStream_t st) {
    GTK_REQUIRE(a && b && out, "null");
    CUtensorMap tmA, tmB;
    {
        const uint64_t dims[2] = {64, 136};
        const uint64_t strides[1] = {128};
        const uint32_t box[2] = {64, 136};
        int rc = make_tmap_bf16(&tmA, a, 2, dims, strides, box, true);
        if (rc) return rc;
    }
    {
        const uint64_t dims[2] = {64, 64};
        const uint64_t strides[1] = {128};
        const uint32_t box[2] = {64, 64};
        int rc = make_tmap_bf16(&tmB, b, 2, dims, strides, box, true);
        if (rc) return rc;
    }
    const size_t smem = 17 * 1024 + 1024 + 8 * 1024 + sizeof(ShiftCtl) + 1024;
    GTK_CUDA(cudaFuncSetAttribute(shift_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    shift_probe_kernel<<<1, 128, smem, st>>>(tmA, tmB, out);
    return check_launch("shift_probe");
}

}  // namespace tc
}  // namespace gtk
#endif  // GTK_INSTRUMENT
