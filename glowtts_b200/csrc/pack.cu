// glowtts_b200/csrc/pack.cu -- varlen packing of the flow decoder's working set.
//
// The reference pads every utterance of a batch to the longest one and runs the whole FlowSpecDecoder on the padding
// (models.py:765-785 with x_mask; 41 % of the squeezed frames at BASELINE configs[1]).  Every flow is per-frame except
// the WaveNet convolutions, whose receptive field per layer is (K/2) dil frames; and everything at a masked frame is
// zero by construction (h = start(x0) * mask, x = (x + res) * mask, output * mask: attentions.py:151, modules.py:166,171).
// So inside the decoder the batch can live as ONE long sequence of P rows:
//
//     [ utt 0 : len_0 rows | G zero rows | utt 1 : len_1 rows | G zero rows | ... | dead rows up to P ]
//
// with G = (K/2) * dilation_rate^(L-1) (the largest halo of any layer): the zero gap rows are exactly what a
// convolution would read from the reference's masked tail, so the results on the valid rows are the reference's, and
// no utterance sees another.  Kernels run on (B = 1, T = P) and skip the tiles behind n_rows = sum(len_b + G).
// P = round_up(B * (T' + G), 128) is static (CUDA-graph friendly); lengths stay on the device.
//
//   pack_plan          lengths from the mask -> off[B+1], row_b[P], row_fr[P], mask_p[P], x_len[B], hdr = {0, n_rows}
//   squeeze_pack       commons.squeeze (commons.py:339-351) fused with the gather into packed rows
//   unsqueeze_unpack   commons.unsqueeze (commons.py:354-364) fused with the scatter back to [B, C, T] (padding zeroed)
//   pack_track         a [B, 1, T] pitch / energy track -> per packed row its n_sqz samples (WNP conditioner input)
#include "common.cuh"
#include "ops.h"

namespace gtk {
namespace flow {

// One block.  mask [B, T] (0/1 prefix mask); squeezed length of b = #{tq : mask[b, n tq + n - 1] != 0} (commons.py:345).
__global__ void __launch_bounds__(1024) pack_plan_kernel(const float* __restrict__ mask, int B, int T, int n, int Tq, int G, int P,
                                                         int* __restrict__ hdr, int* __restrict__ off, int* __restrict__ row_b,
                                                         int* __restrict__ row_fr, float* __restrict__ mask_p,
                                                         float* __restrict__ x_len) {
    extern __shared__ int sm_off[];                         // [B + 1]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    for (int b = warp; b < B; b += nw) {
        int cnt = 0;
        for (int tq = lane; tq < Tq; tq += 32) cnt += (mask ? mask[(size_t)b * T + (size_t)n * tq + n - 1] : 1.f) != 0.f;
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        if (lane == 0) sm_off[b + 1] = cnt;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        sm_off[0] = 0;
        for (int b = 0; b < B; ++b) {
            const int len = sm_off[b + 1];
            x_len[b] = (float)len;
            sm_off[b + 1] = sm_off[b] + len + G;
        }
        hdr[0] = 0;
        hdr[1] = sm_off[B];                                 // n_rows
    }
    __syncthreads();
    for (int b = threadIdx.x; b <= B; b += blockDim.x) off[b] = sm_off[b];
    const int n_rows = sm_off[B];
    for (int r = threadIdx.x; r < P; r += blockDim.x) {
        int b = 0, fr = 0;
        float m = 0.f;
        if (r < n_rows) {
            int lo = 0, hi = B;                             // largest b with off[b] <= r
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (sm_off[mid] <= r) lo = mid; else hi = mid;
            }
            b = lo;
            const int t = r - sm_off[b], len = sm_off[b + 1] - sm_off[b] - G;
            if (t < len) { m = 1.f; fr = b * Tq + t; }
        }
        row_b[r] = b;
        row_fr[r] = fr;
        mask_p[r] = m;
    }
}

int pack_plan(const float* mask, int B, int T, int n, int Tq, int G, int P, int* hdr, int* off, int* row_b, int* row_fr,
              float* mask_p, float* x_len, cudaStream_t st) {
    GTK_REQUIRE(hdr && off && row_b && row_fr && mask_p && x_len && B >= 1 && n >= 1 && G >= 0, "bad args");
    GTK_REQUIRE((size_t)(B + 1) * sizeof(int) <= 48 * 1024, "batch too large for the packing plan");
    GTK_REQUIRE((long long)B * (Tq + G) <= (long long)P, "packed row capacity");
    pack_plan_kernel<<<1, 1024, (size_t)(B + 1) * sizeof(int), st>>>(mask, B, T, n, Tq, G, P, hdr, off, row_b, row_fr, mask_p, x_len);
    return check_launch("pack_plan");
}

// xp[s * C + c][r] = x[b][c][n t + s] for valid rows, 0 elsewhere (all P rows are written).  x [B, C, T_in] with row
// pitch T_in; grid (ceil(P / 256), C).
__global__ void __launch_bounds__(256) squeeze_pack_kernel(const float* __restrict__ x, const int* __restrict__ off,
                                                           const int* __restrict__ row_b, const float* __restrict__ mask_p,
                                                           float* __restrict__ xp, int C, int T_in, int n, int P) {
    const int c = blockIdx.y;
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= P) return;
    const bool ok = mask_p[r] != 0.f;
    const int b = row_b[r];
    const int t = r - off[b];
    const float* src = x + ((size_t)b * C + c) * T_in + (size_t)n * t;
    if (n == 2 && (T_in & 1) == 0 && (reinterpret_cast<uintptr_t>(x) & 7) == 0) {
        float2 v = make_float2(0.f, 0.f);
        if (ok) v = *reinterpret_cast<const float2*>(src);
        xp[(size_t)c * P + r] = v.x;
        xp[(size_t)(C + c) * P + r] = v.y;
    } else {
        for (int s = 0; s < n; ++s) xp[(size_t)(s * C + c) * P + r] = ok ? src[s] : 0.f;
    }
}

int squeeze_pack(const float* x, const int* off, const int* row_b, const float* mask_p, float* xp, int C, int T_in, int n, int P,
                 cudaStream_t st) {
    GTK_REQUIRE(x && off && row_b && mask_p && xp && C >= 1 && C <= 65535, "bad args");
    squeeze_pack_kernel<<<dim3(ceil_div(P, 256), C), 256, 0, st>>>(x, off, row_b, mask_p, xp, C, T_in, n, P);
    return check_launch("squeeze_pack");
}

// out[b][c][n tq + s] = xp[s * C + c][off[b] + tq] for tq < len_b, 0 elsewhere (every element of out [B, C, T_out] is
// written, the frames squeeze dropped included).  grid (ceil(ceil(T_out / n) / 256), C, B).
__global__ void __launch_bounds__(256) unsqueeze_unpack_kernel(const float* __restrict__ xp, const int* __restrict__ off,
                                                               float* __restrict__ out, int C, int T_out, int n, int G, int P) {
    const int c = blockIdx.y, b = blockIdx.z;
    const int tq = blockIdx.x * blockDim.x + threadIdx.x;
    if ((long long)tq * n >= T_out) return;
    const int o = off[b], len = off[b + 1] - o - G;
    float* dst = out + ((size_t)b * C + c) * T_out + (size_t)n * tq;
    const bool ok = tq < len;
    if (n == 2 && (T_out & 1) == 0 && (reinterpret_cast<uintptr_t>(out) & 7) == 0) {
        float2 v = make_float2(0.f, 0.f);
        if (ok) { v.x = xp[(size_t)c * P + o + tq]; v.y = xp[(size_t)(C + c) * P + o + tq]; }
        *reinterpret_cast<float2*>(dst) = v;
    } else {
        for (int s = 0; s < n && n * tq + s < T_out; ++s) dst[s] = ok ? xp[(size_t)(s * C + c) * P + o + tq] : 0.f;
    }
}

int unsqueeze_unpack(const float* xp, const int* off, float* out, int B, int C, int T_out, int n, int G, int P, cudaStream_t st) {
    GTK_REQUIRE(xp && off && out && C >= 1 && C <= 65535 && B <= 65535, "bad args");
    if ((long long)B * C * T_out == 0) return GTK_OK;
    const int tqs = ceil_div(T_out, n);
    unsqueeze_unpack_kernel<<<dim3(ceil_div(tqs, 256), C, B), 256, 0, st>>>(xp, off, out, C, T_out, n, G, P);
    return check_launch("unsqueeze_unpack");
}

// trk[r * n + s] = track[b][n t + s] (valid rows) or 0: the n_sqz samples of a [B, T] pitch / energy track that condition
// packed row r (modules.py:320-322, 353-362: cond_layer1 is 1x1 on the un-squeezed track, then squeezed).
__global__ void __launch_bounds__(256) pack_track_kernel(const float* __restrict__ track, const int* __restrict__ off,
                                                         const int* __restrict__ row_b, const float* __restrict__ mask_p,
                                                         float* __restrict__ trk, int T_in, int n, int P) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= P) return;
    const bool ok = mask_p[r] != 0.f;
    const int b = row_b[r], t = r - off[b];
    for (int s = 0; s < n; ++s) trk[(size_t)r * n + s] = ok ? track[(size_t)b * T_in + (size_t)n * t + s] : 0.f;
}
int pack_track(const float* track, const int* off, const int* row_b, const float* mask_p, float* trk, int T_in, int n, int P,
               cudaStream_t st) {
    GTK_REQUIRE(track && off && row_b && mask_p && trk, "bad args");
    pack_track_kernel<<<ceil_div(P, 256), 256, 0, st>>>(track, off, row_b, mask_p, trk, T_in, n, P);
    return check_launch("pack_track");
}

// d track[b][n t + s] = dtrk[r * n + s] (valid rows), 0 in the padding: every element of dtrack [B, T_out] is written.
__global__ void __launch_bounds__(256) unpack_track_kernel(const float* __restrict__ dtrk, const int* __restrict__ off,
                                                           float* __restrict__ dtrack, int T_out, int n, int G) {
    const int b = blockIdx.y;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T_out) return;
    const int o = off[b], len = off[b + 1] - o - G;
    const int tq = t / n, s = t - tq * n;
    dtrack[(size_t)b * T_out + t] = tq < len ? dtrk[(size_t)(o + tq) * n + s] : 0.f;
}
int unpack_track(const float* dtrk, const int* off, float* dtrack, int B, int T_out, int n, int G, cudaStream_t st) {
    GTK_REQUIRE(dtrk && off && dtrack && B <= 65535, "bad args");
    if ((long long)B * T_out == 0) return GTK_OK;
    unpack_track_kernel<<<dim3(ceil_div(T_out, 256), B), 256, 0, st>>>(dtrk, off, dtrack, T_out, n, G);
    return check_launch("unpack_track");
}

}  // namespace flow
}  // namespace gtk
