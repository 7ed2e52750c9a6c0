// glowtts_b200/csrc/ops.h -- internal declarations of the launchers (one per kernel family).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace gtk {
namespace mas {
size_t workspace_bytes(int B, int T_x, int T_y);
int forward(const float* value, const int* t_xs, const int* t_ys, float* path, int* dur_out, int* idx_out, int B, int T_x,
            int T_y, void* workspace, size_t workspace_bytes_given, int force_generic, cudaStream_t stream);
size_t static_scratch_bytes();
int forward_static(const float* value, const int* t_xs, const int* t_ys, float* path, int* dur_out, int* idx_out, int B, int T_x,
                   int T_y, cudaStream_t stream);
int fused_supported(int D, int T_x, int T_y, int has_logs);
int fused_forward(const float* z, const float* x_m, const int* t_xs, const int* t_ys, float* path, int* dur_out, int* idx_out,
                  float* logp_out, int B, int D, int T_x, int T_y, cudaStream_t stream);
int lengths_from_mask(const float* mask, int* t_xs, int* t_ys, int B, int T_x, int T_y, cudaStream_t stream);
int trace_read(unsigned long long* out, int n_ctas);
int wait_read(long long* out, int n_ctas);
}  // namespace mas

namespace logp {
int forward(const float* x_m, const float* x_logs, const float* z, float* out, int B, int D, int T_x, int T_y, cudaStream_t st);
int generate_path(const float* dur, const float* mask, float* path, int B, int T_x, int T_y, cudaStream_t st);
}  // namespace logp

namespace flow {
int squeeze(const float* x, const float* mask, float* out, float* mask_out, int B, int C, int T, int n, int mask_is_squeezed,
            cudaStream_t st);
int unsqueeze(const float* x, const float* mask, float* out, float* mask_out, int B, int C, int T, int n, int out_T,
              cudaStream_t st);
int mask_len(const float* mask, float* x_len, int B, int T, cudaStream_t st);
// varlen packing of the decoder's working set (pack.cu)
int pack_plan(const float* mask, int B, int T, int n, int Tq, int G, int P, int* hdr, int* off, int* row_b, int* row_fr,
              float* mask_p, float* x_len, cudaStream_t st);
int squeeze_pack(const float* x, const int* off, const int* row_b, const float* mask_p, float* xp, int C, int T_in, int n, int P,
                 cudaStream_t st);
int unsqueeze_unpack(const float* xp, const int* off, float* out, int B, int C, int T_out, int n, int G, int P, cudaStream_t st);
int pack_track(const float* track, const int* off, const int* row_b, const float* mask_p, float* trk, int T_in, int n, int P,
               cudaStream_t st);
int unpack_track(const float* dtrk, const int* off, float* dtrack, int B, int T_out, int n, int G, cudaStream_t st);
int invconv_prep(const float* w, int ns, float* winv, float* logdet, cudaStream_t st);
// per-call constants of up to FLOW_CONST_MAX_BLOCKS flow blocks in one launch (flow_pointwise.cu)
constexpr int FLOW_CONST_MAX_BLOCKS = 32;
struct FlowConstJobs {
    const float* w[FLOW_CONST_MAX_BLOCKS];       // InvConvNear weight [ns][ns] of block k
    const float* logs[FLOW_CONST_MAX_BLOCKS];    // ActNorm logs [C] of block k (or null)
    float* winv;                                 // [n][64] inverses, or null
    float* icld;                                 // [n] log|det W_k|, or null
    int n;
};
struct FlowConstGrads {
    float* dlogs[FLOW_CONST_MAX_BLOCKS];
    float* dW[FLOW_CONST_MAX_BLOCKS];
    int n;
};
int flow_consts(const FlowConstJobs& J, int ns, int C, const float* x_len, float* logdet, int B, int accumulate, cudaStream_t st);
int logdet_const_bwd_all(const float* dlogdet, const float* x_len, int B, const FlowConstGrads& G, int C, const float* winv, int ns,
                         cudaStream_t st);
// n_rows: null, or (varlen packing, B == 1) device pointer to the number of rows in use: rows behind it are skipped
int mix(const float* x, const float* mask, const float* logs, const float* bias, const float* M, int ns, float* z, int B,
        int C, int T, int reverse, cudaStream_t st, const int* n_rows = nullptr);
int mix_bwd(const float* x, const float* mask, const float* logs, const float* bias, const float* M, int ns, const float* dz,
            float* dx, float* dlogs, float* dbias, float* dM, int B, int C, int T, cudaStream_t st, const int* n_rows = nullptr);
int logdet_const(const float* logs, int C, const float* logdetW, float group_count, const float* x_len, float* logdet, int B,
                 int accumulate, cudaStream_t st);
int logdet_const_bwd(const float* dlogdet, const float* x_len, int B, float* dlogs, int C, const float* winv, int ns,
                     float group_count, float* dW, cudaStream_t st);
int actnorm_ddi_stats(const float* x, const float* mask, float* stats, int B, int C, int T, cudaStream_t st);
int actnorm_ddi_finalize(const float* stats, float* logs, float* bias, int C, cudaStream_t st);
int gate_fwd(const float* a, const float* cond, int cond_mode, size_t cond_bstride, float* out, int B, int H, int T,
             float drop_p, DropSeed seed, cudaStream_t st);
int gate_bwd(const float* a, const float* cond, int cond_mode, size_t cond_bstride, const float* dout, float* da, int B, int H,
             int T, float drop_p, DropSeed seed, int scale_out, cudaStream_t st);
int dropout_scale(unsigned long long seed, float p, long long n, float* out, cudaStream_t st);
int row_sum(const float* in, float* out, int R, int T, int accumulate, int rows_per_group, size_t out_group_stride,
            cudaStream_t st);
int channel_sum(const float* in, const float* mask, float* out, int B, int M, int T, int accumulate, cudaStream_t st);
int affine(const float* x, const float* end_out, const float* mask, float* z, float* logdet, int B, int C, int T, int reverse,
           int sigmoid_scale, cudaStream_t st);
int affine_bwd(const float* x, const float* end_out, const float* mask, const float* dz, const float* dlogdet, float* dend,
               float* dx, int B, int C, int T, int sigmoid_scale, cudaStream_t st);
int resskip(const float* rs, const float* x, const float* mask, float* x_out, float* skip, int B, int H, int RS, int T,
            int first, int mask_skip, cudaStream_t st);
int cat_mask(const float* a, const float* b, const float* mask, float* out, int B, int Ca, int Cb, int T, int mask_b,
             cudaStream_t st);
int weightnorm_fwd(const float* g, const float* v, float* w, float* norm_out, int M, int R, cudaStream_t st);
int weightnorm_bwd(const float* g, const float* v, const float* dw, float* dg, float* dv, int M, int R, int accumulate,
                   cudaStream_t st);
int clip_grad_value(float* grad, long long n, float clip, int do_clip, float norm_type, float* acc, cudaStream_t st);
int mle_loss_fwd(const float* z, const float* m, const float* logs, const float* logdet, const float* mask, float* acc4,
                 float* loss, int B, int C, int T, cudaStream_t st);
int mle_loss_bwd(const float* z, const float* m, const float* logs, const float* acc4, const float* gloss, float* dz,
                 float* dm, float* dlogs, float* dlogdet, int B, int C, int T, cudaStream_t st);
int gather_tokens(const float* src, const int* idx, float* out, int B, int C, int T_x, int T_y, cudaStream_t st);
// WN.cond_layer(g) of every block in one launch (cond.cu)
constexpr int COND_MAX_BLOCKS = 32;
struct CondJobs { const float* g[COND_MAX_BLOCKS]; const float* v[COND_MAX_BLOCKS]; const float* b[COND_MAX_BLOCKS]; int n; };
struct CondGrads { float* g[COND_MAX_BLOCKS]; float* v[COND_MAX_BLOCKS]; float* b[COND_MAX_BLOCKS]; int n; };
int cond_all(const CondJobs& J, const float* g, float* cond, int B, int G, int N, cudaStream_t st);
int cond_all_bwd(const CondJobs& J, const CondGrads& D, const float* g, const float* dcond, float* dg, int B, int G, int N,
                 cudaStream_t st);
}  // namespace flow

namespace conv {
struct ConvParams {
    const float* in; const float* wp; const float* bias; const float* mask; const float* resid; float* out;
    int B, Cin, M, T, K, dil;
    int mask_out, mask_resid, mask_in;
    size_t in_bstride, out_bstride, resid_bstride;   // elements between batches; 0 = contiguous
};
struct WgradParams {
    const float* dout; const float* in; const float* mask; float* dw;
    int B, Cin, M, T, dil;
    int mask_dout, mask_in;
    size_t dout_bstride, in_bstride;                 // 0 = contiguous
};
int pack_weight(const float* w, float* wp, int M, int Cin, int K, int transpose, cudaStream_t st);
int conv1d_f32(const ConvParams& p, cudaStream_t st);
int wgrad_f32(const WgradParams& p, int K, cudaStream_t st);
}  // namespace conv
}  // namespace gtk
