/*
 * glowtts_b200.h -- C ABI of the B200-native Glow-TTS hot path (libglowtts_b200.so).
 *
 * Every entry point is stream-ordered on the cudaStream_t passed as `void* stream`, never
 * synchronises the host, never allocates (workspaces are passed in) and never throws: it
 * returns 0 on success or a negative GTK_ERR_* code, with a message in gtk_last_error().
 * All tensors are device pointers, fp32, C-contiguous, laid out as the reference lays them
 * out ([B, C, T] with t fastest; [B, T_x, T_y] with y fastest) unless stated otherwise.
 * Masks are the reference's 0/1 float masks; a [B,1,T] mask is passed as [B,T].
 *
 * Each function names the reference interface (file:line under /root/reference) it replaces.
 * The Python host side (glowtts_b200/) binds these with ctypes and registers them as
 * torch.library ops; INTEGRATION.md shows the binding a reference maintainer would add.
 */
#ifndef GLOWTTS_B200_H
#define GLOWTTS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GTK_OK 0
#define GTK_ERR_INVALID (-1)
#define GTK_ERR_LAUNCH (-2)
#define GTK_ERR_WORKSPACE (-3)
#define GTK_ERR_UNSUPPORTED (-4)

#define GTK_MAX_LAYERS 8

/* ---- library ---------------------------------------------------------------------------- */
int gtk_abi_version(void);
const char* gtk_last_error(void); /* thread-local, valid until the next failing call */
/* Kernels launched by this library in this process so far (bench.py's gpu_launches). */
unsigned long long gtk_launch_count(void);
/* Per-kernel device timing for the roofline report: while enabled, the hot kernels' launches are
 * bracketed by CUDA events on their stream, grouped under a tag.  gtk_prof_read synchronises on the
 * recorded events of tag i and returns the number of launches and their summed duration. */
int gtk_prof_enable(int on);
int gtk_prof_reset(void);
int gtk_prof_tag_count(void);
int gtk_prof_read(int i, char* name, int name_cap, int* launches, float* total_ms);

/* ---- monotonic alignment search --------------------------------------------------------- */
/* Replaces monotonic_align/__init__.py:18-19 (t_x = mask.sum(1)[:,0], t_y = mask.sum(2)[:,0])
 * without the device->host copy.  mask [B,T_x,T_y]; t_xs, t_ys int32 [B]. */
int gtk_mas_lengths_from_mask(const float* mask, int32_t* t_xs, int32_t* t_ys, int B, int T_x, int T_y, void* stream);

/* Bytes of workspace gtk_mas_forward needs for this shape (0 when the decision bits fit in
 * shared memory, which covers T_x <= 512 at the reference's data sizes). */
size_t gtk_mas_workspace_bytes(int B, int T_x, int T_y);

/* Replaces maximum_path_c / maximum_path_each (monotonic_align/core.pyx:9-45) and the host
 * round trip of monotonic_align/__init__.py:11-21.  value [B,T_x,T_y] is NOT modified.
 * path [B,T_x,T_y] is fully written (0/1); path may be NULL when only the alignment's consumers are wanted (durations
 * and the per-frame index; the training step gathers the prior with the index instead of the attn^T matmul,
 * models.py:1118-1119), which halves the kernel's HBM traffic to the 4 B/cell read.  dur_out [B,T_x] (frames per token = attn.sum(2),
 * models.py:1085) and idx_out [B,T_y] (token per frame, -1 beyond t_y) are optional (NULL).
 * Bit-exact with core.pyx for the same fp32 value; utterances with t_x > t_y (undefined
 * behaviour in the reference) or empty lengths produce an all-zero path. */
int gtk_mas_forward(const float* value, const int32_t* t_xs, const int32_t* t_ys, float* path, int32_t* dur_out,
                    int32_t* idx_out, int B, int T_x, int T_y, void* workspace, size_t workspace_bytes,
                    int force_generic, void* stream);

/* Allocation-free variant of gtk_mas_forward.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  Same inputs, outputs and results as gtk_mas_forward(force_generic = 1), but the
 * decision-bit scratch the generic kernel needs is a file-scope device array inside the library (csrc/mas.cu) instead
 * of a caller workspace; the call performs no cudaMalloc / cudaFree and calls no allocating library routine.
 * Limit: B * T_y * ceil(T_x / 32) * 4 <= gtk_mas_static_scratch_bytes() (8 MiB, held by every GPU that loads the
 * library until exit); larger shapes return an error (gtk_last_error() says so) -- use gtk_mas_forward.  All calls share
 * the one array (it keeps the last call's contents; nothing in it needs to start at zero): calls that overlap on the
 * GPU, e.g. on different streams, race. */
size_t gtk_mas_static_scratch_bytes(void);
int gtk_mas_forward_static(const float* value, const int32_t* t_xs, const int32_t* t_ys, float* path, int32_t* dur_out,
                           int32_t* idx_out, int B, int T_x, int T_y, void* stream);

/* Fused variant: models.py:1076-1083 (logp construction + maximum_path) in ONE kernel -- the [B,T_x,T_y] value tensor
 * is computed tile by tile on the tensor cores inside the MAS CTA (3xTF32 mma.sync) and never exists in HBM.
 * Covers mean_only models (x_logs == NULL; all configs/base*.json) with D == 80, T_x <= 256, T_y % 4 == 0;
 * gtk_mas_fused_supported() tells, anything else returns GTK_ERR_UNSUPPORTED (use gtk_logp + gtk_mas_forward).
 * z [B,D,T_y], x_m [B,D,T_x]; path/dur_out/idx_out as gtk_mas_forward; logp_out [B,T_x,T_y] is optional (NULL):
 * when given, the cells inside [0,t_x) x [0,t_y) receive logp (the rest is left untouched).  Not bit-identical to
 * logp -> MAS (fp32 summation order differs): reported as path agreement + max |delta logp|. */
int gtk_mas_fused_supported(int D, int T_x, int T_y, int has_x_logs);
int gtk_mas_fused_logp(const float* z, const float* x_m, const float* x_logs, const int32_t* t_xs, const int32_t* t_ys,
                       float* path, int32_t* dur_out, int32_t* idx_out, float* logp_out, int B, int D, int T_x, int T_y,
                       void* stream);

/* ---- logp (models.py:1076-1082) ---------------------------------------------------------- */
/* logp[b,x,y] = log N(z[b,:,y]; x_m[b,:,x], exp(x_logs[b,:,x])).  x_logs may be NULL (mean_only:
 * x_logs == 0, models.py:712-713).  x_m,x_logs [B,D,T_x]; z [B,D,T_y]; logp [B,T_x,T_y]. */
int gtk_logp(const float* x_m, const float* x_logs, const float* z, float* logp, int B, int D, int T_x, int T_y,
             void* stream);

/* ---- glue (commons.py) ------------------------------------------------------------------- */
/* commons.squeeze (commons.py:339-351): x [B,C,T] -> out [B,C*n,T/n] (masked), mask_out [B,T/n]. mask/mask_out may be NULL. */
int gtk_squeeze(const float* x, const float* mask, float* out, float* mask_out, int B, int C, int T, int n_sqz, void* stream);
/* commons.unsqueeze (commons.py:354-364): x [B,C,T] -> out [B,C/n,T*n] (masked), mask_out [B,T*n]. */
int gtk_unsqueeze(const float* x, const float* mask, float* out, float* mask_out, int B, int C, int T, int n_sqz, void* stream);
/* commons.fused_add_tanh_sigmoid_multiply (commons.py:61-68): a [B,2H,T]; cond_mode 0: no b,
 * 1: b is [B,2H] (broadcast over t, the cond_layer(g) case), 2: b is [B,2H,T]. out [B,H,T]. */
int gtk_gate_fwd(const float* a, const float* b, int cond_mode, float* out, int B, int H, int T, void* stream);
/* d(a) [B,2H,T] from d(out); d(b) is d(a) (mode 2) or its sum over t (mode 1, use gtk_row_sum). */
int gtk_gate_bwd(const float* a, const float* b, int cond_mode, const float* dout, float* da, int B, int H, int T, void* stream);
/* Dropout mask of the decoder kernels (modules.py:120,153): out[i] = keep-scale (0 or 1/(1-p)) of element i of a
 * [B, T', 2H] frame-major in_layer output for layer seed `seed`; gtk_decoder_layer_seed maps the per-call seed of
 * gtk_decoder_desc to the seed of (block, layer).  For tests and for the module-level (unfused) WN. */
int gtk_dropout_scale(unsigned long long seed, float p, long long n, float* out, void* stream);
unsigned long long gtk_decoder_layer_seed(unsigned long long seed, int block, int layer);
int gtk_row_sum(const float* in, float* out, int R, int T, int accumulate, void* stream);
int gtk_channel_sum(const float* in, float* out, int B, int M, int T, int accumulate, void* stream);
/* commons.clip_grad_value_ (commons.py:320-336) over one contiguous range of n gradients: *acc += sum |g|^norm_type and,
 * when do_clip, g = clamp(g, -clip_value, clip_value) in place.  Stream-ordered, no host sync: the caller zeroes acc,
 * calls once per contiguous range (once for a flat gradient buffer) and takes acc^(1/norm_type) on the device. */
int gtk_clip_grad_value(float* grad, long long n, float clip_value, int do_clip, float norm_type, float* acc, void* stream);
/* commons.mle_loss (commons.py:28-33).  m/logs may be NULL (= 0).  acc4: 4 floats of scratch that
 * must be kept for the backward.  loss: 1 float. */
int gtk_mle_loss_fwd(const float* z, const float* m, const float* logs, const float* logdet, const float* mask, float* acc4,
                     float* loss, int B, int C, int T, void* stream);
int gtk_mle_loss_bwd(const float* z, const float* m, const float* logs, const float* acc4, const float* gloss, float* dz,
                     float* dm, float* dlogs, float* dlogdet, int B, int C, int T, void* stream);
/* Prior expansion models.py:1118-1119 as the gather it is: out[b,c,y] = src[b,c,idx[b,y]] (0 if idx<0). */
int gtk_gather_tokens(const float* src, const int32_t* idx, float* out, int B, int C, int T_x, int T_y, void* stream);
/* commons.generate_path (commons.py:127-143): durations [B,T_x] (float) -> path [B,T_x,T_y] * mask. */
int gtk_generate_path(const float* duration, const float* mask, float* path, int B, int T_x, int T_y, void* stream);

/* ---- ActNorm / InvConvNear (modules.py:575-668) ------------------------------------------ */
/* x_len[b] = sum_t mask[b,t]  (modules.py:587,642) */
int gtk_mask_len(const float* mask, float* x_len, int B, int T, void* stream);
/* W^-1 (modules.py:651,668) and logdet(W) (modules.py:658) of the n_split x n_split weight, on device. */
int gtk_invconv_prep(const float* w, int n_split, float* winv, float* logdet_w, void* stream);
/* Fused ActNorm+InvConvNear.  forward: z = mask * M (bias + exp(logs) x); reverse: z = mask * (M x - bias) exp(-logs).
 * logs/bias NULL => InvConvNear only (modules.py:635-665; pass M = W or W^-1); M NULL => ActNorm only
 * (modules.py:584-599; n_split = 1 allows any channel count). */
int gtk_flow_mix(const float* x, const float* mask, const float* logs, const float* bias, const float* M, int n_split,
                 float* z, int B, int C, int T, int reverse, void* stream);
/* Backward of the forward mix; dlogs/dbias [C] and dM [n_split^2] are accumulated into (may be NULL). */
int gtk_flow_mix_bwd(const float* x, const float* mask, const float* logs, const float* bias, const float* M, int n_split,
                     const float* dz, float* dx, float* dlogs, float* dbias, float* dM, int B, int C, int T, void* stream);
/* logdet[b] (+)= (sum_c logs_c + logdet_w * group_count) * x_len[b]; logs / logdet_w may be NULL. */
int gtk_flow_logdet_const(const float* logs, int C, const float* logdet_w, float group_count, const float* x_len,
                          float* logdet, int B, int accumulate, void* stream);
int gtk_flow_logdet_const_bwd(const float* dlogdet, const float* x_len, int B, float* dlogs, int C, const float* winv,
                              int n_split, float group_count, float* dW, void* stream);
/* ActNorm.initialize (modules.py:607-619): stats [3,C] = (sum mask, sum x, sum x^2) then finalize.
 * The split lets a multi-GPU caller all-reduce `stats` in between. */
int gtk_actnorm_ddi_stats(const float* x, const float* mask, float* stats, int B, int C, int T, void* stream);
int gtk_actnorm_ddi_finalize(const float* stats, float* logs, float* bias, int C, void* stream);

/* ---- coupling block pieces (attentions.py:132-186, modules.py:144-171) ------------------- */
/* torch.nn.utils.weight_norm: w[m,:] = g[m] v[m,:] / ||v[m,:]||, R = elements per output channel. */
int gtk_weightnorm_fwd(const float* g, const float* v, float* w, float* norm_out, int M, int R, void* stream);
int gtk_weightnorm_bwd(const float* g, const float* v, const float* dw, float* dg, float* dv, int M, int R, int accumulate,
                       void* stream);
/* W [M,Cin,K] (torch layout) -> packed operand of gtk_conv1d: transpose=0 forward, 1 data-gradient. */
int gtk_conv1d_pack_weight(const float* w, float* wp, int M, int Cin, int K, int transpose, void* stream);
/* F.conv1d(in, W, bias, padding=(K*dil-dil)/2, dilation=dil) with K odd, on a packed weight, plus the fused
 * tail out = (conv + bias + resid*(mask_resid?mask:1)) * (mask_out?mask:1).  flags: bit0 mask_out,
 * bit1 mask_resid, bit2 mask_in. */
int gtk_conv1d(const float* in, const float* wp, const float* bias, const float* mask, const float* resid, float* out,
               int B, int Cin, int M, int T, int K, int dil, int flags, void* stream);
/* dW [M,Cin,K] += sum_{b,t} dout[b,m,t] in[b,c,t+(j-(K-1)/2)dil]; flags: bit0 mask dout, bit1 mask in. */
int gtk_conv1d_wgrad(const float* dout, const float* in, const float* mask, float* dw, int B, int Cin, int M, int T, int K,
                     int dil, int flags, void* stream);
/* rs [B,RS,T], RS==2H: x_out=(x+rs[:, :H])*mask, skip(+)=rs[:,H:]; RS==H: skip(+)=rs (modules.py:166-171). */
int gtk_wn_resskip(const float* rs, const float* x, const float* mask, float* x_out, float* skip, int B, int H, int RS,
                   int T, int first, int mask_skip, void* stream);
/* Affine coupling tail (attentions.py:172-183): end_out=[m|logs]; z=[x0 | (m+exp(logs)x1)mask] and
 * logdet[b] += sum logs*mask (forward), z=[x0 | (x1-m)exp(-logs)mask] (reverse). */
int gtk_coupling_affine(const float* x, const float* end_out, const float* mask, float* z, float* logdet, int B, int C,
                        int T, int reverse, int sigmoid_scale, void* stream);
int gtk_coupling_affine_bwd(const float* x, const float* end_out, const float* mask, const float* dz, const float* dlogdet,
                            float* dend, float* dx, int B, int C, int T, int sigmoid_scale, void* stream);

/* ---- bf16 tensor-core path (tcgen05 + TMA), frame-major activations ----------------------- */
/* "Frames" are bf16 [B, T, ld] tensors with the channel fastest and ld a multiple of 64 (zero padded);
 * they are the K-major UMMA operands of the tensor-core convolutions. */
/* x fp32 [B,C,T] channels [c0, c0+nc) -> frames bf16 [B,T,ld] (columns nc..ld zeroed). */
int gtk_tc_frames_from_tmajor(const float* x, void* frames_bf16, int B, int C, int T, int c0, int nc, int ld, void* stream);
/* W fp32 [M,Cin,K] (torch) -> bf16 [K][rows][cols], zero padded; transpose=0: out[j][m][c] (forward operand, rows>=M,
 * cols>=Cin); transpose=1: out[j][c][m] = W[m][c][K-1-j] (data-gradient operand, rows>=Cin, cols>=M). */
int gtk_tc_pack_weight(const float* w, void* wp_bf16, int M, int Cin, int K, int transpose, int rows, int cols, void* stream);
/* out[b,t,n] = bias[n] + sum_{j,c} in[b, t+(j-K/2)dil, c] wp[j][n][c]  (F.conv1d of modules.py:152,165 /
 * attentions.py:151,170 on frames).  in: frames [B,T,ld_in]; wp: [K][N][ld_in]; N % 16 == 0, N <= 512.
 * flags bit0: multiply the output by mask[b,t].  out_f32 (ld_f32) and out_bf16 (ld_bf16) frame-major, either may be NULL. */
int gtk_tc_conv1d(const void* in_bf16, int ld_in, const void* wp_bf16, const float* bias, const float* mask, int B, int T,
                  int K, int dil, int N, float* out_f32, int ld_f32, void* out_bf16, int ld_bf16, int flags, void* stream);
/* Weight gradient of gtk_tc_conv1d: dw[m][c][j] += sum_{b,t} dy[b,t,m] x[b, t+(j-K/2)dil, c]; dy frames [B,T,ld_y]
 * (M <= ld_y valid columns), x frames [B,T,ld_x] (Cin <= ld_x valid columns, Cin <= 256); dw fp32 [M,Cin,K] (torch
 * layout), ACCUMULATED into with atomics. */
int gtk_tc_wgrad(const void* dy_bf16, int ld_y, const void* x_bf16, int ld_x, float* dw, int B, int T, int M, int Cin, int K,
                 int dil, void* stream);
/* out[g*out_gstride + n] += sum of rows [g*rows_per_group, (g+1)*rows_per_group) of frames in[.., n], n < n_cols
 * (bias gradients: groups=1, rows=B*T; conditioning gradients: groups=B, rows=T). */
int gtk_tc_colsum(const void* in_bf16, int ld, int n_cols, int groups, long long rows_per_group, float* out,
                  size_t out_gstride, void* stream);
/* One fused WaveNet layer (modules.py:151-170) on frames: dilated conv -> (dropout) -> +cond -> tanh*sigmoid ->
 * res/skip 1x1 -> residual / skip update, in one tcgen05 kernel.  h: layer input as bf16 frames [B,T,H] (operand)
 * and fp32 frames (residual master).  win_p = gtk_tc_pack_weight(in_layers[i].weight) [K][2H][H]; wrs_p =
 * gtk_tc_pack_weight(res_skip_layers[i].weight) [1][RS][H].  cond: this layer's 2H slice of cond_layer(g), batch
 * stride cond_bstride floats, or NULL.  first: skip is overwritten; last (RS == H): skip*mask -> skip_out_bf16 and no
 * h_out.  ts_save [B,T,2H] / acts_save [B,T,H] (bf16) are the activations the backward needs, or NULL. */
int gtk_tc_wn_layer(const void* h_bf16, const float* h_f32, const void* win_p, const void* wrs_p, const float* bias_in,
                    const float* bias_rs, const float* cond, size_t cond_bstride, const float* mask, int B, int T, int H,
                    int RS, int K, int dil, int first, int last, float drop_p, unsigned long long drop_seed,
                    float* h_out_f32, void* h_out_bf16, float* skip_f32, void* skip_out_bf16, void* ts_save,
                    void* acts_save, void* stream);

/* Backward of gtk_tc_wn_layer in one tcgen05 kernel (autograd of modules.py:151-170): res/skip data gradient -> gate
 * backward (from the saved tanh|sigmoid, dropout mask regenerated) -> dilated-conv data gradient.
 * drs: bf16 frames [B,T,ld_drs], columns [0,RS) = [d(h_next) * mask | d(skip)] (RS == H: d(skip) only);
 * wrs_t = gtk_tc_pack_weight(res_skip weight, transpose=1) [1][H][pad64(RS)]; win_t = gtk_tc_pack_weight(in_layer
 * weight, transpose=1) [K][H][2H].  ts_tp / resid_tp / out_tp use the TILE-PRIVATE layout the decoder keeps between
 * epilogues: 16-byte vector v of frame t of utterance b lives at (((b * ceil(T/128) + t/128) * n_vec + v) * 128 + t%128)
 * * 16 bytes (n_vec = row bytes / 16).  ts_tp: saved [tanh|sigmoid] bf16 (2H columns); resid_tp: fp32 d(h) carried from
 * the layer above (H columns) or NULL; out_tp: fp32 d(h) of this layer (H columns).  dxin: bf16 frames [B,T,2H] out
 * (operand of the weight gradient); out_bf16: bf16 frames [B,T,ld_out], columns [0,H) = d(h) * mask. */
int gtk_tc_wn_bwd_layer(const void* drs_bf16, int ld_drs, const void* wrs_t, const void* win_t, const void* ts_tp,
                        const float* mask, const float* resid_tp, float* out_tp, void* dxin_bf16, void* out_bf16,
                        int ld_out, int B, int T, int H, int RS, int K, int dil, float drop_p,
                        unsigned long long drop_seed, void* stream);

/* WN.cond_layer(g) (modules.py:122-124, 148-149: a weight-normed 1x1 conv gin_channels -> 2*H*L applied to the speaker
 * embedding g [B, G, 1]) of n_blocks coupling blocks in ONE launch: cond[k][b][n] = bias_k[n] + sum_c W_k[n][c] g[b][c],
 * W_k = weight_g_k[n] * weight_v_k[n][:] / ||weight_v_k[n][:]|| (pass weight_g == NULL, or a NULL entry, for a conv whose
 * weight norm was removed).  weight_g / weight_v / bias: host arrays of n_blocks (<= 32) DEVICE pointers ([N], [N, G],
 * [N]); g [B, G]; cond [n_blocks, B, N] (the layout gtk_block_params.cond expects per block).  G <= 512. */
int gtk_cond_layers(const float* const* weight_g, const float* const* weight_v, const float* const* bias, int n_blocks,
                    const float* g, float* cond, int B, int G, int N, void* stream);
/* Its backward: dcond [n_blocks, B, N] -> d_weight_g / d_weight_v / d_bias (host arrays of device pointers, WRITTEN) and
 * dg [B, G] (ACCUMULATED with atomics: zero it first; may be NULL). */
int gtk_cond_layers_bwd(const float* const* weight_g, const float* const* weight_v, int n_blocks, const float* g,
                        const float* dcond, float* const* d_weight_g, float* const* d_weight_v, float* const* d_bias, float* dg,
                        int B, int G, int N, void* stream);

/* ---- whole decoder (models.py:719-789) --------------------------------------------------- */
typedef struct gtk_decoder_desc {
    int B;              /* batch */
    int in_channels;    /* mel channels (80) */
    int T;              /* mel frames of the input tensor (un-squeezed) */
    int n_sqz;          /* 2 */
    int hidden;         /* 192 */
    int kernel_size;    /* 5 */
    int dilation_rate;  /* 1 */
    int n_layers;       /* 4 */
    int n_blocks;       /* 12 */
    int n_split;        /* 4 */
    int sigmoid_scale;  /* 0 */
    int precision;      /* 0: fp32 CUDA-core convolutions (exact-parity path); 1: bf16 tcgen05 path */
    float p_dropout;    /* nn.Dropout on every WN in_layer output (modules.py:120,153); 0 = off (eval) */
    unsigned long long seed;   /* per-call dropout seed; decoder_backward must be given the forward's value */
    const unsigned long long* seed_dev;   /* if not NULL the per-call seed is read from this DEVICE address by the
                                           * kernels instead (`seed` is ignored): a captured CUDA graph then picks up a
                                           * fresh seed on every replay */
    /* The fork's pitch / energy conditioned coupling blocks (attentions.py:155-160: wn -> wn_energy -> wn_pitch, modules.WNP
     * modules.py:272-362), bf16 path.  energy / pitch: [B, T] tracks on the un-squeezed time axis, or NULL = that WaveNet is
     * the identity (modules.py:323-324).  d_energy / d_pitch: [B, T] gradients of the tracks, WRITTEN by the backward, or NULL. */
    const float* energy;
    const float* pitch;
    float* d_energy;
    float* d_pitch;
} gtk_decoder_desc;

/* Parameters of one modules.WNP (wn_energy / wn_pitch of a coupling block); all NULL when unused. */
typedef struct gtk_wnp_params {
    const float* in_g[GTK_MAX_LAYERS];    /* ....in_layers.{i}.weight_g [2H] */
    const float* in_v[GTK_MAX_LAYERS];    /* ....in_layers.{i}.weight_v [2H, H, K] */
    const float* in_b[GTK_MAX_LAYERS];    /* ....in_layers.{i}.bias [2H] */
    const float* rs_g[GTK_MAX_LAYERS];    /* ....res_skip_layers.{i}.weight_g [RS_i] */
    const float* rs_v[GTK_MAX_LAYERS];    /* ....res_skip_layers.{i}.weight_v [RS_i, H, 1] */
    const float* rs_b[GTK_MAX_LAYERS];    /* ....res_skip_layers.{i}.bias [RS_i] */
    const float* c1_g;                    /* ....cond_layer1.weight_g [2HL/n_sqz] (NULL after remove_weight_norm) */
    const float* c1_v;                    /* ....cond_layer1.weight_v (or .weight) [2HL/n_sqz, 1, 1] */
    const float* c1_b;                    /* ....cond_layer1.bias [2HL/n_sqz] */
} gtk_wnp_params;
typedef struct gtk_wnp_grads {
    float* in_g[GTK_MAX_LAYERS]; float* in_v[GTK_MAX_LAYERS]; float* in_b[GTK_MAX_LAYERS];
    float* rs_g[GTK_MAX_LAYERS]; float* rs_v[GTK_MAX_LAYERS]; float* rs_b[GTK_MAX_LAYERS];
    float* c1_g; float* c1_v; float* c1_b;
} gtk_wnp_grads;

/* Parameters of one (ActNorm, InvConvNear, CouplingBlock) triple; state_dict names in comments.
 * Weight-normed convs take (weight_g, weight_v); after remove_weight_norm pass weight as *_v and NULL as *_g. */
typedef struct gtk_block_params {
    const float* an_logs;                 /* flows.{3b}.logs   [C]  */
    const float* an_bias;                 /* flows.{3b}.bias   [C]  */
    const float* ic_weight;               /* flows.{3b+1}.weight [n_split^2] */
    const float* start_g;                 /* flows.{3b+2}.start.weight_g [H] */
    const float* start_v;                 /* ....start.weight_v [H, C/2, 1]  */
    const float* start_b;                 /* ....start.bias [H] */
    const float* in_g[GTK_MAX_LAYERS];    /* ....wn.in_layers.{i}.weight_g [2H] */
    const float* in_v[GTK_MAX_LAYERS];    /* ....wn.in_layers.{i}.weight_v [2H, H, K] */
    const float* in_b[GTK_MAX_LAYERS];    /* ....wn.in_layers.{i}.bias [2H] */
    const float* rs_g[GTK_MAX_LAYERS];    /* ....wn.res_skip_layers.{i}.weight_g [RS_i] */
    const float* rs_v[GTK_MAX_LAYERS];    /* ....wn.res_skip_layers.{i}.weight_v [RS_i, H, 1] */
    const float* rs_b[GTK_MAX_LAYERS];    /* ....wn.res_skip_layers.{i}.bias [RS_i] */
    const float* end_w;                   /* ....end.weight [C, H, 1] */
    const float* end_b;                   /* ....end.bias [C] */
    const float* cond;                    /* wn.cond_layer(g) output [B, 2H*L] or NULL (modules.py:148-149) */
    gtk_wnp_params wn_energy;             /* flows.{3b+2}.wn_energy.* (used when gtk_decoder_desc.energy != NULL) */
    gtk_wnp_params wn_pitch;              /* flows.{3b+2}.wn_pitch.*  (used when gtk_decoder_desc.pitch != NULL) */
} gtk_block_params;

/* Gradient destinations, same shapes as the parameters; ACCUMULATED into (zero them once per step). */
typedef struct gtk_block_grads {
    float* an_logs; float* an_bias; float* ic_weight;
    float* start_g; float* start_v; float* start_b;
    float* in_g[GTK_MAX_LAYERS]; float* in_v[GTK_MAX_LAYERS]; float* in_b[GTK_MAX_LAYERS];
    float* rs_g[GTK_MAX_LAYERS]; float* rs_v[GTK_MAX_LAYERS]; float* rs_b[GTK_MAX_LAYERS];
    float* end_w; float* end_b;
    float* cond;                          /* [B, 2H*L] or NULL */
    gtk_wnp_grads wn_energy;
    gtk_wnp_grads wn_pitch;
} gtk_block_grads;

/* train != 0: activations are kept in the workspace for gtk_decoder_backward. */
size_t gtk_decoder_workspace_bytes(const gtk_decoder_desc* d, int train);
/* FlowSpecDecoder.forward(reverse=False) (models.py:765-785): x [B,in,T], mask [B,T] -> z [B,in,T'] with
 * T' = (T/n_sqz)*n_sqz, logdet [B]. */
int gtk_decoder_forward(const gtk_decoder_desc* d, const gtk_block_params* blocks, const float* x, const float* mask,
                        float* z, float* logdet, void* workspace, size_t workspace_bytes, int train, void* stream);
/* FlowSpecDecoder.forward(reverse=True): z -> x.  W^-1 is recomputed on device per call (it is a 4x4). */
int gtk_decoder_reverse(const gtk_decoder_desc* d, const gtk_block_params* blocks, const float* z, const float* mask,
                        float* x, void* workspace, size_t workspace_bytes, void* stream);
/* Backward of gtk_decoder_forward(train=1) with the same workspace: dz [B,in,T'], dlogdet [B] -> dx [B,in,T]
 * and parameter gradients. */
int gtk_decoder_backward(const gtk_decoder_desc* d, const gtk_block_params* blocks, const gtk_block_grads* grads,
                         const float* mask, const float* dz, const float* dlogdet, float* dx, void* workspace,
                         size_t workspace_bytes, void* stream);
/* The same backward issued as block ranges blk_hi .. blk_lo (inclusive, descending; first call blk_hi = n_blocks-1,
 * last call blk_lo = 0, consecutive ranges in between).  The running gradient stays in the workspace between calls, so
 * the parameter gradients of a finished range are final: data-parallel training all-reduces them (NCCL, DDP at
 * train_ms_emo_lang_pitch.py:166) while the next range computes. */
int gtk_decoder_backward_range(const gtk_decoder_desc* d, const gtk_block_params* blocks, const gtk_block_grads* grads,
                               const float* mask, const float* dz, const float* dlogdet, float* dx, void* workspace,
                               size_t workspace_bytes, int blk_hi, int blk_lo, void* stream);
/* gtk_decoder_backward_range for a range that is NOT the last one, returning without joining the library's internal side
 * stream: the range's last weight-gradient launch and its reduce keep running there while the caller enqueues the next
 * range, which carries on as the un-split backward would (a range boundary then costs what a block boundary costs).  A
 * consumer of this range's parameter gradients -- the gradient all-reduce of DDP, train_ms_emo_lang_pitch.py:166 -- is
 * ordered with gtk_decoder_backward_join(consumer_stream) AND an ordinary dependency on `stream` (bias / ActNorm gradients
 * are written there).  The last range (blk_lo = 0, any of the two entry points) joins everything.  With the side stream off
 * (GTK_WGRAD_OVERLAP=0/1, fp32 path) this is gtk_decoder_backward_range. */
int gtk_decoder_backward_range_deferred(const gtk_decoder_desc* d, const gtk_block_params* blocks, const gtk_block_grads* grads,
                                        const float* mask, const float* dz, const float* dlogdet, float* dx, void* workspace,
                                        size_t workspace_bytes, int blk_hi, int blk_lo, void* stream);
/* `stream` waits for the side-stream work of the last deferred range on the current device (no-op when there is none). */
int gtk_decoder_backward_join(void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GLOWTTS_B200_H */
