// glowtts_b200/csrc/abi.cu -- the extern "C" surface declared in include/glowtts_b200.h.
#include <stdlib.h>
#include <stdarg.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <string>
#include <vector>

#include "tc.cuh"
#include "ops.h"
#include "../../include/glowtts_b200.h"

namespace gtk {

static thread_local char g_err[512] = "";

bool pdl_enabled() {
    static int on = -1;
    if (on < 0) {
        const char* e = getenv("GTK_PDL");
        on = (e && e[0] == '0') ? 0 : 1;
    }
    return on == 1;
}

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static std::atomic<unsigned long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

// ---- per-kernel event timing -------------------------------------------------------------------
struct ProfTag {
    std::string name;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> spans;
    size_t used = 0;
};
static std::mutex g_prof_mu;
static std::vector<ProfTag> g_prof;
static std::atomic<int> g_prof_on{0};

static ProfTag* prof_tag(const char* tag) {
    for (auto& t : g_prof)
        if (t.name == tag) return &t;
    g_prof.emplace_back();
    g_prof.back().name = tag;
    return &g_prof.back();
}
void prof_begin(const char* tag, cudaStream_t st) {
    if (!g_prof_on.load(std::memory_order_relaxed)) return;
    std::lock_guard<std::mutex> lk(g_prof_mu);
    ProfTag* t = prof_tag(tag);
    if (t->used == t->spans.size()) {
        cudaEvent_t a, b;
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        t->spans.emplace_back(a, b);
    }
    cudaEventRecord(t->spans[t->used].first, st);
}
void prof_end(const char* tag, cudaStream_t st) {
    if (!g_prof_on.load(std::memory_order_relaxed)) return;
    std::lock_guard<std::mutex> lk(g_prof_mu);
    ProfTag* t = prof_tag(tag);
    if (t->used < t->spans.size()) cudaEventRecord(t->spans[t->used++].second, st);
}

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

namespace tc {
int conv1d_plain(const __nv_bfloat16* in, int ld_in, const __nv_bfloat16* wp, const float* bias, const float* mask, int B, int T,
                 int K, int dil, int N, float* out_f32, int ld_f32, __nv_bfloat16* out_bf16, int ld_bf16, int mask_out,
                 cudaStream_t st);
}  // namespace tc

namespace decoder {
size_t workspace_bytes(const gtk_decoder_desc* d, int train);
int forward(const gtk_decoder_desc* d, const gtk_block_params* blocks, const float* x, const float* mask, float* z,
            float* logdet, void* workspace, size_t ws_bytes, int train, cudaStream_t st);
int reverse(const gtk_decoder_desc* d, const gtk_block_params* blocks, const float* z, const float* mask, float* x,
            void* workspace, size_t ws_bytes, cudaStream_t st);
int backward(const gtk_decoder_desc* d, const gtk_block_params* blocks, const gtk_block_grads* grads, const float* mask,
             const float* dz, const float* dlogdet, float* dx, void* workspace, size_t ws_bytes, int blk_hi, int blk_lo,
             cudaStream_t st);
int backward_deferred(const gtk_decoder_desc* d, const gtk_block_params* blocks, const gtk_block_grads* grads, const float* mask,
                      const float* dz, const float* dlogdet, float* dx, void* workspace, size_t ws_bytes, int blk_hi, int blk_lo,
                      cudaStream_t st);
int backward_join_stream(cudaStream_t st);
}  // namespace decoder
}  // namespace gtk

using namespace gtk;
#define ST static_cast<cudaStream_t>(stream)

extern "C" {

int gtk_abi_version(void) { return 1; }
const char* gtk_last_error(void) { return g_err; }
unsigned long long gtk_launch_count(void) { return g_launches.load(); }

int gtk_prof_enable(int on) {
    g_prof_on.store(on ? 1 : 0);
    return GTK_OK;
}
int gtk_prof_reset(void) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    for (auto& t : g_prof) t.used = 0;
    return GTK_OK;
}
int gtk_prof_tag_count(void) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    return (int)g_prof.size();
}
int gtk_prof_read(int i, char* name, int name_cap, int* launches, float* total_ms) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    if (i < 0 || i >= (int)g_prof.size() || !name || !launches || !total_ms) return GTK_ERR_INVALID;
    ProfTag& t = g_prof[i];
    snprintf(name, name_cap, "%s", t.name.c_str());
    float tot = 0.f;
    for (size_t k = 0; k < t.used; ++k) {
        GTK_CUDA(cudaEventSynchronize(t.spans[k].second));
        float ms = 0.f;
        GTK_CUDA(cudaEventElapsedTime(&ms, t.spans[k].first, t.spans[k].second));
        tot += ms;
    }
    *launches = (int)t.used;
    *total_ms = tot;
    return GTK_OK;
}

int gtk_mas_lengths_from_mask(const float* mask, int32_t* t_xs, int32_t* t_ys, int B, int T_x, int T_y, void* stream) {
    return mas::lengths_from_mask(mask, t_xs, t_ys, B, T_x, T_y, ST);
}
size_t gtk_mas_workspace_bytes(int B, int T_x, int T_y) { return mas::workspace_bytes(B, T_x, T_y); }
int gtk_mas_forward(const float* value, const int32_t* t_xs, const int32_t* t_ys, float* path, int32_t* dur_out,
                    int32_t* idx_out, int B, int T_x, int T_y, void* workspace, size_t workspace_bytes, int force_generic,
                    void* stream) {
    return mas::forward(value, t_xs, t_ys, path, dur_out, idx_out, B, T_x, T_y, workspace, workspace_bytes, force_generic, ST);
}

size_t gtk_mas_static_scratch_bytes(void) { return mas::static_scratch_bytes(); }
int gtk_mas_forward_static(const float* value, const int32_t* t_xs, const int32_t* t_ys, float* path, int32_t* dur_out,
                           int32_t* idx_out, int B, int T_x, int T_y, void* stream) {
    return mas::forward_static(value, t_xs, t_ys, path, dur_out, idx_out, B, T_x, T_y, ST);
}
int gtk_mas_fused_supported(int D, int T_x, int T_y, int has_x_logs) { return mas::fused_supported(D, T_x, T_y, has_x_logs); }
int gtk_mas_fused_logp(const float* z, const float* x_m, const float* x_logs, const int32_t* t_xs, const int32_t* t_ys,
                       float* path, int32_t* dur_out, int32_t* idx_out, float* logp_out, int B, int D, int T_x, int T_y,
                       void* stream) {
    if (x_logs) {
        set_error("mas_fused_logp: x_logs != NULL (not a mean_only model) is not fused; use gtk_logp + gtk_mas_forward");
        return GTK_ERR_UNSUPPORTED;
    }
    return mas::fused_forward(z, x_m, t_xs, t_ys, path, dur_out, idx_out, logp_out, B, D, T_x, T_y, ST);
}

int gtk_logp(const float* x_m, const float* x_logs, const float* z, float* logp, int B, int D, int T_x, int T_y, void* stream) {
    return logp::forward(x_m, x_logs, z, logp, B, D, T_x, T_y, ST);
}

int gtk_squeeze(const float* x, const float* mask, float* out, float* mask_out, int B, int C, int T, int n_sqz, void* stream) {
    return flow::squeeze(x, mask, out, mask_out, B, C, T, n_sqz, 0, ST);
}
int gtk_unsqueeze(const float* x, const float* mask, float* out, float* mask_out, int B, int C, int T, int n_sqz, void* stream) {
    return flow::unsqueeze(x, mask, out, mask_out, B, C, T, n_sqz, 0, ST);
}
int gtk_gate_fwd(const float* a, const float* b, int cond_mode, float* out, int B, int H, int T, void* stream) {
    return flow::gate_fwd(a, b, cond_mode, 0, out, B, H, T, 0.f, DropSeed{0ull, nullptr}, ST);
}
int gtk_gate_bwd(const float* a, const float* b, int cond_mode, const float* dout, float* da, int B, int H, int T, void* stream) {
    return flow::gate_bwd(a, b, cond_mode, 0, dout, da, B, H, T, 0.f, DropSeed{0ull, nullptr}, 1, ST);
}
int gtk_dropout_scale(unsigned long long seed, float p, long long n, float* out, void* stream) {
    return flow::dropout_scale(seed, p, n, out, ST);
}
unsigned long long gtk_decoder_layer_seed(unsigned long long seed, int block, int layer) {
    return drop_layer_seed(seed, block, layer);
}
int gtk_row_sum(const float* in, float* out, int R, int T, int accumulate, void* stream) {
    return flow::row_sum(in, out, R, T, accumulate, 0, 0, ST);
}
int gtk_channel_sum(const float* in, float* out, int B, int M, int T, int accumulate, void* stream) {
    return flow::channel_sum(in, nullptr, out, B, M, T, accumulate, ST);
}
int gtk_clip_grad_value(float* grad, long long n, float clip_value, int do_clip, float norm_type, float* acc, void* stream) {
    return flow::clip_grad_value(grad, n, clip_value, do_clip, norm_type, acc, ST);
}
int gtk_mle_loss_fwd(const float* z, const float* m, const float* logs, const float* logdet, const float* mask, float* acc4,
                     float* loss, int B, int C, int T, void* stream) {
    return flow::mle_loss_fwd(z, m, logs, logdet, mask, acc4, loss, B, C, T, ST);
}
int gtk_mle_loss_bwd(const float* z, const float* m, const float* logs, const float* acc4, const float* gloss, float* dz,
                     float* dm, float* dlogs, float* dlogdet, int B, int C, int T, void* stream) {
    return flow::mle_loss_bwd(z, m, logs, acc4, gloss, dz, dm, dlogs, dlogdet, B, C, T, ST);
}
int gtk_gather_tokens(const float* src, const int32_t* idx, float* out, int B, int C, int T_x, int T_y, void* stream) {
    return flow::gather_tokens(src, idx, out, B, C, T_x, T_y, ST);
}
int gtk_generate_path(const float* duration, const float* mask, float* path, int B, int T_x, int T_y, void* stream) {
    return logp::generate_path(duration, mask, path, B, T_x, T_y, ST);
}

int gtk_mask_len(const float* mask, float* x_len, int B, int T, void* stream) { return flow::mask_len(mask, x_len, B, T, ST); }
int gtk_invconv_prep(const float* w, int n_split, float* winv, float* logdet_w, void* stream) {
    return flow::invconv_prep(w, n_split, winv, logdet_w, ST);
}
int gtk_flow_mix(const float* x, const float* mask, const float* logs, const float* bias, const float* M, int n_split, float* z,
                 int B, int C, int T, int reverse, void* stream) {
    return flow::mix(x, mask, logs, bias, M, n_split, z, B, C, T, reverse, ST);
}
int gtk_flow_mix_bwd(const float* x, const float* mask, const float* logs, const float* bias, const float* M, int n_split,
                     const float* dz, float* dx, float* dlogs, float* dbias, float* dM, int B, int C, int T, void* stream) {
    return flow::mix_bwd(x, mask, logs, bias, M, n_split, dz, dx, dlogs, dbias, dM, B, C, T, ST);
}
int gtk_flow_logdet_const(const float* logs, int C, const float* logdet_w, float group_count, const float* x_len,
                          float* logdet, int B, int accumulate, void* stream) {
    return flow::logdet_const(logs, C, logdet_w, group_count, x_len, logdet, B, accumulate, ST);
}
int gtk_flow_logdet_const_bwd(const float* dlogdet, const float* x_len, int B, float* dlogs, int C, const float* winv,
                              int n_split, float group_count, float* dW, void* stream) {
    return flow::logdet_const_bwd(dlogdet, x_len, B, dlogs, C, winv, n_split, group_count, dW, ST);
}
int gtk_actnorm_ddi_stats(const float* x, const float* mask, float* stats, int B, int C, int T, void* stream) {
    return flow::actnorm_ddi_stats(x, mask, stats, B, C, T, ST);
}
int gtk_actnorm_ddi_finalize(const float* stats, float* logs, float* bias, int C, void* stream) {
    return flow::actnorm_ddi_finalize(stats, logs, bias, C, ST);
}

int gtk_weightnorm_fwd(const float* g, const float* v, float* w, float* norm_out, int M, int R, void* stream) {
    return flow::weightnorm_fwd(g, v, w, norm_out, M, R, ST);
}
int gtk_weightnorm_bwd(const float* g, const float* v, const float* dw, float* dg, float* dv, int M, int R, int accumulate,
                       void* stream) {
    return flow::weightnorm_bwd(g, v, dw, dg, dv, M, R, accumulate, ST);
}
int gtk_conv1d_pack_weight(const float* w, float* wp, int M, int Cin, int K, int transpose, void* stream) {
    return conv::pack_weight(w, wp, M, Cin, K, transpose, ST);
}
int gtk_conv1d(const float* in, const float* wp, const float* bias, const float* mask, const float* resid, float* out, int B,
               int Cin, int M, int T, int K, int dil, int flags, void* stream) {
    conv::ConvParams p;
    memset(&p, 0, sizeof(p));
    p.in = in; p.wp = wp; p.bias = bias; p.mask = mask; p.resid = resid; p.out = out;
    p.B = B; p.Cin = Cin; p.M = M; p.T = T; p.K = K; p.dil = dil;
    p.mask_out = flags & 1; p.mask_resid = (flags >> 1) & 1; p.mask_in = (flags >> 2) & 1;
    return conv::conv1d_f32(p, ST);
}
int gtk_conv1d_wgrad(const float* dout, const float* in, const float* mask, float* dw, int B, int Cin, int M, int T, int K,
                     int dil, int flags, void* stream) {
    conv::WgradParams p;
    memset(&p, 0, sizeof(p));
    p.dout = dout; p.in = in; p.mask = mask; p.dw = dw;
    p.B = B; p.Cin = Cin; p.M = M; p.T = T; p.dil = dil;
    p.mask_dout = flags & 1; p.mask_in = (flags >> 1) & 1;
    return conv::wgrad_f32(p, K, ST);
}
int gtk_wn_resskip(const float* rs, const float* x, const float* mask, float* x_out, float* skip, int B, int H, int RS, int T,
                   int first, int mask_skip, void* stream) {
    return flow::resskip(rs, x, mask, x_out, skip, B, H, RS, T, first, mask_skip, ST);
}
int gtk_coupling_affine(const float* x, const float* end_out, const float* mask, float* z, float* logdet, int B, int C, int T,
                        int reverse, int sigmoid_scale, void* stream) {
    return flow::affine(x, end_out, mask, z, logdet, B, C, T, reverse, sigmoid_scale, ST);
}
int gtk_coupling_affine_bwd(const float* x, const float* end_out, const float* mask, const float* dz, const float* dlogdet,
                            float* dend, float* dx, int B, int C, int T, int sigmoid_scale, void* stream) {
    return flow::affine_bwd(x, end_out, mask, dz, dlogdet, dend, dx, B, C, T, sigmoid_scale, ST);
}

int gtk_tc_frames_from_tmajor(const float* x, void* frames_bf16, int B, int C, int T, int c0, int nc, int ld, void* stream) {
    return tc::frames_from_tmajor(x, static_cast<__nv_bfloat16*>(frames_bf16), B, C, T, c0, nc, ld, ST);
}
int gtk_tc_pack_weight(const float* w, void* wp_bf16, int M, int Cin, int K, int transpose, int rows, int cols, void* stream) {
    return tc::pack_weight_bf16(w, static_cast<__nv_bfloat16*>(wp_bf16), M, Cin, K, transpose, rows, cols, ST);
}
int gtk_tc_conv1d(const void* in_bf16, int ld_in, const void* wp_bf16, const float* bias, const float* mask, int B, int T,
                  int K, int dil, int N, float* out_f32, int ld_f32, void* out_bf16, int ld_bf16, int flags, void* stream) {
    return tc::conv1d_plain(static_cast<const __nv_bfloat16*>(in_bf16), ld_in, static_cast<const __nv_bfloat16*>(wp_bf16), bias,
                            mask, B, T, K, dil, N, out_f32, ld_f32, static_cast<__nv_bfloat16*>(out_bf16), ld_bf16, flags & 1, ST);
}

int gtk_tc_wgrad(const void* dy_bf16, int ld_y, const void* x_bf16, int ld_x, float* dw, int B, int T, int M, int Cin, int K,
                 int dil, void* stream) {
    return tc::wgrad_tc_launch(static_cast<const __nv_bfloat16*>(dy_bf16), ld_y, ld_y, static_cast<const __nv_bfloat16*>(x_bf16),
                               ld_x, ld_x, dw, B, T, M, Cin, K, dil, ST);
}
int gtk_tc_colsum(const void* in_bf16, int ld, int n_cols, int groups, long long rows_per_group, float* out,
                  size_t out_gstride, void* stream) {
    return tc::colsum_frames(static_cast<const __nv_bfloat16*>(in_bf16), ld, n_cols, groups, rows_per_group, out, out_gstride, ST);
}
#ifdef GTK_INSTRUMENT
int gtk_debug_mas_trace(unsigned long long* out, int n_ctas) { return mas::trace_read(out, n_ctas); }
int gtk_debug_mas_wait(long long* out, int n_ctas) { return mas::wait_read(out, n_ctas); }
int gtk_debug_wn_trace(unsigned long long* out, int n_ctas) { return tc::wn_trace_read(out, n_ctas); }
int gtk_debug_wn_wait(long long* out, int n_ctas) { return tc::wn_wait_read(out, n_ctas); }
int gtk_debug_wb_wait(long long* out, int n_ctas) { return tc::wb_wait_read(out, n_ctas); }
int gtk_debug_wb_trace(unsigned long long* out, int n_ctas) { return tc::wb_trace_read(out, n_ctas); }
int gtk_debug_ws_trace(unsigned long long* out, int n_ctas) { return tc::ws_trace_read(out, n_ctas); }
int gtk_debug_ws_wait(long long* out, int n_ctas) { return tc::ws_wait_read(out, n_ctas); }
int gtk_debug_gg_wait(long long* out, int n_ctas) { return tc::gg_wait_read(out, n_ctas); }
int gtk_debug_conv_trace_select(int epilogue, int taps) { tc::conv_trace_select(epilogue, taps); return GTK_OK; }
int gtk_debug_conv_trace(unsigned long long* out, int n_ctas) { return tc::conv_trace_read(out, n_ctas); }
int gtk_debug_conv_wait(long long* out, int n_ctas) { return tc::conv_wait_read(out, n_ctas); }
int gtk_debug_shift_probe(const void* a, const void* b, float* out, void* stream) {
    return tc::shift_probe(static_cast<const __nv_bfloat16*>(a), static_cast<const __nv_bfloat16*>(b), out, ST);
}
#endif  // GTK_INSTRUMENT
int gtk_tc_wn_layer(const void* h_bf16, const float* h_f32, const void* win_p, const void* wrs_p, const float* bias_in,
                    const float* bias_rs, const float* cond, size_t cond_bstride, const float* mask, int B, int T, int H,
                    int RS, int K, int dil, int first, int last, float drop_p, unsigned long long drop_seed,
                    float* h_out_f32, void* h_out_bf16, float* skip_f32, void* skip_out_bf16, void* ts_save,
                    void* acts_save, void* stream) {
    tc::WnLayerParams p;
    memset(&p, 0, sizeof(p));
    p.B = B; p.T = T; p.H = H; p.RS = RS; p.taps = K; p.dil = dil; p.first = first; p.last = last;
    p.bias_in = bias_in; p.bias_rs = bias_rs; p.cond = cond; p.cond_bstride = cond_bstride; p.mask = mask;
    p.drop_p = drop_p; p.drop_seed.value = drop_seed; p.drop_seed.dev = nullptr;
    p.h_f32 = h_f32; p.h_out_f32 = h_out_f32; p.h_out_bf16 = static_cast<__nv_bfloat16*>(h_out_bf16);
    p.skip_f32 = skip_f32; p.skip_out_bf16 = static_cast<__nv_bfloat16*>(skip_out_bf16);
    p.ts_save = static_cast<__nv_bfloat16*>(ts_save); p.acts_save = static_cast<__nv_bfloat16*>(acts_save);
    return tc::wn_layer_launch(static_cast<const __nv_bfloat16*>(h_bf16), static_cast<const __nv_bfloat16*>(win_p),
                               static_cast<const __nv_bfloat16*>(wrs_p), p, ST);
}

int gtk_tc_wn_bwd_layer(const void* drs_bf16, int ld_drs, const void* wrs_t, const void* win_t, const void* ts_tp,
                        const float* mask, const float* resid_tp, float* out_tp, void* dxin_bf16, void* out_bf16,
                        int ld_out, int B, int T, int H, int RS, int K, int dil, float drop_p,
                        unsigned long long drop_seed, void* stream) {
    tc::WnBwdParams p;
    memset(&p, 0, sizeof(p));
    p.B = B; p.T = T; p.H = H; p.RS = RS; p.taps = K; p.dil = dil;
    p.mask = mask; p.ts = static_cast<const __nv_bfloat16*>(ts_tp);
    p.drop_p = drop_p; p.drop_seed.value = drop_seed; p.drop_seed.dev = nullptr;
    p.resid_f32 = resid_tp; p.out_f32 = out_tp;
    return tc::wn_bwd_layer_launch(static_cast<const __nv_bfloat16*>(drs_bf16), ld_drs, static_cast<const __nv_bfloat16*>(wrs_t),
                                   static_cast<const __nv_bfloat16*>(win_t), static_cast<__nv_bfloat16*>(dxin_bf16),
                                   static_cast<__nv_bfloat16*>(out_bf16), ld_out, p, ST);
}

int gtk_cond_layers(const float* const* weight_g, const float* const* weight_v, const float* const* bias, int n_blocks,
                    const float* g, float* cond, int B, int G, int N, void* stream) {
    GTK_REQUIRE(weight_v && n_blocks >= 1 && n_blocks <= flow::COND_MAX_BLOCKS, "1..32 blocks per call");
    flow::CondJobs J;
    memset(&J, 0, sizeof(J));
    J.n = n_blocks;
    for (int k = 0; k < n_blocks; ++k) {
        GTK_REQUIRE(weight_v[k] != nullptr, "null weight");
        J.g[k] = weight_g ? weight_g[k] : nullptr; J.v[k] = weight_v[k]; J.b[k] = bias ? bias[k] : nullptr;
    }
    return flow::cond_all(J, g, cond, B, G, N, ST);
}
int gtk_cond_layers_bwd(const float* const* weight_g, const float* const* weight_v, int n_blocks, const float* g,
                        const float* dcond, float* const* d_weight_g, float* const* d_weight_v, float* const* d_bias, float* dg,
                        int B, int G, int N, void* stream) {
    GTK_REQUIRE(weight_v && n_blocks >= 1 && n_blocks <= flow::COND_MAX_BLOCKS, "1..32 blocks per call");
    flow::CondJobs J;
    flow::CondGrads D;
    memset(&J, 0, sizeof(J));
    memset(&D, 0, sizeof(D));
    J.n = D.n = n_blocks;
    for (int k = 0; k < n_blocks; ++k) {
        GTK_REQUIRE(weight_v[k] != nullptr, "null weight");
        J.g[k] = weight_g ? weight_g[k] : nullptr; J.v[k] = weight_v[k];
        D.g[k] = d_weight_g ? d_weight_g[k] : nullptr; D.v[k] = d_weight_v ? d_weight_v[k] : nullptr;
        D.b[k] = d_bias ? d_bias[k] : nullptr;
    }
    return flow::cond_all_bwd(J, D, g, dcond, dg, B, G, N, ST);
}

size_t gtk_decoder_workspace_bytes(const gtk_decoder_desc* d, int train) { return decoder::workspace_bytes(d, train); }
int gtk_decoder_forward(const gtk_decoder_desc* d, const gtk_block_params* blocks, const float* x, const float* mask, float* z,
                        float* logdet, void* workspace, size_t workspace_bytes, int train, void* stream) {
    return decoder::forward(d, blocks, x, mask, z, logdet, workspace, workspace_bytes, train, ST);
}
int gtk_decoder_reverse(const gtk_decoder_desc* d, const gtk_block_params* blocks, const float* z, const float* mask, float* x,
                        void* workspace, size_t workspace_bytes, void* stream) {
    return decoder::reverse(d, blocks, z, mask, x, workspace, workspace_bytes, ST);
}
int gtk_decoder_backward(const gtk_decoder_desc* d, const gtk_block_params* blocks, const gtk_block_grads* grads,
                         const float* mask, const float* dz, const float* dlogdet, float* dx, void* workspace,
                         size_t workspace_bytes, void* stream) {
    GTK_REQUIRE(d != nullptr, "null descriptor");
    return decoder::backward(d, blocks, grads, mask, dz, dlogdet, dx, workspace, workspace_bytes, d->n_blocks - 1, 0, ST);
}
int gtk_decoder_backward_range(const gtk_decoder_desc* d, const gtk_block_params* blocks, const gtk_block_grads* grads,
                               const float* mask, const float* dz, const float* dlogdet, float* dx, void* workspace,
                               size_t workspace_bytes, int blk_hi, int blk_lo, void* stream) {
    return decoder::backward(d, blocks, grads, mask, dz, dlogdet, dx, workspace, workspace_bytes, blk_hi, blk_lo, ST);
}
int gtk_decoder_backward_range_deferred(const gtk_decoder_desc* d, const gtk_block_params* blocks, const gtk_block_grads* grads,
                                        const float* mask, const float* dz, const float* dlogdet, float* dx, void* workspace,
                                        size_t workspace_bytes, int blk_hi, int blk_lo, void* stream) {
    return decoder::backward_deferred(d, blocks, grads, mask, dz, dlogdet, dx, workspace, workspace_bytes, blk_hi, blk_lo, ST);
}
int gtk_decoder_backward_join(void* stream) { return decoder::backward_join_stream(ST); }

}  // extern "C"
