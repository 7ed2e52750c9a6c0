// glowtts_b200/csrc/decoder.cu -- FlowSpecDecoder executor (models.py:719-789): one C call enqueues the
// whole 12 x (ActNorm, InvConvNear, CouplingBlock) stack forward, reverse or backward on a stream.
//
// Design notes (DESIGN.md has the long form):
//  * ActNorm and InvConvNear are ONE memory-bound kernel (flow::mix) - no permute copies, no cuSOLVER.
//  * Activations needed by the backward are kept in the caller's workspace (180 GB HBM: store, don't recompute).
//  * Weight norm is folded on device once per call; its backward is fused behind each weight gradient.
//  * precision 0 runs every contraction in fp32 on CUDA cores (exact-parity path);
//    precision 1 routes the WN / start / end contractions to the bf16 tcgen05 kernels (conv_tc.cu).
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "tc.cuh"
#include "ops.h"
#include "../../include/glowtts_b200.h"

namespace gtk {
namespace decoder {

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct Dims {
    int B, Cin, T, n, C, Tq, H, K, L, NB, NS, dr;
    size_t nq;   // B * Tq
    int n_wn, wn_id[3];   // WaveNets chained in every coupling block (attentions.py:155-160): 0 = wn, 1 = wn_energy, 2 = wn_pitch
};

static int make_dims(const gtk_decoder_desc* d, Dims* o) {
    GTK_REQUIRE(d != nullptr, "null desc");
    GTK_REQUIRE(d->B >= 0 && d->T >= 0 && d->in_channels > 0, "bad shape");
    GTK_REQUIRE(d->n_sqz >= 1, "n_sqz >= 1");
    GTK_REQUIRE(d->n_layers >= 1 && d->n_layers <= GTK_MAX_LAYERS, "n_layers out of range");
    GTK_REQUIRE(d->kernel_size % 2 == 1, "kernel_size must be odd (modules.py:108)");
    GTK_REQUIRE(d->hidden % 2 == 0, "hidden_channels must be even (modules.py:109)");
    GTK_REQUIRE(d->n_split == 2 || d->n_split == 4 || d->n_split == 8, "n_split must be 2, 4 or 8");
    GTK_REQUIRE(d->p_dropout >= 0.f && d->p_dropout < 1.f, "p_dropout must be in [0, 1)");
    o->B = d->B; o->Cin = d->in_channels; o->T = d->T; o->n = d->n_sqz;
    o->C = d->in_channels * d->n_sqz; o->Tq = d->T / d->n_sqz;
    o->H = d->hidden; o->K = d->kernel_size; o->L = d->n_layers; o->NB = d->n_blocks; o->NS = d->n_split;
    o->dr = d->dilation_rate;
    o->nq = (size_t)o->B * o->Tq;
    GTK_REQUIRE(o->C % o->NS == 0 && o->C % 2 == 0, "channels % n_split != 0 (modules.py:637)");
    o->n_wn = 1; o->wn_id[0] = 0; o->wn_id[1] = o->wn_id[2] = -1;
    if (d->energy || d->pitch) {
        GTK_REQUIRE(d->precision == 1, "pitch / energy conditioned blocks run on the bf16 path (precision 1); the fp32 path composes them from the module-level ops");
        GTK_REQUIRE((2 * o->H * o->L) % (16 * o->n) == 0, "2 * hidden * n_layers / n_sqz must be a multiple of 16");
        GTK_REQUIRE(o->n <= 4, "n_sqz <= 4 with pitch / energy tracks");
        if (d->energy) o->wn_id[o->n_wn++] = 1;
        if (d->pitch) o->wn_id[o->n_wn++] = 2;
    }
    return GTK_OK;
}

// Workspace carve-up, in floats.  `train` keeps one activation set per block; otherwise one set is reused.
struct Layout {
    size_t mask_sq, x_len, winv, icld, weff, wp, X, act, rs, bwd, total;
    size_t weff_block, act_block;
    // inside one block's effective-weight set
    size_t we_start, we_in[GTK_MAX_LAYERS], we_rs[GTK_MAX_LAYERS];
    // inside one block's activation set
    size_t a_xmix, a_h[GTK_MAX_LAYERS], a_xin[GTK_MAX_LAYERS], a_acts[GTK_MAX_LAYERS], a_skip, a_end;
    // backward temporaries
    size_t b_dxa, b_dxb, b_dend, b_dskip, b_dha, b_dhb, b_drs, b_dacts, b_dxin, b_dweff;
    int n_x, n_act;
};

static void make_layout(const Dims& D, int train, Layout* L) {
    size_t off = 0;
    auto take = [&](size_t n) { size_t o = off; off = align_up(off + n, 64); return o; };
    const size_t nq = D.nq;
    L->mask_sq = take(nq);
    L->x_len = take(D.B);
    L->winv = take((size_t)D.NB * 64);
    L->icld = take(D.NB);
    // effective weights of one block
    size_t w = 0;
    auto wtake = [&](size_t n) { size_t o = w; w = align_up(w + n, 64); return o; };
    L->we_start = wtake((size_t)D.H * (D.C / 2));
    for (int i = 0; i < D.L; ++i) L->we_in[i] = wtake((size_t)2 * D.H * D.H * D.K);
    for (int i = 0; i < D.L; ++i) L->we_rs[i] = wtake((size_t)((i < D.L - 1) ? 2 * D.H : D.H) * D.H);
    L->weff_block = w;
    L->weff = take(w * D.NB);
    size_t wmax = (size_t)2 * D.H * D.H * D.K;
    if ((size_t)D.C * D.H > wmax) wmax = (size_t)D.C * D.H;
    L->wp = take(wmax);
    L->n_x = train ? D.NB + 1 : 2;
    L->X = take((size_t)L->n_x * D.C * nq);
    size_t a = 0;
    auto atake = [&](size_t n) { size_t o = a; a = align_up(a + n, 64); return o; };
    L->a_xmix = atake((size_t)D.C * nq);
    for (int i = 0; i < D.L; ++i) L->a_h[i] = atake((size_t)D.H * nq);
    for (int i = 0; i < D.L; ++i) L->a_xin[i] = atake((size_t)2 * D.H * nq);
    for (int i = 0; i < D.L; ++i) L->a_acts[i] = atake((size_t)D.H * nq);
    L->a_skip = atake((size_t)D.H * nq);
    L->a_end = atake((size_t)D.C * nq);
    L->act_block = a;
    L->n_act = train ? D.NB : 1;
    L->act = take(a * L->n_act);
    L->rs = take((size_t)2 * D.H * nq);
    L->bwd = off;
    if (train) {
        L->b_dxa = take((size_t)D.C * nq);
        L->b_dxb = take((size_t)D.C * nq);
        L->b_dend = take((size_t)D.C * nq);
        L->b_dskip = take((size_t)D.H * nq);
        L->b_dha = take((size_t)D.H * nq);
        L->b_dhb = take((size_t)D.H * nq);
        L->b_drs = take((size_t)2 * D.H * nq);
        L->b_dacts = take((size_t)D.H * nq);
        L->b_dxin = take((size_t)2 * D.H * nq);
        L->b_dweff = take(wmax);
    }
    L->total = off;
}

struct TcLayout;
static size_t tc_workspace_bytes(const Dims& D, int train);

size_t workspace_bytes(const gtk_decoder_desc* d, int train) {
    Dims D;
    if (make_dims(d, &D) != GTK_OK) return 0;
    if (d->precision == 1) return tc_workspace_bytes(D, train);
    Layout L;
    make_layout(D, train, &L);
    return L.total * sizeof(float) + 256;
}

#define GTK_TRY(expr)            \
    do {                         \
        int rc__ = (expr);       \
        if (rc__ != GTK_OK) return rc__; \
    } while (0)

static int conv(const float* in, size_t in_bs, const float* wp, const float* bias, const float* mask, const float* resid,
                size_t resid_bs, float* out, size_t out_bs, int B, int Cin, int M, int T, int K, int dil, int mask_out,
                int mask_resid, int mask_in, cudaStream_t st) {
    conv::ConvParams p;
    p.in = in; p.wp = wp; p.bias = bias; p.mask = mask; p.resid = resid; p.out = out;
    p.B = B; p.Cin = Cin; p.M = M; p.T = T; p.K = K; p.dil = dil;
    p.mask_out = mask_out; p.mask_resid = mask_resid; p.mask_in = mask_in;
    p.in_bstride = in_bs; p.out_bstride = out_bs; p.resid_bstride = resid_bs;
    return conv::conv1d_f32(p, st);
}

static int ipow(int b, int e) { int r = 1; while (e-- > 0) r *= b; return r; }

// Effective (weight-norm folded) weights of one block -> we[...]
static int fold_weights(const Dims& D, const Layout& L, const gtk_block_params& P, float* we, cudaStream_t st) {
    const int Ch = D.C / 2;
    if (P.start_g) GTK_TRY(flow::weightnorm_fwd(P.start_g, P.start_v, we + L.we_start, nullptr, D.H, Ch, st));
    for (int i = 0; i < D.L; ++i) {
        if (P.in_g[i]) GTK_TRY(flow::weightnorm_fwd(P.in_g[i], P.in_v[i], we + L.we_in[i], nullptr, 2 * D.H, D.H * D.K, st));
        const int RS = (i < D.L - 1) ? 2 * D.H : D.H;
        if (P.rs_g[i]) GTK_TRY(flow::weightnorm_fwd(P.rs_g[i], P.rs_v[i], we + L.we_rs[i], nullptr, RS, D.H, st));
    }
    return GTK_OK;
}
static inline const float* w_start(const Layout& L, const gtk_block_params& P, const float* we) {
    return P.start_g ? we + L.we_start : P.start_v;
}
static inline const float* w_in(const Layout& L, const gtk_block_params& P, const float* we, int i) {
    return P.in_g[i] ? we + L.we_in[i] : P.in_v[i];
}
static inline const float* w_rs(const Layout& L, const gtk_block_params& P, const float* we, int i) {
    return P.rs_g[i] ? we + L.we_rs[i] : P.rs_v[i];
}

static int check_block(const Dims& D, const gtk_block_params& P) {
    GTK_REQUIRE(P.an_logs && P.an_bias && P.ic_weight && P.start_v && P.start_b && P.end_w && P.end_b, "null block parameter");
    for (int i = 0; i < D.L; ++i) GTK_REQUIRE(P.in_v[i] && P.in_b[i] && P.rs_v[i] && P.rs_b[i], "null WN layer parameter");
    return GTK_OK;
}

// start -> WN -> end of one coupling block on xc[:, :C/2]; leaves end_out in act.a_end (attentions.py:151-170).
static int coupling_net(const Dims& D, const Layout& L, const gtk_block_params& P, const float* we, const float* xc,
                        const float* mask, float* act, float* rs, float* wp, const gtk_decoder_desc* d, int blk,
                        cudaStream_t st) {
    const int B = D.B, T = D.Tq, H = D.H, C = D.C, Ch = D.C / 2;
    const size_t cbs = (size_t)C * T;
    GTK_TRY(conv::pack_weight(w_start(L, P, we), wp, H, Ch, 1, 0, st));
    GTK_TRY(conv(xc, cbs, wp, P.start_b, mask, nullptr, 0, act + L.a_h[0], 0, B, Ch, H, T, 1, 1, 1, 0, 0, st));
    for (int i = 0; i < D.L; ++i) {
        const int RS = (i < D.L - 1) ? 2 * H : H;
        GTK_TRY(conv::pack_weight(w_in(L, P, we, i), wp, 2 * H, H, D.K, 0, st));
        GTK_TRY(conv(act + L.a_h[i], 0, wp, P.in_b[i], mask, nullptr, 0, act + L.a_xin[i], 0, B, H, 2 * H, T, D.K,
                     ipow(D.dr, i), 0, 0, 0, st));
        GTK_TRY(flow::gate_fwd(act + L.a_xin[i], P.cond ? P.cond + (size_t)i * 2 * H : nullptr, P.cond ? 1 : 0,
                               (size_t)2 * H * D.L, act + L.a_acts[i], B, H, T, d->p_dropout, make_drop_seed(d->seed, d->seed_dev, blk, i), st));
        GTK_TRY(conv::pack_weight(w_rs(L, P, we, i), wp, RS, H, 1, 0, st));
        GTK_TRY(conv(act + L.a_acts[i], 0, wp, P.rs_b[i], mask, nullptr, 0, rs, 0, B, H, RS, T, 1, 1, 0, 0, 0, st));
        GTK_TRY(flow::resskip(rs, act + L.a_h[i], mask, (i < D.L - 1) ? act + L.a_h[i + 1] : nullptr, act + L.a_skip, B, H, RS,
                              T, i == 0, i == D.L - 1, st));
    }
    GTK_TRY(conv::pack_weight(P.end_w, wp, C, H, 1, 0, st));
    GTK_TRY(conv(act + L.a_skip, 0, wp, P.end_b, mask, nullptr, 0, act + L.a_end, 0, B, H, C, T, 1, 1, 0, 0, 0, st));
    return GTK_OK;
}

// Internal side stream of the bf16 backward: the weight gradients of block k (one persistent stream-K launch + reduce) run
// on it while the data-gradient chain of block k - 1 runs on the caller's stream.  The chain launches fewer CTAs than the
// GPU has SMs (one 128-row tile each), so the two fill the machine together.  Created once per device, lowest priority (the
// chain is the critical path); joined back into the caller's stream before the backward call returns, so the call stays
// stream-ordered for the caller -- and capturable: the side stream enters a CUDA-graph capture through the fork event
// and leaves it through the join.  Three modes (GTK_WGRAD_OVERLAP): 2 = the default, "semi" overlap: the weight gradients of
// block k share the GPU only with the small memory-bound kernels between two blocks (mix_bwd of block k, affine_bwd_frames of
// block k - 1) and the main stream joins before the first tensor-core kernel of block k - 1: 4.28 -> 4.20 ms per step.
// 1 = full overlap with the next block's chain: measured on B200 at configs[1] it LOSES -- 5.29 ms / step against 4.75 ms
// (profiles/wgrad_overlap_r02.txt; re-measured 4.60-4.96 vs 4.40) -- the stream-K weight gradient takes the idle SMs the
// chain's next kernel needs for its early-resident CTAs, and both contend in L2.  0 = everything on the caller's stream.
struct SideStream {
    cudaStream_t s;
    cudaEvent_t fork[2], done[2];
    // deferred join of a backward RANGE (gtk_decoder_backward_range_deferred): the call leaves its last weight-gradient launch
    // and its reduce running on the side stream; the next range picks the state up where an un-split backward would be
    cudaEvent_t red_done;                          // behind the reduce of the last deferred range
    bool carry, carry_used[2], red_pending;
};
static thread_local int g_defer_join = 0;          // set by backward_range_deferred() for the duration of one call
static int side_stream(SideStream** out) {
    constexpr int MAX_DEV = 64;
    static SideStream* tab[MAX_DEV] = {nullptr};
    static std::mutex mu;
    *out = nullptr;
    static int enabled = -1;
    if (enabled < 0) { const char* e = getenv("GTK_WGRAD_OVERLAP"); enabled = e ? ((e[0] == '1' || e[0] == '2') ? (e[0] - '0') : 0) : 2; }
    if (!enabled) return GTK_OK;
    int dev = 0;
    GTK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEV) return GTK_OK;
    std::lock_guard<std::mutex> lock(mu);
    if (!tab[dev]) {
        SideStream* ss = new SideStream();
        int lo = 0, hi = 0;
        GTK_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));          // lo = least priority (numerically greatest)
        GTK_CUDA(cudaStreamCreateWithPriority(&ss->s, cudaStreamNonBlocking, lo));
        for (int i = 0; i < 2; ++i) {
            GTK_CUDA(cudaEventCreateWithFlags(&ss->fork[i], cudaEventDisableTiming));
            GTK_CUDA(cudaEventCreateWithFlags(&ss->done[i], cudaEventDisableTiming));
        }
        GTK_CUDA(cudaEventCreateWithFlags(&ss->red_done, cudaEventDisableTiming));
        ss->carry = false; ss->carry_used[0] = ss->carry_used[1] = false; ss->red_pending = false;
        tab[dev] = ss;
    }
    *out = tab[dev];
    return GTK_OK;
}
// Auxiliary stream of the bf16 forward / reverse: the weight-norm fold + bf16 pack of blocks 1 .. NB-1 runs on it while the
// caller's stream packs block 0, builds the packing plan and runs block 0 -- the packs are ~70 us of one launch that only
// block k >= 1 needs.  Forked from and joined back into the caller's stream with events (capturable, stream-ordered for the
// caller).  GTK_AUX_STREAM=0 keeps everything on the caller's stream.
static int aux_stream(SideStream** out) {
    constexpr int MAX_DEV = 64;
    static SideStream* tab[MAX_DEV] = {nullptr};
    static std::mutex mu;
    *out = nullptr;
    static int enabled = -1;
    if (enabled < 0) { const char* e = getenv("GTK_AUX_STREAM"); enabled = (e && e[0] == '0') ? 0 : 1; }
    if (!enabled) return GTK_OK;
    int dev = 0;
    GTK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEV) return GTK_OK;
    std::lock_guard<std::mutex> lock(mu);
    if (!tab[dev]) {
        SideStream* ss = new SideStream();
        GTK_CUDA(cudaStreamCreateWithFlags(&ss->s, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) {
            GTK_CUDA(cudaEventCreateWithFlags(&ss->fork[i], cudaEventDisableTiming));
            GTK_CUDA(cudaEventCreateWithFlags(&ss->done[i], cudaEventDisableTiming));
        }
        tab[dev] = ss;
    }
    *out = tab[dev];
    return GTK_OK;
}
// GTK_WGRAD_OVERLAP=2: "semi" overlap -- the weight gradients of block k run on the side stream only while the main stream
// runs the small memory-bound kernels between two blocks (mix_bwd of block k, affine_bwd_frames of block k - 1); the main
// stream joins before the first tensor-core kernel of block k - 1.
static bool wgrad_semi_overlap() {
    const char* e = getenv("GTK_WGRAD_OVERLAP");
    return !e || e[0] == '2';
}
static int wgrad_overlap_ctas() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("GTK_WGRAD_CTAS");
        v = e ? atoi(e) : 0;
        if (v <= 0) v = max(16, sm_count() * 2 / 5);     // ~60 of 148: what a ragged B=32 batch's 128-row tiles leave idle
    }
    return v;
}

#include "decoder_tc.inl"

// gtk_decoder_backward_join: `stream` waits for what a deferred range left on the side stream (its weight gradients + reduce)
static int backward_join(cudaStream_t stream) {
    SideStream* side = nullptr;
    GTK_TRY(side_stream(&side));
    if (side && side->red_pending) GTK_CUDA(cudaStreamWaitEvent(stream, side->red_done, 0));
    return GTK_OK;
}

static size_t tc_workspace_bytes(const Dims& D, int train) {
    Layout L32;
    make_layout(D, 0, &L32);
    TcLayout L;
    make_tc_layout(D, L32, train, &L);
    return L.total + 1024;
}

int forward(const gtk_decoder_desc* d, const gtk_block_params* blocks, const float* x, const float* mask, float* z,
            float* logdet, void* workspace, size_t ws_bytes, int train, cudaStream_t st) {
    Dims D;
    GTK_TRY(make_dims(d, &D));
    GTK_REQUIRE(blocks && x && mask && z && logdet && workspace, "null pointer");
    if (d->precision == 1) return tc_forward(d, D, blocks, x, mask, z, logdet, workspace, ws_bytes, train, st);
    if (d->precision != 0) { set_error("decoder: precision %d not available in this build", d->precision); return GTK_ERR_UNSUPPORTED; }
    Layout L;
    make_layout(D, train, &L);
    if (ws_bytes < L.total * sizeof(float)) { set_error("decoder_forward: workspace %zu < %zu", ws_bytes, L.total * sizeof(float)); return GTK_ERR_WORKSPACE; }
    if (D.B == 0) return GTK_OK;
    GTK_CUDA(cudaMemsetAsync(logdet, 0, sizeof(float) * D.B, st));
    if (D.Tq == 0) return GTK_OK;
    float* ws = static_cast<float*>(workspace);
    float* msq = ws + L.mask_sq;
    const size_t xsz = (size_t)D.C * D.nq;
    GTK_TRY(flow::squeeze(x, mask, ws + L.X, msq, D.B, D.Cin, D.T, D.n, 0, st));
    GTK_TRY(flow::mask_len(msq, ws + L.x_len, D.B, D.Tq, st));
    for (int blk = 0; blk < D.NB; ++blk) {
        const gtk_block_params& P = blocks[blk];
        GTK_TRY(check_block(D, P));
        float* we = ws + L.weff + (size_t)blk * L.weff_block;
        float* act = ws + L.act + (size_t)(train ? blk : 0) * L.act_block;
        const float* xin = ws + L.X + (size_t)(train ? blk : (blk & 1)) * xsz;
        float* xout = ws + L.X + (size_t)(train ? blk + 1 : ((blk + 1) & 1)) * xsz;
        float* winv = ws + L.winv + (size_t)blk * 64;
        float* icld = ws + L.icld + blk;
        GTK_TRY(fold_weights(D, L, P, we, st));
        GTK_TRY(flow::invconv_prep(P.ic_weight, D.NS, winv, icld, st));
        GTK_TRY(flow::mix(xin, msq, P.an_logs, P.an_bias, P.ic_weight, D.NS, act + L.a_xmix, D.B, D.C, D.Tq, 0, st));
        GTK_TRY(flow::logdet_const(P.an_logs, D.C, icld, (float)(D.C / D.NS), ws + L.x_len, logdet, D.B, 1, st));
        GTK_TRY(coupling_net(D, L, P, we, act + L.a_xmix, msq, act, ws + L.rs, ws + L.wp, d, blk, st));
        GTK_TRY(flow::affine(act + L.a_xmix, act + L.a_end, msq, xout, logdet, D.B, D.C, D.Tq, 0, d->sigmoid_scale, st));
    }
    const float* xfin = ws + L.X + (size_t)(train ? D.NB : (D.NB & 1)) * xsz;
    GTK_TRY(flow::unsqueeze(xfin, msq, z, nullptr, D.B, D.C, D.Tq, D.n, 0, st));
    return GTK_OK;
}

int reverse(const gtk_decoder_desc* d, const gtk_block_params* blocks, const float* z, const float* mask, float* x,
            void* workspace, size_t ws_bytes, cudaStream_t st) {
    Dims D;
    GTK_TRY(make_dims(d, &D));
    GTK_REQUIRE(blocks && z && mask && x && workspace, "null pointer");
    if (d->precision == 1) return tc_reverse(d, D, blocks, z, mask, x, workspace, ws_bytes, st);
    if (d->precision != 0) { set_error("decoder: precision %d not available in this build", d->precision); return GTK_ERR_UNSUPPORTED; }
    Layout L;
    make_layout(D, 0, &L);
    if (ws_bytes < L.total * sizeof(float)) { set_error("decoder_reverse: workspace %zu < %zu", ws_bytes, L.total * sizeof(float)); return GTK_ERR_WORKSPACE; }
    if (D.B == 0 || D.Tq == 0) return GTK_OK;
    float* ws = static_cast<float*>(workspace);
    float* msq = ws + L.mask_sq;
    const size_t xsz = (size_t)D.C * D.nq;
    float* xa = ws + L.X;
    float* xb = ws + L.X + xsz;
    float* act = ws + L.act;
    float* we = ws + L.weff;
    GTK_TRY(flow::squeeze(z, mask, xa, msq, D.B, D.Cin, D.T, D.n, 0, st));
    for (int blk = D.NB - 1; blk >= 0; --blk) {
        const gtk_block_params& P = blocks[blk];
        GTK_TRY(check_block(D, P));
        float* winv = ws + L.winv + (size_t)blk * 64;
        GTK_TRY(fold_weights(D, L, P, we, st));
        GTK_TRY(flow::invconv_prep(P.ic_weight, D.NS, winv, nullptr, st));
        GTK_TRY(coupling_net(D, L, P, we, xa, msq, act, ws + L.rs, ws + L.wp, d, blk, st));
        GTK_TRY(flow::affine(xa, act + L.a_end, msq, xb, nullptr, D.B, D.C, D.Tq, 1, d->sigmoid_scale, st));
        GTK_TRY(flow::mix(xb, msq, P.an_logs, P.an_bias, winv, D.NS, xa, D.B, D.C, D.Tq, 1, st));
    }
    GTK_TRY(flow::unsqueeze(xa, msq, x, nullptr, D.B, D.C, D.Tq, D.n, 0, st));
    return GTK_OK;
}

#include "decoder_bwd.inl"

// backward() of a range that is not the last one, without the join of the internal side stream at its end (bf16 path with
// the default side-stream weight gradients; any other configuration joins as backward() does)
int backward_deferred(const gtk_decoder_desc* d, const gtk_block_params* blocks, const gtk_block_grads* grads, const float* mask,
                      const float* dz, const float* dlogdet, float* dx, void* workspace, size_t ws_bytes, int blk_hi, int blk_lo,
                      cudaStream_t st) {
    g_defer_join = 1;
    const int rc = backward(d, blocks, grads, mask, dz, dlogdet, dx, workspace, ws_bytes, blk_hi, blk_lo, st);
    g_defer_join = 0;
    return rc;
}
int backward_join_stream(cudaStream_t st) { return backward_join(st); }

}  // namespace decoder
}  // namespace gtk
