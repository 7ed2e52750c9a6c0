// glowtts_b200/csrc/decoder_bwd.inl -- backward of decoder::forward(train=1); included by decoder.cu.
//
// Per block, in reverse order (names as in decoder.cu / attentions.py:132-186):
//   affine_bwd : dZ -> d(end_out), dx1
//   end        : dW_end, db_end ; dskip = W_end^T d(end_out), masked
//   layer i=L-1..0:
//       drs   = (i == L-1) ? dskip : cat(dh_{i+1} * mask, dskip)
//       res_skip : dW, db ; dacts = W_rs^T drs
//       gate_bwd : dacts -> dxin ; dcond += sum_t dxin
//       in_layer : dW, db ; dh_i = conv^T(dxin) + (i < L-1 ? dh_{i+1} * mask : 0)
//   start      : dW, db ; dx0 = dZ0 + W_start^T (dh_0 * mask)
//   mix_bwd    : d(ActNorm, InvConvNear) ; dX_prev ; + the parameter-only logdet terms

static int wgrad(const float* dout, size_t dout_bs, const float* in, size_t in_bs, const float* mask, float* dw, int B,
                 int Cin, int M, int T, int K, int dil, int mask_dout, int mask_in, cudaStream_t st) {
    conv::WgradParams p;
    p.dout = dout; p.in = in; p.mask = mask; p.dw = dw;
    p.B = B; p.Cin = Cin; p.M = M; p.T = T; p.dil = dil;
    p.mask_dout = mask_dout; p.mask_in = mask_in;
    p.dout_bstride = dout_bs; p.in_bstride = in_bs;
    return conv::wgrad_f32(p, K, st);
}

// Weight gradient of a (possibly weight-normed) conv: into (dg, dv) through dweff scratch, or straight into dv.
static int wgrad_param(const float* g, const float* v, float* dg, float* dv, float* dweff, const float* dout, size_t dout_bs,
                       const float* in, size_t in_bs, const float* mask, int B, int Cin, int M, int T, int K, int dil,
                       int mask_dout, cudaStream_t st) {
    if (!dv) return GTK_OK;
    if (g) {
        GTK_REQUIRE(dg != nullptr, "weight_g gradient pointer missing");
        GTK_CUDA(cudaMemsetAsync(dweff, 0, sizeof(float) * (size_t)M * Cin * K, st));
        GTK_TRY(wgrad(dout, dout_bs, in, in_bs, mask, dweff, B, Cin, M, T, K, dil, mask_dout, 0, st));
        return flow::weightnorm_bwd(g, v, dweff, dg, dv, M, Cin * K, 1, st);
    }
    return wgrad(dout, dout_bs, in, in_bs, mask, dv, B, Cin, M, T, K, dil, mask_dout, 0, st);
}

// Blocks blk_hi .. blk_lo (inclusive, descending).  The running gradient dZ lives in the workspace between calls, so the
// whole backward may be issued as consecutive ranges (NB-1 .. a, a-1 .. b, ..., .. 0): the gradients of a finished range
// are final and can be all-reduced while the next range computes.  The first range (blk_hi == NB-1) reads dz, the last
// (blk_lo == 0) writes dx.
int backward(const gtk_decoder_desc* d, const gtk_block_params* blocks, const gtk_block_grads* grads, const float* mask,
             const float* dz, const float* dlogdet, float* dx, void* workspace, size_t ws_bytes, int blk_hi, int blk_lo,
             cudaStream_t st) {
    Dims D;
    GTK_TRY(make_dims(d, &D));
    GTK_REQUIRE(blocks && grads && mask && dz && dlogdet && dx && workspace, "null pointer");
    GTK_REQUIRE(blk_lo >= 0 && blk_hi >= blk_lo && blk_hi < D.NB, "block range");
    const bool first = blk_hi == D.NB - 1, last = blk_lo == 0;
    if (d->precision == 1)
        return tc_backward(d, D, blocks, grads, mask, dz, dlogdet, dx, workspace, ws_bytes, blk_hi, blk_lo, st);
    if (d->precision != 0) { set_error("decoder: precision %d not available in this build", d->precision); return GTK_ERR_UNSUPPORTED; }
    Layout L;
    make_layout(D, 1, &L);
    if (ws_bytes < L.total * sizeof(float)) { set_error("decoder_backward: workspace %zu < %zu", ws_bytes, L.total * sizeof(float)); return GTK_ERR_WORKSPACE; }
    if (D.B == 0 || D.T == 0) return GTK_OK;
    if (first && (D.Tq * D.n != D.T || D.Tq == 0)) GTK_CUDA(cudaMemsetAsync(dx, 0, sizeof(float) * (size_t)D.B * D.Cin * D.T, st));
    if (D.Tq == 0) return GTK_OK;

    float* ws = static_cast<float*>(workspace);
    const float* msq = ws + L.mask_sq;
    const float* x_len = ws + L.x_len;
    const int B = D.B, T = D.Tq, H = D.H, C = D.C, Ch = D.C / 2;
    const size_t xsz = (size_t)C * D.nq, cbs = (size_t)C * T;
    float* dZ = ws + L.b_dxa;       // gradient w.r.t. the current block's output
    float* dM = ws + L.b_dxb;       // gradient w.r.t. xmix, then reused as next dZ
    float* dend = ws + L.b_dend;
    float* dskip = ws + L.b_dskip;
    float* dh_next = ws + L.b_dha;
    float* dh_cur = ws + L.b_dhb;
    float* drs_buf = ws + L.b_drs;
    float* dacts = ws + L.b_dacts;
    float* dxin = ws + L.b_dxin;
    float* dweff = ws + L.b_dweff;
    float* wp = ws + L.wp;

    // unsqueeze^T: dz [B, Cin, Tq*n] -> dZ [B, C, Tq], masked by the squeezed mask
    if (first) GTK_TRY(flow::squeeze(dz, msq, dZ, nullptr, B, D.Cin, D.Tq * D.n, D.n, 1, st));

    for (int blk = blk_hi; blk >= blk_lo; --blk) {
        const gtk_block_params& P = blocks[blk];
        const gtk_block_grads& G = grads[blk];
        const float* we = ws + L.weff + (size_t)blk * L.weff_block;
        const float* act = ws + L.act + (size_t)blk * L.act_block;
        const float* xin_blk = ws + L.X + (size_t)blk * xsz;
        const float* winv = ws + L.winv + (size_t)blk * 64;
        const float* xmix = act + L.a_xmix;

        GTK_TRY(flow::affine_bwd(xmix, act + L.a_end, msq, dZ, dlogdet, dend, dM, B, C, T, d->sigmoid_scale, st));
        // end conv (plain weight, attentions.py:105-110)
        if (G.end_w) GTK_TRY(wgrad(dend, 0, act + L.a_skip, 0, msq, G.end_w, B, H, C, T, 1, 1, 0, 0, st));
        if (G.end_b) GTK_TRY(flow::channel_sum(dend, nullptr, G.end_b, B, C, T, 1, st));
        GTK_TRY(conv::pack_weight(P.end_w, wp, C, H, 1, 1, st));
        GTK_TRY(conv(dend, 0, wp, nullptr, msq, nullptr, 0, dskip, 0, B, C, H, T, 1, 1, 1, 0, 0, st));

        bool have_dh = false;
        for (int i = D.L - 1; i >= 0; --i) {
            const int RS = (i < D.L - 1) ? 2 * H : H;
            const int dil = ipow(D.dr, i);
            const float* drs = dskip;
            if (i < D.L - 1) {
                GTK_TRY(flow::cat_mask(dh_next, dskip, msq, drs_buf, B, H, H, T, 0, st));
                drs = drs_buf;
            }
            GTK_TRY(wgrad_param(P.rs_g[i], P.rs_v[i], G.rs_g[i], G.rs_v[i], dweff, drs, 0, act + L.a_acts[i], 0, msq, B, H, RS,
                                T, 1, 1, 0, st));
            if (G.rs_b[i]) GTK_TRY(flow::channel_sum(drs, nullptr, G.rs_b[i], B, RS, T, 1, st));
            GTK_TRY(conv::pack_weight(w_rs(L, P, we, i), wp, RS, H, 1, 1, st));
            GTK_TRY(conv(drs, 0, wp, nullptr, msq, nullptr, 0, dacts, 0, B, RS, H, T, 1, 1, 0, 0, 0, st));
            const float* cond_i = P.cond ? P.cond + (size_t)i * 2 * H : nullptr;
            const DropSeed lseed = make_drop_seed(d->seed, d->seed_dev, blk, i);
            if (P.cond && G.cond) {
                // the conditioning enters after the dropout: its gradient is the un-scaled gate gradient
                if (d->p_dropout > 0.f)
                    GTK_TRY(flow::gate_bwd(act + L.a_xin[i], cond_i, 1, (size_t)2 * H * D.L, dacts, dxin, B, H, T, d->p_dropout,
                                           lseed, 0, st));
                else
                    GTK_TRY(flow::gate_bwd(act + L.a_xin[i], cond_i, 1, (size_t)2 * H * D.L, dacts, dxin, B, H, T, 0.f, DropSeed{0ull, nullptr}, 1, st));
                GTK_TRY(flow::row_sum(dxin, G.cond + (size_t)i * 2 * H, B * 2 * H, T, 1, 2 * H, (size_t)2 * H * D.L, st));
            }
            if (!(P.cond && G.cond) || d->p_dropout > 0.f)
                GTK_TRY(flow::gate_bwd(act + L.a_xin[i], cond_i, P.cond ? 1 : 0, (size_t)2 * H * D.L, dacts, dxin, B, H, T,
                                       d->p_dropout, lseed, 1, st));
            GTK_TRY(wgrad_param(P.in_g[i], P.in_v[i], G.in_g[i], G.in_v[i], dweff, dxin, 0, act + L.a_h[i], 0, msq, B, H, 2 * H,
                                T, D.K, dil, 0, st));
            if (G.in_b[i]) GTK_TRY(flow::channel_sum(dxin, nullptr, G.in_b[i], B, 2 * H, T, 1, st));
            GTK_TRY(conv::pack_weight(w_in(L, P, we, i), wp, 2 * H, H, D.K, 1, st));
            GTK_TRY(conv(dxin, 0, wp, nullptr, msq, have_dh ? dh_next : nullptr, 0, dh_cur, 0, B, 2 * H, H, T, D.K, dil, 0, 1, 0,
                         st));
            float* t = dh_next; dh_next = dh_cur; dh_cur = t;
            have_dh = true;
        }
        // start conv: h0 = (W x0 + b) * mask
        GTK_TRY(wgrad_param(P.start_g, P.start_v, G.start_g, G.start_v, dweff, dh_next, 0, xmix, cbs, msq, B, Ch, H, T, 1, 1, 1,
                            st));
        if (G.start_b) GTK_TRY(flow::channel_sum(dh_next, msq, G.start_b, B, H, T, 1, st));
        GTK_TRY(conv::pack_weight(w_start(L, P, we), wp, H, Ch, 1, 1, st));
        GTK_TRY(conv(dh_next, 0, wp, nullptr, msq, dZ, cbs, dM, cbs, B, H, Ch, T, 1, 1, 0, 0, 1, st));
        // ActNorm + InvConvNear
        GTK_TRY(flow::mix_bwd(xin_blk, msq, P.an_logs, P.an_bias, P.ic_weight, D.NS, dM, dZ, G.an_logs, G.an_bias, G.ic_weight,
                              B, C, T, st));
        GTK_TRY(flow::logdet_const_bwd(dlogdet, x_len, B, G.an_logs, C, winv, D.NS, (float)(C / D.NS), G.ic_weight, st));
    }
    // squeeze^T: dZ [B, C, Tq] -> dx [B, Cin, T] (row length T; an odd tail frame keeps its zero)
    if (last) GTK_TRY(flow::unsqueeze(dZ, msq, dx, nullptr, B, C, T, D.n, D.T, st));
    return GTK_OK;
}
