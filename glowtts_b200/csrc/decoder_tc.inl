// glowtts_b200/csrc/decoder_tc.inl -- precision 1: the coupling networks on the bf16 tcgen05 kernels
// (conv_tc.cu, wn_tc.cu); ActNorm / InvConvNear / logdet bookkeeping stay fp32.  Included by decoder.cu.
//
// Per block:  mix (fp32, t-major) -> frames(x0) -> start 1x1 [plain epilogue, mask] -> L fused WN layers ->
//             end 1x1 [affine-coupling epilogue: z1, logdet, x0 copy].  Activations inside the coupling network are
//             frame-major ([B, T', H], channel fastest): bf16 operand copies + an fp32 master of the residual stream.

typedef __nv_bfloat16 bf16;

static inline int pad64(int n) { return (n + 63) / 64 * 64; }

constexpr int TC_MAX_WN = 3;                       // wn, wn_energy, wn_pitch
static inline int LI(int w, int i) { return w * GTK_MAX_LAYERS + i; }      // (chain slot of the WaveNet, layer)
struct TcPack {            // bf16 element offsets inside one block's packed-weight set; per-layer entries indexed LI(w, i)
    size_t start, in[TC_MAX_WN * GTK_MAX_LAYERS], rs[TC_MAX_WN * GTK_MAX_LAYERS], end;          // forward operands
    size_t t_start, t_in[TC_MAX_WN * GTK_MAX_LAYERS], t_rs[TC_MAX_WN * GTK_MAX_LAYERS], t_end;  // data-gradient operands
    size_t total;
};
struct TcAct {             // byte offsets inside one block's activation set
    size_t xmix, a0, h[TC_MAX_WN * GTK_MAX_LAYERS], ts[TC_MAX_WN * GTK_MAX_LAYERS], acts[TC_MAX_WN * GTK_MAX_LAYERS],
        skip_bf[TC_MAX_WN], logs, total;
};
// Parameters / gradient destinations of the WaveNet in chain slot w of a block
struct WnW {
    const float* const* in_g; const float* const* in_v; const float* const* in_b;
    const float* const* rs_g; const float* const* rs_v; const float* const* rs_b;
    const float* c1_g; const float* c1_v; const float* c1_b;
};
struct WnG {
    float* const* in_g; float* const* in_v; float* const* in_b; float* const* rs_g; float* const* rs_v; float* const* rs_b;
    float* c1_g; float* c1_v; float* c1_b;
};
static WnW wn_weights(const gtk_block_params& P, int id) {
    if (id == 0) return WnW{P.in_g, P.in_v, P.in_b, P.rs_g, P.rs_v, P.rs_b, nullptr, nullptr, nullptr};
    const gtk_wnp_params& Q = id == 1 ? P.wn_energy : P.wn_pitch;
    return WnW{Q.in_g, Q.in_v, Q.in_b, Q.rs_g, Q.rs_v, Q.rs_b, Q.c1_g, Q.c1_v, Q.c1_b};
}
static WnG wn_grads(const gtk_block_grads& P, int id) {
    if (id == 0) return WnG{P.in_g, P.in_v, P.in_b, P.rs_g, P.rs_v, P.rs_b, nullptr, nullptr, nullptr};
    const gtk_wnp_grads& Q = id == 1 ? P.wn_energy : P.wn_pitch;
    return WnG{Q.in_g, Q.in_v, Q.in_b, Q.rs_g, Q.rs_v, Q.rs_b, Q.c1_g, Q.c1_v, Q.c1_b};
}
struct TcLayout {          // byte offsets in the workspace
    size_t mask_sq, x_len, winv, icld, we, pack, X, act, hf[2], hx, skipf, bwd, total;
    int stack, st_lead, st_halo, st_out;  // wn_stack.cu covers this WaveNet shape: window geometry
    size_t hdr, off, row_b, row_fr;      // varlen packing plan (pack.cu): {0, n_rows}, off[B+1], row_b[P], row_fr[P]; mask_sq = mask_p[P]
    size_t trk[TC_MAX_WN], dtrk[TC_MAX_WN];   // packed pitch / energy samples [P, n_sqz] of chain slots >= 1, and their gradients
    int G, P;                            // gap rows behind every utterance, packed row capacity (multiple of 128)
    size_t we_floats, x_bytes;
    int n_pack, n_act, n_x;
    TcPack P_;
    TcAct A;
    // backward temporaries
    size_t b_dZ, b_dM, b_dend, b_drs, b_dxin, b_lstride, b_dgate, b_dhf[2], b_dh0, b_wpart;   // b_drs / b_dxin: one per layer, b_lstride apart
    int wpart_cap;                       // stream-K launches whose partial tiles fit b_wpart (deferred reduce)
    size_t b_pstride;                    // b_dend .. b_dh0 exist twice (block parity), b_pstride bytes apart: the weight
                                         // gradients of block k read them on the side stream while block k - 1 writes its own
};

static int tc_supported(const Dims& D) {
    if (D.H % 64 != 0 || D.H > 192) { set_error("bf16 path: hidden_channels must be 64, 128 or 192 (got %d)", D.H); return GTK_ERR_UNSUPPORTED; }
    if ((D.C / 2) % 16 != 0) { set_error("bf16 path: in_channels*n_sqz/2 must be a multiple of 16 (got %d)", D.C / 2); return GTK_ERR_UNSUPPORTED; }
    return GTK_OK;
}

static void make_tc_layout(const Dims& D, const Layout& L32, int train, TcLayout* L) {
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 1024); return o; };
    // Inside the bf16 path the batch is ONE packed sequence of P rows (pack.cu): valid frames of every utterance, each
    // followed by G zero rows (G = the widest halo of any WN layer), dead rows up to P.
    L->G = (D.K / 2) * ipow(D.dr, D.L - 1);
    L->P = (int)(((size_t)(D.B > 0 ? D.B : 1) * (D.Tq + L->G) + 127) / 128 * 128);
    const size_t nq = (size_t)L->P;
    const size_t nqp = nq;                                           // rows of the tile-private buffers (whole 128-frame tiles)
    L->stack = tc::wn_stack_geometry(D.H, D.K, D.dr, D.L, pad64(D.C / 2) / 64, &L->st_lead, &L->st_halo, &L->st_out) ? 1 : 0;
    // window-private fp32 masters of the stack kernel: overlapping 128-row windows, tile_out owned rows each
    const size_t nqm = L->stack ? max(nqp, tc::wn_stack_master_rows(L->P, L->st_out)) : nqp;
    const int H = D.H, C = D.C, Ch = D.C / 2, K = D.K;
    L->mask_sq = take(nq * 4);
    L->x_len = take((size_t)D.B * 4);
    L->hdr = take(16);
    L->off = take((size_t)(D.B + 1) * 4);
    L->row_b = take(nq * 4);
    L->row_fr = take(nq * 4);
    for (int w = 1; w < D.n_wn; ++w) { L->trk[w] = take(nq * D.n * 4); L->dtrk[w] = train ? take(nq * D.n * 4) : 0; }
    L->winv = take((size_t)D.NB * 64 * 4);
    L->icld = take((size_t)D.NB * 4);
    L->we_floats = L32.weff_block;
    L->we = take(L32.weff_block * 4);
    // packed weights
    size_t w = 0;
    auto wtake = [&](size_t n) { size_t o = w; w = align_up(w + n, 512); return o; };
    TcPack& P = L->P_;
    P.start = wtake((size_t)H * pad64(Ch));
    for (int w = 0; w < D.n_wn; ++w) {
        for (int i = 0; i < D.L; ++i) P.in[LI(w, i)] = wtake((size_t)K * 2 * H * H);
        for (int i = 0; i < D.L; ++i) P.rs[LI(w, i)] = wtake((size_t)((i < D.L - 1) ? 2 * H : H) * H);
    }
    P.end = wtake((size_t)C * H);
    P.t_start = wtake((size_t)Ch * H);
    for (int w = 0; w < D.n_wn; ++w) {
        for (int i = 0; i < D.L; ++i) P.t_in[LI(w, i)] = wtake((size_t)K * H * 2 * H);
        for (int i = 0; i < D.L; ++i) P.t_rs[LI(w, i)] = wtake((size_t)H * pad64((i < D.L - 1) ? 2 * H : H));
    }
    P.t_end = wtake((size_t)H * pad64(C));
    P.total = w;
    L->n_pack = D.NB;            // all blocks are packed by one launch up front
    L->pack = take(w * 2 * L->n_pack);
    // block inputs / outputs (fp32, t-major)
    L->x_bytes = (size_t)C * nq * 4;
    L->n_x = train ? D.NB + 1 : 2;
    L->X = take(L->x_bytes * L->n_x);
    // activations
    size_t a = 0;
    auto atake = [&](size_t bytes) { size_t o = a; a = align_up(a + bytes, 1024); return o; };
    TcAct& A = L->A;
    A.xmix = atake((size_t)C * nq * 4);
    A.a0 = atake((size_t)pad64(Ch) * nq * 2);
    for (int w = 0; w < D.n_wn; ++w) {
        // h[w][0] of a chained WaveNet IS the masked output of the one before it (skip_bf[w - 1])
        for (int i = 0; i < D.L; ++i) A.h[LI(w, i)] = atake((size_t)H * nq * 2);
        for (int i = 0; i < D.L; ++i) A.ts[LI(w, i)] = atake((size_t)2 * H * nqp * 2);     // tile-private layout: whole tiles
        for (int i = 0; i < D.L; ++i) A.acts[LI(w, i)] = atake((size_t)H * nq * 2);
        A.skip_bf[w] = atake((size_t)H * nq * 2);
    }
    A.logs = atake((size_t)Ch * nq * 4);
    A.total = a;
    L->n_act = train ? D.NB : 1;
    L->act = take(a * L->n_act);
    L->hf[0] = take((size_t)H * nqm * 4);
    L->hf[1] = take((size_t)H * nqp * 4);
    L->hx = (L->stack && D.n_wn > 2) ? take((size_t)H * nqp * 4) : 0;   // second fp32 exchange buffer between chained WaveNets
    L->skipf = take((size_t)H * nqm * 4);
    L->bwd = off;
    if (train) {
        L->b_dZ = take((size_t)C * nq * 4);
        L->b_dM = take((size_t)C * nq * 4);
        L->b_lstride = align_up((size_t)2 * H * nq * 2, 1024);      // per-layer copies: the grouped weight-gradient
        for (int par = 0; par < 2; ++par) {                          // launch at the end of the block reads them all
            const size_t o0 = off;
            const size_t o_dend = take((size_t)pad64(C) * nq * 2);
            const size_t o_drs = take(L->b_lstride * D.L * D.n_wn);         // layer i of chain slot w: index w * L + i
            const size_t o_dxin = take(L->b_lstride * D.L * D.n_wn);
            const size_t o_dgate = take(L->b_lstride * D.L * D.n_wn);       // per layer too: their column sums run once per block
            const size_t o_dh0 = take((size_t)H * nq * 2);
            if (par == 0) { L->b_dend = o_dend; L->b_drs = o_drs; L->b_dxin = o_dxin; L->b_dgate = o_dgate; L->b_dh0 = o_dh0; }
            else L->b_pstride = o0 - L->b_dend;
        }
        L->b_dhf[0] = take((size_t)H * nqp * 4);
        L->b_dhf[1] = take((size_t)H * nqp * 4);
        // partial tiles of the stream-K weight-gradient launches of one backward call, reduced by ONE launch at its end
        L->wpart_cap = min(tc::WGG_MAX_DEFER, D.NB * D.n_wn);
        L->b_wpart = take(tc::wgrad_group_workspace_bytes() * (size_t)L->wpart_cap);
    }
    L->total = off;
}

// weight-norm fold + bf16 operand packs (zero padding included) of EVERY block in one launch (a second one only
// if the model has more than FOLD_PACK_MAX_JOBS conv weights).
static int tc_pack_all(const Dims& D, const TcLayout& L, const gtk_block_params* blocks, bf16* pk_base, int with_transposed,
                       cudaStream_t st, int blk_lo = 0, int blk_hi = -1) {
    const int H = D.H, C = D.C, Ch = D.C / 2, K = D.K;
    const TcPack& Q = L.P_;
    static thread_local tc::FoldPackJobs J;
    J.n = 0;
    bf16* pk = pk_base;
    auto add = [&](const float* g, const float* v, size_t fwd, size_t tr, int M, int Cin, int Kk, int cols_f, int cols_t) {
        tc::FoldPackJob& j = J.job[J.n++];
        j.g = g; j.v = v; j.fwd = pk + fwd; j.tr = with_transposed ? pk + tr : nullptr;
        j.M = M; j.Cin = Cin; j.K = Kk; j.rows_f = M; j.cols_f = cols_f; j.rows_t = Cin; j.cols_t = cols_t; j.groups = 0;
    };
    if (blk_hi < 0) blk_hi = D.NB;
    for (int blk = blk_lo; blk < blk_hi; ++blk) {
        const gtk_block_params& P = blocks[blk];
        GTK_TRY(check_block(D, P));
        if (J.n + 2 + 2 * D.L * D.n_wn > tc::FOLD_PACK_MAX_JOBS) { GTK_TRY(tc::fold_pack(J, st)); J.n = 0; }
        pk = pk_base + (size_t)blk * Q.total;
        add(P.start_g, P.start_v, Q.start, Q.t_start, H, Ch, 1, pad64(Ch), H);
        for (int w = 0; w < D.n_wn; ++w) {
            const WnW W = wn_weights(P, D.wn_id[w]);
            for (int i = 0; i < D.L; ++i) {
                const int RS = (i < D.L - 1) ? 2 * H : H;
                GTK_REQUIRE(W.in_v[i] && W.in_b[i] && W.rs_v[i] && W.rs_b[i] && (w == 0 || (W.c1_v && W.c1_b)), "null WaveNet parameter");
                add(W.in_g[i], W.in_v[i], Q.in[LI(w, i)], Q.t_in[LI(w, i)], 2 * H, H, K, H, 2 * H);
                add(W.rs_g[i], W.rs_v[i], Q.rs[LI(w, i)], Q.t_rs[LI(w, i)], RS, H, 1, H, pad64(RS));
            }
        }
        add(nullptr, P.end_w, Q.end, Q.t_end, C, H, 1, H, pad64(C));
    }
    return J.n ? tc::fold_pack(J, st) : GTK_OK;
}

// Packs of the first block to run on the caller's stream, the others on the auxiliary stream (decoder.cu: aux_stream);
// *join is set to the event the caller's stream must wait for before it runs any other block (null: nothing to wait for).
static int tc_pack_split(const Dims& D, const TcLayout& L, const gtk_block_params* blocks, bf16* pk_base, int with_transposed,
                         int first_blk, cudaStream_t st, cudaEvent_t** join) {
    *join = nullptr;
    SideStream* ax = nullptr;
    GTK_TRY(aux_stream(&ax));
    if (!ax || D.NB < 2) return tc_pack_all(D, L, blocks, pk_base, with_transposed, st);
    GTK_CUDA(cudaEventRecord(ax->fork[0], st));
    GTK_CUDA(cudaStreamWaitEvent(ax->s, ax->fork[0], 0));
    if (first_blk > 0) GTK_TRY(tc_pack_all(D, L, blocks, pk_base, with_transposed, ax->s, 0, first_blk));
    if (first_blk + 1 < D.NB) GTK_TRY(tc_pack_all(D, L, blocks, pk_base, with_transposed, ax->s, first_blk + 1, D.NB));
    GTK_CUDA(cudaEventRecord(ax->done[0], ax->s));
    *join = &ax->done[0];
    return tc_pack_all(D, L, blocks, pk_base, with_transposed, st, first_blk, first_blk + 1);
}

// W^-1 / log|det W| of every block (+ the constant part of logdet when logdet != null), one launch per 32 blocks.
static int tc_flow_consts(const Dims& D, const TcLayout& L, const gtk_block_params* blocks, unsigned char* ws, const float* x_len,
                          float* logdet, cudaStream_t st) {
    static thread_local flow::FlowConstJobs J;
    for (int k0 = 0; k0 < D.NB; k0 += flow::FLOW_CONST_MAX_BLOCKS) {
        J.n = min(flow::FLOW_CONST_MAX_BLOCKS, D.NB - k0);
        for (int k = 0; k < J.n; ++k) {
            GTK_TRY(check_block(D, blocks[k0 + k]));
            J.w[k] = blocks[k0 + k].ic_weight;
            J.logs[k] = blocks[k0 + k].an_logs;
        }
        J.winv = reinterpret_cast<float*>(ws + L.winv) + (size_t)k0 * 64;
        J.icld = reinterpret_cast<float*>(ws + L.icld) + k0;
        GTK_TRY(flow::flow_consts(J, D.NS, D.C, x_len, logdet, D.B, 1, st));
    }
    return GTK_OK;
}

// Device views of the packing plan inside the workspace.
struct TcPlan {
    int* hdr; int* off; int* row_b; int* row_fr; float* mask_p; float* x_len;
    const int* n_rows() const { return hdr + 1; }
};
static TcPlan tc_plan(const TcLayout& L, unsigned char* ws) {
    TcPlan q;
    q.hdr = reinterpret_cast<int*>(ws + L.hdr); q.off = reinterpret_cast<int*>(ws + L.off);
    q.row_b = reinterpret_cast<int*>(ws + L.row_b); q.row_fr = reinterpret_cast<int*>(ws + L.row_fr);
    q.mask_p = reinterpret_cast<float*>(ws + L.mask_sq); q.x_len = reinterpret_cast<float*>(ws + L.x_len);
    return q;
}

// start -> WN -> end(+affine) of one block on the packed rows.  xc = mixed block input (t-major fp32 [C, P]),
// zout = block output.
static int tc_coupling(const gtk_decoder_desc* d, const Dims& D, const TcLayout& L, const gtk_block_params& P, const bf16* pk,
                       const float* xc, float* zout, float* logdet, const TcPlan& pl, unsigned char* act, unsigned char* ws,
                       int reverse, int save, int blk, cudaStream_t st) {
    const int B = 1, T = L.P, H = D.H, C = D.C, Ch = D.C / 2;
    const float* mask = pl.mask_p;
    const TcPack& Q = L.P_;
    const TcAct& A = L.A;
    bf16* a0 = reinterpret_cast<bf16*>(act + A.a0);
    float* hf[2] = {reinterpret_cast<float*>(ws + L.hf[0]), reinterpret_cast<float*>(ws + L.hf[1])};
    float* skipf = reinterpret_cast<float*>(ws + L.skipf);
    GTK_TRY(tc::frames_from_tmajor(xc, a0, B, C, T, 0, Ch, pad64(Ch), st, pl.n_rows()));
    const bool stack = L.stack && tc::wn_stack_enabled();
    const bool end_in_stack = stack && tc::wn_stack_end_ok(H, C) && !(getenv("GTK_WN_STACK_END") && getenv("GTK_WN_STACK_END")[0] == '0');
    if (stack) {
        // start conv + every layer of a WaveNet in ONE launch per WaveNet (wn_stack.cu); a chained WaveNet reads the
        // masked output of the one before it (bf16 operand + fp32 master through the frame-addressed exchange buffers)
        float* hxb[2] = {hf[1], L.hx ? reinterpret_cast<float*>(ws + L.hx) : hf[1]};
        for (int w = 0; w < D.n_wn; ++w) {
            const WnW W = wn_weights(P, D.wn_id[w]);
            tc::WnStackIo io;
            memset(&io, 0, sizeof(io));
            tc::WnStackParams p;
            memset(&p, 0, sizeof(p));
            p.B = B; p.T = T; p.H = H; p.taps = D.K; p.L = D.L;
            p.save = save;
            p.n_rows = pl.n_rows(); p.row_b = pl.row_b; p.row_fr = pl.row_fr; p.mask = mask;
            p.drop_p = d->p_dropout;
            p.hf = hf[0]; p.skipf = skipf;
            if (w == 0) {
                p.has_start = 1; p.n_ks = pad64(Ch) / 64; p.bias_start = P.start_b;
                io.a0 = a0; io.ld_a0 = pad64(Ch); io.wstart = pk + Q.start;
                p.cond_bstride = (size_t)2 * H * D.L;
            } else {                                      // modules.WNP: cond_layer1 on the un-squeezed track, squeezed (modules.py:320-322)
                io.hin = reinterpret_cast<const bf16*>(act + A.skip_bf[w - 1]);
                p.h_in_f32 = hxb[(w - 1) & 1];
                p.trk = reinterpret_cast<const float*>(ws + L.trk[w]);
                p.cw_g = W.c1_g; p.cw_v = W.c1_v; p.cb = W.c1_b;
                p.n_sqz = D.n; p.cs_div = 2 * H * D.L / D.n;
            }
            if (w + 1 < D.n_wn) p.h_out_f32_last = hxb[w & 1];
            else if (end_in_stack) {                      // `end` + affine coupling on the resident WN output
                p.has_end = 1; p.C_end = C; p.bias_end = P.end_b; io.wend = pk + Q.end;
                p.aff_x = xc; p.aff_z = zout; p.aff_logdet = reverse ? nullptr : logdet;
                p.aff_logs_save = save ? reinterpret_cast<float*>(act + A.logs) : nullptr;
                p.aff_reverse = reverse; p.aff_sigmoid_scale = d->sigmoid_scale;
            }
            io.skip_out = reinterpret_cast<bf16*>(act + A.skip_bf[w]);
            for (int i = 0; i < D.L; ++i) {
                tc::WnStackLayer& ly = p.layer[i];
                ly.bias_in = W.in_b[i]; ly.bias_rs = W.rs_b[i];
                ly.cond = (w == 0 && P.cond) ? P.cond + (size_t)i * 2 * H : nullptr;
                ly.drop_seed = make_drop_seed(d->seed, d->seed_dev, blk, D.wn_id[w] * D.L + i);
                ly.ts_save = save ? reinterpret_cast<bf16*>(act + A.ts[LI(w, i)]) : nullptr;
                ly.RS = (i < D.L - 1) ? 2 * H : H; ly.dil = ipow(D.dr, i); ly.cs_base = i * 2 * H;
                io.win[i] = pk + Q.in[LI(w, i)]; io.wrs[i] = pk + Q.rs[LI(w, i)];
                io.hsave[i] = (i == 0 && w > 0) ? nullptr : reinterpret_cast<bf16*>(act + A.h[LI(w, i)]);
                io.acts[i] = reinterpret_cast<bf16*>(act + A.acts[LI(w, i)]);
            }
            GTK_TRY(tc::wn_stack_launch(io, p, st));
        }
    } else {
    {
        tc::ConvTcParams p;
        memset(&p, 0, sizeof(p));
        p.B = B; p.T = T; p.taps = 1; p.dil = 1; p.n_k = pad64(Ch) / 64; p.N = H;
        p.bias = P.start_b; p.mask = mask; p.mask_out = 1;
        p.n_rows = pl.n_rows();
        p.tile_layout = 1;
        p.out_f32 = hf[0]; p.ld_f32 = H;
        p.out_bf16 = reinterpret_cast<bf16*>(act + A.h[0]); p.ld_bf16 = H;
        GTK_TRY(tc::conv_tc_launch(a0, pad64(Ch), pk + Q.start, p, tc::EPI_PLAIN, st));
    }
    // wn -> wn_energy -> wn_pitch (attentions.py:155-160): a chained WaveNet reads the masked output of the one before it
    // (bf16 operand = skip_bf of that slot, fp32 master = the residual buffer its last layer left).
    int cur = 0;                                          // which of the two fp32 residual buffers holds the layer input
    for (int w = 0; w < D.n_wn; ++w) {
        const WnW W = wn_weights(P, D.wn_id[w]);
        for (int i = 0; i < D.L; ++i) {
            tc::WnLayerParams p;
            memset(&p, 0, sizeof(p));
            p.B = B; p.T = T; p.H = H; p.RS = (i < D.L - 1) ? 2 * H : H; p.taps = D.K; p.dil = ipow(D.dr, i);
            p.first = (i == 0); p.last = (i == D.L - 1);
            p.tile_layout = 1;
            p.bias_in = W.in_b[i]; p.bias_rs = W.rs_b[i];
            if (w == 0) { p.cond = P.cond ? P.cond + (size_t)i * 2 * H : nullptr; p.cond_bstride = (size_t)2 * H * D.L; }
            else {                                        // modules.WNP: cond_layer1 on the un-squeezed track, squeezed (modules.py:320-322)
                p.trk = reinterpret_cast<const float*>(ws + L.trk[w]);
                p.cw_g = W.c1_g; p.cw_v = W.c1_v; p.cb = W.c1_b;
                p.n_sqz = D.n; p.cs_div = 2 * H * D.L / D.n; p.cs_base = i * 2 * H;
            }
            p.n_rows = pl.n_rows(); p.row_b = pl.row_b; p.row_fr = pl.row_fr;
            p.mask = mask;
            p.drop_p = d->p_dropout; p.drop_seed = make_drop_seed(d->seed, d->seed_dev, blk, D.wn_id[w] * D.L + i);
            p.h_f32 = hf[cur]; p.h_out_f32 = hf[cur ^ 1];
            p.h_out_bf16 = (i < D.L - 1) ? reinterpret_cast<bf16*>(act + A.h[LI(w, i + 1)]) : nullptr;
            p.skip_f32 = skipf;
            p.skip_out_bf16 = reinterpret_cast<bf16*>(act + A.skip_bf[w]);
            if (p.last && w + 1 < D.n_wn) p.h_out_f32_last = hf[cur ^ 1];
            p.ts_save = save ? reinterpret_cast<bf16*>(act + A.ts[LI(w, i)]) : nullptr;
            p.acts_save = save ? reinterpret_cast<bf16*>(act + A.acts[LI(w, i)]) : nullptr;
            const bf16* h_in = (i == 0 && w > 0) ? reinterpret_cast<const bf16*>(act + A.skip_bf[w - 1])
                                                 : reinterpret_cast<const bf16*>(act + A.h[LI(w, i)]);
            GTK_TRY(tc::wn_layer_launch(h_in, pk + Q.in[LI(w, i)], pk + Q.rs[LI(w, i)], p, st));
            cur ^= 1;
        }
    }
    }   // layer-by-layer path
    if (!end_in_stack) {
        tc::ConvTcParams p;
        memset(&p, 0, sizeof(p));
        p.B = B; p.T = T; p.taps = 1; p.dil = 1; p.n_k = H / 64; p.N = C;
        p.bias = P.end_b; p.mask = mask;
        p.n_rows = pl.n_rows(); p.row_b = pl.row_b;
        p.aff_x = xc; p.aff_z = zout; p.aff_logdet = reverse ? nullptr : logdet;
        p.aff_logs_save = save ? reinterpret_cast<float*>(act + A.logs) : nullptr;
        p.aff_reverse = reverse; p.aff_sigmoid_scale = d->sigmoid_scale;
        GTK_TRY(tc::conv_tc_launch(reinterpret_cast<const bf16*>(act + A.skip_bf[D.n_wn - 1]), H, pk + Q.end, p, tc::EPI_AFFINE, st));
    }
    return GTK_OK;
}

static int tc_forward(const gtk_decoder_desc* d, const Dims& D, const gtk_block_params* blocks, const float* x, const float* mask,
                      float* z, float* logdet, void* workspace, size_t ws_bytes, int train, cudaStream_t st) {
    GTK_TRY(tc_supported(D));
    Layout L32;
    make_layout(D, 0, &L32);
    TcLayout L;
    make_tc_layout(D, L32, train, &L);
    if (ws_bytes < L.total) { set_error("decoder_forward(bf16): workspace %zu < %zu", ws_bytes, L.total); return GTK_ERR_WORKSPACE; }
    if (D.B == 0) return GTK_OK;
    GTK_CUDA(cudaMemsetAsync(logdet, 0, sizeof(float) * D.B, st));
    if (D.Tq == 0) return GTK_OK;
    unsigned char* ws = static_cast<unsigned char*>(workspace);
    const TcPlan pl = tc_plan(L, ws);
    cudaEvent_t* packs_done = nullptr;
    GTK_TRY(tc_pack_split(D, L, blocks, reinterpret_cast<bf16*>(ws + L.pack), train, 0, st, &packs_done));
    GTK_TRY(flow::pack_plan(mask, D.B, D.T, D.n, D.Tq, L.G, L.P, pl.hdr, pl.off, pl.row_b, pl.row_fr, pl.mask_p, pl.x_len, st));
    GTK_TRY(flow::squeeze_pack(x, pl.off, pl.row_b, pl.mask_p, reinterpret_cast<float*>(ws + L.X), D.Cin, D.T, D.n, L.P, st));
    for (int w = 1; w < D.n_wn; ++w)
        GTK_TRY(flow::pack_track(D.wn_id[w] == 1 ? d->energy : d->pitch, pl.off, pl.row_b, pl.mask_p,
                                 reinterpret_cast<float*>(ws + L.trk[w]), D.T, D.n, L.P, st));
    GTK_TRY(tc_flow_consts(D, L, blocks, ws, pl.x_len, logdet, st));   // W^-1, log|det W|, the constant logdet terms
    if (train)     // guard rows behind the last tile in use of the conv inputs the weight gradient reads with a tap shift
        for (int w = 0; w < D.n_wn; ++w) {
            GTK_TRY(tc::zero_guard_rows(reinterpret_cast<bf16*>(ws + L.act + L.A.h[LI(w, 0)]), D.L > 1 ? L.A.h[1] - L.A.h[0] : 0, D.L,
                                        L.A.total, D.NB, D.H, pl.n_rows(), 64, L.P, st));
            if (w + 1 < D.n_wn)
                GTK_TRY(tc::zero_guard_rows(reinterpret_cast<bf16*>(ws + L.act + L.A.skip_bf[w]), 0, 1, L.A.total, D.NB, D.H,
                                            pl.n_rows(), 64, L.P, st));
        }
    for (int blk = 0; blk < D.NB; ++blk) {
        const gtk_block_params& P = blocks[blk];
        GTK_TRY(check_block(D, P));
        if (blk == 1 && packs_done) GTK_CUDA(cudaStreamWaitEvent(st, *packs_done, 0));
        unsigned char* act = ws + L.act + (size_t)(train ? blk : 0) * L.A.total;
        bf16* pk = reinterpret_cast<bf16*>(ws + L.pack) + (size_t)blk * L.P_.total;
        const float* xin = reinterpret_cast<float*>(ws + L.X + (size_t)(train ? blk : (blk & 1)) * L.x_bytes);
        float* xout = reinterpret_cast<float*>(ws + L.X + (size_t)(train ? blk + 1 : ((blk + 1) & 1)) * L.x_bytes);
        float* xmix = reinterpret_cast<float*>(act + L.A.xmix);
        GTK_TRY(flow::mix(xin, pl.mask_p, P.an_logs, P.an_bias, P.ic_weight, D.NS, xmix, 1, D.C, L.P, 0, st, pl.n_rows()));
        GTK_TRY(tc_coupling(d, D, L, P, pk, xmix, xout, logdet, pl, act, ws, 0, train, blk, st));
    }
    const float* xfin = reinterpret_cast<float*>(ws + L.X + (size_t)(train ? D.NB : (D.NB & 1)) * L.x_bytes);
    GTK_TRY(flow::unsqueeze_unpack(xfin, pl.off, z, D.B, D.Cin, D.Tq * D.n, D.n, L.G, L.P, st));
    return GTK_OK;
}

static int tc_reverse(const gtk_decoder_desc* d, const Dims& D, const gtk_block_params* blocks, const float* z, const float* mask,
                      float* x, void* workspace, size_t ws_bytes, cudaStream_t st) {
    GTK_TRY(tc_supported(D));
    Layout L32;
    make_layout(D, 0, &L32);
    TcLayout L;
    make_tc_layout(D, L32, 0, &L);
    if (ws_bytes < L.total) { set_error("decoder_reverse(bf16): workspace %zu < %zu", ws_bytes, L.total); return GTK_ERR_WORKSPACE; }
    if (D.B == 0 || D.Tq == 0) return GTK_OK;
    unsigned char* ws = static_cast<unsigned char*>(workspace);
    const TcPlan pl = tc_plan(L, ws);
    float* xa = reinterpret_cast<float*>(ws + L.X);
    float* xb = reinterpret_cast<float*>(ws + L.X + L.x_bytes);
    unsigned char* act = ws + L.act;
    cudaEvent_t* packs_done = nullptr;
    GTK_TRY(tc_pack_split(D, L, blocks, reinterpret_cast<bf16*>(ws + L.pack), 0, D.NB - 1, st, &packs_done));
    GTK_TRY(flow::pack_plan(mask, D.B, D.T, D.n, D.Tq, L.G, L.P, pl.hdr, pl.off, pl.row_b, pl.row_fr, pl.mask_p, pl.x_len, st));
    GTK_TRY(flow::squeeze_pack(z, pl.off, pl.row_b, pl.mask_p, xa, D.Cin, D.T, D.n, L.P, st));
    for (int w = 1; w < D.n_wn; ++w)
        GTK_TRY(flow::pack_track(D.wn_id[w] == 1 ? d->energy : d->pitch, pl.off, pl.row_b, pl.mask_p,
                                 reinterpret_cast<float*>(ws + L.trk[w]), D.T, D.n, L.P, st));
    GTK_TRY(tc_flow_consts(D, L, blocks, ws, nullptr, nullptr, st));
    for (int blk = D.NB - 1; blk >= 0; --blk) {
        const gtk_block_params& P = blocks[blk];
        GTK_TRY(check_block(D, P));
        if (blk == D.NB - 2 && packs_done) GTK_CUDA(cudaStreamWaitEvent(st, *packs_done, 0));
        float* winv = reinterpret_cast<float*>(ws + L.winv) + (size_t)blk * 64;
        const bf16* pk = reinterpret_cast<const bf16*>(ws + L.pack) + (size_t)blk * L.P_.total;
        GTK_TRY(tc_coupling(d, D, L, P, pk, xa, xb, nullptr, pl, act, ws, 1, 0, blk, st));
        GTK_TRY(flow::mix(xb, pl.mask_p, P.an_logs, P.an_bias, winv, D.NS, xa, 1, D.C, L.P, 1, st, pl.n_rows()));
    }
    GTK_TRY(flow::unsqueeze_unpack(xa, pl.off, x, D.B, D.Cin, D.Tq * D.n, D.n, L.G, L.P, st));
    return GTK_OK;
}

// ------------------------------------------------------------------------------------------ backward
static int tc_backward(const gtk_decoder_desc* d, const Dims& D, const gtk_block_params* blocks, const gtk_block_grads* grads,
                       const float* mask, const float* dz, const float* dlogdet, float* dx, void* workspace, size_t ws_bytes,
                       int blk_hi, int blk_lo, cudaStream_t st) {
    const bool first = blk_hi == D.NB - 1, last = blk_lo == 0;
    GTK_TRY(tc_supported(D));
    Layout L32;
    make_layout(D, 0, &L32);
    TcLayout L;
    make_tc_layout(D, L32, 1, &L);
    if (ws_bytes < L.total) { set_error("decoder_backward(bf16): workspace %zu < %zu", ws_bytes, L.total); return GTK_ERR_WORKSPACE; }
    if (D.B == 0 || D.T == 0) return GTK_OK;
    if (D.Tq == 0) {
        if (first) GTK_CUDA(cudaMemsetAsync(dx, 0, sizeof(float) * (size_t)D.B * D.Cin * D.T, st));
        return GTK_OK;
    }
    (void)mask;                                          // the plan the forward left in the workspace has the lengths

    unsigned char* ws = static_cast<unsigned char*>(workspace);
    const TcPlan pl = tc_plan(L, ws);
    const float* msq = pl.mask_p;
    const int B = 1, T = L.P, H = D.H, C = D.C, Ch = D.C / 2, K = D.K;
    const int ldC = pad64(C), ldCh = pad64(Ch);
    float* dZ = reinterpret_cast<float*>(ws + L.b_dZ);
    float* dM = reinterpret_cast<float*>(ws + L.b_dM);
    float* dhf[2] = {reinterpret_cast<float*>(ws + L.b_dhf[0]), reinterpret_cast<float*>(ws + L.b_dhf[1])};
    float* wpart = reinterpret_cast<float*>(ws + L.b_wpart);   // stream-K partial tiles of the weight gradients
    SideStream* side = nullptr;
    GTK_TRY(side_stream(&side));
    bool side_used[2] = {false, false};
    // main-stream weight gradients: all stream-K launches of this call share ONE reduce launch at its end
    const bool semi = side && wgrad_semi_overlap();
    // Deferred join (gtk_decoder_backward_range_deferred, semi overlap only): a range that is not the last one returns with
    // its last weight-gradient launch and its reduce still running on the side stream; the next range carries on as the
    // un-split backward would -- it waits for that weight gradient before its first tensor-core kernel and never for the
    // reduce (a range boundary then costs what a block boundary costs, not an exposed join: profiles/scaling_r02.txt)
    const bool defer = semi && g_defer_join && !last;
    const bool carried = semi && side->carry && !first;
    if (carried) { side_used[0] = side->carry_used[0]; side_used[1] = side->carry_used[1]; }
    if (side) side->carry = false;
    if (side && first) side->red_pending = false;
    tc::WgradReduceBatch* rbatch = (side && !semi) ? nullptr : tc::wgrad_reduce_batch(wpart, L.wpart_cap);

    if (first) GTK_TRY(flow::squeeze_pack(dz, pl.off, pl.row_b, pl.mask_p, dZ, D.Cin, D.Tq * D.n, D.n, L.P, st));
    if (first)
        for (int w = 1; w < D.n_wn; ++w) GTK_CUDA(cudaMemsetAsync(ws + L.dtrk[w], 0, (size_t)L.P * D.n * sizeof(float), st));
    if (first) {   // gradients of the constant logdet terms of every block, one launch (they only need dlogdet and the saved W^-1)
        static thread_local flow::FlowConstGrads FG;
        for (int k0 = 0; k0 < D.NB; k0 += flow::FLOW_CONST_MAX_BLOCKS) {
            FG.n = min(flow::FLOW_CONST_MAX_BLOCKS, D.NB - k0);
            for (int k = 0; k < FG.n; ++k) { FG.dlogs[k] = grads[k0 + k].an_logs; FG.dW[k] = grads[k0 + k].ic_weight; }
            GTK_TRY(flow::logdet_const_bwd_all(dlogdet, pl.x_len, D.B, FG, C, reinterpret_cast<const float*>(ws + L.winv) + (size_t)k0 * 64,
                                               D.NS, st));
        }
    }

    for (int blk = blk_hi; blk >= blk_lo; --blk) {
        const gtk_block_params& P = blocks[blk];
        const gtk_block_grads& G = grads[blk];
        unsigned char* act = ws + L.act + (size_t)blk * L.A.total;
        const bf16* pk = reinterpret_cast<const bf16*>(ws + L.pack) + (size_t)blk * L.P_.total;
        const TcPack& Q = L.P_;
        const TcAct& A = L.A;
        const float* xin_blk = reinterpret_cast<const float*>(ws + L.X + (size_t)blk * L.x_bytes);
        const float* xmix = reinterpret_cast<const float*>(act + A.xmix);
        const bf16* a0 = reinterpret_cast<const bf16*>(act + A.a0);
        const bf16* skip_bf = reinterpret_cast<const bf16*>(act + A.skip_bf[D.n_wn - 1]);     // input of the `end` conv
        // the gradient temporaries the weight-gradient launch reads: one set per block parity
        const int par = blk & 1;
        unsigned char* wp = ws + (size_t)par * L.b_pstride;
        bf16* dend = reinterpret_cast<bf16*>(wp + L.b_dend);
        auto drs_of = [&](int w, int i) { return reinterpret_cast<bf16*>(wp + L.b_drs + (size_t)(w * D.L + i) * L.b_lstride); };
        auto dxin_of = [&](int w, int i) { return reinterpret_cast<bf16*>(wp + L.b_dxin + (size_t)(w * D.L + i) * L.b_lstride); };
        auto dgate_of = [&](int w, int i) { return reinterpret_cast<bf16*>(wp + L.b_dgate + (size_t)(w * D.L + i) * L.b_lstride); };
        bf16* dh0 = reinterpret_cast<bf16*>(wp + L.b_dh0);
        // the weight gradients of block blk + 2 (same parity) must have finished reading this set
        if (side && side_used[par]) GTK_CUDA(cudaStreamWaitEvent(st, side->done[par], 0));

        // Weight / bias gradients of the whole block are queued here and run as ONE stream-K launch after the
        // data-gradient chain (which is the critical path and keeps every dY it produced) -- on the side stream, i.e.
        // concurrently with the NEXT block's chain; the last block of the call has nothing to hide behind and gets all SMs.
        tc::WgradGroupJob jobs[tc::WGG_MAX_JOBS];
        int n_jobs = 0, n_tiles = 0;
        cudaStream_t wst = side ? side->s : st;
        auto flush = [&]() -> int {
            if (!n_jobs) return GTK_OK;
            if (side) {
                GTK_CUDA(cudaEventRecord(side->fork[par], st));
                GTK_CUDA(cudaStreamWaitEvent(side->s, side->fork[par], 0));
            }
            int rc = tc::wgrad_group_launch(jobs, n_jobs, B, T, wpart, pl.n_rows(), wst,
                                            (side && !semi && blk > blk_lo) ? wgrad_overlap_ctas() : 0, rbatch);
            n_jobs = 0; n_tiles = 0;
            return rc;
        };
        auto queue = [&](const float* g, const float* v, float* dg, float* dv, float* db, const bf16* dy, int ld_y, int cols_y,
                         const bf16* x, int ld_x, int cols_x, int M, int Cin, int Kk, int dil) -> int {
            if (!dv) return db ? tc::colsum_frames_seg(dy, ld_y, M, 1, pl.hdr, db, 0, st) : GTK_OK;   // (no caller leaves dv null)
            if (g && !dg) { set_error("decoder_backward(bf16): weight_g gradient pointer missing"); return GTK_ERR_INVALID; }
            const int tiles = ((M + 127) / 128) * Kk;
            if (n_jobs == tc::WGG_MAX_JOBS || n_tiles + tiles > tc::WGG_MAX_TILES) GTK_TRY(flush());
            tc::WgradGroupJob& J = jobs[n_jobs++];
            n_tiles += tiles;
            J.dy = dy; J.ld_y = ld_y; J.cols_y = cols_y; J.x = x; J.ld_x = ld_x; J.cols_x = cols_x;
            J.g = g; J.v = v; J.dg = dg; J.dv = dv; J.db = db; J.M = M; J.Cin = Cin; J.K = Kk; J.dil = dil;
            return GTK_OK;
        };

        // affine coupling + end conv
        GTK_TRY(tc::affine_bwd_frames(xmix, reinterpret_cast<const float*>(act + A.logs), msq, dZ, dlogdet, dend, ldC, dM, B, C, T,
                                      d->sigmoid_scale, pl.row_b, pl.n_rows(), st));
        GTK_TRY(queue(nullptr, P.end_w, nullptr, G.end_w, G.end_b, dend, ldC, ldC, skip_bf, H, H, C, H, 1, 1));
        // semi overlap: the weight gradients of the block above have had mix_bwd and affine_bwd_frames to hide behind; the
        // tensor-core chain of this block wants the SMs to itself
        if (semi && (blk < blk_hi || carried) && side_used[(blk + 1) & 1]) GTK_CUDA(cudaStreamWaitEvent(st, side->done[(blk + 1) & 1], 0));
        {
            // d(skip) feeds the res_skip gradient of EVERY layer: written into columns [H, 2H) of all per-layer buffers
            tc::ConvTcParams p;
            memset(&p, 0, sizeof(p));
            p.B = B; p.T = T; p.taps = 1; p.dil = 1; p.n_k = ldC / 64; p.N = H;
            p.mask = msq; p.mask_out = 1;
            p.n_rows = pl.n_rows();
            p.out_bf16 = drs_of(D.n_wn - 1, 0) + H; p.ld_bf16 = 2 * H;           // ... of the LAST WaveNet of the chain
            p.rep_bf16 = D.L; p.rep_stride = L.b_lstride / 2;
            GTK_TRY(tc::conv_tc_launch(dend, ldC, pk + Q.t_end, p, tc::EPI_PLAIN, st));
        }
        const bool post_fused = tc::wn_bwd_supported(H, K, 1) && tc::wn_bwd_post_ok(H, Ch) &&
                                !(getenv("GTK_WB_POST") && getenv("GTK_WB_POST")[0] == '0');
        bool post_done = false;
        for (int w = D.n_wn - 1; w >= 0; --w) {
        const WnW W = wn_weights(P, D.wn_id[w]);
        const WnG Gw = wn_grads(G, D.wn_id[w]);
        for (int i = D.L - 1; i >= 0; --i) {
            const int RS = (i < D.L - 1) ? 2 * H : H;
            const int dil = ipow(D.dr, i);
            const bf16* drs_i = (i < D.L - 1) ? drs_of(w, i) : drs_of(w, i) + H;     // [dh_{i+1} mask | dskip]  or  dskip
            bf16* dxin = dxin_of(w, i);
            const bf16* ts = reinterpret_cast<const bf16*>(act + A.ts[LI(w, i)]);
            const bf16* acts = reinterpret_cast<const bf16*>(act + A.acts[LI(w, i)]);
            const bf16* h_i = (i == 0 && w > 0) ? reinterpret_cast<const bf16*>(act + A.skip_bf[w - 1])
                                                : reinterpret_cast<const bf16*>(act + A.h[LI(w, i)]);
            const bool want_dcond = w > 0 || (P.cond && G.cond);
            // the conditioning enters after the dropout: with both active the un-scaled gate gradient is kept as well
            bf16* dpre = (want_dcond && d->p_dropout > 0.f) ? dgate_of(w, i) : nullptr;
            GTK_TRY(queue(W.rs_g[i], W.rs_v[i], Gw.rs_g[i], Gw.rs_v[i], Gw.rs_b[i], drs_i, 2 * H, RS, acts, H, H, RS, H, 1, 1));
            GTK_TRY(queue(W.in_g[i], W.in_v[i], Gw.in_g[i], Gw.in_v[i], Gw.in_b[i], dxin, 2 * H, 2 * H, h_i, H, H, 2 * H, H, K, dil));
            if (tc::wn_bwd_supported(H, K, dil)) {
                // res_skip data gradient -> gate backward -> in_layer data gradient in ONE kernel (wn_bwd_tc.cu)
                tc::WnBwdParams p;
                memset(&p, 0, sizeof(p));
                p.B = B; p.T = T; p.H = H; p.RS = RS; p.taps = K; p.dil = dil;
                p.mask = msq; p.ts = ts;
                p.n_rows = pl.n_rows(); p.row_fr = pl.row_fr; p.dpre = dpre;
                p.drop_p = d->p_dropout; p.drop_seed = make_drop_seed(d->seed, d->seed_dev, blk, D.wn_id[w] * D.L + i);
                p.resid_f32 = (i < D.L - 1) ? dhf[(i + 1) & 1] : nullptr;
                p.out_f32 = dhf[i & 1];
                if (i == 0 && w == 0 && post_fused) {      // data gradient of `start` in the same kernel (GEMM c)
                    p.post_w = pk + Q.t_start; p.post_N = Ch; p.post_C = C; p.post_resid_tm = dZ; p.post_out_tm = dM;
                    post_done = true;
                }
                GTK_TRY(tc::wn_bwd_layer_launch(drs_i, 2 * H, pk + Q.t_rs[LI(w, i)], pk + Q.t_in[LI(w, i)], dxin,
                                                i > 0 ? drs_of(w, i - 1) : dh0, i > 0 ? 2 * H : H, p, st));
            } else {
                {   // res_skip 1x1: data gradient fused with the gate backward
                    tc::ConvTcParams p;
                    memset(&p, 0, sizeof(p));
                    p.B = B; p.T = T; p.taps = 1; p.dil = 1; p.n_k = pad64(RS) / 64; p.N = H;
                    p.mask = msq;
                    p.n_rows = pl.n_rows(); p.row_fr = pl.row_fr;
                    p.tile_layout = 1;
                    p.ts = ts; p.out_bf16 = dxin; p.ld_bf16 = 2 * H;
                    p.drop_p = d->p_dropout; p.drop_seed = make_drop_seed(d->seed, d->seed_dev, blk, D.wn_id[w] * D.L + i);
                    p.out2_bf16 = dpre;
                    GTK_TRY(tc::conv_tc_launch(drs_i, 2 * H, pk + Q.t_rs[LI(w, i)], p, tc::EPI_GATE_BWD, st));
                }
                {   // in_layer dilated conv
                    tc::ConvTcParams p;
                    memset(&p, 0, sizeof(p));
                    p.B = B; p.T = T; p.taps = K; p.dil = dil; p.n_k = 2 * H / 64; p.N = H;
                    p.mask = msq; p.mask_out = 1;
                    p.n_rows = pl.n_rows();
                    p.tile_layout = 1;
                    if (i < D.L - 1) { p.resid_f32 = dhf[(i + 1) & 1]; p.ld_resid = H; p.mask_resid = 1; }
                    p.out_f32 = dhf[i & 1]; p.ld_f32 = H;
                    if (i > 0) { p.out_bf16 = drs_of(w, i - 1); p.ld_bf16 = 2 * H; }
                    else       { p.out_bf16 = dh0; p.ld_bf16 = H; }
                    GTK_TRY(tc::conv_tc_launch(dxin, 2 * H, pk + Q.t_in[LI(w, i)], p, tc::EPI_PLAIN, st));
                }
            }
        }
        if (w > 0) {
            // modules.WNP: gradients of cond_layer1 and of the pitch / energy track from the layers' gate-input gradients,
            // then d(input of this WaveNet) = d(masked output of the one before it) = the d(skip) operand of all ITS layers
            GTK_TRY(tc::wnp_cond_bwd(d->p_dropout > 0.f ? dgate_of(w, 0) : dxin_of(w, 0), L.b_lstride / 2, D.L, 2 * H,
                                     reinterpret_cast<const float*>(ws + L.trk[w]), D.n, 2 * H * D.L / D.n, W.c1_g, W.c1_v, Gw.c1_g,
                                     Gw.c1_v, Gw.c1_b, reinterpret_cast<float*>(ws + L.dtrk[w]), pl.n_rows(), T, st));
            GTK_TRY(tc::frames_fanout(dh0, H, drs_of(w - 1, 0) + H, 2 * H, L.b_lstride / 2, D.L, H, pl.n_rows(), T, st));
        }
        }
        // start conv: h0 = (W x0 + b) * mask   (dh0 is already masked)
        GTK_TRY(queue(P.start_g, P.start_v, G.start_g, G.start_v, G.start_b, dh0, H, H, a0, ldCh, ldCh, H, Ch, 1, 1));
        if (!post_done) {
            tc::ConvTcParams p;
            memset(&p, 0, sizeof(p));
            p.B = B; p.T = T; p.taps = 1; p.dil = 1; p.n_k = H / 64; p.N = Ch;
            p.n_rows = pl.n_rows();
            p.resid_tm = dZ; p.out_tm = dM; p.tm_C = C; p.tm_c0 = 0;
            GTK_TRY(tc::conv_tc_launch(dh0, H, pk + Q.t_start, p, tc::EPI_PLAIN, st));
        }
        GTK_TRY(flush());
        if (P.cond && G.cond)   // d cond_layer(g)[b, layer, n] = sum over the utterance's rows of the gate-input gradient, all layers
            GTK_TRY(tc::colsum_frames_seg(d->p_dropout > 0.f ? dgate_of(0, 0) : dxin_of(0, 0), 2 * H, 2 * H, D.B, pl.off, G.cond,
                                          (size_t)2 * H * D.L, wst, D.L, L.b_lstride / 2, (size_t)2 * H));
        if (side) { GTK_CUDA(cudaEventRecord(side->done[par], side->s)); side_used[par] = true; }
        // ActNorm + InvConvNear (fp32)
        GTK_TRY(flow::mix_bwd(xin_blk, msq, P.an_logs, P.an_bias, P.ic_weight, D.NS, dM, dZ, G.an_logs, G.an_bias, G.ic_weight, B,
                              C, T, st, pl.n_rows()));
    }
    // weight norm + accumulation of every weight gradient of this call (semi overlap: behind the launches on the side stream)
    GTK_TRY(tc::wgrad_group_reduce_flush(rbatch, semi ? side->s : st));
    if (defer) {
        GTK_CUDA(cudaEventRecord(side->red_done, side->s));
        side->red_pending = true; side->carry = true;
        side->carry_used[0] = side_used[0]; side->carry_used[1] = side_used[1];
        return GTK_OK;                                 // no join: the caller orders consumers with gtk_decoder_backward_join
    }
    if (semi) { GTK_CUDA(cudaEventRecord(side->done[0], side->s)); side_used[0] = true; }
    if (last) GTK_TRY(flow::unsqueeze_unpack(dZ, pl.off, dx, D.B, D.Cin, D.T, D.n, L.G, L.P, st));
    if (last)
        for (int w = 1; w < D.n_wn; ++w) {
            float* dtrack = D.wn_id[w] == 1 ? d->d_energy : d->d_pitch;
            if (dtrack) GTK_TRY(flow::unpack_track(reinterpret_cast<const float*>(ws + L.dtrk[w]), pl.off, dtrack, D.B, D.T, D.n, L.G, st));
        }
    for (int par = 0; par < 2; ++par)      // join: every gradient of this call is complete in the caller's stream order
        if (side && side_used[par]) GTK_CUDA(cudaStreamWaitEvent(st, side->done[par], 0));
    return GTK_OK;
}
