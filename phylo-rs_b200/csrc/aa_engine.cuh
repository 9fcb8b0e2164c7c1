// aa_engine.cuh — host side of the 20-state (AminoAcid) engines of a context created with PHB_FLAG_AMINO_ACID.
// Included by phylo_b200.cu after phb_ctx, DevBuf, TreeSched and the CU / CHECK_LAUNCH macros; kernels: aa.cuh.
// One device (the context's first); site patterns are tiled so that the CLVs of a tile fit the CLV budget.
#pragma once

namespace {

inline AaState &aa_of(phb_ctx *ctx) { return ctx->aa; }

int aa_set_alignment(phb_ctx *ctx, int32_t n_rows, int64_t P, const uint8_t *bytes, int32_t layout, const uint64_t *multiplicity)
{
    AaState &A = aa_of(ctx);
    Device &d = ctx->devs[0];
    CU(cudaSetDevice(d.ordinal));
    ctx->have_aln = false;
    ctx->n_rows = n_rows;
    ctx->n_patterns = P;
    A.P_pad = round_up(std::max<int64_t>(P, 1), AA_THREADS);
    ctx->row_invalid_host.assign(std::max(n_rows, 1), 0);
    ctx->row_first_invalid.assign(std::max(n_rows, 1), 0);
    // ASCII -> tip codes on the device, chunk by chunk (the caller's bytes travel as they are; pageable memory through the
    // bounce buffers): aa_pack_kernel
    const int nr1 = std::max(n_rows, 1);
    CU(A.tips.ensure((size_t)nr1 * (size_t)A.P_pad));
    CU(A.row_flags.ensure(sizeof(int) * (size_t)nr1));
    CU(cudaMemsetAsync(A.row_flags.p, 0, sizeof(int) * (size_t)nr1, d.stream));
    if (n_rows > 0 && A.P_pad > P) {
        const int64_t cells = (int64_t)n_rows * (A.P_pad - P);
        PHB_LAUNCH(aa_fill_pad_kernel, dim3((unsigned)((cells + 255) / 256)), dim3(256), d.stream, A.tips.as<uint8_t>(), n_rows, P, A.P_pad);
        CHECK_LAUNCH();
    }
    if (n_rows > 0 && P > 0) {
        const bool pageable = host_is_pageable(bytes);
        const size_t CHUNK = (size_t)128 << 20;
        if (layout == PHB_LAYOUT_PATTERN_MAJOR) {
            const int64_t pp = std::max<int64_t>(32, (int64_t)(CHUNK / (size_t)n_rows) / 32 * 32);
            CU(d.staging.ensure((size_t)std::min<int64_t>(pp, P) * n_rows));
            for (int64_t p0 = 0; p0 < P; p0 += pp) {
                const int64_t np = std::min<int64_t>(pp, P - p0);
                int rc = h2d_from_host(ctx, d, d.staging.as<uint8_t>(), bytes + (size_t)p0 * n_rows, (size_t)np * n_rows, d.stream, pageable);
                if (rc) return rc;
                PHB_LAUNCH(aa_pack_kernel, dim3((unsigned)((np + 31) / 32), (unsigned)((n_rows + 31) / 32)), dim3(32, 8), d.stream,
                           d.staging.as<uint8_t>(), 1, n_rows, np, p0, 0, n_rows, A.tips.as<uint8_t>(), A.P_pad, A.row_flags.as<int>());
                CHECK_LAUNCH();
            }
        } else {
            const int rr = (int)std::max<int64_t>(1, std::min<int64_t>(n_rows, (int64_t)(CHUNK / (size_t)P)));
            CU(d.staging.ensure((size_t)rr * (size_t)P));
            for (int r0 = 0; r0 < n_rows; r0 += rr) {
                const int nr = std::min(rr, n_rows - r0);
                int rc = h2d_from_host(ctx, d, d.staging.as<uint8_t>(), bytes + (size_t)r0 * P, (size_t)nr * P, d.stream, pageable);
                if (rc) return rc;
                PHB_LAUNCH(aa_pack_kernel, dim3((unsigned)((P + 31) / 32), (unsigned)((nr + 31) / 32)), dim3(32, 8), d.stream,
                           d.staging.as<uint8_t>(), 0, n_rows, P, (int64_t)0, r0, nr, A.tips.as<uint8_t>(), A.P_pad, A.row_flags.as<int>());
                CHECK_LAUNCH();
            }
        }
    }
    {
        std::vector<int> flags((size_t)nr1, 0);
        CU(cudaMemcpyAsync(flags.data(), A.row_flags.p, sizeof(int) * (size_t)nr1, cudaMemcpyDeviceToHost, d.stream));
        CU(cudaStreamSynchronize(d.stream));
        for (int r = 0; r < n_rows; ++r) {
            ctx->row_invalid_host[r] = (flags[r] & 1) != 0;
            ctx->row_first_invalid[r] = (flags[r] & 2) != 0;
        }
    }
    std::vector<double> mult((size_t)A.P_pad, 0.0);
    for (int64_t p = 0; p < P; ++p) mult[p] = multiplicity ? (double)multiplicity[p] : 1.0;
    CU(A.mult.ensure(sizeof(double) * mult.size()));
    CU(cudaMemcpyAsync(A.mult.p, mult.data(), sizeof(double) * mult.size(), cudaMemcpyHostToDevice, d.stream));
    CU(cudaStreamSynchronize(d.stream)); // the host vectors go out of scope
    A.tip_const_valid = false;
    ctx->have_aln = true;
    return PHB_OK;
}

// phb_compress_columns / phb_compress_rows in a 20-state context: the unique columns found on the device (compress_impl,
// alphabet-agnostic as Alignment::compress_columns is, alignment.rs:95-139) become the context's alignment without a
// trip through the host.  raw / replist / mult_u64 are compress_impl's device buffers; the work is queued on d.stream.
int aa_adopt_compressed(phb_ctx *ctx, int32_t n_rows, int64_t P, const uint8_t *raw, int64_t pitch, const uint32_t *replist,
                        const uint64_t *mult_u64)
{
    AaState &A = aa_of(ctx);
    Device &d = ctx->devs[0];
    CU(cudaSetDevice(d.ordinal));
    ctx->n_rows = n_rows;
    ctx->n_patterns = P;
    A.P_pad = round_up(std::max<int64_t>(P, 1), AA_THREADS);
    const int nr1 = std::max(n_rows, 1);
    ctx->row_invalid_host.assign(nr1, 0);
    ctx->row_first_invalid.assign(nr1, 0);
    CU(A.tips.ensure((size_t)nr1 * (size_t)A.P_pad));
    CU(A.row_flags.ensure(sizeof(int) * (size_t)nr1));
    CU(A.mult.ensure(sizeof(double) * (size_t)A.P_pad));
    CU(cudaMemsetAsync(A.row_flags.p, 0, sizeof(int) * (size_t)nr1, d.stream));
    if (n_rows > 0 && A.P_pad > P) {
        const int64_t cells = (int64_t)n_rows * (A.P_pad - P);
        PHB_LAUNCH(aa_fill_pad_kernel, dim3((unsigned)((cells + 255) / 256)), dim3(256), d.stream, A.tips.as<uint8_t>(), n_rows, P, A.P_pad);
        CHECK_LAUNCH();
    }
    for (int r0 = 0; r0 < n_rows && P > 0; r0 += 32768) {
        dim3 grid((unsigned)std::min<int64_t>((P + 255) / 256, 1024), (unsigned)std::min(32768, n_rows - r0));
        PHB_LAUNCH(aa_gather_patterns_kernel, grid, dim3(256), d.stream, raw, pitch, replist, P, r0, A.tips.as<uint8_t>(), A.P_pad,
                   A.row_flags.as<int>());
        CHECK_LAUNCH();
    }
    if (P > 0) {
        PHB_LAUNCH(mult_u64_to_double_kernel, dim3((unsigned)std::min<int64_t>((A.P_pad + 255) / 256, 4096)), dim3(256), d.stream,
                   mult_u64, (int64_t)0, P, A.P_pad, A.mult.as<double>());
        CHECK_LAUNCH();
    } else
        CU(cudaMemsetAsync(A.mult.p, 0, sizeof(double) * (size_t)A.P_pad, d.stream));
    std::vector<int> flags((size_t)nr1, 0);
    CU(cudaMemcpyAsync(flags.data(), A.row_flags.p, sizeof(int) * (size_t)nr1, cudaMemcpyDeviceToHost, d.stream));
    CU(cudaStreamSynchronize(d.stream));
    for (int r = 0; r < n_rows; ++r) {
        ctx->row_invalid_host[r] = (flags[r] & 1) != 0;
        ctx->row_first_invalid[r] = (flags[r] & 2) != 0;
    }
    A.tip_const_valid = false;
    ctx->have_aln = true;
    return PHB_OK;
}

int aa_get_alignment(phb_ctx *ctx, uint8_t *bytes, int32_t layout, int64_t first, int64_t cnt)
{
    AaState &A = aa_of(ctx);
    Device &d = ctx->devs[0];
    const int n_rows = ctx->n_rows;
    if (cnt == 0 || n_rows == 0) return PHB_OK;
    CU(cudaSetDevice(d.ordinal));
    CU(d.staging.ensure((size_t)cnt * n_rows));
    for (int r0 = 0; r0 < n_rows; r0 += 32768) {
        dim3 grid((unsigned)std::min<int64_t>((cnt + 255) / 256, 4096), (unsigned)std::min(32768, n_rows - r0));
        PHB_LAUNCH(aa_unpack_kernel, grid, dim3(256), d.stream, A.tips.as<uint8_t>(), n_rows, r0, A.P_pad, first, cnt, (int)layout,
                   d.staging.as<uint8_t>());
        CHECK_LAUNCH();
    }
    CU(cudaMemcpyAsync(bytes, d.staging.p, (size_t)cnt * n_rows, cudaMemcpyDeviceToHost, d.stream));
    CU(cudaStreamSynchronize(d.stream));
    return PHB_OK;
}

int aa_set_model(phb_ctx *ctx, const double *lambda, const double *V, const double *pi, const double *sqrt_pi,
                 const double *inv_sqrt_pi, int32_t n_cat, const double *rates, const double *weights)
{
    AaState &A = aa_of(ctx);
    Device &d = ctx->devs[0];
    CU(cudaSetDevice(d.ordinal));
    AaModelDev &M = A.M;
    memset(&M, 0, sizeof M);
    memcpy(M.lambda, lambda, sizeof M.lambda);
    memcpy(M.V, V, sizeof M.V);
    memcpy(M.pi, pi, sizeof M.pi);
    memcpy(M.sqrt_pi, sqrt_pi, sizeof M.sqrt_pi);
    memcpy(M.inv_sqrt_pi, inv_sqrt_pi, sizeof M.inv_sqrt_pi);
    for (int c = 0; c < n_cat; ++c) {
        M.rates[c] = rates[c];
        M.weights[c] = weights[c];
        M.ln_weights[c] = std::log(weights[c]);
    }
    M.n_cat = n_cat;
    CU(A.model.ensure(sizeof M));
    CU(cudaMemcpyAsync(A.model.p, &M, sizeof M, cudaMemcpyHostToDevice, d.stream));
    CU(cudaStreamSynchronize(d.stream));
    ctx->have_model = true;
    return PHB_OK;
}

// device tables of the tree: any-arity ops by level (TreeSched::uops), pre-order nodes by depth, branch lengths
int aa_upload_tree(phb_ctx *ctx)
{
    AaState &A = aa_of(ctx);
    const TreeSched &T = ctx->tree;
    Device &d = ctx->devs[0];
    CU(cudaSetDevice(d.ordinal));
    std::vector<std::vector<int>> by_depth(T.max_depth + 1);
    for (int v = 0; v < T.n; ++v) by_depth[T.depth[v]].push_back(v);
    A.down_nodes.clear();
    A.down_off.assign(1, 0);
    for (int dep = 0; dep <= T.max_depth; ++dep) {
        for (int v : by_depth[dep]) {
            AaDown nd;
            nd.node = v;
            nd.slot = T.slot[v];
            nd.row = T.slot[v] < 0 ? T.leaf_row[v] : -1;
            nd.parent_slot = v == T.root ? -1 : T.slot[T.parent[v]];
            A.down_nodes.push_back(nd);
        }
        A.down_off.push_back((int)A.down_nodes.size());
    }
    auto up = [&](DevBuf &b, const void *src, size_t bytes) -> cudaError_t {
        cudaError_t e = b.ensure(bytes);
        if (e != cudaSuccess) return e;
        if (bytes) return cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, d.stream);
        return cudaSuccess;
    };
    CU(up(A.blen, T.blen.data(), sizeof(double) * T.n));
    CU(up(A.uops, T.uops.data(), sizeof(GenOp) * T.uops.size()));
    CU(up(A.ukids, T.ukids.data(), sizeof(GenChild) * T.ukids.size()));
    CU(up(A.down, A.down_nodes.data(), sizeof(AaDown) * A.down_nodes.size()));
    CU(up(A.leaf_rows, T.leaf_rows.data(), sizeof(int) * T.leaf_rows.size()));
    CU(cudaStreamSynchronize(d.stream));
    A.tip_const_valid = false;
    ctx->tree_uploaded = true;
    return PHB_OK;
}

// compute_log_likelihood (likelihood.rs:156-209) / compute_marginal_asr (:224-408) at 20 states
int aa_evaluate(phb_ctx *ctx, bool asr, uint8_t *states_out, double *post_out, double *total)
{
    AaState &A = aa_of(ctx);
    const TreeSched &T = ctx->tree;
    Device &d = ctx->devs[0];
    if (!ctx->have_aln || !ctx->have_model || !T.valid)
        return ctx->fail(PHB_ERR_STATE, "set_alignment, set_model and set_tree must precede the evaluation");
    int rc = check_leaves(ctx);
    if (rc) return rc;
    if (!ctx->tree_uploaded && (rc = aa_upload_tree(ctx))) return rc; // after phb_score_nni / phb_score_trees
    CU(cudaSetDevice(d.ordinal));
    const int64_t P = ctx->n_patterns;
    const int K = A.M.n_cat;
    const int n_int = std::max(T.n_internal, 1);
    // tile of patterns: CLVs (160 K bytes per node and pattern) + exponents, states and posteriors when asked for
    size_t free_b = 0, total_b = 0;
    CU(cudaMemGetInfo(&free_b, &total_b));
    size_t budget = ctx->clv_budget ? (size_t)ctx->clv_budget : (free_b + A.clv.cap + A.exps.cap + A.states.cap + A.post.cap) / 10 * 7;
    size_t per_pattern_bytes = (size_t)n_int * K * (AA_NS * sizeof(double) + sizeof(int));
    if (asr) per_pattern_bytes += (size_t)T.n * (1 + (post_out ? AA_NS * sizeof(double) : 0));
    int64_t tile_pad = (int64_t)(budget / per_pattern_bytes) / AA_THREADS * AA_THREADS;
    tile_pad = std::max<int64_t>(AA_THREADS, std::min<int64_t>(tile_pad, A.P_pad));
    const size_t slot_stride = (size_t)K * AA_NS * (size_t)tile_pad;
    CU(A.clv.ensure(sizeof(double) * slot_stride * n_int));
    CU(A.exps.ensure(sizeof(int) * (size_t)n_int * K * tile_pad));
    CU(A.cat_ll.ensure(sizeof(double) * (size_t)(K + 1) * tile_pad));
    CU(A.per_pattern.ensure(sizeof(double) * (size_t)A.P_pad));
    CU(A.leaf_ls.ensure(sizeof(double) * (size_t)A.P_pad));
    CU(A.Pmat.ensure(sizeof(double) * AA_NS * AA_NS * (size_t)T.n * K));
    const int max_blocks = (int)(tile_pad / AA_THREADS);
    CU(A.partials.ensure(sizeof(double) * max_blocks));
    if (asr) {
        CU(A.states.ensure((size_t)T.n * tile_pad));
        if (post_out) CU(A.post.ensure(sizeof(double) * AA_NS * (size_t)T.n * tile_pad));
    }
    const AaModelDev *Md = A.model.as<AaModelDev>();
    PHB_LAUNCH(aa_transition_kernel, dim3((unsigned)(T.n * K)), dim3(AA_THREADS), d.stream, A.blen.as<double>(), T.n, Md, A.Pmat.as<double>());
    CHECK_LAUNCH();
    if (!A.tip_const_valid && P > 0) {
        PHB_LAUNCH(aa_tip_const_kernel, dim3((unsigned)((P + 255) / 256)), dim3(256), d.stream, A.leaf_rows.as<int>(),
                   (int)T.leaf_rows.size(), A.tips.as<uint8_t>(), A.P_pad, P, A.leaf_ls.as<double>());
        CHECK_LAUNCH();
        A.tip_const_valid = true;
    }
    const int root_slot = T.slot[T.root];
    const int root_row = root_slot < 0 ? T.leaf_row[T.root] : -1;
    std::vector<double> h_part;
    double sum = 0.0;
    ctx->stats.n_tiles = 0;
    for (int64_t off = 0; off < P; off += tile_pad) {
        const int tile_len = (int)std::min<int64_t>(tile_pad, P - off);
        const unsigned bx = (unsigned)((tile_len + AA_THREADS - 1) / AA_THREADS);
        for (int l = 1; l <= T.n_levels; ++l) {
            const int b = T.ulevel_off[l - 1], n_ops = T.ulevel_off[l] - b;
            for (int o = 0; o < n_ops; o += 65535) { // grid.y limit
                const int cnt = std::min(65535, n_ops - o);
                PHB_LAUNCH(aa_clv_kernel<false>, dim3(bx, (unsigned)cnt, (unsigned)((K + AA_CSTAGE - 1) / AA_CSTAGE)), dim3(AA_THREADS), d.stream, A.uops.as<GenOp>() + b + o,
                           A.ukids.as<GenChild>(), K, A.Pmat.as<double>(), A.clv.as<double>(), A.exps.as<int>(), slot_stride,
                           (size_t)tile_pad, tile_len, A.tips.as<uint8_t>(), A.P_pad, off, (const AaModelDev *)nullptr, (const double *)nullptr,
                           (const double *)nullptr, (double *)nullptr);
                CHECK_LAUNCH();
            }
        }
        PHB_LAUNCH(aa_root_kernel, dim3(bx), dim3(AA_THREADS), d.stream, root_slot, root_row, Md, A.clv.as<double>(), A.exps.as<int>(),
                   slot_stride, (size_t)tile_pad, tile_len, A.tips.as<uint8_t>(), A.P_pad, off, A.leaf_ls.as<double>(),
                   A.mult.as<double>(), A.per_pattern.as<double>(), A.cat_ll.as<double>(), A.partials.as<double>());
        CHECK_LAUNCH();
        if (asr) {
            for (int dep = 0; dep <= T.max_depth; ++dep) {
                const int b = A.down_off[dep], n_nd = A.down_off[dep + 1] - b;
                for (int o = 0; o < n_nd; o += 65535) {
                    const int cnt = std::min(65535, n_nd - o);
                    PHB_LAUNCH(aa_down_kernel, dim3((bx + AA_R - 1) / AA_R, (unsigned)cnt), dim3(AA_THREADS), d.stream, A.down.as<AaDown>() + b + o, Md,
                               A.Pmat.as<double>(), A.clv.as<double>(), slot_stride, (size_t)tile_pad, tile_len, A.tips.as<uint8_t>(),
                               A.P_pad, off, A.cat_ll.as<double>(), A.states.as<uint8_t>(), post_out ? A.post.as<double>() : (double *)nullptr);
                    CHECK_LAUNCH();
                }
            }
            // pageable destinations of some size go through the pinned bounce buffers (phylo_b200.cu d2h_rows_to_pageable)
            const bool bounce = states_out && (size_t)T.n * (size_t)P >= bounce_min_bytes() && host_is_pageable(states_out);
            if (states_out && bounce) {
                rc = d2h_rows_to_pageable(ctx, d, states_out + off, (size_t)P, A.states.as<uint8_t>(), (size_t)tile_pad, (size_t)tile_len,
                                          T.n, d.stream);
                if (rc) return rc;
            } else if (states_out)
                CU(cudaMemcpy2DAsync(states_out + off, (size_t)P, A.states.p, (size_t)tile_pad, (size_t)tile_len, (size_t)T.n,
                                     cudaMemcpyDeviceToHost, d.stream));
            if (post_out && bounce) {
                rc = d2h_rows_to_pageable(ctx, d, reinterpret_cast<uint8_t *>(post_out + (size_t)off * AA_NS), (size_t)P * AA_NS * sizeof(double),
                                          A.post.as<uint8_t>(), (size_t)tile_len * AA_NS * sizeof(double),
                                          (size_t)tile_len * AA_NS * sizeof(double), T.n, d.stream);
                if (rc) return rc;
            } else if (post_out)
                CU(cudaMemcpy2DAsync(post_out + (size_t)off * AA_NS, (size_t)P * AA_NS * sizeof(double), A.post.p,
                                     (size_t)tile_len * AA_NS * sizeof(double), (size_t)tile_len * AA_NS * sizeof(double), (size_t)T.n,
                                     cudaMemcpyDeviceToHost, d.stream));
        }
        h_part.resize(bx);
        CU(cudaMemcpyAsync(h_part.data(), A.partials.p, sizeof(double) * bx, cudaMemcpyDeviceToHost, d.stream));
        CU(cudaStreamSynchronize(d.stream));
        for (unsigned i = 0; i < bx; ++i) sum += h_part[i]; // fixed order: deterministic
        ctx->stats.n_tiles++;
    }
    CU(cudaStreamSynchronize(d.stream));
    ctx->stats.n_patterns = P;
    ctx->stats.tile_patterns = tile_pad;
    ctx->stats.n_levels = T.n_levels;
    ctx->stats.n_internal = T.n_internal;
    ctx->stats.n_clv_categories = K;
    ctx->stats.kernel_launches = ctx->launches;
    if (total) *total = sum;
    return PHB_OK;
}

// phb_score_nni at 20 states (caller loop: clone, NNI::nni src/tree.rs:1213-1242, log_likelihood).  An NNI at v re-links
// v and its parent p and changes no branch length.  With the base tree's CLVs of a pattern tile resident and the UPPER
// partial of every parent (aa_upper_kernel: one pre-order pass), a neighbour costs two node updates -- v without the
// moved child but with its former sibling, p without that sibling but with the moved child -- and one dot product of
// the new p with up(p).  The updates of a BATCH of neighbours run side by side through aa_clv_kernel (any arity, the
// child lists in the order the re-linked tree has them: the moved child / sibling appended last) into two scratch slots
// per neighbour; aa_nni_root_kernel forms each neighbour's value.
// PHB_AA_NNI_PATH=1 (cross-check): no upper partials, every neighbour re-evaluates its whole path v -> root instead --
// the same kernels on the same operands in the same order as a full re-scoring of the rebuilt tree (PHB_FLAG_NNI_FULL),
// bit-identical to it.
constexpr int AA_NNI_BATCH = 256;
int aa_score_nni(phb_ctx *ctx, int n_moves, const int32_t *node_id, const uint8_t *left_ch, double *totals)
{
    AaState &A = aa_of(ctx);
    const TreeSched &T = ctx->tree;
    Device &d = ctx->devs[0];
    int rc = check_leaves(ctx);
    if (rc) return rc;
    if (!ctx->tree_uploaded && (rc = aa_upload_tree(ctx))) return rc;
    CU(cudaSetDevice(d.ordinal));
    const int64_t P = ctx->n_patterns;
    const int K = A.M.n_cat;
    const int n_int = std::max(T.n_internal, 1);
    const char *path_env = getenv("PHB_AA_NNI_PATH");
    const bool use_upper = !(path_env && path_env[0] == '1');
    auto base_child = [&](int c) {
        GenChild g;
        g.is_tip = T.slot[c] < 0;
        g.src = g.is_tip ? T.leaf_row[c] : T.slot[c];
        g.edge = c;
        g.pad = 0;
        return g;
    };
    // ---- the valid moves
    std::vector<int> valid;
    for (int m = 0; m < n_moves; ++m) {
        totals[m] = NAN;
        const int v = node_id[m];
        if (v < 0 || v >= T.n || v == T.root || T.child_off[v + 1] - T.child_off[v] <= (left_ch[m] ? 1 : 0)) continue;
        const int p = T.parent[v];
        bool has_sib = false;
        for (int k = T.child_off[p]; k < T.child_off[p + 1]; ++k) has_sib |= T.children[k] != v;
        if (has_sib) valid.push_back(m);
    }
    ctx->stats.nni_path_moves = (int64_t)valid.size();
    if (valid.empty() || P == 0) {
        for (int m : valid) totals[m] = 0.0;
        return PHB_OK;
    }
    const int B = (int)std::min<size_t>(valid.size(), AA_NNI_BATCH);
    // slots: [0, n_int) the base tree's CLVs | pi | the upper partials by the node's slot | two per neighbour of a batch
    const int pi_slot = n_int, upper_base = n_int + 1, scratch_base = upper_base + (use_upper ? n_int : 0);
    // With upper partials and all categories in one block, v, p and the dot product are ONE launch per batch
    // (aa_nni_pair_kernel) and nothing of a neighbour is written: no scratch slots.  PHB_AA_NNI_NO_PAIR=1 (cross-check)
    // keeps v in a scratch slot and fuses only p with the dot product (aa_clv_kernel<true>).
    const char *nopair_env = getenv("PHB_AA_NNI_NO_PAIR");
    const bool fuse = use_upper && K <= AA_CSTAGE;
    const bool pair = fuse && !(nopair_env && nopair_env[0] == '1');
    const int n_slots = scratch_base + (pair ? 0 : 2 * B);
    // ---- upper ops by depth, for the parents the moves name and their ancestors
    std::vector<GenOp> ops;
    std::vector<GenChild> kids;
    std::vector<int> upper_off(1, 0);
    if (use_upper) {
        std::vector<char> need(T.n, 0);
        for (int m : valid)
            for (int a = T.parent[node_id[m]]; a != T.root && !need[a]; a = T.parent[a]) need[a] = 1;
        std::vector<std::vector<int>> by_depth(T.max_depth + 1);
        for (int x = 0; x < T.n; ++x)
            if (need[x]) by_depth[T.depth[x]].push_back(x);
        for (int dep = 1; dep <= T.max_depth; ++dep) {
            for (int x : by_depth[dep]) {
                const int g = T.parent[x];
                GenOp op;
                op.out_slot = upper_base + T.slot[x];
                op.child_begin = (int)kids.size();
                op.pad = x + 1;
                GenChild up;
                up.src = g == T.root ? pi_slot : upper_base + T.slot[g];
                up.edge = 0;
                up.is_tip = 0;
                up.pad = 1;
                kids.push_back(up);
                for (int k = T.child_off[g]; k < T.child_off[g + 1]; ++k)
                    if (T.children[k] != x) kids.push_back(base_child(T.children[k]));
                op.child_end = (int)kids.size();
                ops.push_back(op);
            }
            if ((int)ops.size() > upper_off.back()) upper_off.push_back((int)ops.size());
        }
    }
    // ---- batch by batch, the neighbours' ops grouped by step
    struct Batch {
        int first, count;           // range in `valid`
        std::vector<int> step_off;  // ops of step j: [step_off[j], step_off[j+1]) (absolute op indices)
    };
    std::vector<Batch> batches;
    std::vector<int> low_slots(valid.size()), up_slots(valid.size());
    const size_t n_upper_ops = ops.size();
    for (int b0 = 0; b0 < (int)valid.size(); b0 += B) {
        Batch bt;
        bt.first = b0;
        bt.count = std::min(B, (int)valid.size() - b0);
        int max_len = 0;
        std::vector<std::vector<int>> path(bt.count); // v, its parent (, ..., the root)
        for (int i = 0; i < bt.count; ++i) {
            const int v = node_id[valid[b0 + i]];
            if (use_upper)
                path[i] = {v, T.parent[v]};
            else
                for (int a = v; a >= 0; a = a == T.root ? -1 : T.parent[a]) path[i].push_back(a);
            max_len = std::max(max_len, (int)path[i].size());
        }
        for (int j = 0; j < max_len; ++j) {
            bt.step_off.push_back((int)ops.size());
            for (int i = 0; i < bt.count; ++i) {
                const int m = valid[b0 + i];
                const int v = node_id[m];
                if ((int)path[i].size() <= j) continue;
                const int p = T.parent[v];
                const int c1 = T.children[T.child_off[v] + (left_ch[m] ? 1 : 0)];
                int sib = -1;
                for (int k = T.child_off[p]; k < T.child_off[p + 1] && sib < 0; ++k)
                    if (T.children[k] != v) sib = T.children[k];
                const int scratch[2] = {scratch_base + 2 * i, scratch_base + 2 * i + 1};
                GenOp op;
                op.out_slot = scratch[j & 1];
                op.child_begin = (int)kids.size();
                op.pad = 0;
                if (j == 0) { // v: its children without c1, then the former sibling
                    for (int k = T.child_off[v]; k < T.child_off[v + 1]; ++k)
                        if (T.children[k] != c1) kids.push_back(base_child(T.children[k]));
                    kids.push_back(base_child(sib));
                } else {
                    const int a = path[i][j], below = path[i][j - 1];
                    for (int k = T.child_off[a]; k < T.child_off[a + 1]; ++k) {
                        const int c = T.children[k];
                        if (j == 1 && c == sib) continue; // the sibling went below v
                        GenChild g = base_child(c);
                        if (c == below) {
                            g.is_tip = 0;
                            g.src = scratch[(j - 1) & 1];
                        }
                        kids.push_back(g);
                    }
                    if (j == 1) kids.push_back(base_child(c1)); // ... and c1 came up
                }
                op.child_end = (int)kids.size();
                ops.push_back(op);
                if (j + 1 == (int)path[i].size()) {
                    low_slots[b0 + i] = op.out_slot;
                    up_slots[b0 + i] = (use_upper && p != T.root) ? upper_base + T.slot[p] : pi_slot;
                    ops.back().pad = up_slots[b0 + i] + 1; // read by aa_clv_kernel<true>
                }
            }
        }
        bt.step_off.push_back((int)ops.size());
        batches.push_back(std::move(bt));
    }
    ctx->stats.nni_path_steps = (int64_t)(ops.size() - n_upper_ops);
    // ---- buffers: the base tree's CLVs of a pattern tile + two scratch slots per neighbour of a batch
    size_t free_b = 0, total_b = 0;
    CU(cudaMemGetInfo(&free_b, &total_b));
    size_t budget = ctx->clv_budget ? (size_t)ctx->clv_budget : (free_b + A.clv.cap + A.exps.cap) / 10 * 7;
    const size_t per_pattern_bytes = (size_t)n_slots * K * (AA_NS * sizeof(double) + sizeof(int));
    int64_t tile_pad = (int64_t)(budget / per_pattern_bytes) / AA_THREADS * AA_THREADS;
    tile_pad = std::max<int64_t>(AA_THREADS, std::min<int64_t>(tile_pad, A.P_pad));
    const size_t slot_stride = (size_t)K * AA_NS * (size_t)tile_pad;
    const int max_blocks = (int)(tile_pad / AA_THREADS);
    CU(A.clv.ensure(sizeof(double) * slot_stride * n_slots));
    CU(A.exps.ensure(sizeof(int) * (size_t)n_slots * K * tile_pad));
    CU(A.leaf_ls.ensure(sizeof(double) * (size_t)A.P_pad));
    CU(A.Pmat.ensure(sizeof(double) * AA_NS * AA_NS * (size_t)T.n * K));
    CU(A.partials.ensure(sizeof(double) * (size_t)max_blocks * B));
    CU(A.nni_ops.ensure(sizeof(GenOp) * ops.size()));
    CU(A.nni_kids.ensure(sizeof(GenChild) * kids.size()));
    CU(A.nni_roots.ensure(sizeof(int) * 2 * valid.size()));
    CU(cudaMemcpyAsync(A.nni_ops.p, ops.data(), sizeof(GenOp) * ops.size(), cudaMemcpyHostToDevice, d.stream));
    CU(cudaMemcpyAsync(A.nni_kids.p, kids.data(), sizeof(GenChild) * kids.size(), cudaMemcpyHostToDevice, d.stream));
    CU(cudaMemcpyAsync(A.nni_roots.p, low_slots.data(), sizeof(int) * valid.size(), cudaMemcpyHostToDevice, d.stream));
    CU(cudaMemcpyAsync(A.nni_roots.as<int>() + valid.size(), up_slots.data(), sizeof(int) * valid.size(), cudaMemcpyHostToDevice, d.stream));
    const AaModelDev *Md = A.model.as<AaModelDev>();
    PHB_LAUNCH(aa_transition_kernel, dim3((unsigned)(T.n * K)), dim3(AA_THREADS), d.stream, A.blen.as<double>(), T.n, Md, A.Pmat.as<double>());
    CHECK_LAUNCH();
    if (!A.tip_const_valid) {
        PHB_LAUNCH(aa_tip_const_kernel, dim3((unsigned)((P + 255) / 256)), dim3(256), d.stream, A.leaf_rows.as<int>(),
                   (int)T.leaf_rows.size(), A.tips.as<uint8_t>(), A.P_pad, P, A.leaf_ls.as<double>());
        CHECK_LAUNCH();
        A.tip_const_valid = true;
    }
    const unsigned kz = (unsigned)((K + AA_CSTAGE - 1) / AA_CSTAGE);
    std::vector<double> sums(valid.size(), 0.0), h_part;
    ctx->stats.nni_tiles = 0;
    ctx->stats.nni_cache_bytes = (int64_t)(sizeof(double) * slot_stride * n_slots + sizeof(int) * (size_t)n_slots * K * tile_pad);
    for (int64_t off = 0; off < P; off += tile_pad) {
        const int tile_len = (int)std::min<int64_t>(tile_pad, P - off);
        const unsigned bx = (unsigned)((tile_len + AA_THREADS - 1) / AA_THREADS);
        auto clv_launch = [&](const GenOp *o, const GenChild *k, int cnt, bool dot = false) -> int {
            if (dot)
                PHB_LAUNCH(aa_clv_kernel<true>, dim3(bx, (unsigned)cnt, 1), dim3(AA_THREADS), d.stream, o, k, K, A.Pmat.as<double>(), A.clv.as<double>(),
                           A.exps.as<int>(), slot_stride, (size_t)tile_pad, tile_len, A.tips.as<uint8_t>(), A.P_pad, off, Md,
                           (const double *)A.leaf_ls.as<double>(), (const double *)A.mult.as<double>(), A.partials.as<double>());
            else
                PHB_LAUNCH(aa_clv_kernel<false>, dim3(bx, (unsigned)cnt, kz), dim3(AA_THREADS), d.stream, o, k, K, A.Pmat.as<double>(), A.clv.as<double>(),
                           A.exps.as<int>(), slot_stride, (size_t)tile_pad, tile_len, A.tips.as<uint8_t>(), A.P_pad, off, (const AaModelDev *)nullptr,
                           (const double *)nullptr, (const double *)nullptr, (double *)nullptr);
            CHECK_LAUNCH();
            return PHB_OK;
        };
        for (int l = 1; l <= T.n_levels; ++l) { // the base tree, as aa_evaluate does
            const int b = T.ulevel_off[l - 1], n_ops = T.ulevel_off[l] - b;
            for (int o = 0; o < n_ops; o += 65535)
                if ((rc = clv_launch(A.uops.as<GenOp>() + b + o, A.ukids.as<GenChild>(), std::min(65535, n_ops - o)))) return rc;
        }
        PHB_LAUNCH(aa_fill_pi_kernel, dim3((unsigned)std::min<size_t>(((size_t)K * AA_NS * tile_pad + 255) / 256, 2048)), dim3(256), d.stream, Md,
                   A.clv.as<double>(), A.exps.as<int>(), pi_slot, slot_stride, (size_t)tile_pad);
        CHECK_LAUNCH();
        for (size_t l = 0; l + 1 < upper_off.size(); ++l) // the upper partials, parents before children
            for (int o = upper_off[l]; o < upper_off[l + 1]; o += 65535) {
                PHB_LAUNCH(aa_upper_kernel, dim3(bx, (unsigned)std::min(65535, upper_off[l + 1] - o), (unsigned)K), dim3(AA_THREADS), d.stream,
                           A.nni_ops.as<GenOp>() + o, A.nni_kids.as<GenChild>(), K, A.Pmat.as<double>(), A.clv.as<double>(), A.exps.as<int>(),
                           slot_stride, (size_t)tile_pad, tile_len, A.tips.as<uint8_t>(), A.P_pad, off);
                CHECK_LAUNCH();
            }
        for (const Batch &bt : batches) {
            // with upper partials every neighbour has exactly two steps; the second one (p') is fused with the dot product
            if (pair) {
                PHB_LAUNCH(aa_nni_pair_kernel, dim3(bx, (unsigned)bt.count), dim3(AA_THREADS), d.stream, A.nni_ops.as<GenOp>() + bt.step_off[0],
                           A.nni_ops.as<GenOp>() + bt.step_off[1], A.nni_kids.as<GenChild>(), K, A.Pmat.as<double>(), A.clv.as<double>(),
                           A.exps.as<int>(), slot_stride, (size_t)tile_pad, tile_len, A.tips.as<uint8_t>(), A.P_pad, off, Md,
                           A.leaf_ls.as<double>(), A.mult.as<double>(), A.partials.as<double>());
                CHECK_LAUNCH();
            }
            for (size_t j = 0; !pair && j + 1 < bt.step_off.size(); ++j)
                if ((rc = clv_launch(A.nni_ops.as<GenOp>() + bt.step_off[j], A.nni_kids.as<GenChild>(), bt.step_off[j + 1] - bt.step_off[j],
                                     fuse && j == 1)))
                    return rc;
            if (!fuse)
                PHB_LAUNCH(aa_nni_root_kernel, dim3(bx, (unsigned)bt.count), dim3(AA_THREADS), d.stream, A.nni_roots.as<int>() + bt.first,
                       A.nni_roots.as<int>() + valid.size() + bt.first, Md,
                       A.clv.as<double>(), A.exps.as<int>(), slot_stride, (size_t)tile_pad, tile_len, off, A.leaf_ls.as<double>(),
                       A.mult.as<double>(), A.partials.as<double>());
            CHECK_LAUNCH();
            h_part.resize((size_t)bx * bt.count);
            CU(cudaMemcpyAsync(h_part.data(), A.partials.p, sizeof(double) * h_part.size(), cudaMemcpyDeviceToHost, d.stream));
            CU(cudaStreamSynchronize(d.stream));
            for (int i = 0; i < bt.count; ++i)
                for (unsigned x = 0; x < bx; ++x) sums[bt.first + i] += h_part[(size_t)i * bx + x]; // fixed order: deterministic
        }
        ctx->stats.nni_tiles++;
    }
    for (size_t i = 0; i < valid.size(); ++i) totals[valid[i]] = sums[i];
    ctx->stats.kernel_launches = ctx->launches;
    return PHB_OK;
}

// compute_joint_asr (likelihood.rs:421-560) at 20 states: states [n_nodes][n_patterns], *total = sum multiplicity * best
int aa_joint(phb_ctx *ctx, uint8_t *states_out, double *total)
{
    AaState &A = aa_of(ctx);
    const TreeSched &T = ctx->tree;
    Device &d = ctx->devs[0];
    if (!ctx->have_aln || !ctx->have_model || !T.valid)
        return ctx->fail(PHB_ERR_STATE, "set_alignment, set_model and set_tree must precede the evaluation");
    int rc = check_leaves(ctx);
    if (rc) return rc;
    if (!ctx->tree_uploaded && (rc = aa_upload_tree(ctx))) return rc; // after phb_score_nni / phb_score_trees
    CU(cudaSetDevice(d.ordinal));
    const int64_t P = ctx->n_patterns;
    const int K = A.M.n_cat;
    const int n_int = std::max(T.n_internal, 1);
    size_t free_b = 0, total_b = 0;
    CU(cudaMemGetInfo(&free_b, &total_b));
    size_t budget = ctx->clv_budget ? (size_t)ctx->clv_budget : (free_b + A.clv.cap + A.states.cap) / 10 * 7;
    const size_t per_pattern_bytes = (size_t)n_int * K * AA_NS * sizeof(double) + (size_t)T.n + 1;
    int64_t tile_pad = (int64_t)(budget / per_pattern_bytes) / AA_THREADS * AA_THREADS;
    tile_pad = std::max<int64_t>(AA_THREADS, std::min<int64_t>(tile_pad, A.P_pad));
    const size_t slot_stride = (size_t)K * AA_NS * (size_t)tile_pad;
    const size_t n_ent = (size_t)AA_NS * AA_NS * (size_t)T.n * K;
    CU(A.clv.ensure(sizeof(double) * slot_stride * n_int));
    CU(A.per_pattern.ensure(sizeof(double) * (size_t)A.P_pad));
    CU(A.Pmat.ensure(sizeof(double) * n_ent));
    CU(A.lnP.ensure(sizeof(double) * n_ent));
    CU(A.jbest.ensure((size_t)tile_pad));
    CU(A.states.ensure((size_t)T.n * tile_pad));
    CU(A.partials.ensure(sizeof(double) * (size_t)(tile_pad / AA_THREADS)));
    CU(A.parent.ensure(sizeof(int) * (size_t)T.n));
    CU(cudaMemcpyAsync(A.parent.p, T.parent.data(), sizeof(int) * (size_t)T.n, cudaMemcpyHostToDevice, d.stream));
    const AaModelDev *Md = A.model.as<AaModelDev>();
    PHB_LAUNCH(aa_transition_kernel, dim3((unsigned)(T.n * K)), dim3(AA_THREADS), d.stream, A.blen.as<double>(), T.n, Md, A.Pmat.as<double>());
    CHECK_LAUNCH();
    PHB_LAUNCH(log_transition_kernel, dim3((unsigned)((n_ent + 255) / 256)), dim3(256), d.stream, A.Pmat.as<double>(), n_ent, A.lnP.as<double>());
    CHECK_LAUNCH();
    const int root_slot = T.slot[T.root];
    const int root_row = root_slot < 0 ? T.leaf_row[T.root] : -1;
    const bool bounce = states_out && (size_t)T.n * (size_t)P >= bounce_min_bytes() && host_is_pageable(states_out);
    std::vector<double> h_part;
    double sum = 0.0;
    ctx->stats.n_tiles = 0;
    for (int64_t off = 0; off < P; off += tile_pad) {
        const int tile_len = (int)std::min<int64_t>(tile_pad, P - off);
        const unsigned bx = (unsigned)((tile_len + AA_THREADS - 1) / AA_THREADS);
        for (int l = 1; l <= T.n_levels; ++l) {
            const int b = T.ulevel_off[l - 1], n_ops = T.ulevel_off[l] - b;
            for (int o = 0; o < n_ops; o += 65535) {
                const int cnt = std::min(65535, n_ops - o);
                PHB_LAUNCH(aa_joint_up_kernel, dim3(bx, (unsigned)cnt, (unsigned)((K + AA_CSTAGE - 1) / AA_CSTAGE)), dim3(AA_THREADS), d.stream,
                           A.uops.as<GenOp>() + b + o, A.ukids.as<GenChild>(), K, A.lnP.as<double>(), A.clv.as<double>(), slot_stride,
                           (size_t)tile_pad, tile_len, A.tips.as<uint8_t>(), A.P_pad, off);
                CHECK_LAUNCH();
            }
        }
        PHB_LAUNCH(aa_joint_root_kernel, dim3(bx), dim3(AA_THREADS), d.stream, root_slot, root_row, T.root, Md, A.clv.as<double>(), slot_stride,
                   (size_t)tile_pad, tile_len, A.tips.as<uint8_t>(), A.P_pad, off, A.mult.as<double>(), A.per_pattern.as<double>(),
                   A.jbest.as<uint8_t>(), A.states.as<uint8_t>(), A.partials.as<double>());
        CHECK_LAUNCH();
        for (int dep = 1; dep <= T.max_depth; ++dep) {
            const int b = A.down_off[dep], n_nd = A.down_off[dep + 1] - b;
            for (int o = 0; o < n_nd; o += 65535) {
                const int cnt = std::min(65535, n_nd - o);
                PHB_LAUNCH(aa_joint_down_kernel, dim3(bx, (unsigned)cnt), dim3(AA_THREADS), d.stream, A.down.as<AaDown>() + b + o,
                           A.parent.as<int>(), K, A.lnP.as<double>(), A.clv.as<double>(), slot_stride, (size_t)tile_pad, tile_len,
                           A.tips.as<uint8_t>(), A.P_pad, off, A.jbest.as<uint8_t>(), A.states.as<uint8_t>());
                CHECK_LAUNCH();
            }
        }
        if (states_out && bounce) {
            rc = d2h_rows_to_pageable(ctx, d, states_out + off, (size_t)P, A.states.as<uint8_t>(), (size_t)tile_pad, (size_t)tile_len, T.n, d.stream);
            if (rc) return rc;
        } else if (states_out)
            CU(cudaMemcpy2DAsync(states_out + off, (size_t)P, A.states.p, (size_t)tile_pad, (size_t)tile_len, (size_t)T.n,
                                 cudaMemcpyDeviceToHost, d.stream));
        h_part.resize(bx);
        CU(cudaMemcpyAsync(h_part.data(), A.partials.p, sizeof(double) * bx, cudaMemcpyDeviceToHost, d.stream));
        CU(cudaStreamSynchronize(d.stream));
        for (unsigned i = 0; i < bx; ++i) sum += h_part[i];
        ctx->stats.n_tiles++;
    }
    CU(cudaStreamSynchronize(d.stream));
    ctx->stats.n_patterns = P;
    ctx->stats.tile_patterns = tile_pad;
    ctx->stats.kernel_launches = ctx->launches;
    if (total) *total = sum;
    return PHB_OK;
}

int aa_get_per_pattern(phb_ctx *ctx, double *per_pattern)
{
    AaState &A = aa_of(ctx);
    Device &d = ctx->devs[0];
    CU(cudaSetDevice(d.ordinal));
    if (A.per_pattern.cap < sizeof(double) * (size_t)std::max<int64_t>(ctx->n_patterns, 1))
        return ctx->fail(PHB_ERR_STATE, "no evaluation has run yet");
    if (ctx->n_patterns > 0)
        CU(cudaMemcpyAsync(per_pattern, A.per_pattern.p, sizeof(double) * (size_t)ctx->n_patterns, cudaMemcpyDeviceToHost, d.stream));
    CU(cudaStreamSynchronize(d.stream));
    return PHB_OK;
}

int aa_transition_matrices(phb_ctx *ctx, int32_t n_t, const double *t, double *out)
{
    AaState &A = aa_of(ctx);
    Device &d = ctx->devs[0];
    CU(cudaSetDevice(d.ordinal));
    const size_t n_out = (size_t)n_t * A.M.n_cat * AA_NS * AA_NS;
    CU(d.scratch.ensure(sizeof(double) * ((size_t)n_t + n_out)));
    double *d_t = d.scratch.as<double>(), *d_P = d_t + n_t;
    CU(cudaMemcpyAsync(d_t, t, sizeof(double) * n_t, cudaMemcpyHostToDevice, d.stream));
    PHB_LAUNCH(aa_transition_kernel, dim3((unsigned)(n_t * A.M.n_cat)), dim3(AA_THREADS), d.stream, d_t, (int)n_t,
               A.model.as<AaModelDev>(), d_P);
    CHECK_LAUNCH();
    CU(cudaMemcpyAsync(out, d_P, sizeof(double) * n_out, cudaMemcpyDeviceToHost, d.stream));
    CU(cudaStreamSynchronize(d.stream));
    return PHB_OK;
}

void aa_release(phb_ctx *ctx)
{
    AaState &A = aa_of(ctx);
    for (DevBuf *b : {&A.model, &A.tips, &A.mult, &A.leaf_ls, &A.per_pattern, &A.Pmat, &A.blen, &A.clv, &A.exps, &A.cat_ll, &A.partials,
                      &A.uops, &A.ukids, &A.down, &A.leaf_rows, &A.states, &A.post, &A.row_flags, &A.lnP, &A.jbest, &A.parent, &A.nni_ops,
                      &A.nni_kids, &A.nni_roots})
        b->release();
}

} // namespace
