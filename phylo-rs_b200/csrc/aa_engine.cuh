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
                PHB_LAUNCH(aa_clv_kernel, dim3(bx, (unsigned)cnt, (unsigned)((K + AA_CSTAGE - 1) / AA_CSTAGE)), dim3(AA_THREADS), d.stream, A.uops.as<GenOp>() + b + o,
                           A.ukids.as<GenChild>(), K, A.Pmat.as<double>(), A.clv.as<double>(), A.exps.as<int>(), slot_stride,
                           (size_t)tile_pad, tile_len, A.tips.as<uint8_t>(), A.P_pad, off);
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
                      &A.uops, &A.ukids, &A.down, &A.leaf_rows, &A.states, &A.post, &A.row_flags, &A.lnP, &A.jbest, &A.parent})
        b->release();
}

} // namespace
