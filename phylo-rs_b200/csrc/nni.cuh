// nni.cuh — path-only rescoring of NNI neighbours against cached conditional likelihoods.
//
// The reference's caller loop for topology search (BASELINE.json configs[4]) is
//     let mut t = base.clone(); t.nni(node_id, left_ch)?; t.log_likelihood(&model, &aln)?
// (NNI::nni, src/tree.rs:1213-1242; compute_log_likelihood, src/tree/likelihood.rs:156-209), i.e. a
// full pruning pass per neighbour.  An NNI around the edge (p -> v) changes the conditional
// likelihood vectors (CLVs) of v, p and the ancestors of p only (likelihood.rs:113-127 is a product
// over a node's children; every other subtree is untouched), so with the base tree's CLVs resident in
// HBM a neighbour costs one walk from v to the root:
//
//   v'  = (P_a L_a) (.) (P_s L_s)          a = the child of v that stays, s = the sibling of v
//   p'  = (P_v v')  (.) (P_c L_c)          c = the child of v that moves up (branch lengths travel
//   g'  = (P_p p')  (.) (P_u L_u) ...      with the child node), u = the sibling on the way up
//
// followed by the reference's root formula (likelihood.rs:135-140, :61-68, :205).
//
//   clv_cache_level_kernel  level-batched up pass that also stores, per (node, category, pattern),
//                           the exponent accumulated in the node's SUBTREE (exact power-of-two
//                           rescaling stands in for Profile::scale, src/tree/likelihood/profile.rs:20-39)
//   nni_score_kernel        one CTA = (neighbour, 128..512 patterns); a warp = one CLV category of 64
//                           patterns, a thread = a pattern pair; the walk's P matrices are staged in
//                           shared memory; reads one sibling CLV per step of the walk.
//                           CTAs are ordered pattern-block-major with neighbours sorted in DFS order
//                           of v, so the siblings near the root, which almost every walk reads, are
//                           served from the 126 MB L2 instead of HBM.
//   nni_reduce_kernel       fixed-order sum of a neighbour's per-block partials into its total.
#pragma once
#include "kernels.cuh"

namespace phb {


// One step of the walk: the CLV of the node being rebuilt is msg (.) (P[sib_edge] L_sib); unless it
// is the root (up_edge < 0) the next message is P[up_edge] times that CLV.
struct NniStep {
    int32_t sib_src;    // CLV slot (internal sibling) or alignment row (tip sibling)
    int32_t sib_edge;   // dense node id of the sibling == index of its edge's P matrix
    int32_t sib_is_tip;
    int32_t up_edge;    // dense node id of the rebuilt node (its edge to its parent), -1 at the root
};
struct NniMove {
    int32_t step_begin, step_end;
    int32_t a_src, a_edge, a_is_tip; // the child of v that stays: the walk's first message
    int32_t out;                     // index in the caller's totals
    int32_t pad[2];
};

__device__ __forceinline__ int rescale_site_local(double &v0, double &v1, double &v2, double &v3)
{
    double m = fmax(fmax(v0, v1), fmax(v2, v3));
    if (m < RESCALE_THRESHOLD && m > 0.0) {
        int s = -ilogb(m);
        if (s > 1000) s = 1000;
        double f = __longlong_as_double((long long)(1023 + s) << 52);
        v0 *= f; v1 *= f; v2 *= f; v3 *= f;
        return s;
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Up pass with subtree exponents (binary nodes).  Same arithmetic and layout as
// clv_update_level_kernel; exps[slot][cat][tile_pad] = exps[child 0] + exps[child 1] + own rescale.
// ---------------------------------------------------------------------------------------------
template <bool TIP0, bool TIP1>
__device__ __forceinline__ void clv_cache_pair(const NodeOp &op, int cat_clv, int cat_model, int n_cat_model, int n_clv,
                                               const double *__restrict__ Pmat, double *__restrict__ clv,
                                               size_t slot_stride, size_t tile_pad,
                                               const uint8_t *__restrict__ tips, int64_t P_pad, int64_t tile_off,
                                               int *__restrict__ exps, int64_t site)
{
    double p[16];
    double2 a[4], b[4];
    const size_t stride2 = tile_pad / 2;
    const size_t cat_off = (size_t)cat_clv * 4 * tile_pad + (size_t)site;
    const size_t exp_stride = (size_t)n_clv * tile_pad;
    const size_t exp_off = (size_t)cat_clv * tile_pad + (size_t)site;
    int ex = 0, ey = 0;
    load_P(Pmat + ((size_t)op.e0 * n_cat_model + cat_model) * 16, p);
    if (TIP0) contrib_tip(p, tips + (int64_t)op.c0 * P_pad + tile_off + site, a);
    else {
        contrib_internal(p, reinterpret_cast<const double2 *>(clv + (size_t)op.c0 * slot_stride + cat_off), stride2, a);
        const int2 e = *reinterpret_cast<const int2 *>(exps + (size_t)op.c0 * exp_stride + exp_off);
        ex += e.x; ey += e.y;
    }
    load_P(Pmat + ((size_t)op.e1 * n_cat_model + cat_model) * 16, p);
    if (TIP1) contrib_tip(p, tips + (int64_t)op.c1 * P_pad + tile_off + site, b);
    else {
        contrib_internal(p, reinterpret_cast<const double2 *>(clv + (size_t)op.c1 * slot_stride + cat_off), stride2, b);
        const int2 e = *reinterpret_cast<const int2 *>(exps + (size_t)op.c1 * exp_stride + exp_off);
        ex += e.x; ey += e.y;
    }
    double2 v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        v[i].x = a[i].x * b[i].x;
        v[i].y = a[i].y * b[i].y;
    }
    ex += rescale_site_local(v[0].x, v[1].x, v[2].x, v[3].x);
    ey += rescale_site_local(v[0].y, v[1].y, v[2].y, v[3].y);
    double2 *dst = reinterpret_cast<double2 *>(clv + (size_t)op.out_slot * slot_stride + cat_off);
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i * stride2] = v[i];
    int2 eo;
    eo.x = ex; eo.y = ey;
    *reinterpret_cast<int2 *>(exps + (size_t)op.out_slot * exp_stride + exp_off) = eo;
}

__global__ void __launch_bounds__(CLV_THREADS)
clv_cache_level_kernel(const NodeOp *__restrict__ ops, int n_ops, ModelDev M,
                       const double *__restrict__ Pmat, double *__restrict__ clv, size_t slot_stride,
                       size_t tile_pad, int tile_len, int blocks_per_row,
                       const uint8_t *__restrict__ tips, int64_t P_pad, int64_t tile_off,
                       int *__restrict__ exps)
{
    int sb = blockIdx.x % blocks_per_row;
    int rest = blockIdx.x / blocks_per_row;
    int cat_clv = rest % M.n_clv;
    int opi = rest / M.n_clv;
    if (opi >= n_ops) return;
    int64_t site = ((int64_t)sb * CLV_THREADS + threadIdx.x) * 2;
    if (site >= tile_len) return;
    const NodeOp op = ops[opi];
    const int cat_model = M.cat_of_clv[cat_clv];
    switch (op.kind & 3) {
    case 0: clv_cache_pair<false, false>(op, cat_clv, cat_model, M.n_cat, M.n_clv, Pmat, clv, slot_stride, tile_pad, tips, P_pad, tile_off, exps, site); break;
    case 1: clv_cache_pair<true, false>(op, cat_clv, cat_model, M.n_cat, M.n_clv, Pmat, clv, slot_stride, tile_pad, tips, P_pad, tile_off, exps, site); break;
    case 2: clv_cache_pair<false, true>(op, cat_clv, cat_model, M.n_cat, M.n_clv, Pmat, clv, slot_stride, tile_pad, tips, P_pad, tile_off, exps, site); break;
    default: clv_cache_pair<true, true>(op, cat_clv, cat_model, M.n_cat, M.n_clv, Pmat, clv, slot_stride, tile_pad, tips, P_pad, tile_off, exps, site); break;
    }
}

// ---------------------------------------------------------------------------------------------
// Neighbour scoring.  grid = n_site_blocks * n_moves, blockIdx.x = site_block * n_moves + move
// (moves pre-sorted by the host in DFS order of v: consecutive CTAs walk nearly the same path).
//
// CTA = NC * G warps (G = NNI_WARP_GROUPS[NC]); warp w works on CLV category w % NC for the 64 patterns
// of group w / NC, one pattern PAIR per thread (16-byte CLV loads).  The walk's P matrices (two per step)
// and step records are staged in shared memory, NNI_CHUNK steps at a time, so that the per-step matrix
// fetches are shared-memory broadcasts instead of L1 traffic; the sibling CLV of step k+1 is requested
// before the arithmetic of step k.  Categories meet in shared memory for the root's log-sum-exp.
// ---------------------------------------------------------------------------------------------
constexpr int NNI_CHUNK = 24;                       // steps staged per pass
__host__ __device__ constexpr int nni_groups(int nc) { return nc == 1 ? 8 : nc == 2 ? 4 : 2; }
__host__ __device__ constexpr int nni_threads(int nc) { return 32 * nc * nni_groups(nc); }
__host__ __device__ constexpr int nni_block_sites(int nc) { return 64 * nni_groups(nc); }

struct NniSib { // one sibling's data for a pattern pair, as loaded
    double2 x[4];
    int2 e;
    uchar2 code;
};

__device__ __forceinline__ void nni_load_sib(NniSib &o, bool is_tip, int src, int cat, int n_clv,
                                             const double *__restrict__ clv, size_t slot_stride, size_t tile_pad,
                                             const int *__restrict__ exps, const uint8_t *__restrict__ tips, int64_t P_pad,
                                             int64_t tile_off, int64_t site)
{
    if (is_tip) {
        o.code = *reinterpret_cast<const uchar2 *>(tips + (int64_t)src * P_pad + tile_off + site);
        o.e = make_int2(0, 0);
    } else {
        const double2 *s = reinterpret_cast<const double2 *>(clv + (size_t)src * slot_stride + (size_t)cat * 4 * tile_pad + (size_t)site);
        const size_t stride2 = tile_pad / 2;
#pragma unroll
        for (int j = 0; j < 4; ++j) o.x[j] = s[j * stride2];
        o.e = *reinterpret_cast<const int2 *>(exps + ((size_t)src * n_clv + cat) * tile_pad + (size_t)site);
    }
}

// out[i] = sum_j P[i][j] L[j] for both patterns; P (16 doubles) in shared memory
__device__ __forceinline__ void nni_apply_sib(const double *__restrict__ Ps, const NniSib &sb, bool is_tip, double2 (&out)[4])
{
    double p[16];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const double2 t = reinterpret_cast<const double2 *>(Ps)[i];
        p[2 * i] = t.x;
        p[2 * i + 1] = t.y;
    }
    if (is_tip) {
        const uint32_t mx = code_mask(sb.code.x), my = code_mask(sb.code.y);
#pragma unroll
        for (int i = 0; i < 4; ++i) out[i] = make_double2(masked_row(p, i, mx), masked_row(p, i, my));
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            out[i].x = p[4 * i] * sb.x[0].x + p[4 * i + 1] * sb.x[1].x + p[4 * i + 2] * sb.x[2].x + p[4 * i + 3] * sb.x[3].x;
            out[i].y = p[4 * i] * sb.x[0].y + p[4 * i + 1] * sb.x[1].y + p[4 * i + 2] * sb.x[2].y + p[4 * i + 3] * sb.x[3].y;
        }
    }
}

#ifdef PHB_WITH_WALK // the earlier walk design: built for the emulation harness (cross-check in the CPU tests) only
template <int NC, int MINB>
__global__ void __launch_bounds__(nni_threads(NC), MINB)
nni_score_kernel(const NniMove *__restrict__ moves, int n_moves, const NniStep *__restrict__ steps, ModelDev M,
                 const double *__restrict__ Pmat, const double *__restrict__ clv, size_t slot_stride, size_t tile_pad,
                 int tile_len, const int *__restrict__ exps, const uint8_t *__restrict__ tips, int64_t P_pad,
                 int64_t tile_off, const double *__restrict__ leaf_ls /* + tile_off */,
                 const uint8_t *__restrict__ inv_mask /* + tile_off */, const double *__restrict__ mult /* + tile_off */,
                 double *__restrict__ partials /* [move][n_site_blocks] */, int n_site_blocks)
{
    constexpr int BS = nni_block_sites(NC), NT = nni_threads(NC);
    __shared__ __align__(16) double sP[(2 * NNI_CHUNK + 1) * NC * 16]; // [0]: first message; then per step: sibling, up
    __shared__ NniStep sSteps[NNI_CHUNK];
    __shared__ double sLL[NC][BS];
    const int mi = blockIdx.x % n_moves;
    const int sb = blockIdx.x / n_moves;
    const NniMove mv = moves[mi];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cat = warp % NC, grp = warp / NC;
    const int cat_model = M.cat_of_clv[cat];
    const int ncm = M.n_cat;
    const int local = grp * 64 + lane * 2;
    const int64_t site = (int64_t)sb * BS + local;
    const bool live = site < tile_len; // tile_pad is even and the tip rows are padded: the pair load is always in bounds
    double2 msg[4];
    int2 ex = make_int2(0, 0);
    NniSib cur;
    // first message: the child of v that stays
    for (int i = threadIdx.x; i < NC * 16; i += NT) {
        const int c = i >> 4;
        sP[i] = Pmat[((size_t)mv.a_edge * ncm + M.cat_of_clv[c]) * 16 + (i & 15)];
    }
    if (live) nni_load_sib(cur, mv.a_is_tip != 0, mv.a_src, cat, NC, clv, slot_stride, tile_pad, exps, tips, P_pad, tile_off, site);
    __syncthreads();
    if (live) {
        nni_apply_sib(sP + cat * 16, cur, mv.a_is_tip != 0, msg);
        ex = cur.e;
    }
    for (int k0 = mv.step_begin; k0 < mv.step_end; k0 += NNI_CHUNK) {
        const int nk = min(NNI_CHUNK, mv.step_end - k0);
        __syncthreads(); // the previous chunk's matrices are no longer read
        if (threadIdx.x < nk) sSteps[threadIdx.x] = steps[k0 + threadIdx.x];
        for (int i = threadIdx.x; i < nk * 2 * NC * 16; i += NT) {
            const int e = i & 15, c = (i >> 4) % NC, which = (i / (16 * NC)) & 1, k = i / (32 * NC);
            const NniStep st = steps[k0 + k];
            const int edge = which ? st.up_edge : st.sib_edge;
            sP[NC * 16 + i] = edge >= 0 ? Pmat[((size_t)edge * ncm + M.cat_of_clv[c]) * 16 + e] : 0.0;
        }
        __syncthreads();
        if (live) {
            NniSib nxt;
            {
                const NniStep st = sSteps[0];
                nni_load_sib(cur, st.sib_is_tip != 0, st.sib_src, cat, NC, clv, slot_stride, tile_pad, exps, tips, P_pad, tile_off, site);
            }
            for (int k = 0; k < nk; ++k) {
                const NniStep st = sSteps[k];
                if (k + 1 < nk) { // request the next sibling before this step's arithmetic
                    const NniStep sn = sSteps[k + 1];
                    nni_load_sib(nxt, sn.sib_is_tip != 0, sn.sib_src, cat, NC, clv, slot_stride, tile_pad, exps, tips, P_pad, tile_off, site);
                }
                const double *Pk = sP + NC * 16 + (size_t)k * 2 * NC * 16;
                double2 sv[4];
                nni_apply_sib(Pk + cat * 16, cur, st.sib_is_tip != 0, sv);
                ex.x += cur.e.x;
                ex.y += cur.e.y;
                double vx[4], vy[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    vx[i] = msg[i].x * sv[i].x;
                    vy[i] = msg[i].y * sv[i].y;
                }
                ex.x += rescale_site_local(vx[0], vx[1], vx[2], vx[3]);
                ex.y += rescale_site_local(vy[0], vy[1], vy[2], vy[3]);
                if (st.up_edge >= 0) {
                    const double2 *Pu = reinterpret_cast<const double2 *>(Pk + NC * 16 + cat * 16);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const double2 a = Pu[2 * i], b = Pu[2 * i + 1];
                        msg[i].x = a.x * vx[0] + a.y * vx[1] + b.x * vx[2] + b.y * vx[3];
                        msg[i].y = a.x * vy[0] + a.y * vy[1] + b.x * vy[2] + b.y * vy[3];
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) msg[i] = make_double2(vx[i], vy[i]);
                }
                if (k + 1 < nk) cur = nxt;
            }
        }
    }
    // root formula, as root_kernel: this thread's category first, then one thread per pattern mixes them
    if (live) {
        double mx = 0.0, my = 0.0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            mx += M.pi[i] * msg[i].x;
            my += M.pi[i] * msg[i].y;
        }
        double lsx = leaf_ls[site], lsy = leaf_ls[site + 1];
        lsx -= (double)ex.x * LN2;
        lsy -= (double)ex.y * LN2;
        sLL[cat][local] = M.ln_weights[cat_model] + log(mx) + lsx;
        sLL[cat][local + 1] = M.ln_weights[cat_model] + log(my) + lsy;
    }
    __syncthreads();
    double contrib = 0.0;
    for (int q = threadIdx.x; q < BS; q += NT) {
        const int64_t s = (int64_t)sb * BS + q;
        if (s >= tile_len) continue;
        double lls[MAX_CAT];
        double mxl = -INFINITY;
        for (int c = 0; c < M.n_cat; ++c) {
            const int cc = M.clv_idx[c];
            double ll;
            if (cc >= 0) ll = sLL[cc][q];
            else {
                double mass = 0.0;
                const uint32_t m = inv_mask[s];
#pragma unroll
                for (int i = 0; i < 4; ++i) mass += M.pi[i] * (((m >> i) & 1u) ? 1.0 : 0.0);
                ll = M.ln_weights[c] + log(mass) + leaf_ls[s];
            }
            lls[c] = ll;
            mxl = fmax(mxl, ll);
        }
        double site_ll;
        if (!isfinite(mxl)) site_ll = mxl;
        else {
            double acc = 0.0;
            for (int c = 0; c < M.n_cat; ++c) acc += exp(lls[c] - mxl);
            site_ll = mxl + log(acc);
        }
        contrib += mult[s] * site_ll;
    }
    const double r = block_reduce_sum(contrib);
    if (threadIdx.x == 0) partials[(size_t)mi * n_site_blocks + sb] = r;
}

// totals[move.out] += sum of the move's partials (fixed order: deterministic); one block per move
__global__ void __launch_bounds__(ROOT_THREADS)
nni_reduce_kernel(const NniMove *__restrict__ moves, const double *__restrict__ partials, int n_site_blocks,
                  double *__restrict__ totals)
{
    const int mi = blockIdx.x;
    double acc = 0.0;
    for (int i = threadIdx.x; i < n_site_blocks; i += blockDim.x) acc += partials[(size_t)mi * n_site_blocks + i];
    const double r = block_reduce_sum(acc);
    if (threadIdx.x == 0) totals[moves[mi].out] += r;
}
#endif // PHB_WITH_WALK


// =============================================================================================
// Edge-local scoring with "upper" partials (the default for binary trees).
//
// upper_v[j] = P(all data outside the subtree of v, v in state j): one pre-order pass over the cached
// CLVs, after which a neighbour costs a constant number of steps irrespective of the depth of v.  For
// v with children k0, k1, parent p and sibling s, the two neighbours of the edge (p -> v) are
//     lc = 0 (k0 moves up): v' = (P_k1 L_k1)(.)(P_s L_s),  site L = sum_i W_p[i] (P_v v')_i (P_k0 L_k0)_i
//     lc = 1 (k1 moves up): v' = (P_k0 L_k0)(.)(P_s L_s),  site L = sum_i W_p[i] (P_v v')_i (P_k1 L_k1)_i
// with W_p = upper_p (pi at the root) — the same sums the reference's full pruning pass forms at p and
// above, regrouped (likelihood.rs:113-140); branch lengths travel with the child node (tree.rs:1213-1242).
// =============================================================================================

// Pre-order pass, one launch per depth.  One block = (parent, category, 512 patterns); a thread = a
// pattern pair; both children of the (binary) parent are served from one read of the parent's upper.
// clv / exps are indexed by the ROW the host gave a node's lower partial (DownChild::src; only the partials some
// scored edge or some upper partial reads have one), upper / uexps by the row of its upper partial: DownChild::pad =
// that row + 1, 0 = nobody reads the child's upper partial (a tip, a cherry, a node off every path from a scored
// edge to the root); DownOp::parent_slot = the parent's upper row, DownOp::pad = the parent's dense node id.
template <int MINB>
__global__ void __launch_bounds__(CLV_THREADS, MINB)
upper_level_kernel(const DownOp *__restrict__ ops, int n_ops, const DownChild *__restrict__ kids, int root_node, ModelDev M,
                   const double *__restrict__ Pmat, const double *__restrict__ clv, double *__restrict__ upper,
                   size_t slot_stride, size_t tile_pad, int tile_len, int blocks_per_row, const int *__restrict__ exps,
                   int *__restrict__ uexps, const uint8_t *__restrict__ tips, int64_t P_pad, int64_t tile_off)
{
    const int sb = blockIdx.x % blocks_per_row;
    const int rest = blockIdx.x / blocks_per_row;
    const int cat = rest % M.n_clv;
    const int opi = rest / M.n_clv;
    if (opi >= n_ops) return;
    const int64_t site = ((int64_t)sb * CLV_THREADS + threadIdx.x) * 2;
    if (site >= tile_len) return;
    const DownOp op = ops[opi];
    if (op.child_end - op.child_begin != 2) return; // binary trees only (the host checks)
    const DownChild c0 = kids[op.child_begin], c1 = kids[op.child_begin + 1];
    const bool want0 = c0.pad != 0, want1 = c1.pad != 0;
    if (!want0 && !want1) return;
    const int cat_model = M.cat_of_clv[cat];
    const size_t stride2 = tile_pad / 2;
    const size_t cat_off = (size_t)cat * 4 * tile_pad + (size_t)site;
    const size_t exp_off = (size_t)cat * tile_pad + (size_t)site;
    const size_t exp_stride = (size_t)M.n_clv * tile_pad;
    double2 base[4];
    int2 eb = make_int2(0, 0);
    if (op.pad == root_node) {
#pragma unroll
        for (int i = 0; i < 4; ++i) base[i] = make_double2(M.pi[i], M.pi[i]);
    } else {
        const double2 *u = reinterpret_cast<const double2 *>(upper + (size_t)op.parent_slot * slot_stride + cat_off);
#pragma unroll
        for (int i = 0; i < 4; ++i) base[i] = u[i * stride2];
        eb = *reinterpret_cast<const int2 *>(uexps + (size_t)op.parent_slot * exp_stride + exp_off);
    }
#pragma unroll
    for (int which = 0; which < 2; ++which) {
        const DownChild me = which ? c1 : c0, sib = which ? c0 : c1;
        if (!(which ? want1 : want0)) continue;
        double p[16];
        double2 sv[4];
        int2 e = eb;
        load_P(Pmat + ((size_t)sib.node * M.n_cat + cat_model) * 16, p);
        if (sib.is_tip) contrib_tip(p, tips + (int64_t)sib.src * P_pad + tile_off + site, sv);
        else {
            contrib_internal(p, reinterpret_cast<const double2 *>(clv + (size_t)sib.src * slot_stride + cat_off), stride2, sv);
            const int2 es = *reinterpret_cast<const int2 *>(exps + (size_t)sib.src * exp_stride + exp_off);
            e.x += es.x; e.y += es.y;
        }
        double wx[4], wy[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            wx[i] = base[i].x * sv[i].x;
            wy[i] = base[i].y * sv[i].y;
        }
        load_P(Pmat + ((size_t)me.node * M.n_cat + cat_model) * 16, p);
        double ox[4], oy[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) { // upper_me[j] = sum_i P_me[i][j] W[i]
            ox[j] = p[j] * wx[0] + p[4 + j] * wx[1] + p[8 + j] * wx[2] + p[12 + j] * wx[3];
            oy[j] = p[j] * wy[0] + p[4 + j] * wy[1] + p[8 + j] * wy[2] + p[12 + j] * wy[3];
        }
        e.x += rescale_site_local(ox[0], ox[1], ox[2], ox[3]);
        e.y += rescale_site_local(oy[0], oy[1], oy[2], oy[3]);
        double2 *dst = reinterpret_cast<double2 *>(upper + (size_t)(me.pad - 1) * slot_stride + cat_off);
#pragma unroll
        for (int j = 0; j < 4; ++j) dst[j * stride2] = make_double2(ox[j], oy[j]);
        *reinterpret_cast<int2 *>(uexps + (size_t)(me.pad - 1) * exp_stride + exp_off) = e;
    }
}

// Subtree exponents from the running counts the storing sweep leaves (sweep.cuh, SweepParams::exp_out):
// exps[row] = cnt[row] - cnt[pred[row]], pred = the row holding the count before the subtree's first op (-1: zero).
__global__ void nni_exp_rows_kernel(const int *__restrict__ cnt, const int *__restrict__ pred, int n_rows, size_t row_elems /* n_clv * tile_pad */,
                                    int *__restrict__ exps)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= row_elems) return;
    for (int r = 0; r < n_rows; ++r) {
        const int pr = pred[r];
        exps[(size_t)r * row_elems + i] = cnt[(size_t)r * row_elems + i] - (pr >= 0 ? cnt[(size_t)pr * row_elems + i] : 0);
    }
}

// One job = the edge (p -> v): up to two neighbours (out0: k0 moves up, out1: k1 moves up; -1 = not asked).
struct NniEdgeJob {
    int32_t k0_src, k0_edge, k0_is_tip;
    int32_t k1_src, k1_edge, k1_is_tip;
    int32_t s_src, s_edge, s_is_tip;
    int32_t v_edge;       // dense id of v (its edge to p)
    int32_t p_upper_slot; // row of p's upper partial, -1 = p is the root (W = pi)
    int32_t out0, out1;
    int32_t pad[3];
};

// 2^k for k <= 0 (0 below the normal range: such a term is far below one ulp of the sum)
__device__ __forceinline__ double pow2_nonpositive(int k)
{
    return k < -1022 ? 0.0 : __longlong_as_double((long long)(1023 + k) << 52);
}

template <int NC>
__global__ void __launch_bounds__(nni_threads(NC), 2)
nni_edge_kernel(const NniEdgeJob *__restrict__ jobs, int n_jobs, ModelDev M, const double *__restrict__ Pmat,
                const double *__restrict__ clv, const double *__restrict__ upper, size_t slot_stride, size_t tile_pad,
                int tile_len, const int *__restrict__ exps, const int *__restrict__ uexps, const uint8_t *__restrict__ tips,
                int64_t P_pad, int64_t tile_off, const double *__restrict__ leaf_ls, const uint8_t *__restrict__ inv_mask,
                const double *__restrict__ mult, double *__restrict__ partials /* [job][2][n_site_blocks] */, int n_site_blocks)
{
    constexpr int BS = nni_block_sites(NC), NT = nni_threads(NC);
    __shared__ __align__(16) double sP[4 * NC * 16]; // k0, k1, s, v
    __shared__ double sV[2][NC][BS]; // root sums per (move, category, site) and their power-of-two exponents:
    __shared__ int sE[2][NC][BS];    // true value = sV * 2^-sE
    const int ji = blockIdx.x % n_jobs;
    const int sb = blockIdx.x / n_jobs;
    const NniEdgeJob jb = jobs[ji];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cat = warp % NC, grp = warp / NC;
    const int cat_model = M.cat_of_clv[cat];
    const int ncm = M.n_cat;
    const int local = grp * 64 + lane * 2;
    const int64_t site = (int64_t)sb * BS + local;
    const bool live = site < tile_len;
    for (int i = threadIdx.x; i < 4 * NC * 16; i += NT) {
        const int e = i & 15, c = (i >> 4) % NC, which = i / (16 * NC);
        const int edge = which == 0 ? jb.k0_edge : which == 1 ? jb.k1_edge : which == 2 ? jb.s_edge : jb.v_edge;
        sP[i] = Pmat[((size_t)edge * ncm + M.cat_of_clv[c]) * 16 + e];
    }
    NniSib d0, d1, ds;
    double2 w[4];
    int2 ew = make_int2(0, 0);
    if (live) {
        nni_load_sib(d0, jb.k0_is_tip != 0, jb.k0_src, cat, NC, clv, slot_stride, tile_pad, exps, tips, P_pad, tile_off, site);
        nni_load_sib(d1, jb.k1_is_tip != 0, jb.k1_src, cat, NC, clv, slot_stride, tile_pad, exps, tips, P_pad, tile_off, site);
        nni_load_sib(ds, jb.s_is_tip != 0, jb.s_src, cat, NC, clv, slot_stride, tile_pad, exps, tips, P_pad, tile_off, site);
        if (jb.p_upper_slot >= 0) {
            const double2 *u = reinterpret_cast<const double2 *>(upper + (size_t)jb.p_upper_slot * slot_stride +
                                                                 (size_t)cat * 4 * tile_pad + (size_t)site);
            const size_t stride2 = tile_pad / 2;
#pragma unroll
            for (int i = 0; i < 4; ++i) w[i] = u[i * stride2];
            ew = *reinterpret_cast<const int2 *>(uexps + ((size_t)jb.p_upper_slot * NC + cat) * tile_pad + (size_t)site);
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) w[i] = make_double2(M.pi[i], M.pi[i]);
        }
    }
    __syncthreads();
    if (live) {
        double2 c0[4], c1[4], cs[4];
        nni_apply_sib(sP + (0 * NC + cat) * 16, d0, jb.k0_is_tip != 0, c0);
        nni_apply_sib(sP + (1 * NC + cat) * 16, d1, jb.k1_is_tip != 0, c1);
        nni_apply_sib(sP + (2 * NC + cat) * 16, ds, jb.s_is_tip != 0, cs);
        const double2 *Pv = reinterpret_cast<const double2 *>(sP + (3 * NC + cat) * 16);
#pragma unroll
        for (int mvi = 0; mvi < 2; ++mvi) {
            // mvi = 0: k0 moves up, k1 stays;  mvi = 1: k1 moves up, k0 stays
            const double2(&stay)[4] = mvi ? c0 : c1;
            const double2(&up)[4] = mvi ? c1 : c0;
            int ex = ew.x + d0.e.x + d1.e.x + ds.e.x, ey = ew.y + d0.e.y + d1.e.y + ds.e.y;
            double vx[4], vy[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                vx[i] = stay[i].x * cs[i].x;
                vy[i] = stay[i].y * cs[i].y;
            }
            ex += rescale_site_local(vx[0], vx[1], vx[2], vx[3]);
            ey += rescale_site_local(vy[0], vy[1], vy[2], vy[3]);
            double mx = 0.0, my = 0.0;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const double2 a = Pv[2 * i], b = Pv[2 * i + 1];
                const double tx = a.x * vx[0] + a.y * vx[1] + b.x * vx[2] + b.y * vx[3];
                const double ty = a.x * vy[0] + a.y * vy[1] + b.x * vy[2] + b.y * vy[3];
                mx += w[i].x * (tx * up[i].x);
                my += w[i].y * (ty * up[i].y);
            }
            sV[mvi][cat][local] = M.weights[cat_model] * mx;
            sV[mvi][cat][local + 1] = M.weights[cat_model] * my;
            sE[mvi][cat][local] = ex;
            sE[mvi][cat][local + 1] = ey;
        }
    }
    __syncthreads();
    // Categories are mixed in the linear domain after aligning their exponents — one logarithm per site and move
    // instead of a log per category plus a log-sum-exp:  ln( sum_c w_c v_c 2^(e_min - e_c) ) - e_min ln 2 + leaf_ls.
    // The rate-0 (+I) categories enter with exponent 0 and value weight x stationary mass of the shared states.
    double contrib0 = 0.0, contrib1 = 0.0;
    for (int q = threadIdx.x; q < BS; q += NT) {
        const int64_t sq = (int64_t)sb * BS + q;
        if (sq >= tile_len) continue;
        double inv_w = 0.0;
        if (M.n_clv < M.n_cat) {
            double mass = 0.0;
            const uint32_t m = inv_mask[sq];
#pragma unroll
            for (int i = 0; i < 4; ++i) mass += M.pi[i] * (((m >> i) & 1u) ? 1.0 : 0.0);
            for (int c = 0; c < M.n_cat; ++c)
                if (M.clv_idx[c] < 0) inv_w += M.weights[c] * mass;
        }
        const double ls = leaf_ls[sq], mu = mult[sq];
#pragma unroll
        for (int mvi = 0; mvi < 2; ++mvi) {
            int e_min = 0x7fffffff;
#pragma unroll
            for (int cc = 0; cc < NC; ++cc)
                if (!(sV[mvi][cc][q] <= 0.0)) e_min = min(e_min, sE[mvi][cc][q]); // positive or NaN
            if (inv_w > 0.0) e_min = min(e_min, 0);
            double acc = inv_w, site_ll;
            if (e_min == 0x7fffffff) { // nothing positive: ln 0 = -inf (a NaN propagates)
#pragma unroll
                for (int cc = 0; cc < NC; ++cc) acc += sV[mvi][cc][q];
                site_ll = log(acc) + ls;
            } else {
                acc = inv_w > 0.0 ? inv_w * pow2_nonpositive(e_min) : 0.0; // e_min <= 0 whenever inv_w > 0
#pragma unroll
                for (int cc = 0; cc < NC; ++cc)
                    if (!(sV[mvi][cc][q] <= 0.0)) acc += sV[mvi][cc][q] * pow2_nonpositive(e_min - sE[mvi][cc][q]);
                site_ll = log(acc) + (ls - (double)e_min * LN2);
            }
            if (mvi) contrib1 += mu * site_ll;
            else contrib0 += mu * site_ll;
        }
    }
    const double r0 = block_reduce_sum(contrib0);
    const double r1 = block_reduce_sum(contrib1);
    if (threadIdx.x == 0) {
        partials[((size_t)ji * 2 + 0) * n_site_blocks + sb] = r0;
        partials[((size_t)ji * 2 + 1) * n_site_blocks + sb] = r1;
    }
}

// totals[out[row]] += sum of row `row` of partials (fixed order); rows with out < 0 are skipped
__global__ void __launch_bounds__(ROOT_THREADS)
nni_reduce_rows_kernel(const int32_t *__restrict__ out, const double *__restrict__ partials, int n_site_blocks,
                       double *__restrict__ totals)
{
    const int row = blockIdx.x;
    const int o = out[row];
    double acc = 0.0;
    for (int i = threadIdx.x; i < n_site_blocks; i += blockDim.x) acc += partials[(size_t)row * n_site_blocks + i];
    const double r = block_reduce_sum(acc);
    if (threadIdx.x == 0 && o >= 0) totals[o] += r;
}

} // namespace phb
