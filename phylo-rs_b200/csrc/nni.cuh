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
//   nni_score_kernel        one CTA = (neighbour, 128 patterns); one thread = one pattern, all CLV
//                           categories in registers; reads one sibling CLV per step of the walk.
//                           CTAs are ordered pattern-block-major with neighbours sorted in DFS order
//                           of v, so the siblings near the root, which almost every walk reads, are
//                           served from the 126 MB L2 instead of HBM.
//   nni_reduce_kernel       fixed-order sum of a neighbour's per-block partials into its total.
#pragma once
#include "kernels.cuh"

namespace phb {

constexpr int NNI_THREADS = 128;

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
// ---------------------------------------------------------------------------------------------
template <int NC>
__device__ __forceinline__ void nni_matvec(const double *__restrict__ Pm, const double (&x)[4], double (&y)[4])
{
    double p[16];
    load_P(Pm, p);
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] = p[4 * i] * x[0] + p[4 * i + 1] * x[1] + p[4 * i + 2] * x[2] + p[4 * i + 3] * x[3];
}

// contribution of a subtree hanging off the walk: out[i] = sum_j P[i][j] L[j]; returns its exponent
__device__ __forceinline__ int nni_contrib(const double *__restrict__ Pm, bool is_tip, int src, int cat_clv, int n_clv,
                                           const double *__restrict__ clv, size_t slot_stride, size_t tile_pad,
                                           const int *__restrict__ exps, const uint8_t *__restrict__ tips,
                                           int64_t P_pad, int64_t tile_off, int64_t site, double (&out)[4])
{
    double p[16];
    load_P(Pm, p);
    if (is_tip) {
        const uint32_t m = code_mask(tips[(int64_t)src * P_pad + tile_off + site]);
#pragma unroll
        for (int i = 0; i < 4; ++i) out[i] = masked_row(p, i, m);
        return 0;
    }
    const double *s = clv + (size_t)src * slot_stride + (size_t)cat_clv * 4 * tile_pad + (size_t)site;
    const double x0 = s[0], x1 = s[tile_pad], x2 = s[2 * tile_pad], x3 = s[3 * tile_pad];
#pragma unroll
    for (int i = 0; i < 4; ++i) out[i] = p[4 * i] * x0 + p[4 * i + 1] * x1 + p[4 * i + 2] * x2 + p[4 * i + 3] * x3;
    return exps[((size_t)src * n_clv + cat_clv) * tile_pad + (size_t)site];
}

template <int NC>
__global__ void __launch_bounds__(NNI_THREADS)
nni_score_kernel(const NniMove *__restrict__ moves, int n_moves, const NniStep *__restrict__ steps, ModelDev M,
                 const double *__restrict__ Pmat, const double *__restrict__ clv, size_t slot_stride, size_t tile_pad,
                 int tile_len, const int *__restrict__ exps, const uint8_t *__restrict__ tips, int64_t P_pad,
                 int64_t tile_off, const double *__restrict__ leaf_ls /* + tile_off */,
                 const uint8_t *__restrict__ inv_mask /* + tile_off */, const double *__restrict__ mult /* + tile_off */,
                 double *__restrict__ partials /* [move][n_site_blocks] */, int n_site_blocks)
{
    const int mi = blockIdx.x % n_moves;
    const int sb = blockIdx.x / n_moves;
    const NniMove mv = moves[mi];
    const int64_t site = (int64_t)sb * NNI_THREADS + threadIdx.x;
    double contrib = 0.0;
    if (site < tile_len) {
        double msg[NC][4];
        int ex[NC];
        const int ncm = M.n_cat;
#pragma unroll
        for (int c = 0; c < NC; ++c)
            ex[c] = nni_contrib(Pmat + ((size_t)mv.a_edge * ncm + M.cat_of_clv[c]) * 16, mv.a_is_tip != 0, mv.a_src, c, NC, clv,
                                slot_stride, tile_pad, exps, tips, P_pad, tile_off, site, msg[c]);
        for (int k = mv.step_begin; k < mv.step_end; ++k) {
            const NniStep st = steps[k];
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                double sbv[4];
                ex[c] += nni_contrib(Pmat + ((size_t)st.sib_edge * ncm + M.cat_of_clv[c]) * 16, st.sib_is_tip != 0, st.sib_src, c,
                                     NC, clv, slot_stride, tile_pad, exps, tips, P_pad, tile_off, site, sbv);
                double v[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) v[i] = msg[c][i] * sbv[i];
                ex[c] += rescale_site_local(v[0], v[1], v[2], v[3]);
                if (st.up_edge >= 0) nni_matvec<NC>(Pmat + ((size_t)st.up_edge * ncm + M.cat_of_clv[c]) * 16, v, msg[c]);
                else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) msg[c][i] = v[i];
                }
            }
        }
        // root formula, as root_kernel
        double lls[MAX_CAT];
        const double lls_leaf = leaf_ls[site];
        double mx = -INFINITY;
        for (int c = 0; c < M.n_cat; ++c) {
            double mass = 0.0, ls = lls_leaf;
            const int cc = M.clv_idx[c];
            if (cc >= 0) {
#pragma unroll
                for (int q = 0; q < NC; ++q)
                    if (q == cc) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) mass += M.pi[i] * msg[q][i];
                        ls -= (double)ex[q] * LN2;
                    }
            } else {
                const uint32_t m = inv_mask[site];
#pragma unroll
                for (int i = 0; i < 4; ++i) mass += M.pi[i] * (((m >> i) & 1u) ? 1.0 : 0.0);
            }
            const double ll = M.ln_weights[c] + log(mass) + ls;
            lls[c] = ll;
            mx = fmax(mx, ll);
        }
        double site_ll;
        if (!isfinite(mx)) site_ll = mx;
        else {
            double s = 0.0;
            for (int c = 0; c < M.n_cat; ++c) s += exp(lls[c] - mx);
            site_ll = mx + log(s);
        }
        contrib = mult[site] * site_ll;
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

} // namespace phb
