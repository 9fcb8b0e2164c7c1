// joint.cuh — joint (Viterbi) ancestral reconstruction, JointAsr::joint_asr (src/tree/asr.rs:30-37,
// impl src/tree.rs:58-66), engine compute_joint_asr (src/tree/likelihood.rs:421-560).
//
// Same level-batched schedule and the same HBM buffers as the marginal pass, in the max-plus
// semiring and in log space (no rescaling is needed):
//
//   leaf      C_v[i] = ln profile_v[i]  (-inf where the profile is 0)                     :477-480
//   internal  C_v[i] = sum_children max_j ( ln P_c[i][j] + C_c[j] )                       :483-509
//   root      per category: first maximal i of ln pi_i + C_root[i]; cat_ll = ln w + max   :512-524
//             the category with the strictly greatest cat_ll wins (first on ties)         :526-527
//   traceback state_c = first maximal j of ln P_c[state_parent][j] + C_c[j]               :531-538
//
// The reference stores back-pointers per (edge, parent state) during the up pass; here the down
// pass re-evaluates the one row it needs (state_parent, winning category) from the child's C values,
// which are still in the CLV buffer — the same f64 additions in the same order, hence the same
// argmax — so no pointer table (n_nodes x K bytes per pattern) is ever written to HBM.
// When no category has a value > -inf, the reference leaves every node at state 0 (:458-459, :543-549).
#pragma once
#include "kernels.cuh"

namespace phb {

constexpr uint8_t JOINT_NO_CATEGORY = 0xFF;

// ln of every entry of the P(t) table (exact identity gives ln 1 = 0 and ln 0 = -inf; a tiny negative
// entry gives NaN, which never wins a `>` comparison — as in the reference).
__global__ void log_transition_kernel(const double *__restrict__ Pmat, size_t n, double *__restrict__ lnP)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) lnP[i] = log(Pmat[i]);
}

__device__ __forceinline__ double joint_tip_value(uint32_t code, int j, const ModelDev &M)
{
    const uint32_t m = code_mask(code);
    if (!((m >> j) & 1u)) return -INFINITY;
    return (code & 16u) ? 0.0 : M.ln_tip[__popc(m)];
}

// max_j (lnp[4*i + j] + c[j]) with the reference's sequential `val > local_max` semantics (NaN never wins)
__device__ __forceinline__ double joint_row_max(const double (&lnp)[16], int i, const double (&c)[4])
{
    double best = -INFINITY;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const double val = lnp[4 * i + j] + c[j];
        if (val > best) best = val;
    }
    return best;
}
__device__ __forceinline__ int joint_row_argmax(const double *__restrict__ lnp_row, const double (&c)[4])
{
    double best = -INFINITY;
    int arg = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const double val = lnp_row[j] + c[j];
        if (val > best) {
            best = val;
            arg = j;
        }
    }
    return arg;
}

// Up pass, any arity.  One block = (node, category, 512 patterns); one thread = a pattern pair.
__global__ void __launch_bounds__(CLV_THREADS)
joint_up_level_kernel(const GenOp *__restrict__ ops, int n_ops, const GenChild *__restrict__ kids, ModelDev M,
                      const double *__restrict__ lnP, double *__restrict__ cbuf, size_t slot_stride,
                      size_t tile_pad, int tile_len, int blocks_per_row, const uint8_t *__restrict__ tips,
                      int64_t P_pad, int64_t tile_off)
{
    const int sb = blockIdx.x % blocks_per_row;
    const int rest = blockIdx.x / blocks_per_row;
    const int cat = rest % M.n_cat;
    const int opi = rest / M.n_cat;
    if (opi >= n_ops) return;
    const int64_t site = ((int64_t)sb * CLV_THREADS + threadIdx.x) * 2;
    if (site >= tile_len) return;
    const GenOp op = ops[opi];
    const size_t stride2 = tile_pad / 2;
    const size_t cat_off = (size_t)cat * 4 * tile_pad + (size_t)site;
    double vx[4] = {0.0, 0.0, 0.0, 0.0}, vy[4] = {0.0, 0.0, 0.0, 0.0};
    for (int k = op.child_begin; k < op.child_end; ++k) {
        const GenChild ch = kids[k];
        double lnp[16];
        load_P(lnP + ((size_t)ch.edge * M.n_cat + cat) * 16, lnp);
        double cx[4], cy[4];
        if (ch.is_tip) {
            const uchar2 code = *reinterpret_cast<const uchar2 *>(tips + (int64_t)ch.src * P_pad + tile_off + site);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                cx[j] = joint_tip_value(code.x, j, M);
                cy[j] = joint_tip_value(code.y, j, M);
            }
        } else {
            const double2 *src = reinterpret_cast<const double2 *>(cbuf + (size_t)ch.src * slot_stride + cat_off);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const double2 t = src[j * stride2];
                cx[j] = t.x;
                cy[j] = t.y;
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            vx[i] += joint_row_max(lnp, i, cx);
            vy[i] += joint_row_max(lnp, i, cy);
        }
    }
    double2 *dst = reinterpret_cast<double2 *>(cbuf + (size_t)op.out_slot * slot_stride + cat_off);
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i * stride2] = make_double2(vx[i], vy[i]);
}

// Binary nodes (the common case): both children's loads are issued before any arithmetic.
template <bool TIP0, bool TIP1>
__device__ __forceinline__ void joint_up_pair(const NodeOp &op, int cat, const ModelDev &M, const double *__restrict__ lnP,
                                              double *__restrict__ cbuf, size_t slot_stride, size_t tile_pad,
                                              const uint8_t *__restrict__ tips, int64_t P_pad, int64_t tile_off, int64_t site)
{
    const size_t stride2 = tile_pad / 2;
    const size_t cat_off = (size_t)cat * 4 * tile_pad + (size_t)site;
    double ax[4], ay[4], bx[4], by[4];
    if (TIP0) {
        const uchar2 code = *reinterpret_cast<const uchar2 *>(tips + (int64_t)op.c0 * P_pad + tile_off + site);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            ax[j] = joint_tip_value(code.x, j, M);
            ay[j] = joint_tip_value(code.y, j, M);
        }
    } else {
        const double2 *src = reinterpret_cast<const double2 *>(cbuf + (size_t)op.c0 * slot_stride + cat_off);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const double2 t = src[j * stride2];
            ax[j] = t.x;
            ay[j] = t.y;
        }
    }
    if (TIP1) {
        const uchar2 code = *reinterpret_cast<const uchar2 *>(tips + (int64_t)op.c1 * P_pad + tile_off + site);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            bx[j] = joint_tip_value(code.x, j, M);
            by[j] = joint_tip_value(code.y, j, M);
        }
    } else {
        const double2 *src = reinterpret_cast<const double2 *>(cbuf + (size_t)op.c1 * slot_stride + cat_off);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const double2 t = src[j * stride2];
            bx[j] = t.x;
            by[j] = t.y;
        }
    }
    double lnp[16];
    double vx[4], vy[4];
    load_P(lnP + ((size_t)op.e0 * M.n_cat + cat) * 16, lnp);
#pragma unroll
    for (int i = 0; i < 4; ++i) { // the reference starts from 0.0 and adds child by child (likelihood.rs:483, :505)
        vx[i] = 0.0 + joint_row_max(lnp, i, ax);
        vy[i] = 0.0 + joint_row_max(lnp, i, ay);
    }
    load_P(lnP + ((size_t)op.e1 * M.n_cat + cat) * 16, lnp);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        vx[i] += joint_row_max(lnp, i, bx);
        vy[i] += joint_row_max(lnp, i, by);
    }
    double2 *dst = reinterpret_cast<double2 *>(cbuf + (size_t)op.out_slot * slot_stride + cat_off);
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i * stride2] = make_double2(vx[i], vy[i]);
}

__global__ void __launch_bounds__(CLV_THREADS)
joint_up_binary_kernel(const NodeOp *__restrict__ ops, int n_ops, ModelDev M, const double *__restrict__ lnP,
                       double *__restrict__ cbuf, size_t slot_stride, size_t tile_pad, int tile_len, int blocks_per_row,
                       const uint8_t *__restrict__ tips, int64_t P_pad, int64_t tile_off)
{
    const int sb = blockIdx.x % blocks_per_row;
    const int rest = blockIdx.x / blocks_per_row;
    const int cat = rest % M.n_cat;
    const int opi = rest / M.n_cat;
    if (opi >= n_ops) return;
    const int64_t site = ((int64_t)sb * CLV_THREADS + threadIdx.x) * 2;
    if (site >= tile_len) return;
    const NodeOp op = ops[opi];
    switch (op.kind & 3) {
    case 0: joint_up_pair<false, false>(op, cat, M, lnP, cbuf, slot_stride, tile_pad, tips, P_pad, tile_off, site); break;
    case 1: joint_up_pair<true, false>(op, cat, M, lnP, cbuf, slot_stride, tile_pad, tips, P_pad, tile_off, site); break;
    case 2: joint_up_pair<false, true>(op, cat, M, lnP, cbuf, slot_stride, tile_pad, tips, P_pad, tile_off, site); break;
    default: joint_up_pair<true, true>(op, cat, M, lnP, cbuf, slot_stride, tile_pad, tips, P_pad, tile_off, site); break;
    }
}

// Root: winning category, root state, per-pattern value, weighted partial sums.
// root_is_tip: the tree is a single leaf (its C values come from the tip code).
__global__ void __launch_bounds__(ROOT_THREADS)
joint_root_kernel(const double *__restrict__ root_c /* slot base */, size_t tile_pad, int tile_len, ModelDev M,
                  const uint8_t *__restrict__ root_tip_row /* + tile_off, or null */,
                  const double *__restrict__ mult /* + tile_off */, double *__restrict__ per_pattern /* + tile_off */,
                  uint8_t *__restrict__ best_cat /* [tile_pad] */, uint8_t *__restrict__ root_states /* + tile_off */,
                  double *__restrict__ partials)
{
    const int64_t site = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    double contrib = 0.0;
    if (site < tile_len) {
        double ln_pi[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) ln_pi[i] = log(M.pi[i]);
        double best = -INFINITY;
        int bcat = JOINT_NO_CATEGORY, bstate = 0;
        for (int c = 0; c < M.n_cat; ++c) {
            double root_max = -INFINITY;
            int root_best = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const double cv = root_tip_row ? joint_tip_value(root_tip_row[site], i, M)
                                               : root_c[((size_t)c * 4 + i) * tile_pad + site];
                const double ll = ln_pi[i] + cv;
                if (ll > root_max) {
                    root_max = ll;
                    root_best = i;
                }
            }
            const double cat_ll = M.ln_weights[c] + root_max;
            if (cat_ll > best) {
                best = cat_ll;
                bcat = c;
                bstate = root_best;
            }
        }
        per_pattern[site] = best;
        best_cat[site] = (uint8_t)bcat;
        root_states[site] = (uint8_t)bstate;
        contrib = mult[site] * best;
    }
    const double r = block_reduce_sum(contrib);
    if (threadIdx.x == 0) partials[blockIdx.x] = r;
}

// Traceback, level-batched from the root.  One block = (parent node, 256 patterns); one thread = a pattern.
// DownOp::pad carries the parent's dense node id (row of `states`).
__global__ void __launch_bounds__(ROOT_THREADS)
joint_down_level_kernel(const DownOp *__restrict__ ops, int n_ops, const DownChild *__restrict__ kids, ModelDev M,
                        const double *__restrict__ lnP, const double *__restrict__ cbuf, size_t slot_stride,
                        size_t tile_pad, int tile_len, int blocks_per_row, const uint8_t *__restrict__ tips,
                        int64_t P_pad, int64_t tile_off, const uint8_t *__restrict__ best_cat,
                        uint8_t *__restrict__ states /* [node][P_pad] */)
{
    const int sb = blockIdx.x % blocks_per_row;
    const int opi = blockIdx.x / blocks_per_row;
    if (opi >= n_ops) return;
    const int64_t site = (int64_t)sb * blockDim.x + threadIdx.x;
    if (site >= tile_len) return;
    const DownOp op = ops[opi];
    const int cat = best_cat[site];
    const int s_v = states[(int64_t)op.pad * P_pad + tile_off + site];
    for (int k = op.child_begin; k < op.child_end; ++k) {
        const DownChild ch = kids[k];
        int st = 0;
        if (cat != JOINT_NO_CATEGORY) {
            double c[4];
            if (ch.is_tip) {
                const uint32_t code = tips[(int64_t)ch.src * P_pad + tile_off + site];
#pragma unroll
                for (int j = 0; j < 4; ++j) c[j] = joint_tip_value(code, j, M);
            } else {
                const double *src = cbuf + (size_t)ch.src * slot_stride + (size_t)cat * 4 * tile_pad + (size_t)site;
#pragma unroll
                for (int j = 0; j < 4; ++j) c[j] = src[(size_t)j * tile_pad];
            }
            st = joint_row_argmax(lnP + ((size_t)ch.node * M.n_cat + cat) * 16 + 4 * s_v, c);
        }
        states[(int64_t)ch.node * P_pad + tile_off + site] = (uint8_t)st;
    }
}

} // namespace phb
