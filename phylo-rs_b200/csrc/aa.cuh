// aa.cuh — the 20-state (AminoAcid, src/alphabet.rs:66-116) instance of the engines: SURVEY §8 f4.
//
// The reference's engines are generic over A: Alphabet (likelihood.rs:156-166, 224-235; GtrModel<A>, gtr.rs:30-58) and
// the nucleotide kernels of kernels.cuh / sweep.cuh are specialised to 4 states down to their register layout, so the
// 20-state instance is a set of kernels of its own, level-batched with every CLV in HBM:
//   aa_transition_kernel   P(r_c t) per (edge, category), 20x20          rate_matrix.rs:113-135
//   aa_tip_const_kernel    sum of the leaves' Profile::scale() log-scales likelihood.rs:108 (B, Z, J: ln 0.5 each)
//   aa_clv_kernel          one level of prune_pattern_category            likelihood.rs:110-131
//   aa_root_kernel         root mass, ln weights, log_sum_exp, weighting  likelihood.rs:135-140, 61-68, 205-206
//   aa_down_kernel         one depth of the marginal pre-order pass       likelihood.rs:303-390
//   aa_joint_*_kernel      joint (Viterbi) reconstruction                 likelihood.rs:421-560
//   aa_pack / aa_gather_patterns / aa_unpack_kernel   bytes <-> tip codes (set_alignment, compress_*, get_alignment)
//   aa_upper / aa_nni_pair / aa_nni_root_kernel, aa_clv_kernel<true>   NNI neighbours against resident CLVs and upper
//                          partials (caller loop: NNI::nni tree.rs:1213-1242 + log_likelihood)
// Work split: one thread = one site pattern of one rate category (all categories in the down pass, where the mixture
// over categories is thread-local); a block works on ONE (node, category), so every P(t) it touches is block-uniform
// and is staged once in shared memory and read by broadcast.  Per internal child that is a 20x20 mat-vec per thread
// (400 DFMA against 160 bytes of CLV read): the kernel is FP64-bound, not HBM-bound, at 20 states.
// CLV layout: clv[slot][category][state][pattern of the tile] (pattern fastest: coalesced), one int32 exponent per
// (slot, category, pattern); a stored CLV has its maximum in [1,2) (exact power-of-two scaling, as everywhere else).
#pragma once
#include "kernels.cuh"

namespace phb {

constexpr int AA_NS = 20;
constexpr int AA_THREADS = 128;
// tip codes: 0..19 = position in AminoAcid::CANONICAL "ACDEFGHIKLMNPQRSTVWY" (alphabet.rs:73-79)
constexpr uint8_t AA_CODE_B = 20, AA_CODE_Z = 21, AA_CODE_J = 22, AA_CODE_X = 23, AA_CODE_GAP = 24;

// AminoAcid::profile's accepted bytes (alphabet.rs:85-115); anything else is CODE_INVALID
__host__ __device__ inline uint8_t aa_encode_byte(uint8_t c)
{
    const char *canonical = "ACDEFGHIKLMNPQRSTVWY";
    for (int i = 0; i < AA_NS; ++i)
        if ((uint8_t)canonical[i] == c) return (uint8_t)i;
    switch (c) {
    case 'B': return AA_CODE_B;
    case 'Z': return AA_CODE_Z;
    case 'J': return AA_CODE_J;
    case 'X': return AA_CODE_X;
    case '-': return AA_CODE_GAP;
    default: return CODE_INVALID;
    }
}
__host__ __device__ inline uint8_t aa_decode_code(uint8_t code)
{
    const char *all = "ACDEFGHIKLMNPQRSTVWYBZJX-";
    return code <= AA_CODE_GAP ? (uint8_t)all[code] : (uint8_t)'?';
}
// The two states of a two-state ambiguity code, exactly as the reference writes them (alphabet.rs:91-112): B sets
// indices 3 and 11, Z 4 and 14, J 8 and 9 -- the reference's comments name D|N, E|Q, I|L, but its CANONICAL order puts
// those letters elsewhere; parity is with the indices the code uses (and its own test asserts: integration_test.rs:90-102).
__host__ __device__ inline void aa_pair_states(uint32_t code, int &a, int &b)
{
    a = code == AA_CODE_B ? 3 : code == AA_CODE_Z ? 4 : 8;
    b = code == AA_CODE_B ? 11 : code == AA_CODE_Z ? 14 : 9;
}

// Model in global memory (4.4 KB: beyond what is comfortable as a by-value kernel parameter).
struct AaModelDev {
    double lambda[AA_NS];
    double V[AA_NS * AA_NS]; // column-major: V[i + 20 k] = component i of eigenvector k
    double pi[AA_NS];
    double sqrt_pi[AA_NS];
    double inv_sqrt_pi[AA_NS];
    double rates[MAX_CAT];
    double weights[MAX_CAT];
    double ln_weights[MAX_CAT];
    int n_cat;
    int pad;
};

// The up pass uses kernels.cuh's any-arity GenOp / GenChild tables (TreeSched::uops, grouped by level).
// One node of one depth of the pre-order pass.
struct AaDown {
    int32_t node;        // dense id: row of the states / posteriors output
    int32_t slot;        // CLV slot of the node (internal), -1 for a leaf
    int32_t row;         // alignment row of a leaf, -1 otherwise
    int32_t parent_slot; // CLV slot holding the parent's posterior, -1 for the root
};

// ---- alignment bytes -> tip codes, transposed to [row][P_pad] -----------------------------------------------------
// raw: a chunk of the caller's bytes as they came ([pattern][row] when pattern_major, else [row][pattern]), np patterns
// starting at pattern p0 (pattern-major) or all P patterns of rows r0.. (row-major: rows_in_chunk rows).  32 x 32 tiles
// through shared memory so that both the reads and the writes are coalesced.  row_flags[r] |= 1 when the row holds a byte
// AminoAcid::profile rejects, |= 2 when that byte is the row's first pattern (error precedence, check_leaves).
__global__ void aa_pack_kernel(const uint8_t *__restrict__ raw, int pattern_major, int n_rows, int64_t np, int64_t p0, int r0,
                               int rows_in_chunk, uint8_t *__restrict__ tips, int64_t P_pad, int *__restrict__ row_flags)
{
    __shared__ uint8_t tile[32][33];
    const int64_t pb = (int64_t)blockIdx.x * 32; // pattern offset inside the chunk
    const int rb = blockIdx.y * 32;              // row offset inside the chunk
    const int tx = threadIdx.x, ty = threadIdx.y; // 32 x 8
    const int nr = pattern_major ? n_rows : rows_in_chunk;
    if (pattern_major) { // consecutive bytes = consecutive rows of one pattern
        for (int k = ty; k < 32; k += 8) {
            const int64_t p = pb + k;
            const int r = rb + tx;
            tile[k][tx] = (p < np && r < nr) ? raw[(size_t)p * n_rows + r] : (uint8_t)'-';
        }
        __syncthreads();
        for (int k = ty; k < 32; k += 8) {
            const int r = rb + k;
            const int64_t p = pb + tx;
            if (r < nr && p < np) {
                const uint8_t c = aa_encode_byte(tile[tx][k]);
                tips[(size_t)r * P_pad + p0 + p] = c;
                if (c == CODE_INVALID) atomicOr(reinterpret_cast<unsigned *>(row_flags + r), (p0 + p == 0) ? 3u : 1u);
            }
        }
    } else { // consecutive bytes = consecutive patterns of one row: no transposition
        for (int k = ty; k < 32; k += 8) {
            const int r = rb + k;
            const int64_t p = pb + tx;
            if (r < nr && p < np) {
                const uint8_t c = aa_encode_byte(raw[(size_t)r * np + p]);
                tips[(size_t)(r0 + r) * P_pad + p] = c;
                if (c == CODE_INVALID) atomicOr(reinterpret_cast<unsigned *>(row_flags + r0 + r), p == 0 ? 3u : 1u);
            }
        }
    }
}
// The unique columns of a compressed alignment (compress.cuh: raw[row][site] at `pitch`, replist[p] = first site of
// pattern p) -> tip codes [row][P_pad].  blockIdx.y = row; consecutive threads = consecutive patterns (first-appearance
// order, so replist ascends and the reads of one row stay near each other).
__global__ void aa_gather_patterns_kernel(const uint8_t *__restrict__ raw, int64_t pitch, const uint32_t *__restrict__ replist,
                                          int64_t P, int r0, uint8_t *__restrict__ tips, int64_t P_pad, int *__restrict__ row_flags)
{
    const int r = r0 + blockIdx.y;
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
        const uint8_t c = aa_encode_byte(raw[(size_t)r * pitch + replist[p]]);
        tips[(size_t)r * P_pad + p] = c;
        if (c == CODE_INVALID) atomicOr(reinterpret_cast<unsigned *>(row_flags + r), p == 0 ? 3u : 1u);
    }
}
// tip codes -> the canonical byte of each code ('?' for a byte the alphabet rejects), either layout (phb_get_alignment)
__global__ void aa_unpack_kernel(const uint8_t *__restrict__ tips, int n_rows, int row0, int64_t P_pad, int64_t first, int64_t n,
                                 int layout, uint8_t *__restrict__ out)
{
    const int r = row0 + blockIdx.y;
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (int64_t)gridDim.x * blockDim.x) {
        const uint8_t ch = aa_decode_code(tips[(size_t)r * P_pad + first + p]);
        if (layout == 0) out[p * n_rows + r] = ch;
        else out[(int64_t)r * n + p] = ch;
    }
}
__global__ void aa_fill_pad_kernel(uint8_t *__restrict__ tips, int n_rows, int64_t P, int64_t P_pad)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t w = P_pad - P;
    if (w <= 0 || i >= (int64_t)n_rows * w) return;
    tips[(size_t)(i / w) * P_pad + P + i % w] = AA_CODE_GAP;
}

// ---- P(t): diag(inv_sqrt_pi) V diag(exp(lambda r t)) V^T diag(sqrt_pi), exact identity at r t == 0 ------------------
// One block per (edge, category).  Same operation order as RateMatrix::transition (rate_matrix.rs:119-133):
// (V_ik e_k) V_jk summed over k ascending, then the two diagonal scalings.
__global__ void __launch_bounds__(AA_THREADS)
aa_transition_kernel(const double *__restrict__ blen, int n_edges, const AaModelDev *__restrict__ M, double *__restrict__ P)
{
    __shared__ double e[AA_NS];
    const int n_cat = M->n_cat;
    const int edge = blockIdx.x / n_cat, cat = blockIdx.x - edge * n_cat;
    if (edge >= n_edges) return;
    const double t = M->rates[cat] * blen[edge];
    double *out = P + (size_t)blockIdx.x * (AA_NS * AA_NS);
    if (t == 0.0) {
        for (int idx = threadIdx.x; idx < AA_NS * AA_NS; idx += AA_THREADS) out[idx] = (idx / AA_NS == idx % AA_NS) ? 1.0 : 0.0;
        return;
    }
    if (threadIdx.x < AA_NS) e[threadIdx.x] = exp(M->lambda[threadIdx.x] * t);
    __syncthreads();
    for (int idx = threadIdx.x; idx < AA_NS * AA_NS; idx += AA_THREADS) {
        const int i = idx / AA_NS, j = idx - i * AA_NS;
        double acc = __dmul_rn(__dmul_rn(M->V[i], e[0]), M->V[j]);
        for (int k = 1; k < AA_NS; ++k) acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(M->V[i + AA_NS * k], e[k]), M->V[j + AA_NS * k])); // no FMA (kernels.cuh transition_kernel)
        out[idx] = M->inv_sqrt_pi[i] * acc * M->sqrt_pi[j];
    }
}

// ---- leaves' log-scales: Profile::scale() of a leaf profile divides by its maximum (likelihood.rs:108, profile.rs:20-39):
// 1 for a canonical letter, X and the gap (log-scale 0), 0.5 for B, Z, J (log-scale ln 0.5 each).
__global__ void aa_tip_const_kernel(const int *__restrict__ leaf_rows, int n_leaves, const uint8_t *__restrict__ tips,
                                    int64_t P_pad, int64_t P, double *__restrict__ leaf_ls)
{
    const int64_t site = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (site >= P) return;
    int halves = 0;
    for (int k = 0; k < n_leaves; ++k) {
        const uint32_t c = tips[(int64_t)leaf_rows[k] * P_pad + site];
        halves += (c >= AA_CODE_B && c <= AA_CODE_J) ? 1 : 0;
    }
    leaf_ls[site] = (double)halves * -0.69314718055994530942; // ln 0.5
}

// power-of-two normalisation of 20 states to a maximum in [1,2); returns the exponent taken out
// (v_true = v_stored * 2^exponent).  Zero / negative maxima are left alone (Profile::scale's max_val <= 0 return).
__device__ __forceinline__ int aa_normalize(double (&v)[AA_NS])
{
    int h = __double2hiint(v[0]);
#pragma unroll
    for (int i = 1; i < AA_NS; ++i) h = max(h, __double2hiint(v[i]));
    if (h <= 0) return 0;
    int s = 1023 - (h >> 20); // multiply by 2^s
    if (s > 1000) s = 1000;   // subnormal maxima: as far as one exact factor goes
    if (s < -1000) s = -1000;
    const double f = __longlong_as_double((long long)(1023 + s) << 52);
#pragma unroll
    for (int i = 0; i < AA_NS; ++i) v[i] *= f;
    return -s;
}

// the vector a tip contributes through P: sum_j P[i][j] profile[j] with the 0/1 profile Profile::scale() leaves
// (zero terms of the reference's dense product are exact zeros, so the masked sums below are the same values)
__device__ __forceinline__ double aa_tip_row(const double *Ps, int i, uint32_t code)
{
    const double *row = Ps + i * AA_NS;
    if (code < (uint32_t)AA_NS) return row[code];
    if (code <= AA_CODE_J) {
        int a, b;
        aa_pair_states(code, a, b);
        return row[a] + row[b];
    }
    double s = row[0];
#pragma unroll
    for (int j = 1; j < AA_NS; ++j) s += row[j];
    return s;
}

// ---- one level of prune_pattern_category (likelihood.rs:110-131), any arity ---------------------------------------
// grid: x = blocks of AA_THREADS patterns of the tile, y = op of the level, z = category.  values = prod over children
// of P_child * profile_child; the product is rescaled ONCE per node, like the reference's single Profile::scale() per
// node, so a node with very many children leaves the f64 range where the reference does.
// The child's P(t) is staged TRANSPOSED (Pt[j][i] = P[i][j]) and the mat-vec runs column by column,
//     y[0..19] += Pt[j][0..19] * x_j     for j = 0..19  (the reference's gemv order: likelihood.rs:121, column j ascending),
// so that a thread holds the 20 sums and the running product but never the child's whole vector (x_j streams in from
// HBM): 100 registers instead of 148, five blocks per SM instead of three -- the kernel waits on HBM latency, and the
// resident warps are what hides it (9.9 -> 13.5 ms at 1,000 x 20k when a rewrite cost a block per SM).  A tip child
// reads ONE column of P = one contiguous row of Pt.
#ifndef PHB_AA_R
#define PHB_AA_R 1
#endif
constexpr int AA_R = PHB_AA_R; // patterns per thread in the pre-order kernel
constexpr int AA_STAGE = 2;    // children whose P(t) the up-pass kernel stages at once
__device__ __forceinline__ void aa_tip_col(const double *Pt, uint32_t code, double (&y)[AA_NS])
{
    if (code < (uint32_t)AA_NS) {
#pragma unroll
        for (int i = 0; i < AA_NS; ++i) y[i] = Pt[code * AA_NS + i];
    } else if (code <= AA_CODE_J) {
        int a, b;
        aa_pair_states(code, a, b);
#pragma unroll
        for (int i = 0; i < AA_NS; ++i) y[i] = Pt[a * AA_NS + i] + Pt[b * AA_NS + i];
    } else {
#pragma unroll
        for (int i = 0; i < AA_NS; ++i) y[i] = Pt[i];
        for (int j = 1; j < AA_NS; ++j) {
#pragma unroll
            for (int i = 0; i < AA_NS; ++i) y[i] += Pt[j * AA_NS + i];
        }
    }
}
#ifndef PHB_AA_MINB
#define PHB_AA_MINB 4
#endif
constexpr int AA_CSTAGE = 4; // categories a block works through (grid.z = ceil(n_cat / AA_CSTAGE))
constexpr int AA_DCATS = 8;  // categories whose P(t) the pre-order kernel stages at once
// A block = one node x 128 patterns x up to AA_CSTAGE categories.  For a node with at most AA_STAGE children (every node
// of a binary tree) all categories' matrices are staged up front behind ONE barrier pair; a polytomy stages per category
// and child batch.  (Measured at 1,000 x 20k: 9.8 -> 9.0 ms against one category per block.  Staging the children's CLV
// tiles in shared memory with TMA bulk copies, double-buffered over the categories, was tried and was slower -- 10.7 ms:
// at 107 KB of shared memory only two blocks fit an SM, and the resident blocks are what hides the HBM latency here.)
// DOT = true (aa_score_nni, the re-linked parent p' of an NNI neighbour; needs n_cat <= AA_CSTAGE, grid.z == 1): the node's
// vector is not stored but multiplied with the upper partial in slot op.pad - 1 straight away -- sum_i up[i] L[i], the
// category values, their log_sum_exp and the block's sum of multiplicity * site_ll into partials[blockIdx.y][blockIdx.x]:
// aa_nni_root_kernel's arithmetic on the values this kernel would have written, without the 12.8 MB round trip.
template <bool DOT>
__global__ void __launch_bounds__(AA_THREADS, PHB_AA_MINB)
aa_clv_kernel(const GenOp *__restrict__ ops, const GenChild *__restrict__ kids, int n_cat, const double *__restrict__ Pmat,
              double *__restrict__ clv, int *__restrict__ exps, size_t slot_stride, size_t tile_pad, int tile_len,
              const uint8_t *__restrict__ tips, int64_t P_pad, int64_t tile_off, const AaModelDev *__restrict__ M,
              const double *__restrict__ leaf_ls, const double *__restrict__ mult, double *__restrict__ partials)
{
    __shared__ __align__(16) double Pt[AA_CSTAGE * AA_STAGE][AA_NS * AA_NS];
    double kept[DOT ? AA_CSTAGE : 1]; // the category values of this thread's pattern (DOT)
    const GenOp op = ops[blockIdx.y];
    const int cat0 = blockIdx.z * AA_CSTAGE;
    const int ncat = min(AA_CSTAGE, n_cat - cat0);
    const int nch = op.child_end - op.child_begin;
    const bool upfront = nch <= AA_STAGE;
    int64_t site = (int64_t)blockIdx.x * AA_THREADS + threadIdx.x;
    const bool active = site < tile_len;
    if (!active) site = 0; // reads stay in range; nothing is written
    if (upfront) {
        for (int idx = threadIdx.x; idx < ncat * nch * AA_NS * AA_NS; idx += AA_THREADS) {
            const int m = idx / (AA_NS * AA_NS), e = idx - m * (AA_NS * AA_NS); // m = c * nch + child
            const int c = m / nch, k = m - c * nch;
            Pt[c * AA_STAGE + k][(e % AA_NS) * AA_NS + e / AA_NS] =
                Pmat[((size_t)kids[op.child_begin + k].edge * n_cat + cat0 + c) * (AA_NS * AA_NS) + e];
        }
        __syncthreads();
    }
    for (int c = 0; c < ncat; ++c) {
        const int cat = cat0 + c;
        double acc[AA_NS];
#pragma unroll
        for (int i = 0; i < AA_NS; ++i) acc[i] = 1.0;
        int ex = 0;
        for (int k0 = 0; k0 < nch; k0 += AA_STAGE) {
            const int nb = min(AA_STAGE, nch - k0);
            if (!upfront) {
                __syncthreads(); // the previous batch is no longer read
                for (int idx = threadIdx.x; idx < nb * AA_NS * AA_NS; idx += AA_THREADS) {
                    const int k = idx / (AA_NS * AA_NS), e = idx - k * (AA_NS * AA_NS);
                    Pt[k][(e % AA_NS) * AA_NS + e / AA_NS] =
                        Pmat[((size_t)kids[op.child_begin + k0 + k].edge * n_cat + cat) * (AA_NS * AA_NS) + e];
                }
                __syncthreads();
            }
            for (int k = 0; k < nb; ++k) {
                const GenChild ch = kids[op.child_begin + k0 + k];
                const double *P_t = upfront ? Pt[c * AA_STAGE + k] : Pt[k];
                double y[AA_NS];
                if (ch.is_tip) {
                    aa_tip_col(P_t, tips[(int64_t)ch.src * P_pad + tile_off + site], y);
                } else {
                    const double *x_src = clv + (size_t)ch.src * slot_stride + (size_t)cat * AA_NS * tile_pad + site;
                    ex += exps[((size_t)ch.src * n_cat + cat) * tile_pad + site];
                    const double x0 = x_src[0];
#pragma unroll
                    for (int i = 0; i < AA_NS; ++i) y[i] = P_t[i] * x0;
#pragma unroll
                    for (int j = 1; j < AA_NS; ++j) {
                        const double xj = x_src[(size_t)j * tile_pad];
#pragma unroll
                        for (int i = 0; i < AA_NS; ++i) y[i] += P_t[j * AA_NS + i] * xj;
                    }
                }
#pragma unroll
                for (int i = 0; i < AA_NS; ++i) acc[i] *= y[i];
            }
        }
        if (DOT) {
            ex += aa_normalize(acc);
            const int up = op.pad - 1;
            const double *u = clv + (size_t)up * slot_stride + (size_t)cat * AA_NS * tile_pad + site;
            ex += exps[((size_t)up * n_cat + cat) * tile_pad + site];
            double mass = u[0] * acc[0];
#pragma unroll
            for (int i = 1; i < AA_NS; ++i) mass += u[(size_t)i * tile_pad] * acc[i];
            kept[c] = M->ln_weights[cat] + log(mass) + (double)ex * 0.69314718055994530942;
        } else if (active) {
            ex += aa_normalize(acc);
            double *dst = clv + (size_t)op.out_slot * slot_stride + (size_t)cat * AA_NS * tile_pad + site;
#pragma unroll
            for (int i = 0; i < AA_NS; ++i) dst[(size_t)i * tile_pad] = acc[i];
            exps[((size_t)op.out_slot * n_cat + cat) * tile_pad + site] = ex;
        }
    }
    if (DOT) {
        __shared__ double red[AA_THREADS];
        double weighted = 0.0;
        if (active) {
            double mx = -INFINITY;
#pragma unroll
            for (int c = 0; c < AA_CSTAGE; ++c)
                if (c < ncat) mx = fmax(mx, kept[c]);
            double lse = mx;
            if (isfinite(mx)) {
                double sum = 0.0;
#pragma unroll
                for (int c = 0; c < AA_CSTAGE; ++c)
                    if (c < ncat) sum += exp(kept[c] - mx);
                lse = mx + log(sum);
            }
            weighted = mult[tile_off + site] * (lse + leaf_ls[tile_off + site]);
        }
        red[threadIdx.x] = weighted;
        __syncthreads();
        for (int s = AA_THREADS / 2; s > 0; s >>= 1) {
            if ((int)threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
            __syncthreads();
        }
        if (threadIdx.x == 0) partials[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = red[0];
    }
}

// ---- one NNI neighbour in ONE pass (aa_score_nni with upper partials, n_cat <= AA_CSTAGE): the update of v (op ops_v[y]),
// the update of its re-linked parent p (op ops_p[y], one of whose children is v's scratch slot) and the dot product with
// up(p) (slot ops_p[y].pad - 1), per category; v's new vector stays in registers -- neither it nor p's is written.
// In p's product the factor coming from v is taken first (the others follow in list order), so that v's vector is dead
// before the next child's is formed: 128 registers as in aa_clv_kernel.  The matrices of a child batch are staged per
// category.  partials[y][block] = sum over the block's patterns of multiplicity * site_ll.
__global__ void __launch_bounds__(AA_THREADS, PHB_AA_MINB)
aa_nni_pair_kernel(const GenOp *__restrict__ ops_v, const GenOp *__restrict__ ops_p, const GenChild *__restrict__ kids, int n_cat,
                   const double *__restrict__ Pmat, const double *__restrict__ clv, const int *__restrict__ exps, size_t slot_stride,
                   size_t tile_pad, int tile_len, const uint8_t *__restrict__ tips, int64_t P_pad, int64_t tile_off,
                   const AaModelDev *__restrict__ M, const double *__restrict__ leaf_ls, const double *__restrict__ mult,
                   double *__restrict__ partials)
{
    __shared__ __align__(16) double Pt[AA_STAGE][AA_NS * AA_NS];
    __shared__ double red[AA_THREADS];
    const GenOp opv = ops_v[blockIdx.y], opp = ops_p[blockIdx.y];
    int64_t site = (int64_t)blockIdx.x * AA_THREADS + threadIdx.x;
    const bool active = site < tile_len;
    if (!active) site = 0;
    int inline_kid = opp.child_begin; // p's child that is v
    for (int k = opp.child_begin; k < opp.child_end; ++k)
        if (!kids[k].is_tip && kids[k].src == opv.out_slot) inline_kid = k;
    const int edge_v = kids[inline_kid].edge;
    double kept[AA_CSTAGE];
    // stage the transposed P(t) of up to AA_STAGE children (edges given by `edge_of`) for category `cat`
    auto stage = [&](int n, auto edge_of, int cat) {
        __syncthreads(); // the previous matrices are no longer read
        for (int idx = threadIdx.x; idx < n * AA_NS * AA_NS; idx += AA_THREADS) {
            const int k = idx / (AA_NS * AA_NS), e = idx - k * (AA_NS * AA_NS);
            Pt[k][(e % AA_NS) * AA_NS + e / AA_NS] = Pmat[((size_t)edge_of(k) * n_cat + cat) * (AA_NS * AA_NS) + e];
        }
        __syncthreads();
    };
    // y = P_child (child's vector); adds the child's exponent
    auto child_vec = [&](const double *P_t, const GenChild &ch, int cat, double (&y)[AA_NS], int &ex) {
        if (ch.is_tip) {
            aa_tip_col(P_t, tips[(int64_t)ch.src * P_pad + tile_off + site], y);
        } else {
            const double *x_src = clv + (size_t)ch.src * slot_stride + (size_t)cat * AA_NS * tile_pad + site;
            ex += exps[((size_t)ch.src * n_cat + cat) * tile_pad + site];
            const double x0 = x_src[0];
#pragma unroll
            for (int i = 0; i < AA_NS; ++i) y[i] = P_t[i] * x0;
#pragma unroll
            for (int j = 1; j < AA_NS; ++j) {
                const double xj = x_src[(size_t)j * tile_pad];
#pragma unroll
                for (int i = 0; i < AA_NS; ++i) y[i] += P_t[j * AA_NS + i] * xj;
            }
        }
    };
#pragma unroll 1
    for (int c = 0; c < n_cat && c < AA_CSTAGE; ++c) {
        double acc[AA_NS];
#pragma unroll
        for (int i = 0; i < AA_NS; ++i) acc[i] = 1.0;
        int ex = 0;
        for (int k0 = opv.child_begin; k0 < opv.child_end; k0 += AA_STAGE) { // v without the moved child, with its former sibling
            const int nb = min(AA_STAGE, opv.child_end - k0);
            stage(nb, [&](int k) { return kids[k0 + k].edge; }, c);
            for (int k = 0; k < nb; ++k) {
                double y[AA_NS];
                child_vec(Pt[k], kids[k0 + k], c, y, ex);
#pragma unroll
                for (int i = 0; i < AA_NS; ++i) acc[i] *= y[i];
            }
        }
        ex += aa_normalize(acc);
        // p: the factor through v first ...
        stage(1, [&](int) { return edge_v; }, c);
        {
            double y[AA_NS];
#pragma unroll
            for (int i = 0; i < AA_NS; ++i) y[i] = Pt[0][i] * acc[0];
#pragma unroll
            for (int j = 1; j < AA_NS; ++j) {
#pragma unroll
                for (int i = 0; i < AA_NS; ++i) y[i] += Pt[0][j * AA_NS + i] * acc[j];
            }
#pragma unroll
            for (int i = 0; i < AA_NS; ++i) acc[i] = y[i];
        }
        // ... then its other children in list order, the inline one skipped
        for (int k0 = opp.child_begin; k0 < opp.child_end; k0 += AA_STAGE) {
            const int nb = min(AA_STAGE, opp.child_end - k0);
            bool any = false;
            for (int k = 0; k < nb; ++k) any |= (k0 + k != inline_kid);
            if (!any) continue;
            stage(nb, [&](int k) { return kids[k0 + k].edge; }, c);
            for (int k = 0; k < nb; ++k) {
                if (k0 + k == inline_kid) continue;
                double y[AA_NS];
                child_vec(Pt[k], kids[k0 + k], c, y, ex);
#pragma unroll
                for (int i = 0; i < AA_NS; ++i) acc[i] *= y[i];
            }
        }
        ex += aa_normalize(acc);
        const int up = opp.pad - 1;
        const double *u = clv + (size_t)up * slot_stride + (size_t)c * AA_NS * tile_pad + site;
        ex += exps[((size_t)up * n_cat + c) * tile_pad + site];
        double mass = u[0] * acc[0];
#pragma unroll
        for (int i = 1; i < AA_NS; ++i) mass += u[(size_t)i * tile_pad] * acc[i];
        kept[c] = M->ln_weights[c] + log(mass) + (double)ex * 0.69314718055994530942;
    }
    double weighted = 0.0;
    if (active) {
        double mx = -INFINITY;
#pragma unroll
        for (int c = 0; c < AA_CSTAGE; ++c)
            if (c < n_cat) mx = fmax(mx, kept[c]);
        double lse = mx;
        if (isfinite(mx)) {
            double sum = 0.0;
#pragma unroll
            for (int c = 0; c < AA_CSTAGE; ++c)
                if (c < n_cat) sum += exp(kept[c] - mx);
            lse = mx + log(sum);
        }
        weighted = mult[tile_off + site] * (lse + leaf_ls[tile_off + site]);
    }
    red[threadIdx.x] = weighted;
    __syncthreads();
    for (int s = AA_THREADS / 2; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) partials[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = red[0];
}

// the root's vector for one (pattern, category): its CLV, or its tip profile when the tree is a single leaf
__device__ __forceinline__ void aa_load_vec(const double *__restrict__ clv, size_t slot_stride, size_t tile_pad, int slot, int cat,
                                            int64_t site, double (&x)[AA_NS])
{
    const double *src = clv + (size_t)slot * slot_stride + (size_t)cat * AA_NS * tile_pad + site;
#pragma unroll
    for (int j = 0; j < AA_NS; ++j) x[j] = src[(size_t)j * tile_pad];
}
__device__ __forceinline__ void aa_tip_vec(uint32_t code, double (&x)[AA_NS])
{
    int a = -1, b = -1;
    if (code >= AA_CODE_B && code <= AA_CODE_J) aa_pair_states(code, a, b);
#pragma unroll
    for (int j = 0; j < AA_NS; ++j) x[j] = (code >= AA_CODE_X || (int)code == j || j == a || j == b) ? 1.0 : 0.0;
}

// ---- root: cat_ll_c = ln w_c + ln(sum_i pi_i v_i) + log-scale (likelihood.rs:135-140); site_ll = log_sum_exp over the
// categories (:61-68: a non-finite maximum is returned as it is); per-block sum of multiplicity * site_ll (:205-206).
// cat_ll_out[c][pattern of the tile] (without the leaves' log-scale, which is common to the categories) feeds the
// mixture weights of the down pass.
__global__ void __launch_bounds__(AA_THREADS)
aa_root_kernel(int root_slot, int root_row, const AaModelDev *__restrict__ M, const double *__restrict__ clv,
               const int *__restrict__ exps, size_t slot_stride, size_t tile_pad, int tile_len, const uint8_t *__restrict__ tips,
               int64_t P_pad, int64_t tile_off, const double *__restrict__ leaf_ls, const double *__restrict__ mult,
               double *__restrict__ per_pattern, double *__restrict__ cat_ll_out, double *__restrict__ partials)
{
    __shared__ double red[AA_THREADS];
    const int64_t site = (int64_t)blockIdx.x * AA_THREADS + threadIdx.x;
    const int n_cat = M->n_cat;
    double weighted = 0.0;
    if (site < tile_len) {
        double mx = -INFINITY;
        for (int c = 0; c < n_cat; ++c) {
            double x[AA_NS];
            int ex = 0;
            if (root_slot >= 0) {
                aa_load_vec(clv, slot_stride, tile_pad, root_slot, c, site, x);
                ex = exps[((size_t)root_slot * n_cat + c) * tile_pad + site];
            } else
                aa_tip_vec(tips[(int64_t)root_row * P_pad + tile_off + site], x);
            double mass = M->pi[0] * x[0];
#pragma unroll
            for (int i = 1; i < AA_NS; ++i) mass += M->pi[i] * x[i];
            const double ll = M->ln_weights[c] + log(mass) + (double)ex * 0.69314718055994530942;
            cat_ll_out[(size_t)c * tile_pad + site] = ll;
            mx = fmax(mx, ll);
        }
        double lse = mx;
        if (isfinite(mx)) {
            double sum = 0.0;
            for (int c = 0; c < n_cat; ++c) sum += exp(cat_ll_out[(size_t)c * tile_pad + site] - mx);
            lse = mx + log(sum);
        }
        cat_ll_out[(size_t)n_cat * tile_pad + site] = lse; // row n_cat: the log_sum_exp itself
        const double site_ll = lse + leaf_ls[tile_off + site];
        per_pattern[tile_off + site] = site_ll;
        weighted = mult[tile_off + site] * site_ll;
    }
    red[threadIdx.x] = weighted;
    __syncthreads();
    for (int s = AA_THREADS / 2; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) partials[blockIdx.x] = red[0];
}

// ---- NNI neighbours (aa_engine.cuh aa_score_nni) --------------------------------------------------------------------
// Upper partials.  up(x)[i] = the likelihood of everything OUTSIDE the subtree of x given state i at x:
//     out[k]   = up(parent)[k] * prod over the siblings s of x of (P_s L_s)[k]        (up(root) = pi)
//     up(x)[i] = sum_k out[k] P_x[k][i]
// so that sum_i up(x)[i] L(x)[i] is the site likelihood for ANY x.  An op lists the parent's upper vector as a DIRECT
// child (GenChild.pad == 1: multiplied in as it is, no P) and the siblings as ordinary children; GenOp.pad = x + 1 names
// the edge of the closing transposed product.  grid: x = pattern blocks, y = op, z = category.  Stored like a CLV
// (maximum in [1,2), exponent beside it).  One matrix staged at a time: the pass runs once per call.
__global__ void __launch_bounds__(AA_THREADS, 3)
aa_upper_kernel(const GenOp *__restrict__ ops, const GenChild *__restrict__ kids, int n_cat, const double *__restrict__ Pmat,
                double *__restrict__ clv, int *__restrict__ exps, size_t slot_stride, size_t tile_pad, int tile_len,
                const uint8_t *__restrict__ tips, int64_t P_pad, int64_t tile_off)
{
    __shared__ __align__(16) double Ps[AA_NS * AA_NS];
    const GenOp op = ops[blockIdx.y];
    const int cat = blockIdx.z;
    int64_t site = (int64_t)blockIdx.x * AA_THREADS + threadIdx.x;
    const bool active = site < tile_len;
    if (!active) site = 0;
    double acc[AA_NS];
#pragma unroll
    for (int i = 0; i < AA_NS; ++i) acc[i] = 1.0;
    int ex = 0;
    for (int k = op.child_begin; k < op.child_end; ++k) {
        const GenChild ch = kids[k];
        if (ch.pad == 1) { // the parent's upper vector
            const double *x_src = clv + (size_t)ch.src * slot_stride + (size_t)cat * AA_NS * tile_pad + site;
            ex += exps[((size_t)ch.src * n_cat + cat) * tile_pad + site];
#pragma unroll
            for (int i = 0; i < AA_NS; ++i) acc[i] *= x_src[(size_t)i * tile_pad];
            continue;
        }
        __syncthreads(); // the previous matrix is no longer read
        for (int e = threadIdx.x; e < AA_NS * AA_NS; e += AA_THREADS)
            Ps[(e % AA_NS) * AA_NS + e / AA_NS] = Pmat[((size_t)ch.edge * n_cat + cat) * (AA_NS * AA_NS) + e]; // transposed
        __syncthreads();
        double y[AA_NS];
        if (ch.is_tip) {
            aa_tip_col(Ps, tips[(int64_t)ch.src * P_pad + tile_off + site], y);
        } else {
            const double *x_src = clv + (size_t)ch.src * slot_stride + (size_t)cat * AA_NS * tile_pad + site;
            ex += exps[((size_t)ch.src * n_cat + cat) * tile_pad + site];
            const double x0 = x_src[0];
#pragma unroll
            for (int i = 0; i < AA_NS; ++i) y[i] = Ps[i] * x0;
#pragma unroll
            for (int j = 1; j < AA_NS; ++j) {
                const double xj = x_src[(size_t)j * tile_pad];
#pragma unroll
                for (int i = 0; i < AA_NS; ++i) y[i] += Ps[j * AA_NS + i] * xj;
            }
        }
#pragma unroll
        for (int i = 0; i < AA_NS; ++i) acc[i] *= y[i];
    }
    __syncthreads();
    for (int e = threadIdx.x; e < AA_NS * AA_NS; e += AA_THREADS)
        Ps[e] = Pmat[((size_t)(op.pad - 1) * n_cat + cat) * (AA_NS * AA_NS) + e]; // row k = P[k][.]
    __syncthreads();
    double up[AA_NS];
#pragma unroll
    for (int i = 0; i < AA_NS; ++i) up[i] = Ps[i] * acc[0];
#pragma unroll
    for (int k = 1; k < AA_NS; ++k) {
#pragma unroll
        for (int i = 0; i < AA_NS; ++i) up[i] += Ps[k * AA_NS + i] * acc[k];
    }
    if (active) {
        ex += aa_normalize(up);
        double *dst = clv + (size_t)op.out_slot * slot_stride + (size_t)cat * AA_NS * tile_pad + site;
#pragma unroll
        for (int i = 0; i < AA_NS; ++i) dst[(size_t)i * tile_pad] = up[i];
        exps[((size_t)op.out_slot * n_cat + cat) * tile_pad + site] = ex;
    }
}
// up(root): pi for every pattern and category, exponent 0
__global__ void aa_fill_pi_kernel(const AaModelDev *__restrict__ M, double *__restrict__ clv, int *__restrict__ exps, int slot,
                                  size_t slot_stride, size_t tile_pad)
{
    const int n_cat = M->n_cat;
    const size_t n = (size_t)n_cat * AA_NS * tile_pad;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        clv[(size_t)slot * slot_stride + i] = M->pi[(i / tile_pad) % AA_NS];
        if (i < (size_t)n_cat * tile_pad) exps[(size_t)slot * n_cat * tile_pad + i] = 0;
    }
}
// The value of neighbour blockIdx.y: sum_i up[i] L[i] with L in slot low_slots[y] and up in slot up_slots[y] (the pi slot
// when the re-linked parent is the root: then this is aa_root_kernel's arithmetic, bit for bit); the category values stay
// in registers instead of being parked in HBM; per-block sums of multiplicity * site_ll go to partials[neighbour][block].
__global__ void __launch_bounds__(AA_THREADS)
aa_nni_root_kernel(const int *__restrict__ low_slots, const int *__restrict__ up_slots, const AaModelDev *__restrict__ M,
                   const double *__restrict__ clv, const int *__restrict__ exps, size_t slot_stride, size_t tile_pad, int tile_len,
                   int64_t tile_off, const double *__restrict__ leaf_ls, const double *__restrict__ mult, double *__restrict__ partials)
{
    __shared__ double red[AA_THREADS];
    const int64_t site = (int64_t)blockIdx.x * AA_THREADS + threadIdx.x;
    const int n_cat = M->n_cat;
    const int low = low_slots[blockIdx.y], upper = up_slots[blockIdx.y];
    double weighted = 0.0;
    if (site < tile_len) {
        auto cat_value = [&](int c) {
            double x[AA_NS], u[AA_NS];
            aa_load_vec(clv, slot_stride, tile_pad, low, c, site, x);
            aa_load_vec(clv, slot_stride, tile_pad, upper, c, site, u);
            const int ex = exps[((size_t)low * n_cat + c) * tile_pad + site] + exps[((size_t)upper * n_cat + c) * tile_pad + site];
            double mass = u[0] * x[0];
#pragma unroll
            for (int i = 1; i < AA_NS; ++i) mass += u[i] * x[i];
            return M->ln_weights[c] + log(mass) + (double)ex * 0.69314718055994530942;
        };
        constexpr int KEEP = 8; // category values kept in registers; beyond that they are formed a second time
        double kept[KEEP];
        double mx = -INFINITY;
        for (int c = 0; c < n_cat; ++c) {
            const double ll = cat_value(c);
            if (c < KEEP) kept[c] = ll;
            mx = fmax(mx, ll);
        }
        double lse = mx;
        if (isfinite(mx)) {
            double sum = 0.0;
#pragma unroll
            for (int c = 0; c < KEEP; ++c)
                if (c < n_cat) sum += exp(kept[c] - mx);
            for (int c = KEEP; c < n_cat; ++c) sum += exp(cat_value(c) - mx);
            lse = mx + log(sum);
        }
        weighted = mult[tile_off + site] * (lse + leaf_ls[tile_off + site]);
    }
    red[threadIdx.x] = weighted;
    __syncthreads();
    for (int s = AA_THREADS / 2; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) partials[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = red[0];
}

// ---- one depth of the marginal pre-order pass (likelihood.rs:303-390) -------------------------------------------------
// grid: x = pattern blocks of the tile, y = node of the depth.  A thread owns one pattern and loops over the categories:
//   root:   post = pi (.) up                                                        (:304-316)
//   else:   msg = P^T post_parent,  post = up (.) msg                               (:318-341)
//   post /= its sum when the sum is > 0; mixed += mix_c post, mix_c = exp(cat_ll_c - site_ll), uniform when site_ll is
//   not finite (:351-362); mixed /= its sum when > 0; state = LAST maximal index (Iterator::max_by, :385-390).
// The posterior of an internal node overwrites its up-pass CLV in place (its children read it one depth further down;
// nobody reads the up-pass values again); a leaf's goes nowhere but into the mixture.
// A thread owns AA_R patterns, as in the up pass.  A leaf showing a canonical letter has a one-hot up vector, so its
// posterior is that letter whenever the parent's message gives it any mass: only msg[letter] is computed (20 DFMA
// instead of 400) -- identical values, since the other 19 products are exact zeros in the reference too.
__global__ void __launch_bounds__(AA_THREADS, AA_R == 1 ? 3 : 1)
aa_down_kernel(const AaDown *__restrict__ nodes, const AaModelDev *__restrict__ M, const double *__restrict__ Pmat,
               double *__restrict__ clv, size_t slot_stride, size_t tile_pad, int tile_len, const uint8_t *__restrict__ tips,
               int64_t P_pad, int64_t tile_off, const double *__restrict__ cat_ll, uint8_t *__restrict__ states,
               double *__restrict__ post_out)
{
    __shared__ double Pall[AA_DCATS][AA_NS * AA_NS];
    const AaDown nd = nodes[blockIdx.y];
    const int n_cat = M->n_cat;
    // the node's P(t) of every category staged at once behind one barrier (up to AA_DCATS categories; beyond that the
    // staging goes category by category)
    const bool upfront = n_cat <= AA_DCATS;
    if (upfront && nd.parent_slot >= 0) {
        const double *src = Pmat + (size_t)nd.node * n_cat * (AA_NS * AA_NS);
        for (int idx = threadIdx.x; idx < n_cat * AA_NS * AA_NS; idx += AA_THREADS) Pall[idx / (AA_NS * AA_NS)][idx % (AA_NS * AA_NS)] = src[idx];
        __syncthreads();
    }
    int64_t site[AA_R];
    bool active[AA_R];
    double lse[AA_R];
    uint32_t code[AA_R];
    double mixed[AA_R][AA_NS];
#pragma unroll
    for (int r = 0; r < AA_R; ++r) {
        site[r] = ((int64_t)blockIdx.x * AA_R + r) * AA_THREADS + threadIdx.x;
        active[r] = site[r] < tile_len;
        if (!active[r]) site[r] = 0;
        lse[r] = cat_ll[(size_t)n_cat * tile_pad + site[r]];
        code[r] = nd.slot < 0 ? tips[(int64_t)nd.row * P_pad + tile_off + site[r]] : 0u;
#pragma unroll
        for (int i = 0; i < AA_NS; ++i) mixed[r][i] = 0.0;
    }
    for (int c = 0; c < n_cat; ++c) {
        if (!upfront && nd.parent_slot >= 0) {
            __syncthreads();
            const double *src = Pmat + ((size_t)nd.node * n_cat + c) * (AA_NS * AA_NS);
            for (int idx = threadIdx.x; idx < AA_NS * AA_NS; idx += AA_THREADS) Pall[0][idx] = src[idx];
            __syncthreads();
        }
        const double *Ps = upfront ? Pall[c] : Pall[0];
        double mix[AA_R];
#pragma unroll
        for (int r = 0; r < AA_R; ++r)
            mix[r] = isfinite(lse[r]) ? exp(cat_ll[(size_t)c * tile_pad + site[r]] - lse[r]) : 1.0 / (double)n_cat;
        const double *pp = nd.parent_slot >= 0 ? clv + (size_t)nd.parent_slot * slot_stride + (size_t)c * AA_NS * tile_pad : nullptr;
        if (nd.slot < 0 && nd.parent_slot >= 0 && code[0] < (uint32_t)AA_NS && (AA_R == 1 || code[AA_R - 1] < (uint32_t)AA_NS)) {
            // canonical leaf (all of the thread's patterns): post = one-hot when msg[letter] > 0, all zeros otherwise
#pragma unroll
            for (int r = 0; r < AA_R; ++r) {
                double m = Ps[code[r]] * pp[site[r]];
#pragma unroll
                for (int i = 1; i < AA_NS; ++i) m += Ps[i * AA_NS + code[r]] * pp[(size_t)i * tile_pad + site[r]];
                const double v = 1.0 * m; // up (.) msg, then / sum: m / m
                const double one = v > 0.0 ? v / v : v;
#pragma unroll
                for (int i = 0; i < AA_NS; ++i) mixed[r][i] += (i == (int)code[r]) ? mix[r] * one : mix[r] * 0.0;
            }
        } else {
        double post[AA_R][AA_NS];
        if (nd.parent_slot < 0) {
#pragma unroll
            for (int r = 0; r < AA_R; ++r)
#pragma unroll
                for (int i = 0; i < AA_NS; ++i) post[r][i] = M->pi[i];
        } else {
            double p_i[AA_R];
#pragma unroll
            for (int r = 0; r < AA_R; ++r) p_i[r] = pp[site[r]];
#pragma unroll
            for (int j = 0; j < AA_NS; ++j) {
                const double pij = Ps[j];
#pragma unroll
                for (int r = 0; r < AA_R; ++r) post[r][j] = pij * p_i[r];
            }
#pragma unroll
            for (int i = 1; i < AA_NS; ++i) {
#pragma unroll
                for (int r = 0; r < AA_R; ++r) p_i[r] = pp[(size_t)i * tile_pad + site[r]];
#pragma unroll
                for (int j = 0; j < AA_NS; ++j) {
                    const double pij = Ps[i * AA_NS + j];
#pragma unroll
                    for (int r = 0; r < AA_R; ++r) post[r][j] += pij * p_i[r];
                }
            }
        }
#pragma unroll
        for (int r = 0; r < AA_R; ++r) {
            double *own = nd.slot >= 0 ? clv + (size_t)nd.slot * slot_stride + (size_t)c * AA_NS * tile_pad + site[r] : nullptr;
            if (own) {
#pragma unroll
                for (int i = 0; i < AA_NS; ++i) post[r][i] *= own[(size_t)i * tile_pad];
            } else {
                double up[AA_NS];
                aa_tip_vec(code[r], up);
#pragma unroll
                for (int i = 0; i < AA_NS; ++i) post[r][i] *= up[i];
            }
            double sum = 0.0;
#pragma unroll
            for (int i = 0; i < AA_NS; ++i) sum += post[r][i];
            if (sum > 0.0) {
#pragma unroll
                for (int i = 0; i < AA_NS; ++i) post[r][i] /= sum;
            }
            if (own && active[r]) {
#pragma unroll
                for (int i = 0; i < AA_NS; ++i) own[(size_t)i * tile_pad] = post[r][i];
            }
#pragma unroll
            for (int i = 0; i < AA_NS; ++i) mixed[r][i] += mix[r] * post[r][i];
        }
        } // general path
    }
#pragma unroll
    for (int r = 0; r < AA_R; ++r) {
        if (!active[r]) continue;
        double sum = 0.0;
#pragma unroll
        for (int i = 0; i < AA_NS; ++i) sum += mixed[r][i];
        if (sum > 0.0) {
#pragma unroll
            for (int i = 0; i < AA_NS; ++i) mixed[r][i] /= sum;
        }
        int best = 0;
        double best_v = mixed[r][0];
#pragma unroll
        for (int i = 1; i < AA_NS; ++i)
            if (mixed[r][i] >= best_v) {
                best_v = mixed[r][i];
                best = i;
            }
        states[(size_t)nd.node * tile_pad + site[r]] = (uint8_t)best;
        if (post_out) {
            double *o = post_out + ((size_t)nd.node * tile_len + site[r]) * AA_NS;
#pragma unroll
            for (int i = 0; i < AA_NS; ++i) o[i] = mixed[r][i];
        }
    }
}

// =====================================================================================================================
// Joint (Viterbi) reconstruction at 20 states: compute_joint_asr (likelihood.rs:421-560), as joint.cuh does it for 4 --
// log space, max-plus, no back-pointer table (the traceback re-evaluates the one row it needs from the child's C values).
//   leaf      C_v[i] = ln profile_v[i]  (the profile as AminoAcid::profile gives it: ln 1, ln 0.5, -inf)      :477-480
//   internal  C_v[i] = sum_children max_j ( ln P_c[i][j] + C_c[j] )                                            :483-509
//   root      per category the FIRST maximal i of ln pi_i + C_root[i]; the category with the strictly greatest
//             ln w + max wins                                                                                   :512-527
//   traceback state_c = FIRST maximal j of ln P_c[state_parent][j] + C_c[j]                                     :531-538
// `val > best` comparisons throughout: a NaN never wins, as in the reference.
// =====================================================================================================================
constexpr uint8_t AA_NO_CATEGORY = 0xFF;

// ln profile[j] of a tip code
__device__ __forceinline__ double aa_joint_tip_value(uint32_t code, int j)
{
    if (code < (uint32_t)AA_NS) return (int)code == j ? 0.0 : -INFINITY;
    if (code <= AA_CODE_J) {
        int a, b;
        aa_pair_states(code, a, b);
        return (j == a || j == b) ? -0.69314718055994530942 : -INFINITY;
    }
    return 0.0; // X, gap: all ones
}

// Up pass, one level.  Same grid and staging as aa_clv_kernel (ln P transposed, all of the block's categories behind one
// barrier for nodes with at most AA_STAGE children); the buffer holds C values instead of CLVs, no exponents.
__global__ void __launch_bounds__(AA_THREADS, PHB_AA_MINB)
aa_joint_up_kernel(const GenOp *__restrict__ ops, const GenChild *__restrict__ kids, int n_cat, const double *__restrict__ lnP,
                   double *__restrict__ cbuf, size_t slot_stride, size_t tile_pad, int tile_len, const uint8_t *__restrict__ tips,
                   int64_t P_pad, int64_t tile_off)
{
    __shared__ __align__(16) double Lt[AA_CSTAGE * AA_STAGE][AA_NS * AA_NS];
    const GenOp op = ops[blockIdx.y];
    const int cat0 = blockIdx.z * AA_CSTAGE;
    const int ncat = min(AA_CSTAGE, n_cat - cat0);
    const int nch = op.child_end - op.child_begin;
    const bool upfront = nch <= AA_STAGE;
    int64_t site = (int64_t)blockIdx.x * AA_THREADS + threadIdx.x;
    const bool active = site < tile_len;
    if (!active) site = 0;
    if (upfront) {
        for (int idx = threadIdx.x; idx < ncat * nch * AA_NS * AA_NS; idx += AA_THREADS) {
            const int m = idx / (AA_NS * AA_NS), e = idx - m * (AA_NS * AA_NS);
            const int c = m / nch, k = m - c * nch;
            Lt[c * AA_STAGE + k][(e % AA_NS) * AA_NS + e / AA_NS] =
                lnP[((size_t)kids[op.child_begin + k].edge * n_cat + cat0 + c) * (AA_NS * AA_NS) + e];
        }
        __syncthreads();
    }
    for (int c = 0; c < ncat; ++c) {
        const int cat = cat0 + c;
        double acc[AA_NS];
#pragma unroll
        for (int i = 0; i < AA_NS; ++i) acc[i] = 0.0;
        for (int k0 = 0; k0 < nch; k0 += AA_STAGE) {
            const int nb = min(AA_STAGE, nch - k0);
            if (!upfront) {
                __syncthreads();
                for (int idx = threadIdx.x; idx < nb * AA_NS * AA_NS; idx += AA_THREADS) {
                    const int k = idx / (AA_NS * AA_NS), e = idx - k * (AA_NS * AA_NS);
                    Lt[k][(e % AA_NS) * AA_NS + e / AA_NS] = lnP[((size_t)kids[op.child_begin + k0 + k].edge * n_cat + cat) * (AA_NS * AA_NS) + e];
                }
                __syncthreads();
            }
            for (int k = 0; k < nb; ++k) {
                const GenChild ch = kids[op.child_begin + k0 + k];
                const double *L_t = upfront ? Lt[c * AA_STAGE + k] : Lt[k];
                double y[AA_NS];
#pragma unroll
                for (int i = 0; i < AA_NS; ++i) y[i] = -INFINITY;
                const uint32_t code = ch.is_tip ? tips[(int64_t)ch.src * P_pad + tile_off + site] : 0u;
                const double *x_src = ch.is_tip ? nullptr : cbuf + (size_t)ch.src * slot_stride + (size_t)cat * AA_NS * tile_pad + site;
                for (int j = 0; j < AA_NS; ++j) { // j ascending with `>`: the reference's sequential maximum
                    const double cj = ch.is_tip ? aa_joint_tip_value(code, j) : x_src[(size_t)j * tile_pad];
#pragma unroll
                    for (int i = 0; i < AA_NS; ++i) {
                        const double val = L_t[j * AA_NS + i] + cj;
                        if (val > y[i]) y[i] = val;
                    }
                }
#pragma unroll
                for (int i = 0; i < AA_NS; ++i) acc[i] += y[i];
            }
        }
        if (active) {
            double *dst = cbuf + (size_t)op.out_slot * slot_stride + (size_t)cat * AA_NS * tile_pad + site;
#pragma unroll
            for (int i = 0; i < AA_NS; ++i) dst[(size_t)i * tile_pad] = acc[i];
        }
    }
}

// Root: winning category, root state, per-pattern value, block sums of multiplicity * value.
__global__ void __launch_bounds__(AA_THREADS)
aa_joint_root_kernel(int root_slot, int root_row, int root_node, const AaModelDev *__restrict__ M, const double *__restrict__ cbuf,
                     size_t slot_stride, size_t tile_pad, int tile_len, const uint8_t *__restrict__ tips, int64_t P_pad,
                     int64_t tile_off, const double *__restrict__ mult, double *__restrict__ per_pattern,
                     uint8_t *__restrict__ best_cat, uint8_t *__restrict__ states, double *__restrict__ partials)
{
    __shared__ double red[AA_THREADS];
    const int64_t site = (int64_t)blockIdx.x * AA_THREADS + threadIdx.x;
    const int n_cat = M->n_cat;
    double contrib = 0.0;
    if (site < tile_len) {
        double best = -INFINITY;
        int bcat = AA_NO_CATEGORY, bstate = 0;
        const uint32_t code = root_slot < 0 ? tips[(int64_t)root_row * P_pad + tile_off + site] : 0u;
        for (int c = 0; c < n_cat; ++c) {
            double root_max = -INFINITY;
            int root_best = 0;
            for (int i = 0; i < AA_NS; ++i) {
                const double cv = root_slot < 0 ? aa_joint_tip_value(code, i)
                                                : cbuf[(size_t)root_slot * slot_stride + ((size_t)c * AA_NS + i) * tile_pad + site];
                const double ll = log(M->pi[i]) + cv;
                if (ll > root_max) {
                    root_max = ll;
                    root_best = i;
                }
            }
            const double cat_ll = M->ln_weights[c] + root_max;
            if (cat_ll > best) {
                best = cat_ll;
                bcat = c;
                bstate = root_best;
            }
        }
        per_pattern[tile_off + site] = best;
        best_cat[site] = (uint8_t)bcat;
        states[(size_t)root_node * tile_pad + site] = (uint8_t)bstate;
        contrib = mult[tile_off + site] * best;
    }
    red[threadIdx.x] = contrib;
    __syncthreads();
    for (int s = AA_THREADS / 2; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) partials[blockIdx.x] = red[0];
}

// Traceback, one depth (the root's own depth is done by the root kernel).  grid: x = pattern blocks, y = node of the depth.
__global__ void __launch_bounds__(AA_THREADS)
aa_joint_down_kernel(const AaDown *__restrict__ nodes, const int *__restrict__ parent_node, int n_cat, const double *__restrict__ lnP,
                     const double *__restrict__ cbuf, size_t slot_stride, size_t tile_pad, int tile_len,
                     const uint8_t *__restrict__ tips, int64_t P_pad, int64_t tile_off, const uint8_t *__restrict__ best_cat,
                     uint8_t *__restrict__ states)
{
    const AaDown nd = nodes[blockIdx.y];
    if (nd.parent_slot < 0) return; // the root
    const int64_t site = (int64_t)blockIdx.x * AA_THREADS + threadIdx.x;
    if (site >= tile_len) return;
    const int cat = best_cat[site];
    int st = 0;
    if (cat != AA_NO_CATEGORY) {
        const int s_p = states[(size_t)parent_node[nd.node] * tile_pad + site];
        const double *row = lnP + ((size_t)nd.node * n_cat + cat) * (AA_NS * AA_NS) + (size_t)s_p * AA_NS;
        const uint32_t code = nd.slot < 0 ? tips[(int64_t)nd.row * P_pad + tile_off + site] : 0u;
        const double *c_src = nd.slot < 0 ? nullptr : cbuf + (size_t)nd.slot * slot_stride + (size_t)cat * AA_NS * tile_pad + site;
        double best = -INFINITY;
        for (int j = 0; j < AA_NS; ++j) {
            const double val = row[j] + (nd.slot < 0 ? aa_joint_tip_value(code, j) : c_src[(size_t)j * tile_pad]);
            if (val > best) {
                best = val;
                st = j;
            }
        }
    }
    states[(size_t)nd.node * tile_pad + site] = (uint8_t)st;
}

} // namespace phb
