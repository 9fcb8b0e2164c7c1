// kernels.cuh — hand-written sm_100a kernels of the Felsenstein-pruning hot path.
//
// Data layout in HBM:
//   tips   u8  [row][P_pad]                     encoded tip codes, sites (patterns) fastest
//   P      f64 [node][model_cat][4][4]          transition matrix of edge (parent -> node), row-major
//   clv    f64 [slot][clv_cat][state][tile_pad] conditional likelihoods of one site tile, sites fastest
//   scale  i32 [clv_cat][tile_pad]              accumulated power-of-two rescaling exponents (bits)
// Every CLV load/store is a 16-byte (two-site) access on the site axis; P matrices are block-uniform.
#pragma once
#include "cuda_compat.h"
#include <stdint.h>
#include <math.h>

namespace phb {

constexpr int MAX_CAT = 32;
constexpr int CLV_THREADS = 256;          // threads per block in the CLV kernels
constexpr int SITES_PER_BLOCK = 2 * CLV_THREADS;

// tip code: bits 0..3 = SLOT of the state mask over A,C,G,T (slot = row of the sweep's tip tables:
// one-hot masks A,C,G,T are slots 0..3 = the state index, the all-ones mask (N, gap) is slot 4 -- so the five
// codes that make up almost all real data are 0..4 and index the sweep's 5x5 cherry tables directly -- and the
// 10 partial ambiguity masks follow); bit 4 = gap ('-': all-ones profile, log-scale 0); 0xFF = byte outside
// Nucleotide::profile's 17 accepted codes (src/alphabet.rs:44-64).
constexpr uint8_t CODE_GAP = 0x14;
constexpr uint8_t CODE_INVALID = 0xFF;

__host__ __device__ __forceinline__ uint32_t mask_to_slot(uint32_t mask)
{
    return (uint32_t)((0x4EDCBA938762510Full >> (4u * (mask & 15u))) & 15ull);
}
// state mask (bit i = state i allowed) of a tip code
__host__ __device__ __forceinline__ uint32_t code_mask(uint32_t code)
{
    return (uint32_t)((0x0EDCBA97653F8421ull >> (4u * (code & 15u))) & 15ull);
}

__host__ __device__ inline uint8_t encode_mask_byte(uint8_t c)
{
    switch (c) {
    case 'A': return 1;
    case 'C': return 2;
    case 'G': return 4;
    case 'T':
    case 'U': return 8;
    case 'R': return 1 | 4;
    case 'Y': return 2 | 8;
    case 'S': return 2 | 4;
    case 'W': return 1 | 8;
    case 'K': return 4 | 8;
    case 'M': return 1 | 2;
    case 'B': return 2 | 4 | 8;
    case 'D': return 1 | 4 | 8;
    case 'H': return 1 | 2 | 8;
    case 'V': return 1 | 2 | 4;
    case 'N': return 15;
    case '-': return 0x1F;
    default: return CODE_INVALID;
    }
}

__host__ __device__ inline uint8_t encode_byte(uint8_t c)
{
    uint8_t m = encode_mask_byte(c);
    if (m == CODE_INVALID) return CODE_INVALID;
    return (uint8_t)((m & 16u) | mask_to_slot(m));
}

__host__ __device__ inline uint8_t decode_code(uint8_t code)
{
    if (code == CODE_GAP) return '-';
    if (code > 15) return '?';
    const char *tab = "?ACMGRSVTWYHKDBN";
    return (uint8_t)tab[code_mask(code)];
}

// Model handed to kernels by value (constant bank).  phb_set_model, include/phylo_b200.h.
struct ModelDev {
    double lambda[4];
    double V[16]; // column-major: V[i + 4k]
    double pi[4];
    double sqrt_pi[4];
    double inv_sqrt_pi[4];
    double rates[MAX_CAT];
    double weights[MAX_CAT];
    double ln_weights[MAX_CAT];
    double ln_tip[5];         // ln(1/popcount) for popcount 0..4: leaf Profile::scale() log-scales
    int clv_idx[MAX_CAT];     // model category -> CLV category index, -1 = closed-form invariant
    int cat_of_clv[MAX_CAT];  // CLV category -> model category
    int n_cat;
    int n_clv;
};

// One binary internal-node update of a level.  kind bit0/bit1: child 0/1 is a tip.
struct NodeOp {
    int32_t out_slot;
    int32_t c0, c1; // CLV slot (internal child) or alignment row (tip child)
    int32_t e0, e1; // dense node id of the child == index of its edge's P matrix
    int32_t kind;
};

// Generic-arity update (unifurcations, polytomies: src/tree/likelihood.rs:113 iterates any arity).
struct GenOp {
    int32_t out_slot;
    int32_t child_begin, child_end; // range in the GenChild array
    int32_t pad;
};
struct GenChild {
    int32_t src;  // slot or row
    int32_t edge; // dense child id
    int32_t is_tip;
    int32_t pad;
};

// ---------------------------------------------------------------------------------------------
// K0: alignment packing.  ASCII bytes -> tip codes, transposed to [row][P_pad].
// Replaces the byte->profile step of src/tree/likelihood.rs:104-108 (Nucleotide::profile).
// ---------------------------------------------------------------------------------------------
// Tile transpose through shared memory: 64 patterns x 128 rows per block, 128 contiguous input bytes per
// pattern (whole 128-byte lines, read once), four codes per 32-bit store on the pattern axis.  Requires
// dst_off % 4 == 0 (the packing loops use chunk offsets that are multiples of 64).
constexpr int PK_P = 64, PK_R = 128;
__global__ void __launch_bounds__(256)
pack_pattern_major_tile_kernel(const uint8_t *__restrict__ in, int n_rows, int64_t n_patterns,
                               uint8_t *__restrict__ tips, int64_t P_pad, int64_t dst_off,
                               int *__restrict__ row_invalid)
{
    __shared__ uint8_t tile[PK_P][PK_R + 4];
    __shared__ uint8_t lut[256]; // byte -> tip code (blockDim.x == 256)
    lut[threadIdx.x] = encode_byte((uint8_t)threadIdx.x);
    __syncthreads();
    const int64_t p0 = (int64_t)blockIdx.x * PK_P;
    const int r0 = blockIdx.y * PK_R;
    for (int idx = threadIdx.x; idx < PK_P * PK_R; idx += 256) {
        const int j = idx >> 7, r = idx & (PK_R - 1);
        const int64_t p = p0 + j;
        const int row = r0 + r;
        uint8_t code = CODE_GAP;
        if (p < n_patterns && row < n_rows) code = lut[in[p * n_rows + row]];
        tile[j][r] = code;
    }
    __syncthreads();
    const int c = threadIdx.x & 15, rr = threadIdx.x >> 4;
    const int64_t p = p0 + 4 * c;
    for (int it = 0; it < PK_R / 16; ++it) {
        const int r = it * 16 + rr, row = r0 + r;
        if (row >= n_rows || p >= n_patterns) continue;
        const uint32_t c0 = tile[4 * c][r], c1 = tile[4 * c + 1][r], c2 = tile[4 * c + 2][r], c3 = tile[4 * c + 3][r];
        uint8_t *dst = tips + (int64_t)row * P_pad + dst_off + p;
        if (p + 3 < n_patterns) {
            *reinterpret_cast<uint32_t *>(dst) = c0 | (c1 << 8) | (c2 << 16) | (c3 << 24);
            if (c0 == CODE_INVALID || c1 == CODE_INVALID || c2 == CODE_INVALID || c3 == CODE_INVALID) row_invalid[row] = 1;
        } else {
            const uint32_t cs[4] = {c0, c1, c2, c3};
            for (int i = 0; i < 4 && p + i < n_patterns; ++i) {
                dst[i] = (uint8_t)cs[i];
                if (cs[i] == CODE_INVALID) row_invalid[row] = 1;
            }
        }
    }
}

__global__ void pack_row_major_kernel(const uint8_t *__restrict__ in, int n_rows, int64_t n_patterns,
                                      int64_t in_stride, int64_t p_first, uint8_t *__restrict__ tips,
                                      int64_t P_pad, int64_t dst_off, int *__restrict__ row_invalid)
{
    int r = blockIdx.y;
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_patterns;
         p += (int64_t)gridDim.x * blockDim.x) {
        uint8_t code = encode_byte(in[(int64_t)r * in_stride + p_first + p]);
        tips[(int64_t)r * P_pad + dst_off + p] = code;
        if (code == CODE_INVALID) row_invalid[r] = 1;
    }
}

__global__ void fill_pad_kernel(uint8_t *__restrict__ tips, int row0, int64_t P, int64_t P_pad)
{
    int r = row0 + blockIdx.y;
    for (int64_t p = P + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P_pad;
         p += (int64_t)gridDim.x * blockDim.x)
        tips[(int64_t)r * P_pad + p] = CODE_GAP;
}

__global__ void mult_to_double_kernel(const uint64_t *__restrict__ in, int64_t n, int64_t n_pad,
                                      double *__restrict__ out)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pad;
         i += (int64_t)gridDim.x * blockDim.x)
        out[i] = (i < n) ? (in ? (double)in[i] : 1.0) : 0.0;
}

__global__ void unpack_kernel(const uint8_t *__restrict__ tips, int n_rows, int row0, int64_t P_pad,
                              int64_t first, int64_t n, int layout, uint8_t *__restrict__ out)
{
    int r = row0 + blockIdx.y;
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n;
         p += (int64_t)gridDim.x * blockDim.x) {
        uint8_t ch = decode_code(tips[(int64_t)r * P_pad + first + p]);
        if (layout == 0) out[p * n_rows + r] = ch;
        else out[(int64_t)r * n + p] = ch;
    }
}

// ---------------------------------------------------------------------------------------------
// K1: per-branch transition matrices, one thread per (node, model category).
// P = diag(1/sqrt(pi)) * (V * diag(exp(lambda * r * t))) * V^T * diag(sqrt(pi));
// exact identity when r * t == 0.0.  Follows GtrModel::category_transition (src/models/gtr.rs:61-64)
// and RateMatrix::transition (src/models/rate_matrix.rs:113-135).
// The products and sums are rounded one by one (__dmul_rn / __dadd_rn: the compiler must not contract them into FMAs),
// as the reference's are: for a branch of ~1e-8 the off-diagonal entries (~1e-10) are what is left of O(1) terms that
// cancel, one ulp of those terms is 3e-7 of the entry, and a pattern that needs a substitution on such a branch carries
// that straight into its log-likelihood (2e-8 measured against the oracle with contracted FMAs; found by the random-tree
// tests).  With the reference's own rounding sequence the entries agree bit for bit wherever exp() does.
// ---------------------------------------------------------------------------------------------
__global__ void transition_kernel(const double *__restrict__ blen, int n_nodes, ModelDev M,
                                  double *__restrict__ P)
{
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_nodes * M.n_cat) return;
    int node = idx / M.n_cat, cat = idx - node * M.n_cat;
    double t = M.rates[cat] * blen[node];
    double *out = P + (size_t)idx * 16;
    if (t == 0.0) {
#pragma unroll
        for (int i = 0; i < 16; ++i) out[i] = ((i >> 2) == (i & 3)) ? 1.0 : 0.0;
        return;
    }
    double e[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) e[k] = exp(M.lambda[k] * t);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            double acc = __dmul_rn(__dmul_rn(M.V[i + 0], e[0]), M.V[j + 0]);
#pragma unroll
            for (int k = 1; k < 4; ++k) acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(M.V[i + 4 * k], e[k]), M.V[j + 4 * k])); // no FMA: see transition_kernel
            out[i * 4 + j] = M.inv_sqrt_pi[i] * acc * M.sqrt_pi[j];
        }
    }
}

// ---------------------------------------------------------------------------------------------
// K7: per-pattern tip constants.  leaf_ls = sum over leaves of the Profile::scale() log-scale of
// the leaf profile (src/tree/likelihood.rs:108, src/tree/likelihood/profile.rs:20-39: R,Y,.. add
// ln 1/2, B,D,.. ln 1/3, N ln 1/4, gap 0);  inv_mask = AND of all leaf masks = the rate-0 (+I)
// category's CLV at the root in closed form (src/models/gtr.rs:94-99).
// ---------------------------------------------------------------------------------------------
// One block = 128 sites x TIPC_SLICES leaf slices (one warp per slice, a lane = 4 sites, the leaf loop
// unrolled so that several 128-byte row segments are in flight per warp: the kernel is also launched on
// the narrow chunks of the host-streaming path, where one thread per site group walking all leaves was
// latency-bound).  The popcount classes are COUNTED (integers, combined with shared-memory atomics), so
// the result does not depend on how the leaves are split: ls = n2 ln(1/2) + n3 ln(1/3) + n4 ln(1/4).
constexpr int TIPC_SLICES = 8;
constexpr int TIPC_THREADS = 32 * TIPC_SLICES;
constexpr int TIPC_SITES = 128;
__global__ void __launch_bounds__(TIPC_THREADS)
tip_constants_kernel(const uint8_t *__restrict__ tips, int64_t P_pad, int64_t n_sites,
                     const int32_t *__restrict__ leaf_rows, int n_leaves, ModelDev M,
                     double *__restrict__ leaf_ls, uint8_t *__restrict__ inv_mask)
{
    // tips / leaf_ls / inv_mask may point at a pattern offset inside the shard; P_pad is the row stride
    __shared__ int s_cnt[3][TIPC_SITES];
    __shared__ int s_inv[TIPC_SITES];
    const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
    if (threadIdx.x < TIPC_SITES) {
        s_cnt[0][threadIdx.x] = 0;
        s_cnt[1][threadIdx.x] = 0;
        s_cnt[2][threadIdx.x] = 0;
        s_inv[threadIdx.x] = 15;
    }
    __syncthreads();
    const int64_t site = (int64_t)blockIdx.x * TIPC_SITES + lane * 4;
    if (site < n_sites) { // n_sites is a multiple of 4
        int cnt[3][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}, {0, 0, 0, 0}};
        uint32_t inv = 0x0F0F0F0Fu;
        const uint8_t *base = tips + site;
#pragma unroll 8
        for (int l = slice; l < n_leaves; l += TIPC_SLICES) {
            const uchar4 c4 = *reinterpret_cast<const uchar4 *>(base + (int64_t)leaf_rows[l] * P_pad);
            const uint32_t c[4] = {c4.x, c4.y, c4.z, c4.w};
            uint32_t m4 = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t m = code_mask(c[k]);
                m4 |= m << (8 * k);
                const int pc = (c[k] & 16u) ? 0 : __popc(m);
                cnt[0][k] += pc == 2;
                cnt[1][k] += pc == 3;
                cnt[2][k] += pc == 4;
            }
            inv &= m4;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int s = lane * 4 + k;
            if (cnt[0][k]) atomicAdd(&s_cnt[0][s], cnt[0][k]);
            if (cnt[1][k]) atomicAdd(&s_cnt[1][s], cnt[1][k]);
            if (cnt[2][k]) atomicAdd(&s_cnt[2][s], cnt[2][k]);
            atomicAnd(&s_inv[s], (int)((inv >> (8 * k)) & 15u));
        }
    }
    __syncthreads();
    if (threadIdx.x < TIPC_SITES) {
        const int64_t s = (int64_t)blockIdx.x * TIPC_SITES + threadIdx.x;
        if (s < n_sites) {
            leaf_ls[s] = (double)s_cnt[0][threadIdx.x] * M.ln_tip[2] + (double)s_cnt[1][threadIdx.x] * M.ln_tip[3] +
                         (double)s_cnt[2][threadIdx.x] * M.ln_tip[4];
            inv_mask[s] = (uint8_t)s_inv[threadIdx.x];
        }
    }
}

// ---------------------------------------------------------------------------------------------
// K2: level-batched CLV update (the HBM-bound hot kernel).
// One block = one (node, CLV category, 512-site block); one thread = one site PAIR, so every CLV
// access is a 16-byte load/store on the coalesced site axis.  Implements the internal-node branch
// of prune_pattern_category (src/tree/likelihood.rs:110-131):
//     v[i] = prod_children ( sum_j P_child[i][j] * clv_child[j] ),
// with Profile::scale()'s divide-by-max replaced by an exact power-of-two rescale whenever the
// site maximum drops below 2^-256; exponents accumulate in `scale` (integer atomics: exact and
// order-independent).  Tip children read one code byte per site and sum the masked columns of P.
// ---------------------------------------------------------------------------------------------
constexpr double RESCALE_THRESHOLD = 8.636168555094445e-78; // 2^-256

__device__ __forceinline__ void rescale_site(double &v0, double &v1, double &v2, double &v3, int *acc)
{
    double m = fmax(fmax(v0, v1), fmax(v2, v3));
    if (m < RESCALE_THRESHOLD && m > 0.0) {
        int s = -ilogb(m);
        if (s > 1000) s = 1000;
        double f = __longlong_as_double((long long)(1023 + s) << 52);
        v0 *= f; v1 *= f; v2 *= f; v3 *= f;
        atomicAdd(acc, s);
    }
}

__device__ __forceinline__ void load_P(const double *__restrict__ Pm, double (&p)[16])
{
    const double2 *q = reinterpret_cast<const double2 *>(Pm);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        double2 t = __ldg(q + i);
        p[2 * i] = t.x;
        p[2 * i + 1] = t.y;
    }
}

// contribution of an internal child for a site pair: out[i] = sum_j P[i][j] * x[j]
__device__ __forceinline__ void contrib_internal(const double (&p)[16], const double2 *__restrict__ src,
                                                 size_t stride2, double2 (&out)[4])
{
    double2 x0 = src[0], x1 = src[stride2], x2 = src[2 * stride2], x3 = src[3 * stride2];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        out[i].x = p[4 * i] * x0.x + p[4 * i + 1] * x1.x + p[4 * i + 2] * x2.x + p[4 * i + 3] * x3.x;
        out[i].y = p[4 * i] * x0.y + p[4 * i + 1] * x1.y + p[4 * i + 2] * x2.y + p[4 * i + 3] * x3.y;
    }
}

// contribution of a tip child: sum of the columns of P selected by the tip's state mask
__device__ __forceinline__ double masked_row(const double (&p)[16], int i, uint32_t m)
{
    double a = (m & 1u) ? p[4 * i] : 0.0;
    a += (m & 2u) ? p[4 * i + 1] : 0.0;
    a += (m & 4u) ? p[4 * i + 2] : 0.0;
    a += (m & 8u) ? p[4 * i + 3] : 0.0;
    return a;
}
__device__ __forceinline__ void contrib_tip(const double (&p)[16], const uint8_t *__restrict__ codes,
                                            double2 (&out)[4])
{
    uchar2 c = *reinterpret_cast<const uchar2 *>(codes);
    const uint32_t mx = code_mask(c.x), my = code_mask(c.y);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        out[i].x = masked_row(p, i, mx);
        out[i].y = masked_row(p, i, my);
    }
}

template <bool TIP0, bool TIP1>
__device__ __forceinline__ void clv_update_pair(const NodeOp &op, int cat_clv, int cat_model, int n_cat_model,
                                                const double *__restrict__ Pmat, double *__restrict__ clv,
                                                size_t slot_stride, size_t tile_pad,
                                                const uint8_t *__restrict__ tips, int64_t P_pad,
                                                int64_t tile_off, int *__restrict__ scale, int64_t site)
{
    double p[16];
    double2 a[4], b[4];
    const size_t stride2 = tile_pad / 2;
    const size_t cat_off = (size_t)cat_clv * 4 * tile_pad + (size_t)site;
    load_P(Pmat + ((size_t)op.e0 * n_cat_model + cat_model) * 16, p);
    if (TIP0) contrib_tip(p, tips + (int64_t)op.c0 * P_pad + tile_off + site, a);
    else contrib_internal(p, reinterpret_cast<const double2 *>(clv + (size_t)op.c0 * slot_stride + cat_off), stride2, a);
    load_P(Pmat + ((size_t)op.e1 * n_cat_model + cat_model) * 16, p);
    if (TIP1) contrib_tip(p, tips + (int64_t)op.c1 * P_pad + tile_off + site, b);
    else contrib_internal(p, reinterpret_cast<const double2 *>(clv + (size_t)op.c1 * slot_stride + cat_off), stride2, b);
    double2 v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        v[i].x = a[i].x * b[i].x;
        v[i].y = a[i].y * b[i].y;
    }
    int *sc = scale + (size_t)cat_clv * tile_pad + site;
    rescale_site(v[0].x, v[1].x, v[2].x, v[3].x, sc);
    rescale_site(v[0].y, v[1].y, v[2].y, v[3].y, sc + 1);
    double2 *dst = reinterpret_cast<double2 *>(clv + (size_t)op.out_slot * slot_stride + cat_off);
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i * stride2] = v[i];
}

__global__ void __launch_bounds__(CLV_THREADS)
clv_update_level_kernel(const NodeOp *__restrict__ ops, int n_ops, ModelDev M,
                        const double *__restrict__ Pmat, double *__restrict__ clv, size_t slot_stride,
                        size_t tile_pad, int tile_len, int blocks_per_row,
                        const uint8_t *__restrict__ tips, int64_t P_pad, int64_t tile_off,
                        int *__restrict__ scale)
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
    case 0: clv_update_pair<false, false>(op, cat_clv, cat_model, M.n_cat, Pmat, clv, slot_stride, tile_pad, tips, P_pad, tile_off, scale, site); break;
    case 1: clv_update_pair<true, false>(op, cat_clv, cat_model, M.n_cat, Pmat, clv, slot_stride, tile_pad, tips, P_pad, tile_off, scale, site); break;
    case 2: clv_update_pair<false, true>(op, cat_clv, cat_model, M.n_cat, Pmat, clv, slot_stride, tile_pad, tips, P_pad, tile_off, scale, site); break;
    default: clv_update_pair<true, true>(op, cat_clv, cat_model, M.n_cat, Pmat, clv, slot_stride, tile_pad, tips, P_pad, tile_off, scale, site); break;
    }
}

// power-of-two normalisation of one site's four states to a maximum in [1,2); returns the exponent added
__device__ __forceinline__ int normalize_site(double &v0, double &v1, double &v2, double &v3)
{
    const int h = max(max(__double2hiint(v0), __double2hiint(v1)), max(__double2hiint(v2), __double2hiint(v3)));
    if (h <= 0) return 0; // zero / negative maxima are left alone (Profile::scale's max_val <= 0 return)
    int s = 1023 - (h >> 20);
    if (s > 1000) s = 1000;
    const double f = __longlong_as_double((long long)(1023 + s) << 52);
    v0 *= f; v1 *= f; v2 *= f; v3 *= f;
    return s;
}

// Generic arity (1 child, or > 2 children): running product over the children (likelihood.rs:113-131).
// The reference multiplies child contributions whose profiles Profile::scale() left at max == 1 and scales
// once per node, so a node with very many children underflows there (a star of >= ~500 tips gives -inf).
// To leave the normal f64 range at the same place, internal children of a node with >= 3 children are first
// brought to a maximum in [1,2) (exact power of two, exponent accumulated) and the product is rescaled only
// once, at the end -- the same arithmetic as the fused sweep's SW_FLAG_NORM.
__global__ void __launch_bounds__(CLV_THREADS)
clv_update_generic_kernel(const GenOp *__restrict__ ops, int n_ops, const GenChild *__restrict__ kids,
                          ModelDev M, const double *__restrict__ Pmat, double *__restrict__ clv,
                          size_t slot_stride, size_t tile_pad, int tile_len, int blocks_per_row,
                          const uint8_t *__restrict__ tips, int64_t P_pad, int64_t tile_off,
                          int *__restrict__ scale)
{
    int sb = blockIdx.x % blocks_per_row;
    int rest = blockIdx.x / blocks_per_row;
    int cat_clv = rest % M.n_clv;
    int opi = rest / M.n_clv;
    if (opi >= n_ops) return;
    int64_t site = ((int64_t)sb * CLV_THREADS + threadIdx.x) * 2;
    if (site >= tile_len) return;
    const GenOp op = ops[opi];
    const int cat_model = M.cat_of_clv[cat_clv];
    const size_t stride2 = tile_pad / 2;
    const size_t cat_off = (size_t)cat_clv * 4 * tile_pad + (size_t)site;
    int *sc = scale + (size_t)cat_clv * tile_pad + site;
    const bool norm = op.child_end - op.child_begin >= 3;
    int ex = 0, ey = 0;
    double2 v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = make_double2(1.0, 1.0);
    for (int k = op.child_begin; k < op.child_end; ++k) {
        const GenChild ch = kids[k];
        double p[16];
        double2 a[4];
        load_P(Pmat + ((size_t)ch.edge * M.n_cat + cat_model) * 16, p);
        if (ch.is_tip) contrib_tip(p, tips + (int64_t)ch.src * P_pad + tile_off + site, a);
        else {
            const double2 *src = reinterpret_cast<const double2 *>(clv + (size_t)ch.src * slot_stride + cat_off);
            double2 x0 = src[0], x1 = src[stride2], x2 = src[2 * stride2], x3 = src[3 * stride2];
            if (norm) {
                ex += normalize_site(x0.x, x1.x, x2.x, x3.x);
                ey += normalize_site(x0.y, x1.y, x2.y, x3.y);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                a[i].x = p[4 * i] * x0.x + p[4 * i + 1] * x1.x + p[4 * i + 2] * x2.x + p[4 * i + 3] * x3.x;
                a[i].y = p[4 * i] * x0.y + p[4 * i + 1] * x1.y + p[4 * i + 2] * x2.y + p[4 * i + 3] * x3.y;
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            v[i].x *= a[i].x;
            v[i].y *= a[i].y;
        }
    }
    if (ex) atomicAdd(sc, ex);
    if (ey) atomicAdd(sc + 1, ey);
    rescale_site(v[0].x, v[1].x, v[2].x, v[3].x, sc);
    rescale_site(v[0].y, v[1].y, v[2].y, v[3].y, sc + 1);
    double2 *dst = reinterpret_cast<double2 *>(clv + (size_t)op.out_slot * slot_stride + cat_off);
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i * stride2] = v[i];
}

// A tree that is a single leaf (root without children): its "CLV" is the tip mask itself.
__global__ void clv_from_tip_kernel(double *__restrict__ clv, size_t slot_stride, int out_slot, int n_clv,
                                    size_t tile_pad, int tile_len, const uint8_t *__restrict__ tips,
                                    int64_t P_pad, int64_t tile_off, int row)
{
    int64_t site = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (site >= tile_len) return;
    uint32_t m = code_mask(tips[(int64_t)row * P_pad + tile_off + site]);
    for (int c = 0; c < n_clv; ++c)
        for (int i = 0; i < 4; ++i)
            clv[(size_t)out_slot * slot_stride + ((size_t)c * 4 + i) * tile_pad + site] = ((m >> i) & 1u) ? 1.0 : 0.0;
}

// ---------------------------------------------------------------------------------------------
// K3: root kernel.  Per pattern: root_mass_c = sum_i pi_i * clv_root[c][i]; cat_ll_c = ln w_c +
// ln root_mass_c + log_scale_c (src/tree/likelihood.rs:135-140); site_ll = log_sum_exp over
// categories (:61-68, :205); partial = sum mult * site_ll by warp shuffle + block reduction into
// one double per block (fixed order => deterministic).  The rate-0 (+I) category is evaluated in
// closed form from inv_mask unless the context carries it as a CLV category.
// ---------------------------------------------------------------------------------------------
constexpr int ROOT_THREADS = 256;
constexpr double LN2 = 0.6931471805599453;

__device__ __forceinline__ double block_reduce_sum(double v)
{
    __shared__ double warp_sums[32];
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) warp_sums[warp] = v;
    __syncthreads();
    double r = 0.0;
    if (threadIdx.x == 0) {
        int nw = (blockDim.x + 31) >> 5;
        for (int w = 0; w < nw; ++w) r += warp_sums[w];
    }
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(ROOT_THREADS)
root_kernel(const double *__restrict__ root_clv /* slot base */, size_t tile_pad, int tile_len,
            const int *__restrict__ scale, const double *__restrict__ leaf_ls /* + tile_off */,
            const uint8_t *__restrict__ inv_mask, const double *__restrict__ mult, ModelDev M,
            double *__restrict__ per_pattern /* + tile_off */, double *__restrict__ cat_ll_out /* nullable [n_cat][tile_pad] */,
            double *__restrict__ partials)
{
    int64_t site = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    double contrib = 0.0;
    if (site < tile_len) {
        double lls[MAX_CAT];
        const double lls_leaf = leaf_ls[site];
        double mx = -INFINITY;
        for (int c = 0; c < M.n_cat; ++c) {
            double mass = 0.0, ls = lls_leaf;
            int cc = M.clv_idx[c];
            if (cc >= 0) {
                const double *r = root_clv + (size_t)cc * 4 * tile_pad + site;
#pragma unroll
                for (int i = 0; i < 4; ++i) mass += M.pi[i] * r[(size_t)i * tile_pad];
                ls -= (double)scale[(size_t)cc * tile_pad + site] * LN2;
            } else {
                uint32_t m = inv_mask[site];
#pragma unroll
                for (int i = 0; i < 4; ++i) mass += M.pi[i] * (((m >> i) & 1u) ? 1.0 : 0.0);
            }
            double ll = M.ln_weights[c] + log(mass) + ls;
            lls[c] = ll;
            if (cat_ll_out) cat_ll_out[(size_t)c * tile_pad + site] = ll;
            mx = fmax(mx, ll);
        }
        double site_ll;
        if (!isfinite(mx)) site_ll = mx;
        else {
            double s = 0.0;
            for (int c = 0; c < M.n_cat; ++c) s += exp(lls[c] - mx);
            site_ll = mx + log(s);
        }
        per_pattern[site] = site_ll;
        contrib = mult[site] * site_ll;
    }
    double r = block_reduce_sum(contrib);
    if (threadIdx.x == 0) partials[blockIdx.x] = r;
}

// K4: final deterministic reduction of the per-block partials to one double per GPU.
// out[1] carries the fused sweep's error word (an op code the kernel does not know: a corrupted or mis-built
// op table), which is re-armed here for the next evaluation.
__global__ void final_reduce_kernel(const double *__restrict__ partials, int n, double *__restrict__ out,
                                    unsigned int *__restrict__ sweep_err)
{
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) acc += partials[i];
    double r = block_reduce_sum(acc);
    if (threadIdx.x == 0) {
        out[0] = r;
        out[1] = sweep_err ? (double)*sweep_err : 0.0;
        if (sweep_err) *sweep_err = 0u;
    }
}

// ---------------------------------------------------------------------------------------------
// K5/K6: marginal-ASR pre-order pass (src/tree/likelihood.rs:303-398), level-batched from the root.
// Posteriors overwrite the up-pass CLVs in place (same buffers, same layout).  One thread = one
// (parent node, pattern): it reads the parent's per-category posterior once, updates EVERY child
//     post_v = normalise( up_v (.) (P^T post_parent) )            (:321-342)
// and folds the category mixture  mixed_v = normalise( sum_c w_c post_v,c ),
// w_c = exp(cat_ll_c - site_ll) or 1/K when site_ll is not finite (:356-382), then the
// last-maximal-index argmax (:385-390) into one state byte per (node, pattern).
// ---------------------------------------------------------------------------------------------
struct DownOp {
    int32_t parent_slot;
    int32_t child_begin, child_end; // range in DownChild
    int32_t pad;
};
struct DownChild {
    int32_t node;  // dense id: P matrix index and row of the states output
    int32_t src;   // CLV slot (internal) or alignment row (leaf)
    int32_t is_tip;
    int32_t pad;   // 1: internal node with an internal child (the NNI upper pass computes partials only for these)
};

__device__ __forceinline__ int argmax_last(const double (&m)[4])
{
    int best = 0;
#pragma unroll
    for (int i = 1; i < 4; ++i)
        if (!(m[best] > m[i])) best = i;
    return best;
}

__device__ __forceinline__ void mix_weights(const double *__restrict__ cat_ll, size_t tile_pad, int64_t site,
                                            double site_ll, int n_cat, double *w)
{
    if (isfinite(site_ll))
        for (int c = 0; c < n_cat; ++c) w[c] = exp(cat_ll[(size_t)c * tile_pad + site] - site_ll);
    else
        for (int c = 0; c < n_cat; ++c) w[c] = 1.0 / (double)n_cat;
}

// Root: post = normalise(pi (.) up_root) per category (:304-316), in place; plus its own mixture.
__global__ void __launch_bounds__(ROOT_THREADS)
asr_root_kernel(double *__restrict__ root_clv, size_t tile_pad, int tile_len, ModelDev M,
                double *__restrict__ cat_ll /* in: per-category lnL; out: mixing weights */,
                const double *__restrict__ site_ll /* + tile_off */,
                uint8_t *__restrict__ states_row /* + tile_off */, double *__restrict__ post_row /* nullable, [site][4] */)
{
    int64_t site = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (site >= tile_len) return;
    double w[MAX_CAT];
    mix_weights(cat_ll, tile_pad, site, site_ll[site], M.n_cat, w);
    for (int c = 0; c < M.n_cat; ++c) cat_ll[(size_t)c * tile_pad + site] = w[c]; // reused by every down level
    double mixed[4] = {0.0, 0.0, 0.0, 0.0};
    for (int c = 0; c < M.n_cat; ++c) {
        double *r = root_clv + (size_t)c * 4 * tile_pad + site;
        double post[4], sum = 0.0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            post[i] = M.pi[i] * r[(size_t)i * tile_pad];
            sum += post[i];
        }
        if (sum > 0.0) {
#pragma unroll
            for (int i = 0; i < 4; ++i) post[i] /= sum;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            r[(size_t)i * tile_pad] = post[i];
            mixed[i] += w[c] * post[i];
        }
    }
    double sum = mixed[0] + mixed[1] + mixed[2] + mixed[3];
    if (sum > 0.0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) mixed[i] /= sum;
    }
    states_row[site] = (uint8_t)argmax_last(mixed);
    if (post_row) {
#pragma unroll
        for (int i = 0; i < 4; ++i) post_row[site * 4 + i] = mixed[i];
    }
}

constexpr int DOWN_MAX_CHILDREN = 2; // register fast path; larger arities loop in chunks

// One block = (parent node, 512 sites); one thread = a site PAIR (16-byte accesses on the site axis).
// `mixw` holds the per-site category weights exp(cat_ll_c - site_ll) (likelihood.rs:356-363), written
// once per tile by asr_root_kernel into the cat_ll buffer, so no exp() runs here.
#ifndef PHB_DOWN_THREADS
#define PHB_DOWN_THREADS 128
#endif
constexpr int DOWN_THREADS = PHB_DOWN_THREADS; // a block covers 2 * DOWN_THREADS patterns of one parent
#ifndef PHB_DOWN_MINB
#define PHB_DOWN_MINB 5 // measured on B200 (5k x 500k): 128 threads x >= 5 blocks 241 ms; 256 x 1: 262; 128 x 6: 251; 128 x 8: 334
#endif
__global__ void __launch_bounds__(DOWN_THREADS, PHB_DOWN_MINB)
asr_down_level_kernel(const DownOp *__restrict__ ops, int n_ops, const DownChild *__restrict__ kids,
                      ModelDev M, const double *__restrict__ Pmat, double *__restrict__ clv,
                      size_t slot_stride, size_t tile_pad, int tile_len, int blocks_per_row,
                      const uint8_t *__restrict__ tips, int64_t P_pad, int64_t tile_off,
                      const double *__restrict__ mixw,
                      uint8_t *__restrict__ states /* [node][P_pad] */, double *__restrict__ posteriors /* nullable [node][P_pad][4] */)
{
    const int sb = blockIdx.x % blocks_per_row;
    const int opi = blockIdx.x / blocks_per_row;
    if (opi >= n_ops) return;
    const int64_t site = ((int64_t)sb * blockDim.x + threadIdx.x) * 2;
    if (site >= tile_len) return;
    const bool has2 = site + 1 < tile_len;
    const DownOp op = ops[opi];
    for (int k0 = op.child_begin; k0 < op.child_end; k0 += DOWN_MAX_CHILDREN) {
        const int nk = min(DOWN_MAX_CHILDREN, op.child_end - k0);
        double2 mixed[DOWN_MAX_CHILDREN][4];
#pragma unroll
        for (int k = 0; k < DOWN_MAX_CHILDREN; ++k)
#pragma unroll
            for (int i = 0; i < 4; ++i) mixed[k][i] = make_double2(0.0, 0.0);
        for (int c = 0; c < M.n_cat; ++c) {
            const double *pp = clv + (size_t)op.parent_slot * slot_stride + (size_t)c * 4 * tile_pad + site;
            double2 par[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) par[i] = *reinterpret_cast<const double2 *>(pp + (size_t)i * tile_pad);
            const double2 w = *reinterpret_cast<const double2 *>(mixw + (size_t)c * tile_pad + site);
#pragma unroll
            for (int k = 0; k < DOWN_MAX_CHILDREN; ++k) {
                if (k >= nk) break;
                const DownChild ch = kids[k0 + k];
                double p[16];
                load_P(Pmat + ((size_t)ch.node * M.n_cat + c) * 16, p);
                double2 up[4];
                double *self = nullptr;
                if (ch.is_tip) {
                    const uint8_t *tp = tips + (int64_t)ch.src * P_pad + tile_off + site;
                    const uint32_t m0 = code_mask(tp[0]), m1 = code_mask(tp[1]); // site + 1 < P_pad always
#pragma unroll
                    for (int i = 0; i < 4; ++i) up[i] = make_double2(((m0 >> i) & 1u) ? 1.0 : 0.0, ((m1 >> i) & 1u) ? 1.0 : 0.0);
                } else {
                    self = clv + (size_t)ch.src * slot_stride + (size_t)c * 4 * tile_pad + site;
#pragma unroll
                    for (int i = 0; i < 4; ++i) up[i] = *reinterpret_cast<const double2 *>(self + (size_t)i * tile_pad);
                }
                double2 post[4], sum = make_double2(0.0, 0.0);
#pragma unroll
                for (int i = 0; i < 4; ++i) { // msg_i = sum_j P[j][i] * par[j]  (P^T * post_parent)
                    const double mx = p[0 * 4 + i] * par[0].x + p[1 * 4 + i] * par[1].x + p[2 * 4 + i] * par[2].x + p[3 * 4 + i] * par[3].x;
                    const double my = p[0 * 4 + i] * par[0].y + p[1 * 4 + i] * par[1].y + p[2 * 4 + i] * par[2].y + p[3 * 4 + i] * par[3].y;
                    post[i] = make_double2(up[i].x * mx, up[i].y * my);
                    sum.x += post[i].x;
                    sum.y += post[i].y;
                }
                // normalise (skipped when the sum is not > 0: the vector stays as it is, likelihood.rs:337-341)
                const double ix = sum.x > 0.0 ? 1.0 / sum.x : 1.0, iy = sum.y > 0.0 ? 1.0 / sum.y : 1.0;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    post[i].x *= ix; // one reciprocal per site instead of four divisions (<= 1 ulp apart)
                    post[i].y *= iy;
                    if (self) *reinterpret_cast<double2 *>(self + (size_t)i * tile_pad) = post[i];
                    mixed[k][i].x += w.x * post[i].x;
                    mixed[k][i].y += w.y * post[i].y;
                }
            }
        }
#pragma unroll
        for (int k = 0; k < DOWN_MAX_CHILDREN; ++k) {
            if (k >= nk) break;
            const DownChild ch = kids[k0 + k];
            double m0[4] = {mixed[k][0].x, mixed[k][1].x, mixed[k][2].x, mixed[k][3].x};
            double m1[4] = {mixed[k][0].y, mixed[k][1].y, mixed[k][2].y, mixed[k][3].y};
            const double s0 = m0[0] + m0[1] + m0[2] + m0[3], s1 = m1[0] + m1[1] + m1[2] + m1[3];
            if (s0 > 0.0) { // renormalised before the argmax, as the reference does (likelihood.rs:376-390)
#pragma unroll
                for (int i = 0; i < 4; ++i) m0[i] /= s0;
            }
            if (s1 > 0.0) {
#pragma unroll
                for (int i = 0; i < 4; ++i) m1[i] /= s1;
            }
            uint8_t *so = states + (int64_t)ch.node * P_pad + tile_off + site;
            so[0] = (uint8_t)argmax_last(m0);
            if (has2) so[1] = (uint8_t)argmax_last(m1);
            if (posteriors) {
                double *o = posteriors + ((int64_t)ch.node * P_pad + tile_off + site) * 4;
#pragma unroll
                for (int i = 0; i < 4; ++i) o[i] = m0[i];
                if (has2) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) o[4 + i] = m1[i];
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Synthetic alignment simulator (benchmark input generator): per site draw a rate
// category ~ weights, a root state ~ pi, then walk the tree in pre-order drawing each child state
// from row P(r_c t)[parent_state].  Counter-based RNG: splitmix64 of (seed, node, global site).
// ---------------------------------------------------------------------------------------------
__host__ __device__ inline uint64_t splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__host__ __device__ inline double u01(uint64_t seed, uint64_t node, uint64_t site)
{
    uint64_t h = splitmix64(seed ^ splitmix64(node * 0xD1342543DE82EF95ull + 0x632BE59BD9B4E019ull) ^ splitmix64(site));
    return (double)(h >> 11) * (1.0 / 9007199254740992.0);
}

__global__ void simulate_kernel(const int32_t *__restrict__ preorder, const int32_t *__restrict__ parent,
                                const int32_t *__restrict__ depth, const int32_t *__restrict__ leaf_row,
                                int n_nodes, int root, ModelDev M, const double *__restrict__ Pmat,
                                uint8_t *__restrict__ stack /* [max_depth+1][P_pad] */,
                                uint8_t *__restrict__ tips, int64_t P_pad, int64_t n_patterns,
                                int64_t global_offset, uint64_t seed)
{
    int64_t site = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (site >= n_patterns) return;
    uint64_t g = (uint64_t)(global_offset + site);
    double u = u01(seed, 0xFFFFFFFFull, g);
    int cat = M.n_cat - 1;
    double acc = 0.0;
    for (int c = 0; c < M.n_cat; ++c) {
        acc += M.weights[c];
        if (u < acc) { cat = c; break; }
    }
    for (int k = 0; k < n_nodes; ++k) {
        int v = preorder[k];
        int s = 3;
        double uu = u01(seed, (uint64_t)v, g);
        if (v == root) {
            double a = 0.0;
            for (int i = 0; i < 4; ++i) { a += M.pi[i]; if (uu < a) { s = i; break; } }
        } else {
            int ps = stack[(int64_t)(depth[v] - 1) * P_pad + site];
            const double *row = Pmat + ((size_t)v * M.n_cat + cat) * 16 + ps * 4;
            double a = 0.0;
            for (int i = 0; i < 4; ++i) { a += row[i]; if (uu < a) { s = i; break; } }
        }
        stack[(int64_t)depth[v] * P_pad + site] = (uint8_t)s;
        int r = leaf_row[v];
        if (r >= 0) tips[(int64_t)r * P_pad + site] = (uint8_t)s; // one-hot mask of state s has slot s
    }
}

// FP64 FMA peak microbenchmark (confirms the CLV kernel sits under the HBM roof, not the DFMA roof).
__global__ void fp64_fma_kernel(double *out, int iters)
{
    double a0 = threadIdx.x * 1e-3, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double b = 1.0000001, c = 1e-9;
    for (int i = 0; i < iters; ++i) {
        a0 = a0 * b + c; a1 = a1 * b + c; a2 = a2 * b + c; a3 = a3 * b + c;
        a4 = a4 * b + c; a5 = a5 * b + c; a6 = a6 * b + c; a7 = a7 * b + c;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

} // namespace phb
