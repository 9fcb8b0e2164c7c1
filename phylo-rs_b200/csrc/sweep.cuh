// sweep.cuh — the fused post-order sweep: the whole Felsenstein pruning of one block of site patterns
// inside one CTA, conditional-likelihood vectors (CLVs) never leaving the SM.
//
// Replaces, for compute_log_likelihood (src/tree/likelihood.rs:156-209), the per-level CLV kernels of
// kernels.cuh: instead of writing every internal node's CLV to HBM and reading it back once
// (32*K*(1+i)+t algorithmic bytes per update) a CTA walks the tree in a depth-first
// order chosen by the host so that at most `need(root)` partial results are alive at any time
// (Sethi–Ullman numbering: the child subtree that needs more slots is evaluated first).  The running
// result lives in registers, pending siblings in a shared-memory stack (overflow levels in an
// L2-resident global stack), and the only HBM traffic left is the tip codes (n bytes per pattern).
//
// Work split: 1 thread = R (=4) consecutive site patterns of ONE rate category; a warp = 128 patterns
// of one category, so every transition matrix is warp-uniform.  A dedicated producer warp streams the
// per-op records (transition matrices / tip tables, op-ordered so the stream is read linearly) and the
// tip-code rows into a shared-memory ring with TMA bulk copies (cp.async.bulk + mbarrier
// complete_tx); consumer warps never touch global memory in the steady state.
//
// Per-op arithmetic follows prune_pattern_category (src/tree/likelihood.rs:110-131):
//     v[i] = prod_children sum_j P_child[i][j] * clv_child[j]
// with a tip child's sum replaced by a lookup of sum_{j in mask} P[i][j] (the 15 non-empty state
// masks of Nucleotide::profile, src/alphabet.rs:44-64, tabulated once per (tip edge, category)), and
// Profile::scale()'s divide-by-max replaced by exact power-of-two rescaling tracked per (site, category).
#pragma once
#include "kernels.cuh"

namespace phb {

// ---- op stream ---------------------------------------------------------------------------------------
enum SweepKind : int32_t {
    SW_TT = 0,      // cur = tip(A) * tip(B)                    [optional push of cur first]
    SW_TI = 1,      // cur = tip(A) * (P_B cur)
    SW_II = 2,      // cur = (P_A pop(level)) * (P_B cur)
    SW_APPLY = 3,   // cur = P_A cur                            (unifurcation / polytomy accumulation)
    SW_MULTIP = 4,  // cur = cur * tip(A)
    SW_MULPOP = 5,  // cur = cur * pop(level)
    SW_LOADTIP = 6, // cur = tip(A)                             [optional push of cur first]
};
constexpr int SW_FLAG_PUSH = 1;    // push cur to `level` before executing (TT / LOADTIP only)
constexpr int SW_FLAG_RESCALE = 2; // node complete: run the underflow check
constexpr int SW_FLAG_U = 4;       // TT / TI: operate on the second register accumulator `u` (cur untouched)
constexpr int SW_FLAG_XU = 8;      // II: the pending operand is `u`, not a stack level
constexpr int SW_FLAG_NORM = 16;   // APPLY of a child of a node with >= 3 children: bring the child's CLV to max in [1,2)
                                   // first (exact power of two), as the reference's Profile::scale() leaves every child
                                   // at max == 1 before the parent's product (profile.rs:20-39, likelihood.rs:113-131)

constexpr int SW_P_BYTES = 128;   // 4x4 f64 row-major
constexpr int SW_TAB_BYTES = 512; // 16 masks x 4 f64, slot order SW_PERM
constexpr int SW_BAR_BYTES = 512; // mbarrier area at the start of dynamic shared memory

// Per-op header, one table for the whole tree in global memory (identical for every site block, so the
// consumers prefetch it two ops ahead into registers: nothing on the op's critical path waits for it).
//   w0 = code | flags << 8 | level << 16      w1 = szA | szB << 16   (per-category bytes of operand A / B)
//        code = kind, + 8 when the op works on / takes its pending operand from the second accumulator `u`
//   w2 = CLV slot of the node this op completes (valid with SW_FLAG_RESCALE; used by the storing variant)
struct SweepOp {
    uint32_t w0, w1, w2, w3;
};
// Producer-side descriptor (global memory): where the record is and which tip rows it needs.
struct SweepDesc {
    uint64_t rec_off;  // byte offset of the record in the stream
    uint32_t rec_bytes;
    int32_t rowA, rowB; // alignment rows of tip operands, -1 = none
    int32_t pad[3];
};
// One operand block the transition kernel has to fill.
struct SweepBlock {
    uint64_t dst_off; // byte offset of category 0's block; category c at dst_off + c * cat_stride
    uint32_t cat_stride;
    int32_t edge;     // dense node id whose branch length parametrises P
    int32_t is_table;
    int32_t pad;
};

// Tip tables are indexed by the SLOT that the packed tip codes already carry (kernels.cuh:
// mask_to_slot): one-hot masks (A,C,G,T) are slots 0..3, so their 32-byte rows fall in four different
// bank groups and gathers on unambiguous data are conflict-free.

// ---- K1s: transition matrices / tip tables straight into the op-ordered stream --------------------
__global__ void sweep_transition_kernel(const SweepBlock *__restrict__ blocks, int n_blocks,
                                        const double *__restrict__ blen, ModelDev M,
                                        unsigned char *__restrict__ stream)
{
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_blocks * M.n_clv) return;
    int b = idx / M.n_clv, cc = idx - b * M.n_clv;
    const SweepBlock blk = blocks[b];
    const int cat = M.cat_of_clv[cc];
    double t = M.rates[cat] * blen[blk.edge];
    double p[16];
    if (t == 0.0) {
#pragma unroll
        for (int i = 0; i < 16; ++i) p[i] = ((i >> 2) == (i & 3)) ? 1.0 : 0.0;
    } else {
        double e[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) e[k] = exp(M.lambda[k] * t);
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                double acc = (M.V[i + 0] * e[0]) * M.V[j + 0];
#pragma unroll
                for (int k = 1; k < 4; ++k) acc += (M.V[i + 4 * k] * e[k]) * M.V[j + 4 * k];
                p[i * 4 + j] = M.inv_sqrt_pi[i] * acc * M.sqrt_pi[j];
            }
    }
    double *out = reinterpret_cast<double *>(stream + blk.dst_off + (size_t)cc * blk.cat_stride);
    if (!blk.is_table) {
#pragma unroll
        for (int i = 0; i < 16; ++i) out[i] = p[i];
    } else {
        for (uint32_t m = 0; m < 16; ++m) {
            double *row = out + 4 * mask_to_slot(m);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                double a = (m & 1u) ? p[4 * i] : 0.0;
                a += (m & 2u) ? p[4 * i + 1] : 0.0;
                a += (m & 4u) ? p[4 * i + 2] : 0.0;
                a += (m & 8u) ? p[4 * i + 3] : 0.0;
                row[i] = a;
            }
        }
    }
}

// ---- mbarrier / TMA bulk-copy primitives -----------------------------------------------------------
#ifdef PHB_HOST_EMU
typedef emu::Mbar sw_mbar_t;
__device__ inline void sw_mbar_init(sw_mbar_t *b, int count) { b->count = b->pending = count; b->tx = 0; b->phase = 0; }
__device__ inline void sw_fence_init() {}
__device__ inline void sw_mbar_expect_tx(sw_mbar_t *b, uint32_t bytes) { b->tx += bytes; b->pending -= 1; emu::mbar_check(b); }
__device__ inline void sw_mbar_arrive(sw_mbar_t *b) { b->pending -= 1; emu::mbar_check(b); }
__device__ inline void sw_mbar_wait(sw_mbar_t *b, uint32_t parity) { while (b->phase == (int)parity) emu::yield(); }
__device__ inline void sw_bulk_g2s(void *dst, const void *src, uint32_t bytes, sw_mbar_t *b)
{
    memcpy(dst, src, bytes);
    b->tx -= bytes;
    emu::mbar_check(b);
}
__device__ inline void sw_consumer_barrier(int n) { emu::named_barrier(1, (unsigned)n); }
#else
typedef unsigned long long sw_mbar_t;
__device__ __forceinline__ uint32_t sw_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void sw_mbar_init(sw_mbar_t *b, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(sw_smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void sw_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void sw_mbar_expect_tx(sw_mbar_t *b, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sw_smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void sw_mbar_arrive(sw_mbar_t *b)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(sw_smem_u32(b)) : "memory");
}
__device__ __forceinline__ void sw_mbar_wait(sw_mbar_t *b, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(sw_smem_u32(b)),
        "r"(parity)
        : "memory");
}
// TMA bulk copy global -> shared, completion signalled on the mbarrier (complete_tx::bytes).
__device__ __forceinline__ void sw_bulk_g2s(void *dst, const void *src, uint32_t bytes, sw_mbar_t *b)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(sw_smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(sw_smem_u32(b))
                 : "memory");
}
__device__ __forceinline__ void sw_consumer_barrier(int n) { asm volatile("bar.sync 1, %0;" ::"r"(n) : "memory"); }
#endif

// ---- kernel parameters -------------------------------------------------------------------------------
constexpr int SW_R = 4;              // site patterns per thread
#ifndef PHB_SW_CONSUMERS
#define PHB_SW_CONSUMERS 320
#endif
constexpr int SW_MAX_CONSUMERS = PHB_SW_CONSUMERS; // consumer threads per CTA (+ one producer warp)

struct SweepParams {
    const SweepOp *ops;
    const SweepDesc *descs;
    const unsigned char *stream;
    int n_ops;
    int Q;            // site quads per CTA; consumer threads = Q * n_clv
    int n_consumers;  // Q * n_clv
    int n_stages;     // ring depth
    int stage_bytes;  // bytes per ring stage (record + two code rows)
    int rec_cap;      // max record bytes (offset of the code rows inside a stage)
    int smem_levels;  // stack levels held in shared memory; deeper ones go to gstack
    int g_levels;     // levels available in gstack per CTA
    double *gstack;   // [cta][g_level][R][2][n_consumers] double2
    const uint8_t *tips;
    int64_t P_pad;    // row stride of tips
    int64_t P;        // valid patterns
    int64_t n_blocks; // ceil(P / (4 Q))
    const double *leaf_ls;
    const uint8_t *inv_mask;
    const double *mult;
    double *per_pattern;
    double *partials; // [n_blocks]
    unsigned int *error_flag;
    // storing variant (marginal-ASR up pass): every completed node's CLV goes to clv_out[slot][cat][state][site]
    // (site relative to the first pattern of this launch), the per-category log-likelihoods to cat_ll_out[cat][site]
    double *clv_out;
    double *cat_ll_out;
    size_t slot_stride; // n_clv * 4 * tile_pad
    size_t tile_pad;
};

__device__ __forceinline__ void sw_matvec(const double (&p)[16], const double (&x)[4], double (&y)[4])
{
#pragma unroll
    for (int i = 0; i < 4; ++i) y[i] = p[4 * i] * x[0] + p[4 * i + 1] * x[1] + p[4 * i + 2] * x[2] + p[4 * i + 3] * x[3];
}

__device__ __forceinline__ void sw_load_P(const unsigned char *blk, double (&p)[16])
{
    const double2 *q = reinterpret_cast<const double2 *>(blk);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        double2 t = q[i];
        p[2 * i] = t.x;
        p[2 * i + 1] = t.y;
    }
}

__device__ __forceinline__ void sw_load_tip(const unsigned char *tab, uint32_t code, double (&t)[4])
{
    const double2 *q = reinterpret_cast<const double2 *>(tab + 32u * (code & 15u));
    double2 a = q[0], b = q[1];
    t[0] = a.x; t[1] = a.y; t[2] = b.x; t[3] = b.y;
}

// exact power-of-two rescale of one site's 4 states when the largest drops below 2^-256 (ALWAYS = false), or
// unconditionally to a maximum in [1,2) (ALWAYS = true: the child of a polytomy, see SW_FLAG_NORM)
template <bool ALWAYS = false>
__device__ __forceinline__ void sw_rescale(double (&v)[4], int &sc)
{
    int h = max(max(__double2hiint(v[0]), __double2hiint(v[1])), max(__double2hiint(v[2]), __double2hiint(v[3])));
    if ((ALWAYS || h < ((1023 - 256) << 20)) && h > 0) { // positive (zero / negative maxima are left alone)
        int e = h >> 20;                     // biased exponent of the maximum (0 = subnormal)
        int s = 1023 - e;
        if (s > 1000) s = 1000;
        double f = __longlong_as_double((long long)(1023 + s) << 52);
#pragma unroll
        for (int i = 0; i < 4; ++i) v[i] *= f;
        sc += s;
    }
}

// The fused sweep.  grid = min(n_blocks, #SMs) persistent CTAs; block = n_consumers (padded to a warp
// multiple) + 32 producer threads; dynamic shared memory = barriers + ring + stack.
// STORE = true is the up pass of compute_marginal_asr (src/tree/likelihood.rs:286-298): the same sweep, but
// every completed node's CLV is also written to HBM ONCE (the level kernels write it and read it back), at
// whatever power-of-two scale it has in registers — the pre-order pass normalises per node and category
// (:330-342), so only the per-category totals, which the root stage writes to cat_ll_out, carry exponents.
template <bool STORE>
__global__ void __launch_bounds__(SW_MAX_CONSUMERS + 32, 1)
sweep_kernel(SweepParams p, ModelDev M)
{
    PHB_DYN_SMEM(smem);
    const int tid = threadIdx.x;
    const int n_cons_warps = (p.n_consumers + 31) >> 5;
    const int prod_tid0 = n_cons_warps * 32;
    sw_mbar_t *full = reinterpret_cast<sw_mbar_t *>(smem);
    sw_mbar_t *empty = full + p.n_stages;
    unsigned char *ring = smem + SW_BAR_BYTES;
    unsigned char *stack = ring + (size_t)p.n_stages * p.stage_bytes;
    if (tid == 0) {
        for (int s = 0; s < p.n_stages; ++s) {
            sw_mbar_init(&full[s], 1);
            sw_mbar_init(&empty[s], n_cons_warps);
        }
        sw_fence_init();
    }
    __syncthreads();
    const int sites_per_block = 4 * p.Q;
    const int code_bytes = 4 * p.Q;

    if (tid >= prod_tid0) {
        // ===================== producer warp: one lane feeds the ring =====================
        if (tid == prod_tid0) {
            int s = 0;
            uint32_t parity = 1; // a fresh barrier passes a parity-1 wait
            const int n_stages = p.n_stages, stage_bytes = p.stage_bytes, rec_cap = p.rec_cap, n_ops = p.n_ops;
            for (int64_t b = blockIdx.x; b < p.n_blocks; b += gridDim.x) {
                const uint8_t *tip_base = p.tips + b * sites_per_block;
                SweepDesc d = p.descs[0];
                for (int k = 0; k < n_ops; ++k) {
                    const SweepDesc nd = p.descs[k + 1 < n_ops ? k + 1 : k]; // prefetch the next descriptor
                    sw_mbar_wait(&empty[s], parity);
                    unsigned char *st = ring + (size_t)s * stage_bytes;
                    uint32_t bytes = d.rec_bytes + (d.rowA >= 0 ? code_bytes : 0) + (d.rowB >= 0 ? code_bytes : 0);
                    sw_mbar_expect_tx(&full[s], bytes);
                    if (d.rec_bytes) sw_bulk_g2s(st, p.stream + d.rec_off, d.rec_bytes, &full[s]); // MULPOP has no record
                    if (d.rowA >= 0) sw_bulk_g2s(st + rec_cap, tip_base + (int64_t)d.rowA * p.P_pad, code_bytes, &full[s]);
                    if (d.rowB >= 0) sw_bulk_g2s(st + rec_cap + code_bytes, tip_base + (int64_t)d.rowB * p.P_pad, code_bytes, &full[s]);
                    d = nd;
                    if (++s == n_stages) {
                        s = 0;
                        parity ^= 1u;
                    }
                }
            }
        }
        return;
    }

    // ===================== consumer warps =====================
    const bool active = tid < p.n_consumers;
    const int t = active ? tid : 0;   // inactive lanes of the last warp shadow thread 0 (no stores)
    const int cc = t / p.Q;           // CLV category of this thread
    const int q = t - cc * p.Q;       // site quad inside the block
    const int lane = tid & 31;
    const int n_cons = p.n_consumers;
    const int n_stages = p.n_stages, stage_bytes = p.stage_bytes, n_ops = p.n_ops, smem_levels = p.smem_levels;
    const int lvl_stride = SW_R * 2 * n_cons; // double2 elements per stack level
    double2 *const sstack = reinterpret_cast<double2 *>(stack) + t;
    double2 *const gstack = reinterpret_cast<double2 *>(p.gstack) + (size_t)blockIdx.x * p.g_levels * lvl_stride + t;
    const int code_off = p.rec_cap + 4 * q;

    int s = 0;
    uint32_t parity = 0;
    const unsigned char *st = ring;
    const uint4 *const optab = reinterpret_cast<const uint4 *>(p.ops);
    for (int64_t b = blockIdx.x; b < p.n_blocks; b += gridDim.x) {
        double v[SW_R][4]; // running result
        double u[SW_R][4]; // second accumulator: a chain-shaped pending sibling is evaluated here (SW_FLAG_U)
        int sc[SW_R];
#pragma unroll
        for (int r = 0; r < SW_R; ++r) {
            sc[r] = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) v[r][i] = u[r][i] = 1.0;
        }
        // op headers run two ops ahead of the arithmetic (same table for every block)
        uint4 hq0 = __ldg(optab), hq1 = __ldg(optab + (n_ops > 1 ? 1 : 0));
        for (int k = 0; k < n_ops; ++k) {
            const uint4 h = hq0;
            hq0 = hq1;
            hq1 = __ldg(optab + (k + 2 < n_ops ? k + 2 : n_ops - 1));
            const int code = h.x & 0xff, kind = code & 7, flags = (h.x >> 8) & 0xff, level = (int)(h.x >> 16);
            const int szA = h.y & 0xffff, szB = (int)(h.y >> 16);
            sw_mbar_wait(&full[s], parity);
            const unsigned char *blkA = st + cc * (szA + szB);
            const unsigned char *blkB = blkA + szA;
            const unsigned char *codesA = st + code_off;
            const unsigned char *codesB = codesA + code_bytes;
            // ---- the op bodies, written once over the accumulator they update ----
            auto push = [&]() { // park the running result before a TT / LOADTIP overwrites it
                double2 *slot = level < smem_levels ? sstack + level * lvl_stride
                                                    : gstack + (size_t)(level - smem_levels) * lvl_stride;
                if (active) {
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) {
                        slot[(2 * r) * n_cons] = make_double2(v[r][0], v[r][1]);
                        slot[(2 * r + 1) * n_cons] = make_double2(v[r][2], v[r][3]);
                    }
                }
            };
            auto op_tt = [&](double (&acc)[SW_R][4]) {
                const uchar4 ca = *reinterpret_cast<const uchar4 *>(codesA);
                const uchar4 cb = *reinterpret_cast<const uchar4 *>(codesB);
                const uint32_t a4[4] = {ca.x, ca.y, ca.z, ca.w}, b4[4] = {cb.x, cb.y, cb.z, cb.w};
#pragma unroll
                for (int r = 0; r < SW_R; ++r) {
                    double ta[4], tb[4];
                    sw_load_tip(blkA, a4[r], ta);
                    sw_load_tip(blkB, b4[r], tb);
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[r][i] = ta[i] * tb[i];
                }
            };
            auto op_ti = [&](double (&acc)[SW_R][4]) {
                const uchar4 ca = *reinterpret_cast<const uchar4 *>(codesA);
                const uint32_t a4[4] = {ca.x, ca.y, ca.z, ca.w};
                double P[16];
                sw_load_P(blkB, P);
#pragma unroll
                for (int r = 0; r < SW_R; ++r) {
                    double y[4], ta[4];
                    sw_load_tip(blkA, a4[r], ta);
                    sw_matvec(P, acc[r], y);
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[r][i] = ta[i] * y[i];
                }
            };
            // v = (P_A x) * (P_B v)
            auto op_ii = [&](const double (&x)[SW_R][4]) {
                double P[16], xa[SW_R][4];
                sw_load_P(blkA, P);
#pragma unroll
                for (int r = 0; r < SW_R; ++r) sw_matvec(P, x[r], xa[r]);
                sw_load_P(blkB, P);
#pragma unroll
                for (int r = 0; r < SW_R; ++r) {
                    double y[4];
                    sw_matvec(P, v[r], y);
#pragma unroll
                    for (int i = 0; i < 4; ++i) v[r][i] = xa[r][i] * y[i];
                }
            };
            auto pop = [&](double (&x)[SW_R][4]) {
                const double2 *slot = level < smem_levels ? sstack + level * lvl_stride
                                                          : gstack + (size_t)(level - smem_levels) * lvl_stride;
#pragma unroll
                for (int r = 0; r < SW_R; ++r) {
                    const double2 a = slot[(2 * r) * n_cons], c = slot[(2 * r + 1) * n_cons];
                    x[r][0] = a.x; x[r][1] = a.y; x[r][2] = c.x; x[r][3] = c.y;
                }
            };
            // underflow guard: one test for the four sites, the smallest per-site maximum decides
            auto guard = [&](double (&acc)[SW_R][4]) {
                int hmin = 0x7fffffff;
#pragma unroll
                for (int r = 0; r < SW_R; ++r)
                    hmin = min(hmin, max(max(__double2hiint(acc[r][0]), __double2hiint(acc[r][1])),
                                         max(__double2hiint(acc[r][2]), __double2hiint(acc[r][3]))));
                if (hmin < ((1023 - 256) << 20)) {
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) sw_rescale(acc[r], sc[r]);
                }
            };
            const bool resc = flags & SW_FLAG_RESCALE;
            switch (code) { // dense dispatch code, precomputed by the host scheduler
            case SW_TT:
                if (flags & SW_FLAG_PUSH) push();
                op_tt(v);
                if (resc) guard(v);
                break;
            case SW_TT | 8:
                op_tt(u);
                if (resc) guard(u);
                break;
            case SW_TI:
                op_ti(v);
                if (resc) guard(v);
                break;
            case SW_TI | 8:
                op_ti(u);
                if (resc) guard(u);
                break;
            case SW_II | 8:
                op_ii(u);
                if (resc) guard(v);
                break;
            case SW_II: {
                double x[SW_R][4];
                pop(x);
                op_ii(x);
                if (resc) guard(v);
                break;
            }
            case SW_APPLY: {
                if (flags & SW_FLAG_NORM) {
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) sw_rescale<true>(v[r], sc[r]);
                }
                double P[16];
                sw_load_P(blkA, P);
#pragma unroll
                for (int r = 0; r < SW_R; ++r) {
                    double y[4];
                    sw_matvec(P, v[r], y);
#pragma unroll
                    for (int i = 0; i < 4; ++i) v[r][i] = y[i];
                }
                if (resc) guard(v);
                break;
            }
            case SW_MULTIP:
            case SW_LOADTIP: {
                if (flags & SW_FLAG_PUSH) push();
                const uchar4 ca = *reinterpret_cast<const uchar4 *>(codesA);
                const uint32_t a4[4] = {ca.x, ca.y, ca.z, ca.w};
#pragma unroll
                for (int r = 0; r < SW_R; ++r) {
                    double ta[4];
                    sw_load_tip(blkA, a4[r], ta);
#pragma unroll
                    for (int i = 0; i < 4; ++i) v[r][i] = (kind == SW_LOADTIP) ? ta[i] : v[r][i] * ta[i];
                }
                if (resc) guard(v);
                break;
            }
            case SW_MULPOP: {
                double x[SW_R][4];
                pop(x);
#pragma unroll
                for (int r = 0; r < SW_R; ++r)
#pragma unroll
                    for (int i = 0; i < 4; ++i) v[r][i] *= x[r][i];
                if (resc) guard(v);
                break;
            }
            default:
                if (lane == 0) atomicOr(p.error_flag, 1u);
                break;
            }
            if (STORE && resc && active) {
                const int64_t rel = b * sites_per_block + 4 * q; // first of this thread's 4 patterns, launch-relative
                double *dst = p.clv_out + (size_t)h.z * p.slot_stride + (size_t)cc * 4 * p.tile_pad + (size_t)rel;
                const bool use_u = flags & SW_FLAG_U;
                if (rel + 3 < p.P) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        double2 *d2 = reinterpret_cast<double2 *>(dst + (size_t)i * p.tile_pad);
                        d2[0] = use_u ? make_double2(u[0][i], u[1][i]) : make_double2(v[0][i], v[1][i]);
                        d2[1] = use_u ? make_double2(u[2][i], u[3][i]) : make_double2(v[2][i], v[3][i]);
                    }
                } else {
#pragma unroll
                    for (int r = 0; r < SW_R; ++r)
                        if (rel + r < p.P) {
#pragma unroll
                            for (int i = 0; i < 4; ++i) dst[(size_t)i * p.tile_pad + r] = use_u ? u[r][i] : v[r][i];
                        }
                }
            }
            __syncwarp();
            if (lane == 0) sw_mbar_arrive(&empty[s]);
            st += stage_bytes;
            if (++s == n_stages) {
                s = 0;
                st = ring;
                parity ^= 1u;
            }
        }

        // ---- root: per-category log-likelihoods, category mix, weights, block reduction ----
        // (src/tree/likelihood.rs:135-140, :61-68, :205).  The stack is idle here; its first bytes hold
        // the [model category][site] exchange buffer.
        sw_consumer_barrier(n_cons_warps * 32); // warps drift: nobody may still be using stack level 0
        double *xch = reinterpret_cast<double *>(stack);
        const int64_t site0 = b * sites_per_block + 4 * q;
        if (active) {
            const int cat = M.cat_of_clv[cc];
#pragma unroll
            for (int r = 0; r < SW_R; ++r) {
                const int64_t site = site0 + r;
                double ll = 0.0;
                if (site < p.P) {
                    double mass = 0.0;
#pragma unroll
                    for (int i = 0; i < 4; ++i) mass += M.pi[i] * v[r][i];
                    ll = M.ln_weights[cat] + log(mass) + (p.leaf_ls[site] - (double)sc[r] * LN2);
                }
                xch[(size_t)cat * sites_per_block + 4 * q + r] = ll;
                if (STORE && site < p.P) p.cat_ll_out[(size_t)cat * p.tile_pad + (size_t)site] = ll;
            }
            if (cc == 0 && M.n_clv < M.n_cat) { // rate-0 (+I) categories in closed form from the tip masks
                for (int c = 0; c < M.n_cat; ++c) {
                    if (M.clv_idx[c] >= 0) continue;
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) {
                        const int64_t site = site0 + r;
                        double ll = 0.0;
                        if (site < p.P) {
                            const uint32_t m = p.inv_mask[site];
                            double mass = 0.0;
#pragma unroll
                            for (int i = 0; i < 4; ++i) mass += M.pi[i] * (((m >> i) & 1u) ? 1.0 : 0.0);
                            ll = M.ln_weights[c] + log(mass) + p.leaf_ls[site];
                        }
                        xch[(size_t)c * sites_per_block + 4 * q + r] = ll;
                        if (STORE && site < p.P) p.cat_ll_out[(size_t)c * p.tile_pad + (size_t)site] = ll;
                    }
                }
            }
        }
        sw_consumer_barrier(n_cons_warps * 32);
        double contrib = 0.0;
        if (active && cc == 0) {
#pragma unroll
            for (int r = 0; r < SW_R; ++r) {
                const int64_t site = site0 + r;
                if (site >= p.P) continue;
                double mx = -INFINITY;
                for (int c = 0; c < M.n_cat; ++c) mx = fmax(mx, xch[(size_t)c * sites_per_block + 4 * q + r]);
                double site_ll;
                if (!isfinite(mx)) site_ll = mx;
                else {
                    double sum = 0.0;
                    for (int c = 0; c < M.n_cat; ++c) sum += exp(xch[(size_t)c * sites_per_block + 4 * q + r] - mx);
                    site_ll = mx + log(sum);
                }
                p.per_pattern[site] = site_ll;
                contrib += p.mult[site] * site_ll;
            }
        }
        // block reduction over the consumer warps (fixed order => deterministic)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) contrib += __shfl_down_sync(0xffffffffu, contrib, o);
        sw_consumer_barrier(n_cons_warps * 32); // everyone is done reading xch
        double *wsum = reinterpret_cast<double *>(stack);
        if (lane == 0) wsum[tid >> 5] = contrib;
        sw_consumer_barrier(n_cons_warps * 32);
        if (tid == 0) {
            double tot = 0.0;
            for (int w = 0; w < n_cons_warps; ++w) tot += wsum[w];
            p.partials[b] = tot;
        }
        sw_consumer_barrier(n_cons_warps * 32); // the stack is reused by the next block
    }
}

} // namespace phb
