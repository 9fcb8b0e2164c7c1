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
// masks of Nucleotide::profile, src/alphabet.rs:44-64, tabulated once per (tip edge, category)), a
// cherry (two tip children) by ONE lookup in the 5x5 table of products over the codes A,C,G,T,N/gap,
// and Profile::scale()'s divide-by-max replaced by exact power-of-two rescaling tracked per (site, category).
#pragma once
#include "kernels.cuh"

namespace phb {

// ---- op stream ---------------------------------------------------------------------------------------
enum SweepKind : int32_t {
    SW_TT = 0,      // cur = tip(A) * tip(B) (cherry table)     [optional push of cur first]
    SW_TI = 1,      // cur = tip(A) * (P_B cur)
    SW_II = 2,      // cur = (P_A pop(level)) * (P_B cur)
    SW_APPLY = 3,   // cur = P_A cur                            (unifurcation / polytomy accumulation)
    SW_MULTIP = 4,  // cur = cur * tip(A)
    SW_MULPOP = 5,  // cur = cur * pop(level)
    SW_LOADTIP = 6, // cur = tip(A)                             [optional push of cur first]
};
constexpr int SW_FLAG_PUSH = 1;    // push cur to `level` before executing (TT / LOADTIP only)
constexpr int SW_FLAG_DONE = 2;    // this op completes a node (the storing variant writes its CLV out)
constexpr int SW_FLAG_U = 4;       // TT / TI: operate on the second register accumulator `u` (cur untouched)
constexpr int SW_FLAG_XU = 8;      // II: the pending operand is `u`, not a stack level
constexpr int SW_FLAG_NORM = 16;   // APPLY of a child of a node with >= 3 children: bring the child's CLV to max in [1,2)
                                   // first (exact power of two), as the reference's Profile::scale() leaves every child
                                   // at max == 1 before the parent's product (profile.rs:20-39, likelihood.rs:113-131)
constexpr int SW_FLAG_GUARD = 32;  // after the op: rescale the sites whose maximum fell below 2^-256 (placed by the host
                                   // where the bound on the accumulated shrink requires it, sweep_sched.h plan_guards)
constexpr int SW_FLAG_NORMOUT = 64; // after the op: rescale every site to max in [1,2) (child of an op whose own
                                    // worst-case shrink leaves no room for a lazily scaled operand)

constexpr int SW_P_BYTES = 128;    // 4x4 f64 row-major
constexpr int SW_TAB_BYTES = 512;  // 16 masks x 4 f64, slot order (kernels.cuh mask_to_slot)
constexpr int SW_TAB2_BYTES = 800; // cherry table: 5 x 5 code pairs (A,C,G,T,N/gap) x 4 f64, row = a + 5 b
constexpr int SW_BAR_BYTES = 512;  // mbarrier area at the start of dynamic shared memory
constexpr int SW_HDR_BYTES = 16;   // every record starts with the header of the NEXT op

// Per-op header.  The table lives in global memory (two padding entries); the transition kernel copies header
// k + 1 to the front of record k, so a consumer picks up the next op's header from the ring stage it is working
// on -- one warp-uniform LDS whose latency the current op's arithmetic hides -- and only the first header of a
// site block comes from global memory.
//   w0 = code | flags << 8 | level << 16      w1 = per-category record bytes | szA << 16 (operand B at +szA)
//        code = kind, + 8 when the op works on / takes its pending operand from the second accumulator `u`
//   w2 = CLV slot of the node this op completes (valid with SW_FLAG_DONE; used by the storing variant)
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
    int32_t is_table; // 0: P(t) 4x4; 1: tip table of `edge`; 2: cherry table of (edge, edge2)
    int32_t edge2;
};

// ---- K1s: transition matrices / tip tables straight into the op-ordered stream --------------------
__device__ __forceinline__ void sw_transition(const ModelDev &M, double t, double (&p)[16])
{
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
}
// sum over the states of `mask` of row i of P, in ascending state order (bit-identical for every table kind)
__device__ __forceinline__ double sw_masked(const double (&p)[16], int i, uint32_t m)
{
    double a = (m & 1u) ? p[4 * i] : 0.0;
    a += (m & 2u) ? p[4 * i + 1] : 0.0;
    a += (m & 4u) ? p[4 * i + 2] : 0.0;
    a += (m & 8u) ? p[4 * i + 3] : 0.0;
    return a;
}

// Threads past the operand blocks copy the op headers into the stream: the record of op k starts with the
// header of op k + 1 (SW_HDR_BYTES), which the consumers read from the ring while op k's arithmetic runs.
__global__ void sweep_transition_kernel(const SweepBlock *__restrict__ blocks, int n_blocks,
                                        const double *__restrict__ blen, ModelDev M,
                                        unsigned char *__restrict__ stream, const SweepOp *__restrict__ ops,
                                        const SweepDesc *__restrict__ descs, int n_ops)
{
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_blocks * M.n_clv) {
        const int k = idx - n_blocks * M.n_clv;
        if (k < n_ops) *reinterpret_cast<uint4 *>(stream + descs[k].rec_off) = *reinterpret_cast<const uint4 *>(ops + k + 1);
        return;
    }
    int b = idx / M.n_clv, cc = idx - b * M.n_clv;
    const SweepBlock blk = blocks[b];
    const int cat = M.cat_of_clv[cc];
    double p[16];
    sw_transition(M, M.rates[cat] * blen[blk.edge], p);
    double *out = reinterpret_cast<double *>(stream + blk.dst_off + (size_t)cc * blk.cat_stride);
    if (blk.is_table == 0) {
#pragma unroll
        for (int i = 0; i < 16; ++i) out[i] = p[i];
    } else if (blk.is_table == 1) {
        for (uint32_t m = 0; m < 16; ++m) {
            double *row = out + 4 * mask_to_slot(m);
#pragma unroll
            for (int i = 0; i < 4; ++i) row[i] = sw_masked(p, i, m);
        }
    } else {
        double q[16];
        sw_transition(M, M.rates[cat] * blen[blk.edge2], q);
        for (uint32_t a = 0; a < 5; ++a)
            for (uint32_t c = 0; c < 5; ++c) { // two planes of 25 x 16 bytes: states 0,1 | states 2,3
                double *lo = out + 2 * (a + 5 * c), *hi = lo + SW_TAB2_BYTES / 16;
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    lo[i] = sw_masked(p, i, code_mask(a)) * sw_masked(q, i, code_mask(c));
                    hi[i] = sw_masked(p, i + 2, code_mask(a)) * sw_masked(q, i + 2, code_mask(c));
                }
            }
    }
}

// ---- shared-memory, mbarrier and TMA bulk-copy primitives -----------------------------------------
// Device build: everything in dynamic shared memory (barriers, ring stages, stack) is addressed by its 32-bit
// shared-window address, derived ONCE per thread from an opaque base, and read / written with explicit
// ld.shared / st.shared.  Left to itself the compiler re-derives the window base (S2R SR_CgaCtaId / SR_SWINHI)
// and the lane id (S2R SR_TID) inside the op loop to save registers, and the loop then waits on those
// special-register reads (ncu: 12 % of all warp samples, profiles/r2b_sweep_source.txt).
#ifdef PHB_HOST_EMU
typedef emu::Mbar sw_mbar_t;
typedef unsigned char *sw_sptr;
__device__ inline sw_sptr sw_sbase(unsigned char *smem) { return smem; }
__device__ inline double2 sw_lds128(sw_sptr a) { return *reinterpret_cast<const double2 *>(a); }
__device__ inline uint32_t sw_lds32(sw_sptr a) { uint32_t v; memcpy(&v, a, 4); return v; }
__device__ inline uint4 sw_lds128u(sw_sptr a) { return *reinterpret_cast<const uint4 *>(a); }
__device__ inline void sw_sts128(sw_sptr a, double2 v) { *reinterpret_cast<double2 *>(a) = v; }
__device__ inline void sw_sts64(sw_sptr a, double v) { *reinterpret_cast<double *>(a) = v; }
__device__ inline double sw_lds64(sw_sptr a) { return *reinterpret_cast<const double *>(a); }
__device__ inline uint32_t sw_laneid() { return threadIdx.x & 31u; }
__device__ inline emu::Mbar *sw_bar(sw_sptr base, int i) { return reinterpret_cast<emu::Mbar *>(base) + i; }
__device__ inline void sw_mbar_init(sw_sptr base, int i, int count) { emu::Mbar *b = sw_bar(base, i); b->count = b->pending = count; b->tx = 0; b->phase = 0; }
__device__ inline void sw_fence_init() {}
__device__ inline void sw_mbar_expect_tx(sw_sptr base, int i, uint32_t bytes) { emu::Mbar *b = sw_bar(base, i); b->tx += bytes; b->pending -= 1; emu::mbar_check(b); }
__device__ inline void sw_mbar_arrive(sw_sptr base, int i) { emu::Mbar *b = sw_bar(base, i); b->pending -= 1; emu::mbar_check(b); }
__device__ inline void sw_mbar_wait(sw_sptr base, int i, uint32_t parity) { emu::Mbar *b = sw_bar(base, i); while (b->phase == (int)parity) emu::yield(); }
__device__ inline void sw_bulk_g2s(sw_sptr dst, const void *src, uint32_t bytes, sw_sptr base, int i)
{
    emu::Mbar *b = sw_bar(base, i);
    memcpy(dst, src, bytes);
    b->tx -= bytes;
    emu::mbar_check(b);
}
__device__ inline void sw_mbar_arrive_elect(sw_sptr base, int i)
{
    __syncwarp();
    if (sw_laneid() == 0u) sw_mbar_arrive(base, i);
}
__device__ inline void sw_consumer_barrier(int n) { emu::named_barrier(1, (unsigned)n); }
#define SW_OPAQUE(x) ((void)0)
#else
#define SW_OPAQUE(x) asm volatile("" : "+r"(x))
typedef unsigned long long sw_mbar_t;
typedef uint32_t sw_sptr;
__device__ __forceinline__ sw_sptr sw_sbase(unsigned char *smem)
{
    uint32_t a = (uint32_t)__cvta_generic_to_shared(smem), r;
    asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(a));
    return r;
}
__device__ __forceinline__ double2 sw_lds128(sw_sptr a)
{
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a));
    return v;
}
__device__ __forceinline__ double sw_lds64(sw_sptr a)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint32_t sw_lds32(sw_sptr a)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint4 sw_lds128u(sw_sptr a)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ void sw_sts128(sw_sptr a, double2 v)
{
    asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(a), "d"(v.x), "d"(v.y) : "memory");
}
__device__ __forceinline__ void sw_sts64(sw_sptr a, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory"); }
__device__ __forceinline__ uint32_t sw_laneid()
{
    uint32_t l;
    asm volatile("mov.u32 %0, %%laneid;" : "=r"(l));
    return l;
}
__device__ __forceinline__ void sw_mbar_init(sw_sptr base, int i, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(base + 8u * (uint32_t)i), "r"(count) : "memory");
}
__device__ __forceinline__ void sw_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void sw_mbar_expect_tx(sw_sptr base, int i, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(base + 8u * (uint32_t)i), "r"(bytes) : "memory");
}
__device__ __forceinline__ void sw_mbar_arrive(sw_sptr base, int i)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(base + 8u * (uint32_t)i) : "memory");
}
__device__ __forceinline__ void sw_mbar_wait(sw_sptr base, int i, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(base + 8u * (uint32_t)i),
        "r"(parity)
        : "memory");
}
// TMA bulk copy global -> shared, completion signalled on the mbarrier (complete_tx::bytes).
__device__ __forceinline__ void sw_bulk_g2s(sw_sptr dst, const void *src, uint32_t bytes, sw_sptr base, int i)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(base + 8u * (uint32_t)i)
                 : "memory");
}
// one elected lane of the (converged) warp arrives: no lane-id register, no separate warp barrier
__device__ __forceinline__ void sw_mbar_arrive_elect(sw_sptr base, int i)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "elect.sync _|p, 0xffffffff;\n"
        "@p mbarrier.arrive.shared::cta.b64 _, [%0];\n"
        "}\n" ::"r"(base + 8u * (uint32_t)i)
        : "memory");
}
__device__ __forceinline__ void sw_consumer_barrier(int n) { asm volatile("bar.sync 1, %0;" ::"r"(n) : "memory"); }
#endif

// ---- kernel parameters -------------------------------------------------------------------------------
constexpr int SW_R = 4;              // site patterns per thread
// 11 consumer warps + the producer warp = 12 warps = 3 per SM sub-partition (16K registers each: 168 per thread);
// 10 + 1 left two schedulers a warp short (tools/ubench.cu: DFMA throughput at 10 warps is 83 % of that at 12)
#ifndef PHB_SW_CONSUMERS
#define PHB_SW_CONSUMERS 352
#endif
constexpr int SW_MAX_CONSUMERS = PHB_SW_CONSUMERS; // consumer threads per CTA (+ one producer warp)

struct SweepParams {
    const SweepOp *ops;
    const SweepDesc *descs;
    const unsigned char *stream;
    int n_ops;
    int Q;            // site quads per CTA; consumer threads = Q * n_clv
    int n_consumers;  // Q * n_clv
    int n_stages;     // ring depth (a power of two)
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

__device__ __forceinline__ void sw_load_P(sw_sptr blk, double (&p)[16])
{
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        double2 t = sw_lds128(blk + 16u * i);
        p[2 * i] = t.x;
        p[2 * i + 1] = t.y;
    }
}

// one 32-byte row of a tip table
__device__ __forceinline__ void sw_load_row(sw_sptr tab, uint32_t row, double (&t)[4])
{
    const sw_sptr a = tab + 32u * row;
    double2 x = sw_lds128(a), y = sw_lds128(a + 16u);
    t[0] = x.x; t[1] = x.y; t[2] = y.x; t[3] = y.y;
}
// one row of a cherry table: states 0,1 in the first plane, 2,3 in the second (16-byte row stride: the four
// rows a == b, which carry most sites, fall in four different bank groups)
__device__ __forceinline__ void sw_load_row2(sw_sptr tab, uint32_t row, double (&t)[4])
{
    const sw_sptr a = tab + 16u * row;
    double2 x = sw_lds128(a), y = sw_lds128(a + (uint32_t)(SW_TAB2_BYTES / 2));
    t[0] = x.x; t[1] = x.y; t[2] = y.x; t[3] = y.y;
}

// exact power-of-two rescale of one site's 4 states when the largest drops below 2^-256 (ALWAYS = false), or
// unconditionally to a maximum in [1,2) (ALWAYS = true: SW_FLAG_NORM / SW_FLAG_NORMOUT)
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
    const int n_stages = p.n_stages, n_ops = p.n_ops;
    const uint32_t stage_bytes = (uint32_t)p.stage_bytes;
    const sw_sptr bars = sw_sbase(smem);            // mbarriers: full[s] = index s, empty[s] = index n_stages + s
    const sw_sptr ring = bars + SW_BAR_BYTES;       // n_stages x stage_bytes
    const sw_sptr stack_s = ring + (uint32_t)n_stages * stage_bytes;
    unsigned char *const stack = smem + SW_BAR_BYTES + (size_t)n_stages * stage_bytes; // same place, generic (root stage)
    if (tid == 0) {
        for (int s = 0; s < n_stages; ++s) {
            sw_mbar_init(bars, s, 1);
            sw_mbar_init(bars, n_stages + s, n_cons_warps);
        }
        sw_fence_init();
    }
    __syncthreads();
    const int sites_per_block = 4 * p.Q;
    const uint32_t code_bytes = 4u * (uint32_t)p.Q;
    const uint32_t smask = (uint32_t)n_stages - 1u;

    if (tid >= prod_tid0) {
        // ===================== producer warp =====================
        // Lane 0 waits for the stage and posts the byte count; lanes 0..2 then issue one bulk copy each (the
        // record and the two tip-code rows).
        const int pl = tid - prod_tid0;
        uint32_t s = 0, parity = 1; // a fresh barrier passes a parity-1 wait
        const uint32_t rec_cap = (uint32_t)p.rec_cap;
        for (int64_t b = blockIdx.x; b < p.n_blocks; b += gridDim.x) {
            const uint8_t *tip_base = p.tips + b * sites_per_block;
            SweepDesc d = p.descs[0];
#pragma unroll 1
            for (int k = 0; k < n_ops; ++k) {
                const SweepDesc nd = p.descs[k + 1]; // the table carries one padding entry
                const sw_sptr st = ring + s * stage_bytes;
                const uint32_t bytes = d.rec_bytes + (d.rowA >= 0 ? code_bytes : 0u) + (d.rowB >= 0 ? code_bytes : 0u);
                if (pl == 0) {
                    sw_mbar_wait(bars, n_stages + (int)s, parity);
                    sw_mbar_expect_tx(bars, (int)s, bytes);
                }
                __syncwarp();
                if (pl == 0) {
                    if (d.rec_bytes) sw_bulk_g2s(st, p.stream + d.rec_off, d.rec_bytes, bars, (int)s); // MULPOP has no record
                } else if (pl == 1) {
                    if (d.rowA >= 0) sw_bulk_g2s(st + rec_cap, tip_base + (int64_t)d.rowA * p.P_pad, code_bytes, bars, (int)s);
                } else if (pl == 2) {
                    if (d.rowB >= 0) sw_bulk_g2s(st + rec_cap + code_bytes, tip_base + (int64_t)d.rowB * p.P_pad, code_bytes, bars, (int)s);
                }
                d = nd;
                s = (s + 1u) & smask;
                parity ^= (s == 0u) ? 1u : 0u;
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
    const int smem_levels = p.smem_levels;
    const int lvl_stride = SW_R * 2 * n_cons; // double2 elements per stack level
    const sw_sptr sstack = stack_s + 16u * (uint32_t)t;
    double2 *const gstack = reinterpret_cast<double2 *>(p.gstack) + (size_t)blockIdx.x * p.g_levels * lvl_stride + t;
    const uint32_t code_off = (uint32_t)p.rec_cap + 4u * (uint32_t)q;
    const uint32_t row_stride0 = 16u * (uint32_t)n_cons; // bytes between the two halves / the sites of a stack entry

    uint32_t s = 0, parity = 0;
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
        uint4 hn = __ldg(optab); // header of the first op; every later one arrives with the previous op's record
#pragma unroll 1
        for (int k = 0; k < n_ops; ++k) {
            const uint4 h = hn;
            const uint32_t code = h.x & 0xffu, flags = h.x >> 8;
            const sw_sptr st = ring + s * stage_bytes;
            const sw_sptr blkA = st + (uint32_t)SW_HDR_BYTES + (uint32_t)cc * (h.y & 0xffffu);
            const sw_sptr blkB = blkA + (h.y >> 16);
            const sw_sptr codesA = st + code_off;
            const sw_sptr codesB = codesA + code_bytes;
            sw_mbar_wait(bars, (int)s, parity);
            hn = sw_lds128u(st);
            // ---- the op bodies, written once over the accumulator they update ----
            const int level = (int)((h.x >> 16) & 0xffu);
            auto push = [&]() { // park the running result before a TT / LOADTIP overwrites it
                if (!active) return;
                if (level < smem_levels) {
                    uint32_t row_stride = row_stride0; // opaque here: keeps the seven multiples out of every op's prologue
                    SW_OPAQUE(row_stride);
                    const sw_sptr slot = sstack + (uint32_t)(level * lvl_stride) * 16u;
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) {
                        sw_sts128(slot + (uint32_t)(2 * r) * row_stride, make_double2(v[r][0], v[r][1]));
                        sw_sts128(slot + (uint32_t)(2 * r + 1) * row_stride, make_double2(v[r][2], v[r][3]));
                    }
                } else {
                    int nc = n_cons;
                    SW_OPAQUE(nc);
                    double2 *slot = gstack + (size_t)(level - smem_levels) * lvl_stride;
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) {
                        slot[(2 * r) * nc] = make_double2(v[r][0], v[r][1]);
                        slot[(2 * r + 1) * nc] = make_double2(v[r][2], v[r][3]);
                    }
                }
            };
            auto pop = [&](double (&x)[SW_R][4]) {
                if (level < smem_levels) {
                    uint32_t row_stride = row_stride0;
                    SW_OPAQUE(row_stride);
                    const sw_sptr slot = sstack + (uint32_t)(level * lvl_stride) * 16u;
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) {
                        const double2 a = sw_lds128(slot + (uint32_t)(2 * r) * row_stride), c = sw_lds128(slot + (uint32_t)(2 * r + 1) * row_stride);
                        x[r][0] = a.x; x[r][1] = a.y; x[r][2] = c.x; x[r][3] = c.y;
                    }
                } else {
                    int nc = n_cons;
                    SW_OPAQUE(nc);
                    const double2 *slot = gstack + (size_t)(level - smem_levels) * lvl_stride;
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) {
                        const double2 a = slot[(2 * r) * nc], c = slot[(2 * r + 1) * nc];
                        x[r][0] = a.x; x[r][1] = a.y; x[r][2] = c.x; x[r][3] = c.y;
                    }
                }
            };
            // cherry: one row of the 5x5 product table per site when both codes are among A,C,G,T,N/gap (slots
            // 0..4); a site with any other ambiguity code sums the masked columns of the two P(t) that follow the
            // table, in the reference's order (rare, divergent)
            auto op_tt = [&](double (&acc)[SW_R][4]) {
                const uint32_t ca = sw_lds32(codesA) & 0x0f0f0f0fu, cb = sw_lds32(codesB) & 0x0f0f0f0fu;
                // a byte is <= 4  <=>  bit 3 is clear and bit 2 does not come with bit 0 or 1
                const uint32_t bad = (ca & 0x08080808u) | (ca & 0x04040404u & ((ca << 1) | (ca << 2))) |
                                     (cb & 0x08080808u) | (cb & 0x04040404u & ((cb << 1) | (cb << 2)));
                const bool fast = bad == 0u;
                if (fast) {
                    const uint32_t idx = ca + 5u * cb; // per byte a + 5 b <= 24: no carries
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) sw_load_row2(blkA, (idx >> (8 * r)) & 0xffu, acc[r]);
                } else {
                    double PA[16], PB[16];
                    sw_load_P(blkB, PA);
                    sw_load_P(blkB + SW_P_BYTES, PB);
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) {
                        const uint32_t ma = code_mask((ca >> (8 * r)) & 15u), mb = code_mask((cb >> (8 * r)) & 15u);
#pragma unroll
                        for (int i = 0; i < 4; ++i) acc[r][i] = sw_masked(PA, i, ma) * sw_masked(PB, i, mb);
                    }
                }
            };
            auto op_ti = [&](double (&acc)[SW_R][4]) {
                const uint32_t ca = sw_lds32(codesA);
                double P[16];
                sw_load_P(blkB, P);
#pragma unroll
                for (int r = 0; r < SW_R; ++r) {
                    double y[4], ta[4];
                    sw_load_row(blkA, (ca >> (8 * r)) & 15u, ta);
                    sw_matvec(P, acc[r], y);
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[r][i] = ta[i] * y[i];
                }
            };
            // v = (P_A u) * (P_B v).  The pending operand is ALWAYS in `u`: either the chain-shaped sibling was
            // evaluated there (SW_FLAG_XU) or it is popped into it -- `u` is dead whenever a stack level is popped (it
            // is live only between a chain's first TT|U and the II|XU that consumes it, and chains hold no pops) -- so
            // the op needs no third set of 16 doubles
            auto op_ii = [&]() {
                double P[16];
                sw_load_P(blkA, P);
#pragma unroll
                for (int r = 0; r < SW_R; ++r) {
                    double y[4];
                    sw_matvec(P, u[r], y);
#pragma unroll
                    for (int i = 0; i < 4; ++i) u[r][i] = y[i];
                }
                sw_load_P(blkB, P);
#pragma unroll
                for (int r = 0; r < SW_R; ++r) {
                    double y[4];
                    sw_matvec(P, v[r], y);
#pragma unroll
                    for (int i = 0; i < 4; ++i) v[r][i] = u[r][i] * y[i];
                }
            };
            // underflow guard: one test for the four sites, the smallest per-site maximum decides
            auto guard = [&](double (&acc)[SW_R][4]) {
                if (flags & SW_FLAG_NORMOUT) {
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) sw_rescale<true>(acc[r], sc[r]);
                    return;
                }
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
            const bool resc = flags & (SW_FLAG_GUARD | SW_FLAG_NORMOUT);
            switch (code) { // dense dispatch code, precomputed by the host scheduler
            case SW_TI:
                op_ti(v);
                if (resc) guard(v);
                break;
            case SW_TI | 8:
                op_ti(u);
                if (resc) guard(u);
                break;
            case SW_II | 8:
                op_ii();
                if (resc) guard(v);
                break;
            case SW_TT | 8:
                op_tt(u);
                if (resc) guard(u);
                break;
            case SW_TT:
                if (flags & SW_FLAG_PUSH) push();
                op_tt(v);
                if (resc) guard(v);
                break;
            case SW_II:
                pop(u);
                op_ii();
                if (resc) guard(v);
                break;
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
                const uint32_t ca = sw_lds32(codesA);
#pragma unroll
                for (int r = 0; r < SW_R; ++r) {
                    double ta[4];
                    sw_load_row(blkA, (ca >> (8 * r)) & 15u, ta);
#pragma unroll
                    for (int i = 0; i < 4; ++i) v[r][i] = (code == SW_LOADTIP) ? ta[i] : v[r][i] * ta[i];
                }
                if (resc) guard(v);
                break;
            }
            case SW_MULPOP:
                pop(u);
#pragma unroll
                for (int r = 0; r < SW_R; ++r)
#pragma unroll
                    for (int i = 0; i < 4; ++i) v[r][i] *= u[r][i];
                if (resc) guard(v);
                break;
            default:
                if (lane == 0) atomicOr(p.error_flag, 1u);
                break;
            }
            if (STORE && (flags & SW_FLAG_DONE) && active) {
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
            sw_mbar_arrive_elect(bars, n_stages + (int)s); // every lane's reads of the stage have been consumed by now
            s = (s + 1u) & smask;
            parity ^= (s == 0u) ? 1u : 0u;
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
