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
    SW_TS = 7,      // cur = tip(A) * super(B)                  [optional push of cur first]
    SW_SI = 8,      // cur = super(A) * (P_B cur)
};
// dispatch code of an op: 2 * kind + 1 when the op works on / takes its pending operand from the second accumulator
// `u` -- dense, so that the op loop's switch becomes one indexed branch
constexpr int SW_CODE_U = 1;
__host__ __device__ constexpr int sw_code(int kind, bool u) { return 2 * kind + (u ? SW_CODE_U : 0); }
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
constexpr int SW_FLAG_FIRST = 128; // first op of a ring stage: wait for the stage's data before touching it
constexpr int SW_FLAG_LAST = 256;  // last op of a ring stage: hand the stage back to the producer afterwards

constexpr uint32_t SW_NO_ROW = 0xffffffffu; // header w2 of a completing op whose node the storing variant does not write

constexpr int SW_P_BYTES = 128;    // 4x4 f64 row-major
// Tables are PLANAR with a 16-byte row stride: states 0,1 of all rows, then states 2,3.  The rows that carry almost
// every site (A,C,G,T; the pairs a == b) then fall in different bank groups of a quarter-warp's LDS.128.
constexpr int SW_TIP_BYTES = 160;  // tip table: the 5 codes A,C,G,T,N/gap x 4 f64 (any other ambiguity code = sum of rows)
constexpr int SW_TAB2_BYTES = 800; // cherry / super-tip table: 5 x 5 code pairs x 4 f64, row = a + 5 b
constexpr int SW_BAR_BYTES = 512;  // mbarrier area at the start of dynamic shared memory
constexpr int SW_HDR_BYTES = 16;   // every record starts with the header of the NEXT op

// Per-op header.  The table lives in global memory (two padding entries); the transition kernel copies header
// k + 1 to the front of record k, so a consumer picks up the next op's header from the ring stage it is working
// on -- one warp-uniform LDS whose latency the current op's arithmetic hides -- and only the first header of a
// site block comes from global memory.
//   w0 = code | flags << 8 | level << 18     w1 = per-category record bytes | szA << 16 (operand B at +szA)
//        code = sw_code(kind, works on / takes its pending operand from the second accumulator `u`)
//   w3 = byte offset of the op's record inside its ring stage | first tip-code row slot of the op << 16
//   w2 = CLV slot of the node this op completes (valid with SW_FLAG_DONE; used by the storing variant)
struct SweepOp {
    uint32_t w0, w1, w2, w3;
};
// Producer-side descriptor of one ring stage (global memory): several consecutive ops share a stage -- their records
// are adjacent in the stream and arrive with ONE bulk copy, their tip-code rows with one copy each -- so the
// full / empty handshake, the stage arithmetic and the producer's issue work are paid once per stage, not per op.
constexpr int SW_STAGE_ROWS = 8;        // tip-code row slots per stage
constexpr int SW_STAGE_REC_BYTES = 8192; // record bytes per stage (at least the largest single record)
struct SweepStage {
    uint64_t rec_off;  // byte offset of the stage's first record in the stream
    uint32_t rec_bytes;
    int32_t n_rows;
    int32_t rows[SW_STAGE_ROWS]; // alignment rows whose tip codes the stage's ops read, in slot order
};
// One operand block the transition kernel has to fill.
struct SweepBlock {
    uint64_t dst_off; // byte offset of category 0's block; category c at dst_off + c * cat_stride
    uint32_t cat_stride;
    int32_t edge;     // dense node id whose branch length parametrises P
    int32_t is_table; // 0: P(t) 4x4; 1: tip table of `edge`; 2: cherry table tip(edge) (.) tip(edge2);
                      // 3: super-tip table P(edge) * (tip(edge2) (.) tip(edge3)); 4: P(t) transposed (down_sweep.cuh)
    int32_t edge2, edge3;
    int32_t pad;
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
                double acc = __dmul_rn(__dmul_rn(M.V[i + 0], e[0]), M.V[j + 0]);
#pragma unroll
                for (int k = 1; k < 4; ++k) acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(M.V[i + 4 * k], e[k]), M.V[j + 4 * k])); // no FMA: see transition_kernel
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
                                        const uint64_t *__restrict__ op_off, int n_ops)
{
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_blocks * M.n_clv) {
        const int k = idx - n_blocks * M.n_clv;
        if (k < n_ops) *reinterpret_cast<uint4 *>(stream + op_off[k]) = *reinterpret_cast<const uint4 *>(ops + k + 1);
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
    } else if (blk.is_table == 4) { // P^T, row-major (the pre-order sweep's message is P^T post_parent)
#pragma unroll
        for (int i = 0; i < 16; ++i) out[i] = p[(i & 3) * 4 + (i >> 2)];
    } else if (blk.is_table == 1) {
        for (uint32_t a = 0; a < 5; ++a) { // two planes of 5 x 16 bytes
            double *lo = out + 2 * a, *hi = lo + SW_TIP_BYTES / 16;
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                lo[i] = sw_masked(p, i, code_mask(a));
                hi[i] = sw_masked(p, i + 2, code_mask(a));
            }
        }
    } else {
        // tip(x) (.) tip(y) for the 25 code pairs; is_table == 3: carried through the cherry's own edge
        double x[16], y[16];
        const bool super = blk.is_table == 3;
        sw_transition(M, M.rates[cat] * blen[super ? blk.edge2 : blk.edge], x);
        sw_transition(M, M.rates[cat] * blen[super ? blk.edge3 : blk.edge2], y);
        for (uint32_t a = 0; a < 5; ++a)
            for (uint32_t c = 0; c < 5; ++c) { // two planes of 25 x 16 bytes: states 0,1 | states 2,3
                double ch[4], w[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) ch[j] = sw_masked(x, j, code_mask(a)) * sw_masked(y, j, code_mask(c));
                if (super) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) w[i] = p[4 * i] * ch[0] + p[4 * i + 1] * ch[1] + p[4 * i + 2] * ch[2] + p[4 * i + 3] * ch[3];
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) w[i] = ch[i];
                }
                double *lo = out + 2 * (a + 5 * c), *hi = lo + SW_TAB2_BYTES / 16;
                lo[0] = w[0]; lo[1] = w[1]; hi[0] = w[2]; hi[1] = w[3];
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
    const SweepStage *stages;
    const unsigned char *stream;
    int n_ops;
    int n_stage_descs; // ring stages one sweep of the tree consists of
    int Q;            // site quads per CTA; consumer threads = Q * n_clv
    int n_consumers;  // Q * n_clv
    int n_stages;     // ring depth (a power of two)
    int stage_bytes;  // bytes per ring stage (records + SW_STAGE_ROWS code rows)
    int rec_cap;      // record bytes per stage (offset of the code rows inside a stage)
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
    size_t slot_stride; // n_clv * 4 * tile_pad; blocked layout: n_blocks * n_clv * 4 * (4 Q)
    size_t tile_pad;
    int blocked;        // clv_out is [slot][site block][cat][state][4 Q]: one node's block is contiguous (down_sweep.cuh)
    // NNI cache pass (phylo_b200.cu score_nni_paths): exp_out != null -- a completed node with row w2 is first brought
    // to a maximum in [1,2) and the RUNNING exponent count of its (site, category) goes to exp_out[row][cat][site]; the
    // exponent of the node's own subtree is the difference to the count before the subtree's first op (a row of its
    // own: the ops of a subtree are consecutive).  Only rows < n_store get their CLV written.
    int *exp_out;
    uint32_t n_store;
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

// one row of a planar table: states 0,1 at tab + 16 row, states 2,3 one plane further
__device__ __forceinline__ void sw_load_prow(sw_sptr tab, uint32_t plane, uint32_t row, double (&t)[4])
{
    const sw_sptr a = tab + 16u * row;
    double2 x = sw_lds128(a), y = sw_lds128(a + plane);
    t[0] = x.x; t[1] = x.y; t[2] = y.x; t[3] = y.y;
}
// a byte of w is > 4 (a code other than A,C,G,T,N/gap)  <=>  bit 3 set, or bit 2 together with bit 0 or 1
__device__ __forceinline__ uint32_t sw_rare(uint32_t w) { return (w & 0x08080808u) | (w & 0x04040404u & ((w << 1) | (w << 2))); }
// ---- partial ambiguity codes (R,Y,S,W,K,M,B,D,H,V): rare, kept OUT OF LINE so that their loops cost the op bodies
// no registers.  Tip vector for such a code: the sum of the table rows of the code's states in ascending order -- the
// order in which the reference's P * profile accumulates them, so the sum is bit-identical to the masked column sum
// (likelihood.rs:121, alphabet.rs:44-64).
struct sw_vec4 {
    double t[4];
};
#ifdef PHB_HOST_EMU
#define SW_NOINLINE __attribute__((noinline))
#else
#define SW_NOINLINE __noinline__
#endif
__device__ SW_NOINLINE sw_vec4 sw_tip_slow(sw_sptr tab, uint32_t code)
{
    sw_vec4 o;
    const uint32_t m = code_mask(code);
    bool first = true;
    for (uint32_t s = 0; s < 4; ++s) {
        if (!((m >> s) & 1u)) continue;
        double r[4];
        sw_load_prow(tab, SW_TIP_BYTES / 2, s, r);
        for (int i = 0; i < 4; ++i) o.t[i] = first ? r[i] : o.t[i] + r[i];
        first = false;
    }
    if (first)
        for (int i = 0; i < 4; ++i) o.t[i] = 0.0;
    return o;
}
// Row of a 5x5 pair table for a pair with a partial code: the table is linear in each tip's profile, so it is the sum
// of the rows of the codes' states (differs from product-of-sums by round-off only; the storing variant never comes
// here for an explicit cherry, its records carry the two P(t)).
__device__ SW_NOINLINE sw_vec4 sw_pair_slow(sw_sptr tab, uint32_t x, uint32_t y)
{
    sw_vec4 o;
    const uint32_t mx = x <= 4u ? (x == 4u ? 15u : 1u << x) : code_mask(x), my = y <= 4u ? (y == 4u ? 15u : 1u << y) : code_mask(y);
    bool first = true;
    // a code among A,C,G,T,N/gap keeps its own row index (N/gap is one row, not four)
    for (uint32_t b = 0; b < 4; ++b) {
        if (!((my >> b) & 1u)) continue;
        for (uint32_t a = 0; a < 4; ++a) {
            if (!((mx >> a) & 1u)) continue;
            double r[4];
            sw_load_prow(tab, SW_TAB2_BYTES / 2, (x <= 4u ? x : a) + 5u * (y <= 4u ? y : b), r);
            for (int i = 0; i < 4; ++i) o.t[i] = first ? r[i] : o.t[i] + r[i];
            first = false;
            if (x <= 4u) break;
        }
        if (y <= 4u) break;
    }
    if (first)
        for (int i = 0; i < 4; ++i) o.t[i] = 0.0;
    return o;
}
__device__ __forceinline__ void sw_tip_any(sw_sptr tab, uint32_t code, double (&t)[4])
{
    if (code <= 4u) sw_load_prow(tab, SW_TIP_BYTES / 2, code, t);
    else {
        const sw_vec4 o = sw_tip_slow(tab, code);
#pragma unroll
        for (int i = 0; i < 4; ++i) t[i] = o.t[i];
    }
}
__device__ __forceinline__ void sw_pair_any(sw_sptr tab, uint32_t x, uint32_t y, double (&t)[4])
{
    if (x <= 4u && y <= 4u) sw_load_prow(tab, SW_TAB2_BYTES / 2, x + 5u * y, t);
    else {
        const sw_vec4 o = sw_pair_slow(tab, x, y);
#pragma unroll
        for (int i = 0; i < 4; ++i) t[i] = o.t[i];
    }
}

struct sw_false { static constexpr bool value = false; };
struct sw_true { static constexpr bool value = true; };

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
        // Lane 0 waits for the stage and posts the byte count; lanes 0..8 then issue one bulk copy each (the
        // records of the stage's ops and up to eight tip-code rows).
        const int pl = tid - prod_tid0;
        uint32_t s = 0, parity = 1; // a fresh barrier passes a parity-1 wait
        const uint32_t rec_cap = (uint32_t)p.rec_cap;
        const int n_sd = p.n_stage_descs;
        const int my_row = pl >= 1 && pl <= SW_STAGE_ROWS ? pl - 1 : 0; // lane i + 1 copies row slot i
        for (int64_t b = blockIdx.x; b < p.n_blocks; b += gridDim.x) {
            const uint8_t *tip_base = p.tips + b * sites_per_block;
            // this lane's view of the stage descriptor, one stage ahead: record (lane 0) or one row (lanes 1..8)
            uint64_t rec_off = p.stages[0].rec_off;
            uint32_t rec_bytes = p.stages[0].rec_bytes;
            int n_rows = p.stages[0].n_rows, row = p.stages[0].rows[my_row];
#pragma unroll 1
            for (int k = 0; k < n_sd; ++k) {
                const SweepStage *nx = p.stages + k + 1; // the table carries one padding entry
                const uint64_t n_rec_off = nx->rec_off;
                const uint32_t n_rec_bytes = nx->rec_bytes;
                const int n_n_rows = nx->n_rows, n_row = nx->rows[my_row];
                const sw_sptr st = ring + s * stage_bytes;
                if (pl == 0) {
                    sw_mbar_wait(bars, n_stages + (int)s, parity);
                    sw_mbar_expect_tx(bars, (int)s, rec_bytes + (uint32_t)n_rows * code_bytes);
                }
                __syncwarp();
                if (pl == 0) sw_bulk_g2s(st, p.stream + rec_off, rec_bytes, bars, (int)s);
                else if (pl <= n_rows) sw_bulk_g2s(st + rec_cap + (uint32_t)(pl - 1) * code_bytes, tip_base + (int64_t)row * p.P_pad, code_bytes, bars, (int)s);
                rec_off = n_rec_off;
                rec_bytes = n_rec_bytes;
                n_rows = n_n_rows;
                row = n_row;
                s = (s + 1u) & smask;
                parity ^= (s == 0u) ? 1u : 0u;
            }
        }
        return;
    }

    // ===================== consumer warps =====================
    const int t = tid;                // n_consumers is a multiple of 32 (SweepGeom::plan): every consumer lane works
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
        for (int k = 0; k < n_ops;) { // one ring stage per trip: wait, run its ops, hand it back
        const sw_sptr st = ring + s * stage_bytes;
        sw_mbar_wait(bars, (int)s, parity);
        uint32_t flags;
#pragma unroll 1
        do {
            ++k;
            const uint4 h = hn;
            const uint32_t code = h.x & 0xffu;
            flags = (h.x >> 8) & 0x3ffu;
            const sw_sptr rec = st + (h.w & 0xffffu);
            const sw_sptr blkA = rec + (uint32_t)SW_HDR_BYTES + (uint32_t)cc * (h.y & 0xffffu);
            const sw_sptr blkB = blkA + (h.y >> 16);
            const sw_sptr codes0 = st + code_off + (h.w >> 16) * code_bytes; // this thread's four codes in the op's first row slot
            hn = sw_lds128u(rec);
            // ---- the op bodies, written once over the accumulator they update ----
            const int level = (int)(h.x >> 18);
            auto push = [&]() { // park the running result before a TT / LOADTIP overwrites it
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
            // explicit cherry: one row of the 5x5 product table per site when both codes are among A,C,G,T,N/gap
            // (slots 0..4).  Any other code: the storing variant's records carry the two P(t) and the masked column
            // sums are formed in the reference's order; the plain records do not, and the table rows are summed.
            auto op_tt = [&](double (&acc)[SW_R][4]) {
                const uint32_t ca = sw_lds32(codes0) & 0x0f0f0f0fu, cb = sw_lds32(codes0 + code_bytes) & 0x0f0f0f0fu;
                if ((sw_rare(ca) | sw_rare(cb)) == 0u) {
                    const uint32_t idx = ca + 5u * cb; // per byte a + 5 b <= 24: no carries
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) sw_load_prow(blkA, SW_TAB2_BYTES / 2, (idx >> (8 * r)) & 0xffu, acc[r]);
                } else if ((h.y & 0xffffu) > (uint32_t)SW_TAB2_BYTES) { // the record carries the two P(t) behind the table
                    double PA[16], PB[16];
                    sw_load_P(blkA + SW_TAB2_BYTES, PA);
                    sw_load_P(blkA + SW_TAB2_BYTES + SW_P_BYTES, PB);
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) {
                        const uint32_t ma = code_mask((ca >> (8 * r)) & 15u), mb = code_mask((cb >> (8 * r)) & 15u);
#pragma unroll
                        for (int i = 0; i < 4; ++i) acc[r][i] = sw_masked(PA, i, ma) * sw_masked(PB, i, mb);
                    }
                } else {
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) sw_pair_any(blkA, (ca >> (8 * r)) & 15u, (cb >> (8 * r)) & 15u, acc[r]);
                }
            };
            // tip x super-tip: two lookups and a product, no matrix-vector product at all
            auto op_ts = [&](double (&acc)[SW_R][4]) {
                const uint32_t ca = sw_lds32(codes0) & 0x0f0f0f0fu, cx = sw_lds32(codes0 + code_bytes) & 0x0f0f0f0fu,
                               cy = sw_lds32(codes0 + 2u * code_bytes) & 0x0f0f0f0fu;
                if ((sw_rare(ca) | sw_rare(cx) | sw_rare(cy)) == 0u) {
                    const uint32_t idx = cx + 5u * cy;
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) {
                        double ta[4], w[4];
                        sw_load_prow(blkA, SW_TIP_BYTES / 2, (ca >> (8 * r)) & 0xffu, ta);
                        sw_load_prow(blkB, SW_TAB2_BYTES / 2, (idx >> (8 * r)) & 0xffu, w);
#pragma unroll
                        for (int i = 0; i < 4; ++i) acc[r][i] = ta[i] * w[i];
                    }
                } else {
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) {
                        double ta[4], w[4];
                        sw_tip_any(blkA, (ca >> (8 * r)) & 15u, ta);
                        sw_pair_any(blkB, (cx >> (8 * r)) & 15u, (cy >> (8 * r)) & 15u, w);
#pragma unroll
                        for (int i = 0; i < 4; ++i) acc[r][i] = ta[i] * w[i];
                    }
                }
            };
            // SUPER = false: tip(A) (.) (P_B acc);  SUPER = true: super-tip(A) (.) (P_B acc)
            auto op_ti = [&](double (&acc)[SW_R][4], auto super_tag) {
                constexpr bool SUPER = decltype(super_tag)::value;
                const uint32_t ca = sw_lds32(codes0) & 0x0f0f0f0fu;
                const uint32_t cb = SUPER ? (sw_lds32(codes0 + code_bytes) & 0x0f0f0f0fu) : 0u;
                double P[16];
                sw_load_P(blkB, P);
                if ((sw_rare(ca) | sw_rare(cb)) == 0u) {
                    const uint32_t idx = ca + 5u * cb;
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) {
                        double y[4], ta[4];
                        sw_load_prow(blkA, SUPER ? SW_TAB2_BYTES / 2 : SW_TIP_BYTES / 2, (idx >> (8 * r)) & 0xffu, ta);
                        sw_matvec(P, acc[r], y);
#pragma unroll
                        for (int i = 0; i < 4; ++i) acc[r][i] = ta[i] * y[i];
                    }
                } else {
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) {
                        double y[4], ta[4];
                        if (SUPER) sw_pair_any(blkA, (ca >> (8 * r)) & 15u, (cb >> (8 * r)) & 15u, ta);
                        else sw_tip_any(blkA, (ca >> (8 * r)) & 15u, ta);
                        sw_matvec(P, acc[r], y);
#pragma unroll
                        for (int i = 0; i < 4; ++i) acc[r][i] = ta[i] * y[i];
                    }
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
            switch (code) { // dense dispatch code, precomputed by the host scheduler
            case sw_code(SW_TI, false):
                op_ti(v, sw_false());
                break;
            case sw_code(SW_TI, true):
                op_ti(u, sw_false());
                break;
            case sw_code(SW_SI, false):
                op_ti(v, sw_true());
                break;
            case sw_code(SW_SI, true):
                op_ti(u, sw_true());
                break;
            case sw_code(SW_II, true):
                op_ii();
                break;
            case sw_code(SW_II, false):
                pop(u);
                op_ii();
                break;
            case sw_code(SW_TS, true):
                op_ts(u);
                break;
            case sw_code(SW_TS, false):
                if (flags & SW_FLAG_PUSH) push();
                op_ts(v);
                break;
            case sw_code(SW_TT, true):
                op_tt(u);
                break;
            case sw_code(SW_TT, false):
                if (flags & SW_FLAG_PUSH) push();
                op_tt(v);
                break;
            case sw_code(SW_APPLY, false): {
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
                break;
            }
            case sw_code(SW_MULTIP, false):
            case sw_code(SW_LOADTIP, false): {
                if (flags & SW_FLAG_PUSH) push();
                const uint32_t ca = sw_lds32(codes0) & 0x0f0f0f0fu;
#pragma unroll
                for (int r = 0; r < SW_R; ++r) {
                    double ta[4];
                    sw_tip_any(blkA, (ca >> (8 * r)) & 15u, ta);
#pragma unroll
                    for (int i = 0; i < 4; ++i) v[r][i] = (code == (uint32_t)sw_code(SW_LOADTIP, false)) ? ta[i] : v[r][i] * ta[i];
                }
                break;
            }
            case sw_code(SW_MULPOP, false):
                pop(u);
#pragma unroll
                for (int r = 0; r < SW_R; ++r)
#pragma unroll
                    for (int i = 0; i < 4; ++i) v[r][i] *= u[r][i];
                break;
            default:
                if (lane == 0) atomicOr(p.error_flag, 1u);
                break;
            }
            if (flags & (SW_FLAG_GUARD | SW_FLAG_NORMOUT)) { // the accumulator the op wrote: u for the U-flagged chain ops
                if (flags & SW_FLAG_U) guard(u);
                else guard(v);
            }
            if (STORE && (flags & SW_FLAG_DONE) && h.z != SW_NO_ROW) {
                const bool use_u = flags & SW_FLAG_U;
                if (p.exp_out) {
#pragma unroll
                    for (int r = 0; r < SW_R; ++r) {
                        if (use_u) sw_rescale<true>(u[r], sc[r]);
                        else sw_rescale<true>(v[r], sc[r]);
                    }
                    const int64_t rel = b * sites_per_block + 4 * q;
                    int *e = p.exp_out + ((size_t)h.z * M.n_clv + cc) * p.tile_pad + (size_t)rel;
#pragma unroll
                    for (int r = 0; r < SW_R; ++r)
                        if (rel + r < p.P) e[r] = sc[r];
                }
                if (h.z >= p.n_store) {
                    // a counter-only row
                } else if (p.blocked) { // whole blocks: the buffer covers n_blocks * 4 Q patterns
                    double *dst = p.clv_out + (size_t)h.z * p.slot_stride + ((size_t)b * M.n_clv + cc) * (size_t)(4 * sites_per_block) + 4 * q;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        double2 *d2 = reinterpret_cast<double2 *>(dst + (size_t)i * sites_per_block);
                        d2[0] = use_u ? make_double2(u[0][i], u[1][i]) : make_double2(v[0][i], v[1][i]);
                        d2[1] = use_u ? make_double2(u[2][i], u[3][i]) : make_double2(v[2][i], v[3][i]);
                    }
                } else {
                    const int64_t rel = b * sites_per_block + 4 * q; // first of this thread's 4 patterns, launch-relative
                    double *dst = p.clv_out + (size_t)h.z * p.slot_stride + (size_t)cc * 4 * p.tile_pad + (size_t)rel;
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
            }
        } while (!(flags & SW_FLAG_LAST));
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
        {
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
        if (cc == 0) {
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
