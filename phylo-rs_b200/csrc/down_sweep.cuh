// down_sweep.cuh — the pre-order pass of compute_marginal_asr (src/tree/likelihood.rs:303-390) as an on-chip sweep.
//
// The level-batched pre-order kernels (kernels.cuh, asr_down_level_kernel) move, per internal node and pattern, the
// parent's posterior (read), the node's up-pass CLV (read) and the node's posterior (write) through HBM: 96*K bytes.
// A node's posterior is needed by its children only, so a depth-first pre-order walk can keep it in registers: the
// CTA walks the tree from the root, the current node's per-category posterior lives in registers, a child's up-pass
// CLV arrives ONCE through a TMA ring (it is stored by the storing sweep, sweep.cuh, in site-block-major layout so
// that one node's tile is one contiguous bulk copy), and the only things written are the state byte per (node,
// pattern) and, if asked for, the mixed posterior.  Where a node has two internal children, one child's posterior
// waits in a shared-memory stack (deep levels: an L2-resident global stack) while the other subtree is walked; the
// child with the smaller stack need goes first, so the depth is the tree's Strahler number, as in the up sweep.
//
// Work split: one thread = ONE site pattern, ALL K rate categories (the up sweep has one category per thread).  The
// category mixture  mixed_v = normalise(sum_c w_c post_v,c)  (likelihood.rs:356-382) is then thread-local, and every
// transition matrix is still warp-uniform (the categories are looped over, not spread over lanes).
//
// Per child v of the current node p, per category c (likelihood.rs:321-342):
//     msg    = P_v^T post_p            (the record carries P^T row-major, so this is the up sweep's mat-vec)
//     post_v = up_v (.) msg, divided by its sum when the sum is > 0
// then  mixed = sum_c w_c post_v,c,  divided by its sum when > 0, and the last maximal index is the state (:385-390).
#pragma once
#include "sweep.cuh"

#include <algorithm>
#include <vector>

namespace phb {

enum DownKind : int32_t {
    DN_INIT = 0,    // cur = normalise(pi (.) up_root); emits the root's state            (likelihood.rs:304-316)
    DN_LEAF = 1,    // leaf child: state from tip profile (.) P^T cur; cur unchanged
    DN_DESCEND = 2, // internal child: cur = its posterior (the walk continues below it)
    DN_DEFER = 3,   // internal child: its posterior goes to stack `level`; cur unchanged
    DN_RESUME = 4,  // cur = pop(level): the walk continues below the deferred child
};

// Per-op header (global table, + padding entry; the next op's header rides at the front of each record).
//   w0 = kind | flags << 8 (SW_FLAG_FIRST / SW_FLAG_LAST of its ring stage) | level << 20
//   w1 = dense node id (row of the states output)
//   w2 = byte offset of the op's record inside its ring stage | tip-code row slot of a LEAF op << 16
// Producer descriptor of one ring stage: a stage opens with at most ONE op that reads a CLV tile (INIT, DESCEND,
// DEFER) and takes the tile-less ops that follow it (LEAF: a tip-code row each; RESUME), so that the handshake and the
// producer's issue work are paid per tile, not per op.  The records of a stage's ops are adjacent in the stream.
constexpr int DN_STAGE_ROWS = 8; // tip-code row slots per stage
constexpr int DN_STAGE_OPS = 10; // ops per stage at most
struct DownDesc {
    uint64_t rec_off;
    uint32_t rec_bytes;
    int32_t slot;   // CLV slot of the node whose up-pass tile the stage's first op reads, -1 = none
    int32_t n_rows;
    int32_t rows[DN_STAGE_ROWS]; // alignment rows of the stage's leaf ops, in slot order
    int32_t pad[3];
};

struct DownSched {
    bool valid = false;
    int n_clv = 0, n_ops = 0, n_levels = 0;
    int rec_bytes = 0; // bytes of a full record
    int n_stage_descs = 0;
    size_t stream_bytes = 0;
    std::vector<SweepOp> ops;
    std::vector<DownDesc> descs;
    std::vector<SweepBlock> blocks; // transposed transition matrices to write into the stream (is_table = 4)
    std::vector<uint64_t> op_off;   // stream offset of every op's record (where the transition kernel drops the headers)
    std::vector<int> built_kid_off, built_kids, built_leaf_row, built_slot;
    int built_root = -1;
    int count_kind[8] = {0};

    bool same_topology(int root, const std::vector<int> &child_off, const std::vector<int> &children,
                       const std::vector<int> &leaf_row, const std::vector<int> &slot, int n_clv_) const
    {
        return valid && n_clv == n_clv_ && root == built_root && child_off == built_kid_off && children == built_kids &&
               leaf_row == built_leaf_row && slot == built_slot;
    }

    void build(int n, int root, const std::vector<int> &child_off, const std::vector<int> &children,
               const std::vector<int> &leaf_row, const std::vector<int> &slot, int n_clv_)
    {
        valid = false;
        n_clv = n_clv_;
        built_kid_off = child_off;
        built_kids = children;
        built_leaf_row = leaf_row;
        built_slot = slot;
        built_root = root;
        for (int &c : count_kind) c = 0;
        auto is_leaf = [&](int v) { return child_off[v + 1] == child_off[v]; };
        // post-order for the stack needs D(v): what walking v's subtree needs while v's posterior is in registers
        std::vector<int> order;
        {
            std::vector<int> st(1, root);
            while (!st.empty()) {
                int v = st.back();
                st.pop_back();
                order.push_back(v);
                for (int k = child_off[v]; k < child_off[v + 1]; ++k) st.push_back(children[k]);
            }
            std::reverse(order.begin(), order.end());
        }
        std::vector<int> D(n, 0);
        std::vector<std::vector<int>> ints(n);      // internal children, ascending need
        std::vector<std::vector<int>> defer_lvl(n); // stack level of ints[v][j], j >= 1
        for (int v : order) {
            if (is_leaf(v)) continue;
            for (int k = child_off[v]; k < child_off[v + 1]; ++k)
                if (!is_leaf(children[k])) ints[v].push_back(children[k]);
            std::stable_sort(ints[v].begin(), ints[v].end(), [&](int a, int b) { return D[a] < D[b]; });
            int need = ints[v].empty() ? 0 : D[ints[v][0]];
            defer_lvl[v].assign(ints[v].size(), -1);
            int prev = -1;
            for (size_t j = 1; j < ints[v].size(); ++j) {
                // pending while the subtrees of ints[v][0..j-1] are walked: above their needs, distinct from earlier ones
                const int lvl = std::max(D[ints[v][j - 1]], prev + 1);
                defer_lvl[v][j] = lvl;
                prev = lvl;
                need = std::max(need, std::max(lvl + 1, D[ints[v][j]]));
            }
            D[v] = need;
        }
        n_levels = D[root];
        struct Rec { int kind, level, node, slot, row, edge; };
        std::vector<Rec> recs;
        auto emit = [&](int kind, int level, int node, int slot_, int row, int edge) {
            recs.push_back({kind, level, node, slot_, row, edge});
            count_kind[kind]++;
        };
        emit(DN_INIT, 0, root, slot[root], -1, -1);
        struct Frame { int v; size_t j; };
        std::vector<Frame> st;
        auto enter = [&](int v) { // cur == post_v: leaves, then deferred children, then the first internal child
            for (int k = child_off[v]; k < child_off[v + 1]; ++k) {
                const int c = children[k];
                if (is_leaf(c)) emit(DN_LEAF, 0, c, -1, leaf_row[c], c);
            }
            for (size_t j = ints[v].size(); j-- > 1;) emit(DN_DEFER, defer_lvl[v][j], ints[v][j], slot[ints[v][j]], -1, ints[v][j]);
            st.push_back({v, 0});
        };
        enter(root);
        while (!st.empty()) {
            Frame &f = st.back();
            const int v = f.v;
            if (f.j >= ints[v].size()) {
                st.pop_back();
                continue;
            }
            const size_t j = f.j++;
            const int c = ints[v][j];
            if (j == 0) emit(DN_DESCEND, 0, c, slot[c], -1, c);
            else emit(DN_RESUME, defer_lvl[v][j], c, -1, -1, -1);
            enter(c);
        }
        // records: [next header][K transposed matrices]; INIT and RESUME carry the header only.  Stages: greedy.
        n_ops = (int)recs.size();
        rec_bytes = SW_HDR_BYTES + n_clv * SW_P_BYTES;
        descs.clear();
        ops.assign(recs.size() + 2, SweepOp{0, 0, 0, 0});
        op_off.assign(recs.size(), 0);
        blocks.clear();
        size_t off = 0;
        DownDesc cur;
        int cur_ops = 0;
        auto reset = [&]() {
            memset(&cur, 0, sizeof cur);
            cur.slot = -1;
            for (int &x : cur.rows) x = -1;
            cur_ops = 0;
        };
        reset();
        for (size_t k = 0; k < recs.size(); ++k) {
            const Rec &r = recs[k];
            const uint32_t bytes = r.edge >= 0 ? (uint32_t)rec_bytes : (uint32_t)SW_HDR_BYTES;
            const bool tile = r.slot >= 0;
            if (cur_ops > 0 && (tile || cur_ops >= DN_STAGE_OPS || (r.row >= 0 && cur.n_rows >= DN_STAGE_ROWS))) {
                ops[k - 1].w0 |= (uint32_t)SW_FLAG_LAST << 8;
                descs.push_back(cur);
                reset();
            }
            uint32_t fl = 0;
            if (cur_ops == 0) {
                cur.rec_off = off;
                fl |= SW_FLAG_FIRST;
            }
            op_off[k] = off;
            ops[k].w2 = cur.rec_bytes | ((uint32_t)(r.row >= 0 ? cur.n_rows : 0) << 16);
            if (tile) cur.slot = r.slot;
            if (r.row >= 0) cur.rows[cur.n_rows++] = r.row;
            cur.rec_bytes += bytes;
            ++cur_ops;
            if (r.edge >= 0) blocks.push_back({off + SW_HDR_BYTES, (uint32_t)SW_P_BYTES, r.edge, 4, -1, -1, 0});
            ops[k].w0 = (uint32_t)r.kind | (fl << 8) | ((uint32_t)r.level << 20);
            ops[k].w1 = (uint32_t)r.node;
            off += bytes;
        }
        if (cur_ops > 0) {
            ops[recs.size() - 1].w0 |= (uint32_t)SW_FLAG_LAST << 8;
            descs.push_back(cur);
        }
        n_stage_descs = (int)descs.size();
        reset();
        descs.push_back(cur); // padding entry: the producer reads one stage ahead
        stream_bytes = off;
        valid = true;
    }
};

struct DownParams {
    const SweepOp *ops;
    const DownDesc *descs;
    const unsigned char *stream;
    int n_ops;
    int n_stage_descs;
    int S;           // patterns per CTA pass == consumer threads == the storing sweep's block width
    int n_stages;
    int stage_bytes; // records + CLV tile + tip-code rows
    int data_off;    // offset of the CLV tile inside a stage (>= DN_STAGE_OPS records, 128-byte aligned)
    int rows_off;    // offset of the tip-code rows inside a stage; row slot j at rows_off + j * row_pitch
    int row_pitch;
    int smem_levels, g_levels;
    double *gstack;  // [cta][g_level][K*4][S]
    const double *clv; // blocked: [slot][block][cat][state][S]
    size_t slot_stride; // doubles between slots: n_blocks * K * 4 * S
    const uint8_t *tips; // + first pattern of the launch
    int64_t P_pad;
    int64_t P;        // valid patterns of the launch
    int64_t n_blocks;
    const double *cat_ll; // [model cat][tile_pad] per-category log-likelihoods written by the storing sweep
    size_t tile_pad;
    const double *site_ll; // + first pattern
    uint8_t *states;       // [node][P_pad] + first pattern
    double *posteriors;    // nullable [node][P_pad][4] + first pattern * 4
};

__device__ __forceinline__ int dn_argmax_last(const double (&m)[4])
{
    int best = 0;
#pragma unroll
    for (int i = 1; i < 4; ++i)
        if (!(m[best] > m[i])) best = i;
    return best;
}

// grid = min(n_blocks, #SMs) persistent CTAs; block = S consumer threads + one producer warp.
template <int K>
__global__ void __launch_bounds__(384, 1)
asr_down_sweep_kernel(DownParams p, ModelDev M)
{
    PHB_DYN_SMEM(smem);
    const int tid = threadIdx.x;
    const int S = p.S;
    const int n_cons_warps = S >> 5;
    const int n_stages = p.n_stages, n_ops = p.n_ops;
    const uint32_t stage_bytes = (uint32_t)p.stage_bytes, data_off = (uint32_t)p.data_off;
    const sw_sptr bars = sw_sbase(smem);
    const sw_sptr ring = bars + SW_BAR_BYTES;
    const sw_sptr stack_s = ring + (uint32_t)n_stages * stage_bytes;
    if (tid == 0) {
        for (int s = 0; s < n_stages; ++s) {
            sw_mbar_init(bars, s, 1);
            sw_mbar_init(bars, n_stages + s, n_cons_warps);
        }
        sw_fence_init();
    }
    __syncthreads();
    const uint32_t tile_bytes = (uint32_t)(K * 4 * S * 8);

    if (tid >= S) {
        // ===================== producer warp: lane 0 posts the byte count and copies the records, lane 1 the CLV
        // tile, lanes 2.. one tip-code row each
        const int pl = tid - S;
        const int my_row = pl >= 2 && pl < 2 + DN_STAGE_ROWS ? pl - 2 : 0;
        uint32_t s = 0, parity = 1;
        const int n_sd = p.n_stage_descs;
        for (int64_t b = blockIdx.x; b < p.n_blocks; b += gridDim.x) {
            const double *clv_blk = p.clv + (size_t)b * (size_t)(K * 4 * S);
            const uint8_t *tip_base = p.tips + b * S;
            uint64_t rec_off = p.descs[0].rec_off;
            uint32_t rec_bytes = p.descs[0].rec_bytes;
            int slot = p.descs[0].slot, n_rows = p.descs[0].n_rows, row = p.descs[0].rows[my_row];
#pragma unroll 1
            for (int k = 0; k < n_sd; ++k) {
                const DownDesc *nx = p.descs + k + 1;
                const uint64_t n_rec_off = nx->rec_off;
                const uint32_t n_rec_bytes = nx->rec_bytes;
                const int n_slot = nx->slot, n_n_rows = nx->n_rows, n_row = nx->rows[my_row];
                const sw_sptr st = ring + s * stage_bytes;
                if (pl == 0) {
                    sw_mbar_wait(bars, n_stages + (int)s, parity);
                    sw_mbar_expect_tx(bars, (int)s, rec_bytes + (slot >= 0 ? tile_bytes : 0u) + (uint32_t)n_rows * (uint32_t)S);
                }
                __syncwarp();
                if (pl == 0) sw_bulk_g2s(st, p.stream + rec_off, rec_bytes, bars, (int)s);
                else if (pl == 1) {
                    if (slot >= 0) sw_bulk_g2s(st + data_off, clv_blk + (size_t)slot * p.slot_stride, tile_bytes, bars, (int)s);
                } else if (pl - 2 < n_rows)
                    sw_bulk_g2s(st + (uint32_t)p.rows_off + (uint32_t)(pl - 2) * (uint32_t)p.row_pitch, tip_base + (int64_t)row * p.P_pad,
                                (uint32_t)S, bars, (int)s);
                rec_off = n_rec_off;
                rec_bytes = n_rec_bytes;
                slot = n_slot;
                n_rows = n_n_rows;
                row = n_row;
                if (++s == (uint32_t)n_stages) {
                    s = 0;
                    parity ^= 1u;
                }
            }
        }
        return;
    }

    // ===================== consumers: thread = one pattern, all K categories =====================
    const int smem_levels = p.smem_levels;
    const uint32_t lvl_bytes = tile_bytes;
    const sw_sptr my_stack = stack_s + 8u * (uint32_t)tid;
    double *const my_gstack = p.gstack + (size_t)blockIdx.x * p.g_levels * (size_t)(K * 4 * S) + tid;
    const bool fast_leaves = p.posteriors == nullptr;
    uint32_t s = 0, parity = 0;
    const uint4 *const optab = reinterpret_cast<const uint4 *>(p.ops);
    for (int64_t b = blockIdx.x; b < p.n_blocks; b += gridDim.x) {
        const int64_t site = b * S + tid; // launch-relative
        const bool live = site < p.P;
        double cur[K][4], w[K];
        {
            // mixture weights of the pattern: exp(cat_ll_c - site_ll), or 1/K when site_ll is not finite (:356-363)
            const double sl = live ? p.site_ll[site] : 0.0;
#pragma unroll
            for (int c = 0; c < K; ++c) {
                w[c] = 1.0 / (double)K;
                if (live && isfinite(sl)) w[c] = exp(p.cat_ll[(size_t)c * p.tile_pad + (size_t)site] - sl);
#pragma unroll
                for (int i = 0; i < 4; ++i) cur[c][i] = 0.0;
            }
        }
        uint4 hn = __ldg(optab);
#pragma unroll 1
        for (int k = 0; k < n_ops;) { // one ring stage per trip
        const sw_sptr st = ring + s * stage_bytes;
        sw_mbar_wait(bars, (int)s, parity);
        uint32_t flags;
#pragma unroll 1
        do {
            ++k;
            const uint4 h = hn;
            const uint32_t kind = h.x & 0xffu;
            flags = (h.x >> 8) & 0xfffu;
            const int level = (int)(h.x >> 20);
            const sw_sptr rec = st + (h.z & 0xffffu);
            hn = sw_lds128u(rec);
            const sw_sptr data = st + data_off + 8u * (uint32_t)tid;
            if (kind == DN_RESUME) {
                if (level < smem_levels) {
#pragma unroll
                    for (int c = 0; c < K; ++c)
#pragma unroll
                        for (int i = 0; i < 4; ++i) cur[c][i] = sw_lds64(my_stack + (uint32_t)level * lvl_bytes + (uint32_t)((c * 4 + i) * S) * 8u);
                } else {
                    const double *g = my_gstack + (size_t)(level - smem_levels) * (size_t)(K * 4 * S);
#pragma unroll
                    for (int c = 0; c < K; ++c)
#pragma unroll
                        for (int i = 0; i < 4; ++i) cur[c][i] = g[(size_t)(c * 4 + i) * S];
                }
                continue;
            }
            uint32_t tipcode = 0;
            if (kind == DN_LEAF) {
                const sw_sptr rowp = st + (uint32_t)p.rows_off + (h.z >> 16) * (uint32_t)p.row_pitch;
                tipcode = (sw_lds32(rowp + ((uint32_t)tid & ~3u)) >> (8u * ((uint32_t)tid & 3u))) & 15u;
                if (fast_leaves && tipcode <= 3u) {
                    // an unambiguous tip: its posterior is one-hot wherever the message reaches its state in every
                    // category (msg_c[a] > 0: post_c = e_a up to rounding, mixed[a] > 0 = the only non-zero entry), so
                    // the state is the tip's own -- four FMAs per category instead of the whole update.  Anything
                    // else (a message that is not positive) takes the general path below.
                    bool all_pos = true;
#pragma unroll
                    for (int c = 0; c < K; ++c) {
                        const sw_sptr rowa = rec + (uint32_t)SW_HDR_BYTES + (uint32_t)(c * SW_P_BYTES) + 32u * tipcode;
                        const double2 a = sw_lds128(rowa), b2 = sw_lds128(rowa + 16u);
                        const double m = a.x * cur[c][0] + a.y * cur[c][1] + b2.x * cur[c][2] + b2.y * cur[c][3];
                        all_pos = all_pos && (m > 0.0);
                    }
                    if (all_pos) {
                        if (live) p.states[(int64_t)h.y * p.P_pad + site] = (uint8_t)tipcode;
                        continue;
                    }
                }
            }
            {
                double mixed[4] = {0.0, 0.0, 0.0, 0.0};
                const uint32_t tipmask = code_mask(tipcode);
#pragma unroll
                for (int c = 0; c < K; ++c) {
                    double post[4];
                    if (kind == DN_INIT) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) post[i] = M.pi[i] * sw_lds64(data + (uint32_t)((c * 4 + i) * S) * 8u);
                    } else {
                        double PT[16], msg[4];
                        sw_load_P(rec + (uint32_t)SW_HDR_BYTES + (uint32_t)(c * SW_P_BYTES), PT);
                        sw_matvec(PT, cur[c], msg);
                        if (kind == DN_LEAF) {
#pragma unroll
                            for (int i = 0; i < 4; ++i) post[i] = (((tipmask >> i) & 1u) ? 1.0 : 0.0) * msg[i];
                        } else {
#pragma unroll
                            for (int i = 0; i < 4; ++i) post[i] = sw_lds64(data + (uint32_t)((c * 4 + i) * S) * 8u) * msg[i];
                        }
                    }
                    const double sum = ((post[0] + post[1]) + post[2]) + post[3];
                    // left as it is when the sum is not > 0 (:311-315, :337-341); one reciprocal per child (<= 1 ulp from the
                    // reference's four divisions, as in asr_down_level_kernel), true divisions at the root (asr_root_kernel)
                    if (kind == DN_INIT) {
                        if (sum > 0.0) {
#pragma unroll
                            for (int i = 0; i < 4; ++i) post[i] /= sum;
                        }
                    } else {
                        const double inv = sum > 0.0 ? 1.0 / sum : 1.0;
#pragma unroll
                        for (int i = 0; i < 4; ++i) post[i] *= inv;
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) mixed[i] += w[c] * post[i];
                    if (kind == DN_DEFER) {
                        if (level < smem_levels) {
#pragma unroll
                            for (int i = 0; i < 4; ++i) sw_sts64(my_stack + (uint32_t)level * lvl_bytes + (uint32_t)((c * 4 + i) * S) * 8u, post[i]);
                        } else {
                            double *g = my_gstack + (size_t)(level - smem_levels) * (size_t)(K * 4 * S);
#pragma unroll
                            for (int i = 0; i < 4; ++i) g[(size_t)(c * 4 + i) * S] = post[i];
                        }
                    } else if (kind != DN_LEAF) { // INIT, DESCEND: the walk continues with this node's posterior
#pragma unroll
                        for (int i = 0; i < 4; ++i) cur[c][i] = post[i];
                    }
                }
                // The reference renormalises the mixture before the argmax (:376-390).  Division by a positive number is
                // monotone, so it can only change the winner by rounding two entries together: when the largest entry
                // beats every other by more than 2^-40 relative, the argmax is that of the unnormalised mixture and the
                // four divisions are skipped (they are the most expensive thing in the op); posteriors on request.
                const double msum = ((mixed[0] + mixed[1]) + mixed[2]) + mixed[3];
                int best = dn_argmax_last(mixed);
                bool clear = msum > 0.0 && !p.posteriors;
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if (i != best && !(mixed[best] * (1.0 - 0x1p-40) > mixed[i])) clear = false;
                if (!clear) {
                    if (msum > 0.0) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) mixed[i] /= msum;
                    }
                    best = dn_argmax_last(mixed);
                }
                if (live) {
                    p.states[(int64_t)h.y * p.P_pad + site] = (uint8_t)best;
                    if (p.posteriors) {
                        double *o = p.posteriors + ((int64_t)h.y * p.P_pad + site) * 4;
                        *reinterpret_cast<double2 *>(o) = make_double2(mixed[0], mixed[1]);
                        *reinterpret_cast<double2 *>(o + 2) = make_double2(mixed[2], mixed[3]);
                    }
                }
            }
        } while (!(flags & SW_FLAG_LAST));
        sw_mbar_arrive_elect(bars, n_stages + (int)s);
        if (++s == (uint32_t)n_stages) {
            s = 0;
            parity ^= 1u;
        }
        }
        // the stack is reused by the next block: nobody may still be reading it (deferred entries are always
        // resumed within the block, so no cross-block hazard exists on the data; the ring needs no barrier)
    }
}

} // namespace phb
