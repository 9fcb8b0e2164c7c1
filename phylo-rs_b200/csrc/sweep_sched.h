// sweep_sched.h — host scheduler of the fused sweep (sweep.cuh): turns the child lists of the
// flattened tree (phb_set_tree) into the op stream a CTA executes for every block of site patterns.
//
// The reference evaluates nodes in DFSPostOrderIterator order (src/iter.rs:112-138) and keeps every
// node's profile in a HashMap (src/tree/likelihood.rs:95).  Any order that visits children before
// parents gives the same per-node arithmetic; this scheduler picks the depth-first order that
// minimises the number of partial results alive at once (Sethi–Ullman: among a node's internal
// children the one whose subtree needs more stack levels goes first), and assigns every pending
// sibling a STATIC stack level = need(subtree evaluated while it waits).  Nested waits have strictly
// smaller needs, so levels never collide and the low (hot) levels can live in shared memory.
#pragma once
#include "sweep.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace phb {

struct SweepSched {
    bool valid = false;
    int n_clv = 0;
    int n_ops = 0;
    int n_levels = 0;   // stack levels used (need(root))
    int rec_cap = 0;    // largest record
    size_t stream_bytes = 0;
    std::vector<SweepDesc> descs;
    std::vector<SweepBlock> blocks;
    std::vector<SweepOp> ops;           // per-op header table (global memory, prefetched by the consumers)
    int count_kind[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int n_push = 0, n_reg_sibling = 0;  // pending siblings parked in the stack / kept in the second accumulator
    // guard planning (plan_guards): the op that completes each node, the flags the schedule itself needs, and the
    // tree in post-order
    std::vector<int> done_op;           // node -> index of the op that completes it (-1: leaf)
    std::vector<uint32_t> base_w0;      // per op: w0 without SW_FLAG_GUARD / SW_FLAG_NORMOUT
    std::vector<int> post_order;        // children before parents
    std::vector<int> kid_off, kids;     // copy of the child lists
    int n_guards = 0, n_normout = 0;

    // child_off/children: CSR child lists; leaf_row: alignment row per leaf; root must be internal.
    // slot: CLV slot per node (-1 for leaves), recorded in the header of the op that completes the node.
    void build(int n, int root, const std::vector<int> &child_off, const std::vector<int> &children,
               const std::vector<int> &leaf_row, int n_clv_, const std::vector<int> &slot = std::vector<int>())
    {
        valid = false;
        n_clv = n_clv_;
        descs.clear();
        blocks.clear();
        for (int &c : count_kind) c = 0;
        auto is_leaf = [&](int v) { return child_off[v + 1] == child_off[v]; };
        // post-order (children before parents) for the need computation
        std::vector<int> order;
        order.reserve(n);
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
        post_order = order;
        kid_off = child_off;
        kids = children;
        std::vector<int> need(n, 0);
        std::vector<std::vector<int>> ints(n), tips(n);
        for (int v : order) {
            if (is_leaf(v)) continue;
            for (int k = child_off[v]; k < child_off[v + 1]; ++k) {
                int c = children[k];
                (is_leaf(c) ? tips[v] : ints[v]).push_back(c);
            }
            std::stable_sort(ints[v].begin(), ints[v].end(), [&](int a, int b) { return need[a] > need[b]; });
            int nd = 0;
            if (!ints[v].empty()) nd = need[ints[v][0]];
            if (ints[v].size() >= 2) nd = std::max(nd, need[ints[v][1]] + 1);
            need[v] = nd;
        }
        n_levels = need[root];
        // chain[v]: the subtree of v is a chain of binary TT / TI updates (no internal x internal node, no
        // polytomy): it runs entirely in ONE accumulator, so as the pending-side sibling of a binary node it
        // can be evaluated in the second register accumulator `u` with no stack traffic at all.
        std::vector<char> chain(n, 0);
        for (int v : order) {
            if (is_leaf(v)) continue;
            const size_t ni = ints[v].size(), nt = tips[v].size();
            chain[v] = ni + nt == 2 && (ni == 0 || (ni == 1 && chain[ints[v][0]]));
        }

        struct Hdr { int kind, flags, level, szA, szB; };
        struct Rec { Hdr h; int edgeA, edgeB, rowA, rowB; bool tabA, tabB; int node; };
        std::vector<Rec> recs;
        int pending_push = -1;
        bool u_mode = false;
        n_push = n_reg_sibling = 0;
        auto emit = [&](int kind, int level, int edgeA, bool tabA, int edgeB, bool tabB) {
            Rec r;
            memset(&r, 0, sizeof r);
            r.h.kind = kind;
            r.h.level = level;
            if (u_mode) r.h.flags |= SW_FLAG_U;
            if (kind == SW_TT || kind == SW_LOADTIP) {
                if (pending_push >= 0) {
                    r.h.flags |= SW_FLAG_PUSH;
                    r.h.level = pending_push;
                    pending_push = -1;
                    ++n_push;
                }
            }
            r.edgeA = edgeA; r.edgeB = edgeB; r.tabA = tabA; r.tabB = tabB;
            r.h.szA = edgeA < 0 ? 0 : (tabA ? SW_TAB_BYTES : SW_P_BYTES);
            r.h.szB = edgeB < 0 ? 0 : (tabB ? SW_TAB_BYTES : SW_P_BYTES);
            if (kind == SW_TT) { // one 5x5 table of products for the cherry instead of two tip tables, followed by
                r.h.szA = SW_TAB2_BYTES; // the two P(t) for the sites that carry a partial ambiguity code
                r.h.szB = 2 * SW_P_BYTES;
            }
            r.rowA = (edgeA >= 0 && tabA) ? leaf_row[edgeA] : -1;
            r.rowB = (edgeB >= 0 && tabB) ? leaf_row[edgeB] : -1;
            recs.push_back(r);
            count_kind[kind]++;
        };
        struct Frame { int v, j; bool ret; };
        std::vector<Frame> st;
        st.push_back({root, 0, false});
        while (!st.empty()) {
            Frame &f = st.back();
            const int v = f.v;
            const std::vector<int> &iv = ints[v], &tv = tips[v];
            const bool binary = iv.size() + tv.size() == 2;
            if (f.ret) { // child iv[j-1] just finished; its CLV is the running value
                f.ret = false;
                if (!binary) {
                    emit(SW_APPLY, 0, iv[f.j - 1], false, -1, false);
                    // a product of >= 3 child contributions: every child enters at max in [1,2), as in the
                    // reference (children are max-normalised, profile.rs:20-39), so the product leaves the
                    // normal range exactly where the reference's own product does
                    if (iv.size() + tv.size() >= 3) recs.back().h.flags |= SW_FLAG_NORM;
                    if (f.j - 1 >= 1) emit(SW_MULPOP, need[iv[f.j - 1]], -1, false, -1, false);
                }
            }
            if (f.j < (int)iv.size()) {
                if (f.j >= 1) {
                    if (binary && chain[iv[f.j]] && !u_mode) {
                        u_mode = true; // the waiting sibling stays in v; this subtree runs in u
                        ++n_reg_sibling;
                    } else
                        pending_push = need[iv[f.j]];
                }
                f.ret = true;
                int c = iv[f.j++];
                st.push_back({c, 0, false});
                continue;
            }
            if (binary) {
                if (iv.size() == 2) {
                    const bool from_u = u_mode; // set when iv[1] was evaluated in u (chains hold no II, so no nesting)
                    u_mode = false;
                    // operand A is the waiting side: the popped CLV of iv[0], or u holding iv[1]
                    if (from_u) emit(SW_II, need[iv[1]], iv[1], false, iv[0], false);
                    else emit(SW_II, need[iv[1]], iv[0], false, iv[1], false);
                    if (from_u) recs.back().h.flags |= SW_FLAG_XU;
                }
                else if (iv.size() == 1) emit(SW_TI, 0, tv[0], true, iv[0], false);
                else emit(SW_TT, 0, tv[0], true, tv[1], true);
            } else if (iv.empty()) {
                size_t k = 0;
                if (tv.size() >= 2) {
                    emit(SW_TT, 0, tv[0], true, tv[1], true);
                    k = 2;
                } else {
                    emit(SW_LOADTIP, 0, tv[0], true, -1, false);
                    k = 1;
                }
                for (; k < tv.size(); ++k) emit(SW_MULTIP, 0, tv[k], true, -1, false);
            } else {
                for (int tnode : tv) emit(SW_MULTIP, 0, tnode, true, -1, false);
            }
            recs.back().h.flags |= SW_FLAG_DONE;
            recs.back().node = v;
            st.pop_back();
        }
        // dense renumbering of the stack levels actually used (the register accumulator usually frees level 0)
        {
            std::vector<int> used(n_levels + 2, 0);
            for (const Rec &r : recs)
                if ((r.h.flags & SW_FLAG_PUSH) || r.h.kind == SW_MULPOP || (r.h.kind == SW_II && !(r.h.flags & SW_FLAG_XU)))
                    used[r.h.level] = 1;
            int next = 0;
            for (int &u : used) u = u ? next++ : -1;
            for (Rec &r : recs)
                if (used[r.h.level] >= 0) r.h.level = used[r.h.level];
            n_levels = next;
        }
        // lay the records out in op order
        n_ops = (int)recs.size();
        size_t off = 0;
        rec_cap = 0;
        descs.assign(recs.size() + 1, SweepDesc()); // + one padding entry (the producer prefetches k + 1)
        memset(descs.data(), 0, sizeof(SweepDesc) * descs.size());
        descs.back().rowA = descs.back().rowB = -1;
        for (size_t k = 0; k < recs.size(); ++k) {
            const Rec &r = recs[k];
            const uint32_t per_cat = (uint32_t)(r.h.szA + r.h.szB);
            const uint32_t bytes = (uint32_t)SW_HDR_BYTES + (uint32_t)n_clv * per_cat; // next op's header + operand blocks
            SweepDesc &d = descs[k];
            memset(&d, 0, sizeof d);
            d.rec_off = off;
            d.rec_bytes = bytes;
            d.rowA = r.rowA;
            d.rowB = r.rowB;
            const size_t b0 = off + SW_HDR_BYTES;
            if (r.h.kind == SW_TT) {
                blocks.push_back({b0, per_cat, r.edgeA, 2, r.edgeB});
                blocks.push_back({b0 + (uint32_t)SW_TAB2_BYTES, per_cat, r.edgeA, 0, -1});
                blocks.push_back({b0 + (uint32_t)(SW_TAB2_BYTES + SW_P_BYTES), per_cat, r.edgeB, 0, -1});
            } else {
                if (r.edgeA >= 0) blocks.push_back({b0, per_cat, r.edgeA, r.tabA ? 1 : 0, -1});
                if (r.edgeB >= 0) blocks.push_back({b0 + (uint32_t)r.h.szA, per_cat, r.edgeB, r.tabB ? 1 : 0, -1});
            }
            rec_cap = std::max(rec_cap, (int)bytes);
            off += bytes;
        }
        stream_bytes = off;
        ops.assign(recs.size() + 2, SweepOp{0, 0, 0, 0}); // + two padding entries (the consumers prefetch k + 2)
        base_w0.assign(recs.size(), 0);
        done_op.assign(n, -1);
        for (size_t k = 0; k < recs.size(); ++k) {
            const Hdr &h = recs[k].h;
            const uint32_t code = (uint32_t)h.kind | ((h.flags & (SW_FLAG_U | SW_FLAG_XU)) ? 8u : 0u);
            base_w0[k] = code | ((uint32_t)h.flags << 8) | ((uint32_t)h.level << 16);
            ops[k].w1 = (uint32_t)(h.szA + h.szB) | ((uint32_t)h.szA << 16);
            const int nd = (h.flags & SW_FLAG_DONE) ? recs[k].node : -1;
            if (nd >= 0) done_op[nd] = (int)k;
            ops[k].w2 = (nd >= 0 && (size_t)nd < slot.size() && slot[nd] >= 0) ? (uint32_t)slot[nd] : 0u;
            ops[k].w3 = 0;
        }
        plan_guards(std::vector<double>(), std::vector<double>()); // until bounds are known: guard after every node
        valid = true;
    }

    // Decides where the underflow guard runs.  A guard after every node costs ~13 dependent integer instructions per
    // op; almost all of them are provably idle.  b[v] / bd[v] bound, in bits, how much the edge above node v can
    // shrink a vector's maximum: every entry of P_v is >= 2^-b[v], every diagonal entry >= 2^-bd[v] (worst rate
    // category; sweep_guard_bounds).  For a binary node p = (a, c):
    //     max(CLV_p) >= max(CLV_a) * max(CLV_c) * 2^-s,   s = min(b[a] + bd[c], bd[a] + b[c])
    // (take the row where one child's maximum sits: its own P contributes a diagonal entry, the other child's P at
    // least its smallest entry; a tip child's table entry is a sum of entries of P).  E[v] is the resulting lower
    // bound on log2 max(CLV_v), accumulated over the unguarded part of v's subtree.  Whenever it would fall below
    // -LIM the children get a guard (their maxima are then >= 2^-256), and if even that is not enough -- the edge
    // pair alone can shrink by more than LIM - 512 bits, i.e. r*t below ~1e-25 -- they are normalised to [1,2) on
    // completion, which is the reference's own operating point (Profile::scale leaves max == 1).  LIM = 600 keeps
    // >= 422 bits of dynamic range below the maximum before an entry turns subnormal.  Nodes that are not binary
    // are always guarded.  Empty bounds: guard after every node (the schedule's state before set_model / blen).
    // Rescaling is an exact power of two, so the placement changes no bit of any CLV.
    void plan_guards(const std::vector<double> &b, const std::vector<double> &bd)
    {
        const int n = (int)done_op.size();
        const double LIM = 600.0, T = 256.0;
        std::vector<char> guard(n, 0), normout(n, 0);
        if ((int)b.size() != n || (int)bd.size() != n || getenv("PHB_SWEEP_GUARD_ALL")) {
            for (int v = 0; v < n; ++v) guard[v] = done_op[v] >= 0;
        } else {
            std::vector<double> Epre(n, 0.0), Epost(n, 0.0);
            auto leaf = [&](int v) { return kid_off[v + 1] == kid_off[v]; };
            auto set_guard = [&](int c) {
                if (leaf(c)) return;
                guard[c] = 1;
                Epost[c] = std::max(Epre[c], -T);
            };
            for (int v : post_order) {
                if (leaf(v)) continue;
                const int nk = kid_off[v + 1] - kid_off[v];
                if (nk != 2) {
                    for (int k = kid_off[v]; k < kid_off[v + 1]; ++k) set_guard(kids[k]);
                    guard[v] = 1;
                    Epre[v] = Epost[v] = -T;
                    continue;
                }
                const int a = kids[kid_off[v]], c = kids[kid_off[v] + 1];
                const double s = std::min(b[a] + bd[c], bd[a] + b[c]);
                double E = Epost[a] + Epost[c] - s;
                if (!(E >= -LIM)) {
                    set_guard(a);
                    set_guard(c);
                    E = Epost[a] + Epost[c] - s;
                    if (!(E >= -LIM)) {
                        for (int x : {a, c})
                            if (!leaf(x)) {
                                normout[x] = 1;
                                Epost[x] = 0.0;
                            }
                        E = Epost[a] + Epost[c] - s;
                        if (!(E >= -LIM)) E = -LIM; // the edge pair alone: same exposure as the reference's own product
                    }
                }
                Epre[v] = Epost[v] = E;
            }
        }
        n_guards = n_normout = 0;
        for (size_t k = 0; k < base_w0.size(); ++k) ops[k].w0 = base_w0[k];
        for (int v = 0; v < n; ++v) {
            if (done_op[v] < 0) continue;
            if (normout[v]) {
                ops[done_op[v]].w0 |= (uint32_t)SW_FLAG_NORMOUT << 8;
                ++n_normout;
            } else if (guard[v]) {
                ops[done_op[v]].w0 |= (uint32_t)SW_FLAG_GUARD << 8;
                ++n_guards;
            }
        }
    }
};

// Per-edge shrink bounds for SweepSched::plan_guards, from the model's eigen-system (phb_set_model) and the branch
// lengths.  Q = diag(1/sqrt(pi)) V diag(lambda) V^T diag(sqrt(pi)); with mu = max_i |Q_ii| and q = min_{i != j} Q_ij,
// uniformisation (P(t) = sum_k Poisson(k; mu t) R^k, R = I + Q/mu >= 0) gives, term k = 1 and k = 0,
//     P_ij(t) >= e^{-mu t} t q  (i != j),      P_ii(t) >= e^{-mu t},
// and because P(t + s) = P(s) P(t) with stochastic P(s), the smallest entry of P(t) never decreases with t: the
// bound is evaluated at t* = min(t, 1/mu).  Worst rate category (smallest positive rate; a rate-0 category has
// P = I, whose off-diagonal zeros bound nothing: b = inf).  One bit of slack covers host/device rounding.
inline void sweep_guard_bounds(const ModelDev &M, const std::vector<double> &blen, std::vector<double> &b, std::vector<double> &bd)
{
    const size_t n = blen.size();
    b.assign(n, INFINITY);
    bd.assign(n, 0.0);
    double Q[16], mu = 0.0, qmin = INFINITY;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            double acc = 0.0;
            for (int k = 0; k < 4; ++k) acc += M.V[i + 4 * k] * M.lambda[k] * M.V[j + 4 * k];
            Q[4 * i + j] = M.inv_sqrt_pi[i] * acc * M.sqrt_pi[j];
            if (i == j) mu = std::max(mu, std::fabs(Q[4 * i + j]));
            else qmin = std::min(qmin, Q[4 * i + j]);
        }
    double rmin = INFINITY, rmax = 0.0;
    bool zero_rate = false;
    for (int c = 0; c < M.n_clv; ++c) {
        const double r = M.rates[M.cat_of_clv[c]];
        if (r > 0.0) {
            rmin = std::min(rmin, r);
            rmax = std::max(rmax, r);
        } else
            zero_rate = true;
    }
    if (!(mu > 0.0) || !(qmin > 0.0) || !std::isfinite(mu) || !std::isfinite(rmin) || zero_rate) return; // b = inf: always guard
    const double LOG2E = 1.4426950408889634;
    for (size_t v = 0; v < n; ++v) {
        const double t = blen[v];
        if (!(t > 0.0) || !std::isfinite(t)) continue;
        // smallest entry: worst at the slowest category; smallest diagonal entry: worst at the fastest
        const double tau = rmin * t, ts = std::min(tau, 1.0 / mu);
        const double floor_bits = mu * ts * LOG2E - std::log2(ts * qmin); // -log2(e^{-mu t*} t* q)
        b[v] = floor_bits + 1.0;
        bd[v] = std::min(mu * rmax * t * LOG2E, floor_bits) + 1.0;
    }
}

// Launch geometry / shared-memory carve-up for a schedule.
struct SweepGeom {
    int Q = 0, n_consumers = 0, threads = 0, n_stages = 0, stage_bytes = 0, smem_levels = 0, g_levels = 0;
    size_t level_bytes = 0, smem_bytes = 0;
    int64_t planned_P = -1;
    // Picks the site quads per CTA pass (Q) for a shard of P patterns on n_sm SMs.  A pass over the op list
    // costs about t(n) = 304 + 2.67 n clocks per op for n consumer threads (measured on B200 at n = 256 and
    // 320), and the persistent grid needs ceil(blocks / n_sm) passes, so a slightly narrower or wider CTA
    // can save a whole pass when the block count lands just above a multiple of the SM count (the strong-
    // scaling shards of 125k-500k patterns).  PHB_SWEEP_Q overrides the choice (experiments).
    bool plan(const SweepSched &S, size_t smem_max, int64_t P = 0, int n_sm = 148)
    {
        const int q_max = (SW_MAX_CONSUMERS / S.n_clv) & ~3;
        if (q_max < 4) return false;
        Q = q_max;
        if (P > 0 && n_sm > 0) {
            double best = 1e300;
            for (int q = q_max; q >= 8 && q >= q_max / 2; q -= 4) {
                const int64_t blocks = (P + 4 * q - 1) / (4 * q);
                const int64_t passes = (blocks + n_sm - 1) / n_sm;
                const double cost = (double)passes * (304.0 + 2.67 * q * S.n_clv);
                if (cost < best - 1e-9) {
                    best = cost;
                    Q = q;
                }
            }
        }
        if (const char *e = getenv("PHB_SWEEP_Q")) {
            const int q = atoi(e) & ~3;
            if (q >= 4 && q <= q_max) Q = q;
        }
        planned_P = P;
        n_consumers = Q * S.n_clv;
        threads = ((n_consumers + 31) / 32) * 32 + 32;
        level_bytes = (size_t)n_consumers * SW_R * 32;
        stage_bytes = (S.rec_cap + 2 * 4 * Q + 127) / 128 * 128;
        const int want_levels = std::max(S.n_levels, 1);
        for (n_stages = 8; n_stages >= 2; n_stages /= 2) { // a power of two: the kernel wraps with a mask
            size_t fixed = SW_BAR_BYTES + (size_t)n_stages * stage_bytes;
            if (fixed + level_bytes > smem_max) continue;
            smem_levels = (int)std::min<size_t>((smem_max - fixed) / level_bytes, (size_t)want_levels);
            if (smem_levels >= std::min(want_levels, 3) || n_stages == 2) break;
        }
        if (n_stages < 2) return false;
        g_levels = std::max(0, S.n_levels - smem_levels);
        smem_bytes = SW_BAR_BYTES + (size_t)n_stages * stage_bytes + (size_t)smem_levels * level_bytes;
        return true;
    }
};

} // namespace phb
