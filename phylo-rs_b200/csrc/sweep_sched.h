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
    int max_rows = 0;   // most tip-code rows any op needs (<= 3: a tip and a super-tip)
    int n_stage_descs = 0;
    // arithmetic one sweep executes per (site pattern, CLV category), fast paths: FP64 instructions (DMUL / DFMA) and
    // flops (DFMA = 2).  A 4x4 matrix-vector product is 4 DMUL + 12 DFMA; lookups cost none.
    double fp64_instr = 0, flops = 0;
    size_t stream_bytes = 0;
    std::vector<SweepStage> stages;      // ring stages (+ one padding entry): the producer's work list
    std::vector<uint64_t> op_off;        // stream offset of every op's record (the transition kernel drops the headers there)
    int stage_rec_cap = 0;               // record bytes a ring stage must hold
    std::vector<SweepBlock> blocks;
    std::vector<SweepOp> ops;           // per-op header table (global memory, prefetched by the consumers)
    int count_kind[16] = {0};
    int n_push = 0, n_reg_sibling = 0;  // pending siblings parked in the stack / kept in the second accumulator
    // guard planning (plan_guards): the op that completes each node, the flags the schedule itself needs, and the
    // tree in post-order
    std::vector<int> done_op;           // node -> index of the op that completes it (-1: leaf)
    std::vector<int> first_op;          // node -> index of the first op of its subtree (the ops of a subtree are
                                        // consecutive: first_op[v] .. done_op[v]); -1: leaf / folded cherry
    std::vector<int> op_node;           // op -> node it completes (-1: not a completing op)
    std::vector<uint32_t> base_w0;      // per op: w0 without SW_FLAG_GUARD / SW_FLAG_NORMOUT
    std::vector<int> post_order;        // children before parents
    std::vector<int> kid_off, kids;     // copy of the child lists
    std::vector<int> built_leaf_row;    // and of the leaf -> alignment row map, root and slots the schedule was built for
    std::vector<int> built_slot;
    int built_root = -1;
    // the schedule depends on the topology only (branch lengths enter through the stream and the guard plan): a
    // set_tree with the same child lists, rows and root -- every step of a branch-length optimisation -- reuses it
    bool same_topology(int root, const std::vector<int> &child_off, const std::vector<int> &children,
                       const std::vector<int> &leaf_row, const std::vector<int> &slot) const
    {
        return valid && root == built_root && child_off == kid_off && children == kids && leaf_row == built_leaf_row && slot == built_slot;
    }
    int n_guards = 0, n_normout = 0;

    // child_off/children: CSR child lists; leaf_row: alignment row per leaf; root must be internal.
    // slot: CLV slot per node (-1 for leaves), recorded in the header of the op that completes the node.
    // collapse: fold every cherry (two leaf children) whose parent is binary into its parent's op as a SUPER-TIP:
    //   the parent reads P_cherry-edge * (tip(x) (.) tip(y)) for the code pair of its two leaves from a 25-row
    //   table, exactly as it reads a tip's masked column sums -- the cherry's own op, its product and its
    //   parent's matrix-vector product all disappear (a third of the ops and of the FP64 work of a Yule tree).
    //   Not for the storing variant (the pre-order pass needs every node's CLV), and not for a cherry whose
    //   `no_collapse` mark is set (plan_guards: branch x rate so small that the unscaled table could underflow).
    bool collapsed_build = false;
    std::vector<char> pseudo;            // node is a cherry folded into its parent
    std::vector<char> built_no_collapse; // the mask the current schedule was built with
    void build(int n, int root, const std::vector<int> &child_off, const std::vector<int> &children,
               const std::vector<int> &leaf_row, int n_clv_, const std::vector<int> &slot = std::vector<int>(),
               bool collapse = false, const std::vector<char> &no_collapse = std::vector<char>())
    {
        valid = false;
        n_clv = n_clv_;
        collapsed_build = collapse;
        stages.clear();
        op_off.clear();
        blocks.clear();
        for (int &c : count_kind) c = 0;
        auto is_leaf = [&](int v) { return child_off[v + 1] == child_off[v]; };
        auto arity = [&](int v) { return child_off[v + 1] - child_off[v]; };
        // post-order (children before parents) for the need computation
        std::vector<int> order, parent(n, -1);
        order.reserve(n);
        {
            std::vector<int> st(1, root);
            while (!st.empty()) {
                int v = st.back();
                st.pop_back();
                order.push_back(v);
                for (int k = child_off[v]; k < child_off[v + 1]; ++k) {
                    parent[children[k]] = v;
                    st.push_back(children[k]);
                }
            }
            std::reverse(order.begin(), order.end());
        }
        post_order = order;
        kid_off = child_off;
        kids = children;
        built_leaf_row = leaf_row;
        built_slot = slot;
        built_root = root;
        pseudo.assign(n, 0);
        if (collapse)
            for (int v = 0; v < n; ++v) {
                if (arity(v) != 2 || v == root || arity(parent[v]) != 2) continue;
                if (!is_leaf(children[child_off[v]]) || !is_leaf(children[child_off[v] + 1])) continue;
                if ((size_t)v < no_collapse.size() && no_collapse[v]) continue;
                pseudo[v] = 1;
            }
        // a node with two cherry children keeps one of them as an ordinary cherry op (the record of an op with two
        // super-tips would be the largest by far and set the size of every ring stage)
        for (int v = 0; v < n; ++v)
            if (arity(v) == 2 && pseudo[children[child_off[v]]] && pseudo[children[child_off[v] + 1]]) pseudo[children[child_off[v]]] = 0;
        auto tipish = [&](int v) { return is_leaf(v) || pseudo[v]; };
        std::vector<int> need(n, 0);
        std::vector<std::vector<int>> ints(n), tips(n);
        for (int v : order) {
            if (tipish(v)) continue;
            for (int k = child_off[v]; k < child_off[v + 1]; ++k) {
                int c = children[k];
                (tipish(c) ? tips[v] : ints[v]).push_back(c);
            }
            std::stable_sort(ints[v].begin(), ints[v].end(), [&](int a, int b) { return need[a] > need[b]; });
            int nd = 0;
            if (!ints[v].empty()) nd = need[ints[v][0]];
            if (ints[v].size() >= 2) nd = std::max(nd, need[ints[v][1]] + 1);
            need[v] = nd;
        }
        n_levels = need[root];
        // chain[v]: the subtree of v is a chain of binary tip x tip / tip x internal updates (no internal x internal
        // node, no polytomy): it runs entirely in ONE accumulator, so as the pending-side sibling of a binary node it
        // can be evaluated in the second register accumulator `u` with no stack traffic at all.
        std::vector<char> chain(n, 0);
        for (int v : order) {
            if (tipish(v)) continue;
            const size_t ni = ints[v].size(), nt = tips[v].size();
            chain[v] = ni + nt == 2 && (ni == 0 || (ni == 1 && chain[ints[v][0]]));
        }

        // operand of an op: a tip (5-row table of its edge, one code row), a super-tip (25-row table of a folded
        // cherry, two code rows), the transition matrix of an edge, or -- SW_TT only -- the product table of the
        // two tip edges of an explicit cherry
        enum { O_NONE = 0, O_TIP, O_SUPER, O_P, O_CHERRY };
        struct Opd { int type, node; };
        struct Hdr { int kind, flags, level, szA, szB, nrowA; };
        struct Rec { Hdr h; Opd A, B; int rows[4]; int node; };
        std::vector<Rec> recs;
        int pending_push = -1;
        bool u_mode = false;
        n_push = n_reg_sibling = 0;
        auto opd_tip = [&](int v) { return Opd{pseudo[v] ? O_SUPER : O_TIP, v}; };
        auto opd_bytes = [&](const Opd &o, bool store_p) {
            switch (o.type) {
            case O_TIP: return SW_TIP_BYTES;
            case O_SUPER: return SW_TAB2_BYTES;
            case O_P: return SW_P_BYTES;
            case O_CHERRY: return SW_TAB2_BYTES + (store_p ? 2 * SW_P_BYTES : 0);
            default: return 0;
            }
        };
        auto emit = [&](int kind, int level, Opd A, Opd B) {
            Rec r;
            memset(&r, 0, sizeof r);
            r.h.kind = kind;
            r.h.level = level;
            if (u_mode) r.h.flags |= SW_FLAG_U;
            if (kind == SW_TT || kind == SW_TS || kind == SW_LOADTIP) { // ops that START a value
                if (pending_push >= 0) {
                    r.h.flags |= SW_FLAG_PUSH;
                    r.h.level = pending_push;
                    pending_push = -1;
                    ++n_push;
                }
            }
            r.A = A;
            r.B = B;
            // the storing variant keeps the two P(t) behind an explicit cherry's table: its ambiguity-code path then
            // follows the reference's summation order bit for bit (posterior ties); the plain variant sums table rows
            r.h.szA = opd_bytes(A, !collapse);
            r.h.szB = opd_bytes(B, false);
            int nr = 0;
            auto add_rows = [&](const Opd &o) {
                if (o.type == O_TIP) r.rows[nr++] = leaf_row[o.node];
                else if (o.type == O_SUPER) {
                    r.rows[nr++] = leaf_row[children[child_off[o.node]]];
                    r.rows[nr++] = leaf_row[children[child_off[o.node] + 1]];
                }
            };
            for (int &x : r.rows) x = -1;
            if (A.type == O_CHERRY) { // the two tip edges ride in A.node (first leaf) and B.node (second leaf)
                r.rows[nr++] = leaf_row[A.node];
                r.rows[nr++] = leaf_row[B.node];
                r.h.nrowA = 1;
            } else {
                add_rows(A);
                r.h.nrowA = nr;
                add_rows(B);
            }
            recs.push_back(r);
            count_kind[kind]++;
        };
        const Opd none{O_NONE, -1};
        struct Frame { int v, j; bool ret; };
        std::vector<Frame> st;
        st.push_back({root, 0, false});
        first_op.assign(n, -1);
        first_op[root] = 0;
        while (!st.empty()) {
            Frame &f = st.back();
            const int v = f.v;
            const std::vector<int> &iv = ints[v], &tv = tips[v];
            const bool binary = iv.size() + tv.size() == 2;
            if (f.ret) { // child iv[j-1] just finished; its CLV is the running value
                f.ret = false;
                if (!binary) {
                    emit(SW_APPLY, 0, Opd{O_P, iv[f.j - 1]}, none);
                    // a product of >= 3 child contributions: every child enters at max in [1,2), as in the
                    // reference (children are max-normalised, profile.rs:20-39), so the product leaves the
                    // normal range exactly where the reference's own product does
                    if (iv.size() + tv.size() >= 3) recs.back().h.flags |= SW_FLAG_NORM;
                    if (f.j - 1 >= 1) emit(SW_MULPOP, need[iv[f.j - 1]], none, none);
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
                first_op[c] = (int)recs.size(); // the next op emitted opens c's subtree
                st.push_back({c, 0, false});
                continue;
            }
            if (binary) {
                if (iv.size() == 2) {
                    const bool from_u = u_mode; // set when iv[1] was evaluated in u (chains hold no II, so no nesting)
                    u_mode = false;
                    // operand A is the waiting side: the popped CLV of iv[0], or u holding iv[1]
                    if (from_u) emit(SW_II, need[iv[1]], Opd{O_P, iv[1]}, Opd{O_P, iv[0]});
                    else emit(SW_II, need[iv[1]], Opd{O_P, iv[0]}, Opd{O_P, iv[1]});
                    if (from_u) recs.back().h.flags |= SW_FLAG_XU;
                } else if (iv.size() == 1) {
                    if (pseudo[tv[0]]) emit(SW_SI, 0, Opd{O_SUPER, tv[0]}, Opd{O_P, iv[0]});
                    else emit(SW_TI, 0, Opd{O_TIP, tv[0]}, Opd{O_P, iv[0]});
                } else {
                    int a = tv[0], c = tv[1];
                    if (pseudo[a]) std::swap(a, c); // at most one of the two is a super-tip: it goes second
                    if (pseudo[c]) emit(SW_TS, 0, Opd{O_TIP, a}, Opd{O_SUPER, c});
                    else emit(SW_TT, 0, Opd{O_CHERRY, a}, Opd{O_NONE, c});
                }
            } else if (iv.empty()) {
                size_t k = 0;
                if (tv.size() >= 2) {
                    emit(SW_TT, 0, Opd{O_CHERRY, tv[0]}, Opd{O_NONE, tv[1]});
                    k = 2;
                } else {
                    emit(SW_LOADTIP, 0, Opd{O_TIP, tv[0]}, none);
                    k = 1;
                }
                for (; k < tv.size(); ++k) emit(SW_MULTIP, 0, Opd{O_TIP, tv[k]}, none);
            } else {
                for (int tnode : tv) emit(SW_MULTIP, 0, Opd{O_TIP, tnode}, none);
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
        // lay the records out in op order and pack consecutive ops into ring stages (greedy: up to
        // SW_STAGE_REC_BYTES of records and SW_STAGE_ROWS tip-code rows per stage)
        n_ops = (int)recs.size();
        size_t off = 0;
        rec_cap = 0;
        max_rows = 0;
        op_off.assign(recs.size(), 0);
        std::vector<uint32_t> op_w3(recs.size(), 0);
        std::vector<char> op_first(recs.size(), 0), op_last(recs.size(), 0);
        auto add_block = [&](size_t at, uint32_t per_cat, const Opd &o, const Opd &other, bool store_p) {
            auto kid = [&](int c, int i) { return children[child_off[c] + i]; };
            switch (o.type) {
            case O_TIP: blocks.push_back({at, per_cat, o.node, 1, -1, -1, 0}); break;
            case O_P: blocks.push_back({at, per_cat, o.node, 0, -1, -1, 0}); break;
            case O_SUPER: blocks.push_back({at, per_cat, o.node, 3, kid(o.node, 0), kid(o.node, 1), 0}); break;
            case O_CHERRY:
                blocks.push_back({at, per_cat, o.node, 2, other.node, -1, 0});
                if (store_p) {
                    blocks.push_back({at + (uint32_t)SW_TAB2_BYTES, per_cat, o.node, 0, -1, -1, 0});
                    blocks.push_back({at + (uint32_t)(SW_TAB2_BYTES + SW_P_BYTES), per_cat, other.node, 0, -1, -1, 0});
                }
                break;
            default: break;
            }
        };
        for (size_t k = 0; k < recs.size(); ++k) {
            const Rec &r = recs[k];
            const uint32_t per_cat = (uint32_t)(r.h.szA + r.h.szB);
            rec_cap = std::max(rec_cap, (int)(SW_HDR_BYTES + n_clv * per_cat));
        }
        stage_rec_cap = std::max(rec_cap, SW_STAGE_REC_BYTES);
        SweepStage cur;
        memset(&cur, 0, sizeof cur);
        auto flush = [&](size_t k_last) {
            op_last[k_last] = 1;
            stages.push_back(cur);
            memset(&cur, 0, sizeof cur);
        };
        for (size_t k = 0; k < recs.size(); ++k) {
            const Rec &r = recs[k];
            const uint32_t per_cat = (uint32_t)(r.h.szA + r.h.szB);
            const uint32_t bytes = (uint32_t)SW_HDR_BYTES + (uint32_t)n_clv * per_cat; // next op's header + operand blocks
            int nr = 0;
            for (int i = 0; i < 4; ++i) nr += r.rows[i] >= 0;
            max_rows = std::max(max_rows, nr);
            if (cur.rec_bytes > 0 && (cur.rec_bytes + bytes > (uint32_t)stage_rec_cap || cur.n_rows + nr > SW_STAGE_ROWS)) flush(k - 1);
            if (cur.rec_bytes == 0) {
                cur.rec_off = off;
                op_first[k] = 1;
            }
            op_off[k] = off;
            op_w3[k] = cur.rec_bytes | ((uint32_t)cur.n_rows << 16);
            for (int i = 0; i < 4; ++i)
                if (r.rows[i] >= 0) cur.rows[cur.n_rows++] = r.rows[i];
            cur.rec_bytes += bytes;
            const size_t b0 = off + SW_HDR_BYTES;
            add_block(b0, per_cat, r.A, r.B, !collapse);
            if (r.A.type != O_CHERRY) add_block(b0 + (uint32_t)r.h.szA, per_cat, r.B, r.A, false);
            off += bytes;
        }
        if (!recs.empty()) flush(recs.size() - 1);
        n_stage_descs = (int)stages.size();
        memset(&cur, 0, sizeof cur);
        stages.push_back(cur); // padding entry: the producer reads one stage ahead
        stream_bytes = off;
        fp64_instr = flops = 0;
        for (const Rec &r : recs) {
            int mv = 0, mul = 0; // matrix-vector products, elementwise products of 4
            switch (r.h.kind) {
            case SW_TI: case SW_SI: mv = 1; mul = 1; break;
            case SW_II: mv = 2; mul = 1; break;
            case SW_TS: case SW_MULTIP: case SW_MULPOP: mul = 1; break;
            case SW_APPLY: mv = 1; break;
            default: break; // TT, LOADTIP: table rows only
            }
            fp64_instr += 16.0 * mv + 4.0 * mul;
            flops += 28.0 * mv + 4.0 * mul;
        }
        ops.assign(recs.size() + 2, SweepOp{0, 0, 0, 0}); // + two padding entries
        base_w0.assign(recs.size(), 0);
        done_op.assign(n, -1);
        op_node.assign(recs.size(), -1);
        for (size_t k = 0; k < recs.size(); ++k) {
            const Hdr &h = recs[k].h;
            const uint32_t code = (uint32_t)sw_code(h.kind, (h.flags & (SW_FLAG_U | SW_FLAG_XU)) != 0);
            const uint32_t fl = (uint32_t)h.flags | (op_first[k] ? (uint32_t)SW_FLAG_FIRST : 0u) | (op_last[k] ? (uint32_t)SW_FLAG_LAST : 0u);
            base_w0[k] = code | (fl << 8) | ((uint32_t)h.level << 18);
            ops[k].w1 = (uint32_t)(h.szA + h.szB) | ((uint32_t)h.szA << 16);
            const int nd = (h.flags & SW_FLAG_DONE) ? recs[k].node : -1;
            if (nd >= 0) done_op[nd] = (int)k;
            op_node[k] = nd;
            // the storing variant skips a node whose row is SW_NO_ROW (a node nobody asked for)
            ops[k].w2 = (nd >= 0 && (size_t)nd < slot.size() && slot[nd] >= 0) ? (uint32_t)slot[nd] : SW_NO_ROW;
            ops[k].w3 = op_w3[k];
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
                if (leaf(c) || done_op[c] < 0) return; // a folded cherry has no op of its own: nothing to guard
                guard[c] = 1;
                Epost[c] = std::max(Epre[c], -T);
            };
            for (int v : post_order) {
                if (leaf(v)) continue;
                const int nk = kid_off[v + 1] - kid_off[v];
                if (nk != 2) {
                    for (int k = kid_off[v]; k < kid_off[v + 1]; ++k) set_guard(kids[k]);
                    guard[v] = done_op[v] >= 0;
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
                            if (!leaf(x) && done_op[x] >= 0) {
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

// Cherries that must stay explicit ops: the folded table holds P_v (tip_x (.) tip_y) UNSCALED, so its entries can be as
// small as 2^-(s + b[v]), s the cherry's own shrink bound; beyond 400 bits (branch x rate below ~1e-60) the explicit
// op with its guard / normalisation is kept.
inline void sweep_no_collapse(const std::vector<int> &child_off, const std::vector<int> &children, const std::vector<double> &b,
                              const std::vector<double> &bd, std::vector<char> &out)
{
    const int n = (int)child_off.size() - 1;
    out.assign(n, 0);
    if ((int)b.size() != n) return;
    for (int v = 0; v < n; ++v) {
        if (child_off[v + 1] - child_off[v] != 2) continue;
        const int a = children[child_off[v]], c = children[child_off[v] + 1];
        const double s = std::min(b[a] + bd[c], bd[a] + b[c]);
        out[v] = !(s + b[v] <= 400.0);
    }
}

// Site quads per CTA: consumer warps are always full, so Q * n_clv is a multiple of 32 (and Q of 4: 16-byte
// tip-code rows for the bulk copies).
inline int sweep_q_step(int n_clv)
{
    int step = 4;
    while ((step * n_clv) % 32) step += 4;
    return step;
}
inline int sweep_q_max(int n_clv) { return SW_MAX_CONSUMERS / n_clv / sweep_q_step(n_clv) * sweep_q_step(n_clv); }

// Block width shared by the storing sweep and the pre-order sweep (down_sweep.cuh): 4 Q patterns, where 4 Q is also
// the pre-order kernel's consumer thread count (a multiple of 32, at most 352 so that 12 warps keep 168 registers).
inline int sweep_q_asr(int n_clv)
{
    int step = sweep_q_step(n_clv);
    while (step % 8) step += sweep_q_step(n_clv);
    const int cap = std::min(88, SW_MAX_CONSUMERS / n_clv);
    return cap / step * step;
}

// Launch geometry / shared-memory carve-up for a schedule.
struct SweepGeom {
    int Q = 0, n_consumers = 0, threads = 0, n_stages = 0, stage_bytes = 0, smem_levels = 0, g_levels = 0;
    size_t level_bytes = 0, smem_bytes = 0;
    int64_t planned_P = -1;
    // Picks the site quads per CTA pass (Q) for a shard of P patterns on n_sm SMs.  A pass of the persistent grid over
    // the op list is latency-bound: its time depends on the resident warps per scheduler, hardly on the width within
    // such a class (B200, 10k-taxon tree, ms per pass at Q = 16 / 24 | 40 / 56 | 72 / 88: 2.79 / 2.87 | 3.53 / 3.59 |
    // 4.10 / 4.14, profiles/r2_pass_time_vs_width.txt), and the grid needs ceil(blocks / n_sm) passes -- so the widest
    // CTA wins unless a narrower one saves a pass.  tail_split: the caller sends the blocks left over after the last full
    // pass as a launch of their own, just wide enough for one pass (run_sweep); the cost model then charges that pass at
    // its own, cheaper class.  PHB_SWEEP_Q overrides the choice (experiments).
    static int warps_per_scheduler(int q, int n_clv) { return (q * n_clv / 32 + 1 + 3) / 4; }
    static double pass_cost(int q, int n_clv)
    {
        const int c = warps_per_scheduler(q, n_clv);
        return (c <= 1 ? 0.69 : c == 2 ? 0.86 : 1.0) * (1.0 + 0.0006 * q);
    }
    bool plan(const SweepSched &S, size_t smem_max, int64_t P = 0, int n_sm = 148, int force_q = 0, bool tail_split = false)
    {
        const int step = sweep_q_step(S.n_clv), q_max = sweep_q_max(S.n_clv);
        if (q_max < step) return false;
        Q = q_max;
        if (force_q > 0) Q = force_q; // the storing sweep writes blocks the pre-order sweep reads back: fixed width
        else if (P > 0 && n_sm > 0) {
            double best = 1e300;
            for (int q = q_max; q >= step && q >= q_max / 2; q -= step) {
                const int64_t blocks = (P + 4 * q - 1) / (4 * q);
                const int64_t full = blocks / n_sm, rem = blocks - full * n_sm;
                double cost;
                if (tail_split && full > 0 && rem > 0) {
                    const int64_t tail_P = P - full * n_sm * 4 * q;
                    int qt = (int)((tail_P + (int64_t)4 * n_sm - 1) / ((int64_t)4 * n_sm));
                    qt = std::min(q, (qt + step - 1) / step * step);
                    cost = (double)full * pass_cost(q, S.n_clv) + pass_cost(qt, S.n_clv);
                } else
                    cost = (double)(full + (rem > 0 ? 1 : 0)) * pass_cost(q, S.n_clv);
                if (cost < best - 1e-9) {
                    best = cost;
                    Q = q;
                }
            }
        }
        if (const char *e = force_q > 0 ? nullptr : getenv("PHB_SWEEP_Q")) {
            const int q = atoi(e) / step * step;
            if (q >= step && q <= q_max) Q = q;
        }
        planned_P = P;
        n_consumers = Q * S.n_clv;
        threads = n_consumers + 32;
        level_bytes = (size_t)n_consumers * SW_R * 32;
        stage_bytes = ((S.stage_rec_cap + 15) / 16 * 16 + SW_STAGE_ROWS * 4 * Q + 127) / 128 * 128;
        const int want_levels = std::max(S.n_levels, 1);
        for (n_stages = 4; n_stages >= 2; n_stages /= 2) { // a power of two: the kernel wraps with a mask
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
