"""Shared parity checks: run a Case through a C-ABI library (real sm_100a build, or the emulation
harness build for the CPU scheduler tests) and compare with the oracle.

Tolerances are BASELINE.json's: total lnL 1e-9 relative, per-pattern lnL 1e-8 absolute, ASR argmax
states bit-exact (sites whose top-2 mixed posteriors differ by < 1e-10 are audited, not hidden)."""
import numpy as np

import oracle
from oracle import workloads
from phylo_rs_b200 import capi

REL_TOL = 1e-9
ABS_TOL = 1e-8


def run_lnl(lib, case, flags=0, devices=None, budget=None):
    with capi.Context(devices=devices, flags=flags, library=lib) as ctx:
        if budget:
            ctx.set_clv_budget(budget)
        workloads.load_into(ctx, case)
        tot, per = ctx.log_likelihood(per_pattern=True)
        st = ctx.stats()
    return tot, per, st


def run_asr(lib, case, want_post=False, devices=None, budget=None, flags=0):
    with capi.Context(devices=devices, flags=flags, library=lib) as ctx:
        if budget:
            ctx.set_clv_budget(budget)
        workloads.load_into(ctx, case)
        tot, states, post = ctx.marginal_asr(want_post)
        per = ctx.per_pattern()
    return tot, per, states, post


def assert_lnl_close(tot, per, ref_tot, ref_per):
    fin = np.isfinite(ref_per)
    assert np.array_equal(np.isfinite(per), fin)
    assert np.array_equal(per[~fin], ref_per[~fin]) or np.all(np.isneginf(per[~fin]) == np.isneginf(ref_per[~fin]))
    if fin.any():
        assert np.max(np.abs(per[fin] - ref_per[fin])) <= ABS_TOL
    if np.isfinite(ref_tot):
        assert abs(tot - ref_tot) <= REL_TOL * max(abs(ref_tot), 1e-300)
    else:
        assert tot == ref_tot or (np.isnan(tot) and np.isnan(ref_tot))


def check_lnl(lib, case, **kw):
    tot, per, st = run_lnl(lib, case, **kw)
    ref_tot, ref_per = oracle.log_likelihood_flat(case.ft, case.model, case.patterns, case.mult, mode=0, n_threads=8)
    assert_lnl_close(tot, per, ref_tot, ref_per)
    return tot, per, st


def check_asr(lib, case, want_post=False, **kw):
    tot, per, states, post = run_asr(lib, case, want_post, **kw)
    ref_tot, ref_per, ref_states, ref_post = oracle.marginal_asr_flat(
        case.ft, case.model, case.patterns, case.mult, want_posteriors=True, n_threads=8)
    assert_lnl_close(tot, per, ref_tot, ref_per)
    diff = states != ref_states
    if diff.any():
        # bit-exact bar; a mismatch is tolerated ONLY at a numerical tie of the oracle's own top two
        srt = np.sort(ref_post, axis=2)
        gap = srt[..., -1] - srt[..., -2]
        assert np.all(gap[diff] < 1e-10), f"{int(diff.sum())} ASR state mismatches away from ties"
    if want_post:
        assert np.max(np.abs(post - ref_post)) <= 1e-9
    return int(diff.sum())


def check_nni(lib, case, moves=None, flags=0, budget=None, devices=None, expect_path=True):
    """phb_score_nni against the oracle run on each rebuilt neighbour tree (the reference's caller loop:
    clone, NNI::nni src/tree.rs:1213-1242, log_likelihood).  Returns (totals, stats)."""
    ft = case.ft
    if moves is None:
        moves = [(v, lc) for v in range(ft["n_nodes"]) for lc in (0, 1)]
    with capi.Context(devices=devices, flags=flags, library=lib) as ctx:
        if budget:
            ctx.set_clv_budget(budget)
        workloads.load_into(ctx, case)
        base = ctx.log_likelihood()
        got = ctx.score_nni([m[0] for m in moves], [m[1] for m in moves])
        st = ctx.stats()
        assert ctx.log_likelihood() == base  # the base tree is untouched
    ids = ft["ids"]
    n_valid = 0
    for (v, lc), val in zip(moves, got):
        t = case.tree.clone()
        try:
            failed = not (0 <= v < ft["n_nodes"]) or t.nni(int(ids[v]), bool(lc))
        except IndexError:  # a unifurcation's missing second child / a missing sibling: the reference's indexing panics
            failed = True
        if failed:
            assert np.isnan(val)
            continue
        n_valid += 1
        ref, _ = oracle.log_likelihood_flat(workloads.flatten_fast(t, case.leaf_order), case.model, case.patterns, case.mult,
                                            mode=2, n_threads=8)
        if np.isfinite(ref):
            assert abs(val - ref) <= REL_TOL * abs(ref), (v, lc, val, ref)
        else:
            assert val == ref
    if expect_path is not None:
        assert st["nni_path_moves"] == (n_valid if expect_path else 0)
    return got, st


def check_score_trees(lib, case, flags=0, seed=72):
    """phb_score_trees: the caller loop `t = base.clone(); t.spr(e1, e2); t.log_likelihood(..)` (src/tree/ops.rs:34-41,
    src/tree.rs:1161-1204) as one call.  The candidates are built by the oracle's tree with the reference's prune /
    graft semantics (the pruned node's parent stays behind as a unifurcation, a new node splits the target edge)."""
    from oracle.refhost import Node

    base = case.tree
    ids = [i for i in base.get_node_ids()]
    rng = np.random.default_rng(seed)
    cands, refs = [], []
    for _ in range(400):
        if len(cands) == 5:
            break
        t = base.clone()
        x, y = int(rng.choice(ids)), int(rng.choice(ids))
        px, py = t.nodes[x].parent, t.nodes[y].parent
        if px is None or py is None or y in set(t.preord_ids(x)):
            continue
        t.delete_edge(px, x)                # SPR::prune: the subtree leaves its parent, which stays behind
        new = Node(t.next_id())             # SPR::graft: a new node splits the target edge (no weight: length 0) ...
        t.split_edge((py, y), new)
        t.set_child(new.id, x)              # ... and takes the pruned subtree as its second child
        ft = workloads.flatten_fast(t, case.leaf_order)
        cands.append((ft["postorder"], ft["child_off"], ft["children"], ft["blen"], ft["leaf_row"], ft["root"]))
        refs.append(oracle.log_likelihood_flat(ft, case.model, case.patterns, case.mult, mode=0, n_threads=4)[0])
    assert len(cands) >= 3
    with capi.Context(flags=flags, library=lib) as ctx:
        workloads.load_into(ctx, case)
        before = ctx.log_likelihood()
        bad = list(cands[0])
        bad[4] = np.full_like(np.asarray(bad[4]), -1)  # no leaf has a row: unschedulable, scores NaN
        got = ctx.score_trees(cands + [tuple(bad)])
        assert ctx.log_likelihood() == before  # the context's own tree is untouched
    assert np.isnan(got[-1])
    for g, r in zip(got[:-1], refs):
        assert abs(g - r) <= 1e-9 * abs(r)
    assert len({round(float(x), 6) for x in got[:-1]}) > 1


def check_joint(lib, case, devices=None, budget=None, flags=0):
    """phb_joint_asr against the oracle's compute_joint_asr restatement.  Per-pattern Viterbi values within
    1e-8 absolute, total within 1e-9 relative, states bit-exact except at audited numerical ties: a differing
    pattern is accepted only if the GPU's state assignment reaches the oracle's best value to 1e-9 under the
    oracle's own arithmetic (two assignments of equal joint probability)."""
    with capi.Context(devices=devices, flags=flags, library=lib) as ctx:
        if budget:
            ctx.set_clv_budget(budget)
        workloads.load_into(ctx, case)
        tot, states = ctx.joint_asr()
        per = ctx.per_pattern()
        st = ctx.stats()
    ref_tot, ref_per, ref_states = oracle.joint_asr_flat(case.ft, case.model, case.patterns, case.mult, n_threads=8)
    assert_lnl_close(tot, per, ref_tot, ref_per)
    diff = np.nonzero((states != ref_states).any(axis=0))[0]
    for p in diff:
        val = oracle.joint_assignment_value(case.ft, case.model, case.patterns[p], states[:, p])
        assert abs(val - ref_per[p]) <= 1e-9 * max(1.0, abs(ref_per[p])), f"pattern {p}: joint states differ away from a tie"
    return len(diff), st


def check_compress(lib, rows, devices=None):
    """phb_compress_columns against the oracle's restatement of Alignment::compress_columns
    (src/alignment.rs:93-125): pattern count, site_to_pattern and multiplicity bit-exact, and the context's
    alignment afterwards holds the patterns in first-appearance order (canonical bytes: U reads back as T)."""
    from oracle import RefAlignment

    rows = np.ascontiguousarray(rows, np.uint8)
    n_rows, width = rows.shape
    aln = RefAlignment({f"r{i}": bytes(rows[i]) for i in range(n_rows)}, width)
    pats, s2p, mult, _lo = aln.compress_columns()
    with capi.Context(devices=devices, library=lib) as ctx:
        n, got_s2p, got_mult = ctx.compress_columns(rows)
        back = ctx.get_alignment() if n and n_rows else None
    assert n == len(pats)
    assert np.array_equal(got_s2p.astype(np.int64), s2p)
    assert np.array_equal(got_mult, mult)
    if back is not None:
        canon = pats.copy()
        canon[canon == ord("U")] = ord("T")
        valid = np.isin(canon, np.frombuffer(b"ACGTRYSWKMBDHVN-", np.uint8))
        assert np.array_equal(back[valid], canon[valid]) and np.all(back[~valid] == ord("?"))
    return n


NONBINARY_NEWICKS = [
    # polytomies of arity 3..6, unifurcations, a polytomy under a polytomy (the oracle itself is checked on these by
    # brute-force enumeration: tests/test_oracle_kat.py)
    "((a:0.1,b:0.2,c:0.3):0.2,(d:0.1,e:0.4,f:0.1,g:0.2):0.3,h:0.5);",
    "(((a:0.2,b:0.1):0.1):0.3,((c:0.3):0.2,d:0.1,e:0.2,f:0.3,g:0.1,h:0.2):0.2,(i:0.1,j:0.3):0.4);",
    "((a:0.05,b:0.4,c:0.1,d:0.3,e:0.2):0.6,(f:0.2,(g:0.1,h:0.1,i:0.3):0.2,j:0.2):0.1,k:0.7,l:0.02);",
    "(a:0.3,b:0.1,c:0.2,d:0.4,e:0.5,f:0.6);",
]


def nonbinary_case(nwk, model, n_sites=300):
    import re

    from oracle import parse_newick

    tree = parse_newick(nwk)
    names = re.findall(r"[(,]([a-z])", nwk)
    ft = workloads.flatten_fast(tree, names)
    rng = np.random.default_rng(len(nwk))
    pats = np.frombuffer(b"ACGTACGTACGTRYN-", np.uint8)[rng.integers(16, size=(n_sites, len(names)))]
    return workloads.Case(tree, ft, workloads.make_model(model), model, np.ascontiguousarray(pats), np.ones(n_sites, np.uint64), names)


def check_nonbinary(lib):
    for nwk in NONBINARY_NEWICKS:
        for model in ("gtr", "gtr_i_g4"):
            case = nonbinary_case(nwk, model)
            check_lnl(lib, case)
            check_lnl(lib, case, flags=capi.FLAG_LEVEL_KERNELS)
            check_asr(lib, case, want_post=True)
            check_joint(lib, case)
        # every (node, left_ch) of the shape: a unifurcation's only child moves (left_ch false), its missing second child
        # and a node without a sibling are the reference's index panics -> NaN
        check_nni(lib, nonbinary_case(nwk, "gtr_g4", n_sites=40), expect_path=False)


def check_nni_cache_variants(lib, case):
    """phb_score_nni on ONE context, several move sets in a row (the rows of the cache and the sweep schedule are re-planned
    per call), each once with the storing-sweep cache and once with the level-kernel cache (PHB_NNI_LEVEL_CACHE): the two
    must agree to round-off, and both with full rescoring of a few neighbours."""
    import os

    ft = case.ft
    co = ft["child_off"]
    internal = [v for v in range(ft["n_nodes"]) if co[v + 1] - co[v] == 2 and v != ft["root"]]
    cherries = [v for v in internal if all(co[c + 1] == co[c] for c in ft["children"][co[v]:co[v] + 2])]
    sets = [[(v, lc) for v in internal[:7] for lc in (0, 1)], [(v, 1) for v in cherries[:5]],
            [(v, lc) for v in internal[::3] for lc in (1, 0)], [(internal[-1], 0)]]
    with capi.Context(library=lib) as ctx:
        workloads.load_into(ctx, case)
        base = ctx.log_likelihood()
        for moves in sets:
            ids, left = [m[0] for m in moves], [m[1] for m in moves]
            a = ctx.score_nni(ids, left)
            os.environ["PHB_NNI_LEVEL_CACHE"] = "1"
            try:
                b = ctx.score_nni(ids, left)
            finally:
                del os.environ["PHB_NNI_LEVEL_CACHE"]
            assert np.all(np.abs(a - b) <= 1e-11 * np.abs(b)), (moves, a, b)
            assert ctx.log_likelihood() == base
    with capi.Context(flags=capi.FLAG_NNI_FULL, library=lib) as ctx:
        workloads.load_into(ctx, case)
        moves = sets[2][:6]
        full = ctx.score_nni([m[0] for m in moves], [m[1] for m in moves])
    with capi.Context(library=lib) as ctx:
        workloads.load_into(ctx, case)
        got = ctx.score_nni([m[0] for m in moves], [m[1] for m in moves])
    assert np.all(np.abs(got - full) <= REL_TOL * np.abs(full))


def random_tree_case(seed, n_leaves_max=40, n_sites=96):
    """A random rooted tree of any shape: arities 1..5 mixed, branch lengths from exactly 0 over 1e-8 to 4 (+ a missing
    one, which the reference reads as 0: likelihood.rs:115-118), a random model of the family and 0-20 % ambiguity codes."""
    from oracle import parse_newick

    rng = np.random.default_rng(seed)
    n_leaves = int(rng.integers(2, n_leaves_max + 1))
    names = [f"t{i}" for i in range(n_leaves)]

    def blen():
        r = rng.random()
        if r < 0.06:
            return ":0"
        if r < 0.10:
            return ""  # no branch length
        if r < 0.18:
            return f":{10.0 ** rng.uniform(-8, -4):.3e}"
        if r < 0.24:
            return f":{rng.uniform(2.0, 4.0):.4f}"
        return f":{rng.exponential(0.15) + 1e-4:.5f}"

    nodes = [n + blen() for n in names]
    while len(nodes) > 1:
        k = int(min(len(nodes), rng.choice([1, 2, 2, 2, 2, 3, 3, 4, 5])))
        if k == 1 and len(nodes) == 1:
            break
        idx = sorted(rng.choice(len(nodes), size=k, replace=False).tolist(), reverse=True)
        kids = [nodes.pop(i) for i in idx]
        last = len(nodes) == 0
        nodes.append("(" + ",".join(kids) + ")" + ("" if last else blen()))
        if last:
            break
    nwk = nodes[0] + ";"
    tree = parse_newick(nwk)
    ft = workloads.flatten_fast(tree, names)
    model = workloads.MODELS[int(rng.integers(len(workloads.MODELS)))]
    m = workloads.make_model(model)
    pats = workloads.simulate_patterns(ft, m, n_sites, rng)
    amb = float(rng.choice([0.0, 0.05, 0.2]))
    if amb > 0:
        mask = rng.random(pats.shape) < amb
        codes = np.frombuffer(workloads.AMBIG, np.uint8)
        pats[mask] = codes[rng.integers(len(codes), size=int(mask.sum()))]
    mult = rng.integers(1, 5, size=n_sites).astype(np.uint64)
    return workloads.Case(tree, ft, m, model, np.ascontiguousarray(pats), mult, names), nwk


def check_random_trees(lib, seeds):
    for seed in seeds:
        case, nwk = random_tree_case(seed)
        try:
            check_lnl(lib, case)
            check_lnl(lib, case, flags=capi.FLAG_LEVEL_KERNELS)
            check_asr(lib, case, want_post=True)
            check_joint(lib, case)
            if seed % 3 == 0:  # NNI neighbours of the random shape: the cached-partials walk when it is binary
                rng = np.random.default_rng(seed + 20_000)
                n = case.ft["n_nodes"]
                moves = [(int(v), int(lc)) for v, lc in zip(rng.integers(n, size=8), rng.integers(2, size=8))]
                check_nni(lib, case, moves=moves, expect_path=None)
        except AssertionError as e:
            raise AssertionError(f"seed {seed}, model {case.model_name}, tree {nwk}: {e}") from e


def check_random_trees_amino_acid(lib, seeds):
    """the same random shapes through the 20-state engines"""
    for seed in seeds:
        case, nwk = random_tree_case(seed, n_leaves_max=25, n_sites=40)
        rng = np.random.default_rng(seed + 10_000)
        model = workloads.AA_MODELS[int(rng.integers(3))]
        m = workloads.make_aa_model(model, seed=0xAA20 + seed % 5)
        pats = workloads.simulate_patterns(case.ft, m, 40, rng)
        if seed % 2:
            mask = rng.random(pats.shape) < 0.15
            codes = np.frombuffer(workloads.AA_AMBIG, np.uint8)
            pats[mask] = codes[rng.integers(len(codes), size=int(mask.sum()))]
        c2 = workloads.Case(case.tree, case.ft, m, model, np.ascontiguousarray(pats), case.mult[:40], case.leaf_order)
        try:
            check_lnl(lib, c2, flags=capi.FLAG_AMINO_ACID)
            assert check_asr(lib, c2, want_post=True, flags=capi.FLAG_AMINO_ACID) == 0
            check_joint(lib, c2, flags=capi.FLAG_AMINO_ACID)
            if seed % 3 == 0:  # a few NNI neighbours of the random shape, each re-scored in full
                n = c2.ft["n_nodes"]
                moves = [(int(v), int(lc)) for v, lc in zip(rng.integers(n, size=6), rng.integers(2, size=6))]
                import os

                a, _ = check_nni(lib, c2, moves=moves, flags=capi.FLAG_AMINO_ACID)
                os.environ["PHB_AA_NNI_PATH"] = "1"  # without upper partials: bit-identical to full re-scoring
                try:
                    w, _ = check_nni(lib, c2, moves=moves, flags=capi.FLAG_AMINO_ACID)
                finally:
                    del os.environ["PHB_AA_NNI_PATH"]
                b, _ = check_nni(lib, c2, moves=moves, flags=capi.FLAG_AMINO_ACID | capi.FLAG_NNI_FULL, expect_path=False)
                assert np.array_equal(w, b, equal_nan=True), (w, b)
                fin = np.isfinite(b)
                assert np.array_equal(np.isnan(a), np.isnan(b)) and np.allclose(a[fin], b[fin], rtol=1e-11, atol=0), (a, b)
                assert np.array_equal(a[~fin], b[~fin], equal_nan=True)
        except AssertionError as e:
            raise AssertionError(f"seed {seed}, model {model}, tree {nwk}: {e}") from e
