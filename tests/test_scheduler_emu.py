"""Host scheduler / tiling / sharding / ABI logic of the product sources, exercised on the CPU through
the emulation-harness build (tests/emu: the same .cu files compiled by g++ against a CUDA-runtime shim).
Numerical parity CLAIMS are made by tests/test_gpu_parity.py on the real sm_100a library; these tests
exist so that indexing and scheduling bugs are found before GPU time is spent."""
import numpy as np
import pytest

import oracle
from oracle import RefAlignment, RefModel, parse_newick, workloads
from phylo_rs_b200 import capi

import _cases


@pytest.mark.parametrize("model", ["jc69", "gtr", "gtr_g4", "gtr_i_g4", "gtr_i"])
def test_lnl_small_models(emu_lib, model):
    case = workloads.make_case(12, 300, model=model, seed=11, ambiguity=0.1)
    _cases.check_lnl(emu_lib, case)


def test_lnl_literal_invariant_flag(emu_lib):
    case = workloads.make_case(10, 200, model="gtr_i_g4", seed=3, ambiguity=0.1)
    a = _cases.check_lnl(emu_lib, case)[0]
    b = _cases.check_lnl(emu_lib, case, flags=capi.FLAG_LITERAL_INVARIANT)[0]
    assert abs(a - b) <= 1e-10 * abs(a)


def test_tiling_is_result_invariant(emu_lib):
    case = workloads.make_case(9, 1500, model="gtr_g4", seed=4)
    full = _cases.check_lnl(emu_lib, case, flags=capi.FLAG_LEVEL_KERNELS)
    per_site = 8 * 4 * 4 * 8  # slots x cats x states x f64
    tiled = _cases.check_lnl(emu_lib, case, budget=per_site * 512, flags=capi.FLAG_LEVEL_KERNELS)
    assert tiled[2]["n_tiles"] >= 3
    assert np.array_equal(full[1], tiled[1])
    sweep = _cases.check_lnl(emu_lib, case)  # the fused sweep agrees with the level kernels
    assert np.max(np.abs(sweep[1] - full[1])) < 1e-11


def test_device_sharding_is_result_invariant(emu_lib):
    case = workloads.make_case(9, 1100, model="gtr_g4", seed=5)
    one = _cases.check_lnl(emu_lib, case)
    for devs in ([0, 1], [0, 1, 2, 3]):
        many = _cases.check_lnl(emu_lib, case, devices=devs)
        assert np.array_equal(one[1], many[1])
        assert abs(one[0] - many[0]) <= 1e-12 * abs(one[0])


@pytest.mark.parametrize("model", ["jc69", "gtr_g4", "gtr_i_g4"])
def test_level_kernels_flag(emu_lib, model):
    case = workloads.make_case(12, 300, model=model, seed=11, ambiguity=0.1)
    _cases.check_lnl(emu_lib, case, flags=capi.FLAG_LEVEL_KERNELS)


def test_sweep_many_blocks_per_cta(emu_lib):
    """More site blocks than (emulated) SMs: every persistent CTA sweeps the tree several times."""
    case = workloads.make_case(7, 2600, model="gtr_g4", seed=14)  # <= 320-site blocks -> >= 9 blocks on 3 CTAs
    st = _cases.check_lnl(emu_lib, case)[2]
    assert 160 <= st["sweep_block_sites"] <= 352 and st["sweep_block_sites"] % 16 == 0
    # 6 internal nodes, the cherries folded into their parents' ops; one launch for the full passes + one, narrower, for the rest
    assert 3 <= st["sweep_ops"] < 6 and st["clv_launches"] in (1, 2)


def test_sweep_global_stack_levels(emu_lib, monkeypatch):
    """A balanced 64-leaf tree: 16 of the 32 cherries are folded into their parents' ops (the other cherry of each
    pair stays an op), the cherry-level pending sibling lives in the second register accumulator, the rest are stack
    levels; with little shared memory most spill to the global stack."""
    def balanced(lo, hi):
        if hi - lo == 1:
            return f"t{lo}:0.{1 + lo % 7}"
        mid = (lo + hi) // 2
        return f"({balanced(lo, mid)},{balanced(mid, hi)}):0.0{1 + (lo + hi) % 9}"

    tree = parse_newick(balanced(0, 64) + ";")
    rng = np.random.default_rng(3)
    lo = [f"t{i}" for i in range(64)]
    ft = workloads.flatten_fast(tree, lo)
    m = workloads.make_model("gtr_g4")
    pats = workloads.simulate_patterns(ft, m, 700, rng)
    case = workloads.Case(tree, ft, m, "gtr_g4", pats, np.ones(700, np.uint64), lo)
    a = _cases.check_lnl(emu_lib, case)
    monkeypatch.setenv("PHB_EMU_SMEM_BYTES", str(512 + 2 * 4736 + 2 * 32768))
    b = _cases.check_lnl(emu_lib, case)
    assert np.array_equal(a[1], b[1])
    assert a[2]["sweep_stack_levels"] == 3 and a[2]["sweep_smem_levels"] == 3 and a[2]["sweep_ops"] == 63 - 16
    assert b[2]["sweep_stack_levels"] == 3 and 1 <= b[2]["sweep_smem_levels"] < 3


def test_sweep_schedule_shapes(emu_lib):
    """Caterpillar (need 0), polytomies and unifurcations through the generic ops."""
    cat = "(((((a:0.1,b:0.2):0.05,c:0.1):0.02,d:0.3):0.01,e:0.2):0.03,f:0.1);"
    poly = "(((a:0.1,b:0.2,c:0.1)x:0.05,(d:0.3,e:0.2)y:0.01,((f:0.1,g:0.2)z:0.1)w:0.2,h:0.3,i:0.1)r:0.1,(j:0.2)u:0.3);"
    rng = np.random.default_rng(5)
    for nwk in (cat, poly):
        tree = parse_newick(nwk)
        names = [tree.nodes[i].taxa for i in tree.get_node_ids() if tree.is_leaf(i)]
        ft = workloads.flatten_fast(tree, names)
        for model in ("jc69", "gtr_i_g4"):
            m = workloads.make_model(model)
            pats = workloads.simulate_patterns(ft, m, 90, rng)
            pats[rng.random(pats.shape) < 0.2] = ord("N")
            case = workloads.Case(tree, ft, m, model, pats, np.ones(90, np.uint64), names)
            _cases.check_lnl(emu_lib, case)


def test_asr_small(emu_lib):
    for model in ("jc69", "gtr_g4", "gtr_i_g4"):
        case = workloads.make_case(10, 200, model=model, seed=6, ambiguity=0.1)
        assert _cases.check_asr(emu_lib, case, want_post=True) == 0


def test_asr_sharded_and_tiled(emu_lib):
    case = workloads.make_case(8, 1300, model="gtr_g4", seed=8)
    assert _cases.check_asr(emu_lib, case, devices=[0, 1], budget=7 * 4 * 4 * 8 * 512) == 0


def test_polytomy_and_unifurcation(emu_lib):
    tree = parse_newick("((A:0.1,B:0.2,C:0.3)X:0.05,(D:0.1)Y:0.2,E:0.3)R;")
    aln = RefAlignment.from_fasta_bytes(b">A\nACGTN\n>B\nACGAR\n>C\nAGGT-\n>D\nTCGTA\n>E\nACCTA\n")
    m = RefModel.gtr([.3, .2, .2, .3], [1, 2.5, 1, 1, 2.5, 1]).with_gamma(0.5, 4)
    pats, _s2p, mult, lo = aln.compress_columns()
    case = workloads.Case(tree, oracle.flatten(tree, lo), m, "gtr_g4", pats, mult, lo)
    _cases.check_lnl(emu_lib, case)
    assert _cases.check_asr(emu_lib, case, want_post=True) == 0


@pytest.mark.parametrize("k,depth", [(4, 110), (6, 110), (16, 110), (300, 1), (800, 1)])
def test_high_arity_with_deep_subtrees(emu_lib, k, depth):
    """a root with k children, each a 110-leaf caterpillar (branches 0.9), and stars of 300 /
    800 tips.  Every child CLV enters the product max-normalised (profile.rs:20-39), so the product leaves the
    f64 range only where the reference's does (the 800-star is -inf there too).  Default sweep, level kernels and
    the marginal-ASR up passes must all agree with the oracle."""
    case = workloads.polytomy_case(k, depth, 12, model="gtr_g4", seed=100 + k)
    tot, per, _ = _cases.check_lnl(emu_lib, case)
    tot_l, per_l, _ = _cases.check_lnl(emu_lib, case, flags=capi.FLAG_LEVEL_KERNELS)
    if (k, depth) == (800, 1):
        assert tot == -np.inf and tot_l == -np.inf  # the reference's own underflow (likelihood.rs:123-125)
    else:
        assert np.isfinite(tot) and np.max(np.abs(per - per_l)) < 1e-9
    assert _cases.check_asr(emu_lib, case) == 0
    assert _cases.check_asr(emu_lib, case, flags=capi.FLAG_LEVEL_KERNELS) == 0


def test_folded_cherries_match_explicit_cherries(emu_lib, monkeypatch):
    """The plain sweep folds every cherry under a binary parent into the parent's op (a 25-row table of
    P_cherry (tip_x (.) tip_y)); with PHB_SWEEP_NO_COLLAPSE the cherries stay ops.  Same per-pattern lnL to round-off,
    with ambiguity codes (the folded tables sum rows for the partial codes) and with gaps."""
    for model, amb in (("jc69", 0.0), ("gtr_i_g4", 0.25)):
        case = workloads.make_case(60, 80, model=model, seed=41, ambiguity=amb)
        a = _cases.check_lnl(emu_lib, case)
        monkeypatch.setenv("PHB_SWEEP_NO_COLLAPSE", "1")
        b = _cases.check_lnl(emu_lib, case)
        monkeypatch.delenv("PHB_SWEEP_NO_COLLAPSE")
        assert a[2]["sweep_ops"] < b[2]["sweep_ops"] == b[2]["n_internal"]
        assert np.max(np.abs(a[1] - b[1])) < 1e-10


def test_score_trees_spr_candidates(emu_lib):
    """phb_score_trees: the caller loop `t = base.clone(); t.spr(e1, e2); t.log_likelihood(..)` as one call"""
    _cases.check_score_trees(emu_lib, workloads.make_case(14, 60, model="gtr_g4", seed=71, ambiguity=0.05))


def test_schedule_reuse_across_branch_length_updates(emu_lib):
    """set_tree with the same topology and new branch lengths (a branch-length optimiser's loop) reuses the sweep
    schedule; the stream and the guard plan follow the new lengths: same result as a fresh context."""
    case = workloads.make_case(40, 120, model="gtr_i_g4", seed=51, ambiguity=0.05)
    ft = case.ft
    rng = np.random.default_rng(52)
    with capi.Context(library=emu_lib) as ctx:
        workloads.load_into(ctx, case)
        first = ctx.log_likelihood()
        for _ in range(3):
            ft["blen"] = np.clip(rng.exponential(0.3, ft["n_nodes"]), 1e-6, 5.0).astype(np.float32).astype(np.float64)
            ctx.set_tree(ft["postorder"], ft["child_off"], ft["children"], ft["blen"], ft["leaf_row"], ft["root"])
            tot, per = ctx.log_likelihood(per_pattern=True)
            with capi.Context(library=emu_lib) as fresh:
                workloads.load_into(fresh, case)
                assert fresh.log_likelihood() == tot
            ref_tot, ref_per = oracle.log_likelihood_flat(ft, case.model, case.patterns, case.mult, mode=0, n_threads=4)
            _cases.assert_lnl_close(tot, per, ref_tot, ref_per)
        assert tot != first


def test_guard_plan_matches_guard_everywhere(emu_lib, monkeypatch):
    """The host places the underflow guard only where its bound on the accumulated shrink needs it
    (sweep_sched.h plan_guards).  Rescaling is an exact power of two, so the placement changes no CLV bit: the
    per-pattern lnL agrees with the guard-after-every-node run to round-off of the final logarithm."""
    case = workloads.make_case(700, 24, model="gtr_i_g4", seed=31)
    a = _cases.check_lnl(emu_lib, case)
    assert 0 < a[2]["sweep_guards"] < 0.25 * a[2]["n_internal"] and a[2]["sweep_normouts"] == 0
    monkeypatch.setenv("PHB_SWEEP_GUARD_ALL", "1")
    b = _cases.check_lnl(emu_lib, case)
    assert b[2]["sweep_guards"] == b[2]["sweep_ops"] < b[2]["n_internal"]  # one per op; folded cherries have none
    assert np.max(np.abs(a[1] - b[1])) < 1e-9 and a[1].min() < -300


def test_guard_plan_extreme_branch_lengths(emu_lib):
    """Branches of 1e-30 .. 1e-3 under a very skewed Gamma (alpha = 0.05: the slowest category's rate is ~1e-12) make
    single edges able to shrink a CLV by hundreds of bits: their children are normalised on completion."""
    rng = np.random.default_rng(77)
    case = workloads.make_case(120, 40, model="gtr_g4", seed=32)
    ft = case.ft
    ft["blen"] = np.where(rng.random(ft["n_nodes"]) < 0.5, 10.0 ** rng.uniform(-30, -3, ft["n_nodes"]), ft["blen"]).astype(np.float32).astype(np.float64)
    case.model = RefModel.gtr(workloads.PI, workloads.GTR_RATES).with_gamma(0.05, 4)
    rngp = np.random.default_rng(78)
    case.patterns = workloads.simulate_patterns(dict(ft, preorder=_preorder(ft)), case.model, 40, rngp)
    case.patterns[rngp.random(case.patterns.shape) < 0.3] = ord("A")  # conflicts across near-zero branches
    tot, per, st = _cases.check_lnl(emu_lib, case)
    assert st["sweep_normouts"] > 0
    _cases.check_lnl(emu_lib, case, flags=capi.FLAG_LEVEL_KERNELS)


def _preorder(ft):
    pre, st = [], [int(ft["root"])]
    co, ch = ft["child_off"], ft["children"]
    while st:
        v = st.pop()
        pre.append(v)
        st.extend(reversed(ch[co[v]:co[v + 1]].tolist()))
    return np.asarray(pre, np.int32)


def test_reference_fixture_and_internal_row(emu_lib):
    # integration_test.rs:14-17 + :663 — row C names an internal node: ignored by pruning, never validated
    tree = parse_newick("((A:0.601,B:0.301)C:0.100,D:0.200)E:0.001;")
    aln = RefAlignment.from_fasta_bytes(b">A\nACGTACGT\n>B\nAGGTAGGT\n>C\nAC?TAC!T\n>D\nTTTTTTTT\n")
    pats, _s2p, mult, lo = aln.compress_columns()
    m = RefModel.jukes_cantor()
    case = workloads.Case(tree, oracle.flatten(tree, lo), m, "jc69", pats, mult, lo)
    _cases.check_lnl(emu_lib, case)


def test_errors(emu_lib):
    case = workloads.make_case(6, 20, model="jc69", seed=2)
    with capi.Context(library=emu_lib) as ctx:
        with pytest.raises(capi.PhbError) as e:
            ctx.log_likelihood()
        assert e.value.variant == "State"
        workloads.load_into(ctx, case)
        ft = case.ft
        lr = ft["leaf_row"].copy()
        leaf = int(np.where(lr >= 0)[0][0])
        lr[leaf] = -1
        ctx.set_tree(ft["postorder"], ft["child_off"], ft["children"], ft["blen"], lr, ft["root"])
        with pytest.raises(capi.PhbError) as e:
            ctx.log_likelihood()
        assert e.value.variant == "InvalidAlignment"
        bad = case.patterns.copy()
        bad[3, ft["leaf_row"][leaf]] = ord("X")
        ctx.set_alignment(bad)
        ctx.set_tree(ft["postorder"], ft["child_off"], ft["children"], ft["blen"], ft["leaf_row"], ft["root"])
        with pytest.raises(capi.PhbError) as e:
            ctx.log_likelihood()
        assert e.value.variant == "AlphabetMismatch"
        with pytest.raises(capi.PhbError) as e:  # node before its child
            ctx.set_tree(ft["postorder"][::-1], ft["child_off"], ft["children"], ft["blen"], ft["leaf_row"], ft["root"])
        assert e.value.variant == "NumericalInstability"


def test_error_precedence_follows_the_reference(emu_lib):
    """likelihood.rs:97-107 walks the leaves in post-order on pattern 0: a leaf without a row and a leaf with an
    unaccepted byte IN PATTERN 0 are reported in the order they are met; an unaccepted byte in a later pattern
    can only surface after pattern 0 went through (so a missing row wins)."""
    case = workloads.make_case(6, 20, model="jc69", seed=2)
    ft = case.ft
    leaves = [int(v) for v in ft["postorder"] if ft["leaf_row"][v] >= 0]
    first, last = leaves[0], leaves[-1]
    lr = ft["leaf_row"].copy()
    lr[last] = -1  # the LAST leaf of the post-order has no row
    for pat, want in ((0, "AlphabetMismatch"), (3, "InvalidAlignment")):
        bad = case.patterns.copy()
        bad[pat, ft["leaf_row"][first]] = ord("X")  # the FIRST leaf holds an unaccepted byte in pattern `pat`
        with capi.Context(library=emu_lib) as ctx:
            workloads.load_into(ctx, case)
            ctx.set_alignment(bad)
            ctx.set_tree(ft["postorder"], ft["child_off"], ft["children"], ft["blen"], lr, ft["root"])
            with pytest.raises(capi.PhbError) as e:
                ctx.log_likelihood()
            assert e.value.variant == want


def test_empty_alignment_is_zero(emu_lib):
    case = workloads.make_case(6, 5, model="jc69", seed=2)
    case.patterns = case.patterns[:0]
    case.mult = case.mult[:0]
    with capi.Context(library=emu_lib) as ctx:
        workloads.load_into(ctx, case)
        assert ctx.log_likelihood() == 0.0


def test_zero_branch_conflict_and_all_gap(emu_lib):
    tree = parse_newick("((A:0,B:0)C:0,D:0)E:0;")
    aln = RefAlignment.from_fasta_bytes(b">A\nAC-\n>B\nCC-\n>D\nGC-\n")
    pats, _s2p, mult, lo = aln.compress_columns()
    case = workloads.Case(tree, oracle.flatten(tree, lo), RefModel.jukes_cantor(), "jc69", pats, mult, lo)
    tot, per, _ = _cases.check_lnl(emu_lib, case)
    assert tot == -np.inf and np.isfinite(per[1]) and abs(per[2]) < 1e-13
    assert _cases.check_asr(emu_lib, case) == 0


def test_multiplicity_weights(emu_lib):
    case = workloads.make_case(8, 400, model="gtr_g4", seed=9, compress=True)
    assert case.mult.max() > 1
    _cases.check_lnl(emu_lib, case)


def test_score_nni_matches_rebuilt_trees(emu_lib):
    """Path-only rescoring (csrc/nni.cuh) == the oracle on every rebuilt neighbour, incl. leaf/root moves (NaN)."""
    case = workloads.make_case(10, 150, model="gtr_g4", seed=12)
    got, st = _cases.check_nni(emu_lib, case)
    assert st["nni_tiles"] == 1 and st["nni_path_steps"] > 0
    full, st_full = _cases.check_nni(emu_lib, case, flags=capi.FLAG_NNI_FULL, expect_path=False)
    walk, _ = _cases.check_nni(emu_lib, case, flags=capi.FLAG_NNI_WALK)  # v -> root walks instead of upper partials
    ok = np.isfinite(got)
    assert np.array_equal(ok, np.isfinite(full)) and np.array_equal(ok, np.isfinite(walk))
    assert np.max(np.abs(got[ok] - full[ok]) / np.abs(full[ok])) < 1e-12
    assert np.max(np.abs(got[ok] - walk[ok]) / np.abs(walk[ok])) < 1e-12
    # repeated and half-asked edges: every request gets its own answer
    internal = [v for v in range(case.ft["n_nodes"]) if case.ft["child_off"][v + 1] > case.ft["child_off"][v] and v != case.ft["root"]]
    a, b = internal[0], internal[-1]
    _cases.check_nni(emu_lib, case, moves=[(a, 1), (a, 1), (b, 0), (a, 0), (a, 1)])


@pytest.mark.parametrize("model,amb", [("jc69", 0.0), ("gtr_i_g4", 0.1)])
def test_score_nni_models_ambiguity_tiling(emu_lib, model, amb):
    case = workloads.make_case(14, 1300, model=model, seed=31, ambiguity=amb, compress=(amb == 0.0))
    n_int = 13
    n_clv = 1 if model == "jc69" else 4
    budget = n_int * n_clv * 72 * 512 + 1  # one 512-pattern tile of cached CLVs + upper partials + exponents
    got, st = _cases.check_nni(emu_lib, case, budget=budget)
    assert st["nni_tiles"] == (len(case.patterns) + 511) // 512
    one, _ = _cases.check_nni(emu_lib, case)
    ok = np.isfinite(got)
    assert np.max(np.abs(got[ok] - one[ok]) / np.abs(one[ok])) < 1e-13
    sharded, _ = _cases.check_nni(emu_lib, case, devices=[0, 0, 0])  # three pattern shards, host sum per neighbour
    assert np.max(np.abs(sharded[ok] - one[ok]) / np.abs(one[ok])) < 1e-13


def test_score_nni_deep_tree_rescaling(emu_lib):
    """600 taxa: subtree likelihoods fall below 2^-256, so the cached subtree exponents and the walk's own
    rescales both take part."""
    case = workloads.make_case(600, 24, model="gtr_g4", seed=33)
    ft = case.ft
    rng = np.random.default_rng(5)
    internal = [v for v in range(ft["n_nodes"]) if ft["child_off"][v + 1] > ft["child_off"][v]]
    moves = [(int(v), int(lc)) for v in rng.choice(internal, 40, replace=False) for lc in (0, 1)]
    _cases.check_nni(emu_lib, case, moves=moves)
    _cases.check_nni(emu_lib, case, moves=moves[:20], flags=capi.FLAG_NNI_WALK)


def test_score_nni_polytomy_falls_back(emu_lib):
    tree = parse_newick("((A:0.1,B:0.2,C:0.05)X:0.1,(D:0.3,E:0.1)Y:0.2)R;")
    aln = RefAlignment.from_fasta_bytes(b">A\nACGTAC\n>B\nACGTTC\n>C\nAGGTAC\n>D\nACCTAC\n>E\nTCGTAC\n")
    m = RefModel.gtr([.3, .2, .2, .3], [1, 2.5, 1, 1, 2.5, 1]).with_gamma(0.5, 4)
    pats, _s2p, mult, lo = aln.compress_columns()
    case = workloads.Case(tree, oracle.flatten(tree, lo), m, "gtr_g4", pats, mult, lo)
    ft = case.ft
    # moves on binary nodes only: the full-rescoring fallback treats nodes of other arity as invalid too
    _cases.check_nni(emu_lib, case, expect_path=False)


@pytest.mark.parametrize("model", ["jc69", "gtr_g4", "gtr_i_g4"])
def test_joint_asr_models(emu_lib, model):
    case = workloads.make_case(12, 300, model=model, seed=11, ambiguity=0.1)
    _cases.check_joint(emu_lib, case)


def test_joint_asr_tiling_sharding_and_shapes(emu_lib):
    case = workloads.make_case(9, 1500, model="gtr_g4", seed=4, compress=True)
    _cases.check_joint(emu_lib, case)
    _, st = _cases.check_joint(emu_lib, case, budget=8 * 4 * 4 * 8 * 512)
    assert st["n_tiles"] >= 2
    _cases.check_joint(emu_lib, case, devices=[0, 0, 0])
    tree = parse_newick("((A:0.1,B:0.2,C:0.3)X:0.05,(D:0.1)Y:0.2,E:0.3)R;")  # polytomy + unifurcation
    aln = RefAlignment.from_fasta_bytes(b">A\nACGTN\n>B\nACGAR\n>C\nAGGT-\n>D\nTCGTA\n>E\nACCTA\n")
    m = RefModel.gtr([.3, .2, .2, .3], [1, 2.5, 1, 1, 2.5, 1]).with_gamma(0.5, 4)
    pats, _s2p, mult, lo = aln.compress_columns()
    _cases.check_joint(emu_lib, workloads.Case(tree, oracle.flatten(tree, lo), m, "gtr_g4", pats, mult, lo))
    zero = parse_newick("((A:0,B:0)C:0,D:0)E:0;")  # conflicting leaves at distance 0: -inf, state 0 everywhere
    zaln = RefAlignment.from_fasta_bytes(b">A\nACGT\n>B\nAGGT\n>D\nTTTT\n")
    pats, _s2p, mult, lo = zaln.compress_columns()
    zc = workloads.Case(zero, oracle.flatten(zero, lo), RefModel.jukes_cantor(), "jc69", pats, mult, lo)
    assert _cases.check_joint(emu_lib, zc)[0] == 0


def test_compress_columns_matches_reference_walk(emu_lib):
    rng = np.random.default_rng(77)
    # few distinct columns, many repeats; T vs U and an invalid byte are distinct keys (raw bytes are compared)
    base = rng.choice(np.frombuffer(b"ACGTUN-R?", np.uint8), size=(7, 40))
    rows = base[:, rng.integers(0, 40, size=1000)]
    n = _cases.check_compress(emu_lib, rows)
    assert n <= 40
    _cases.check_compress(emu_lib, rows, devices=[0, 0, 0])             # shards of the unique patterns
    with capi.Context(library=emu_lib) as ctx:                          # one pointer per row (Alignment.seqs layout)
        a = ctx.compress_columns(rows)
        b = ctx.compress_rows([bytes(r) for r in rows])
        assert a[0] == b[0] and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
    _cases.check_compress(emu_lib, rng.choice(np.frombuffer(b"ACGT", np.uint8), size=(30, 333)))  # (almost) all unique
    _cases.check_compress(emu_lib, np.full((3, 17), ord("A"), np.uint8))  # one pattern
    _cases.check_compress(emu_lib, rng.choice(np.frombuffer(b"AC", np.uint8), size=(1, 9)))       # a single row
    with capi.Context(library=emu_lib) as ctx:
        n, s2p, mult = ctx.compress_columns(np.zeros((4, 0), np.uint8))
        assert n == 0 and len(s2p) == 0 and len(mult) == 0


def test_compress_then_score_equals_host_compression(emu_lib):
    """compress on the device, then evaluate: same lnL as handing over the host-compressed patterns."""
    case = workloads.make_case(9, 700, model="gtr_i_g4", seed=21)
    rows = np.ascontiguousarray(case.patterns.T)  # [row][site]: uncompressed simulated sites
    host = workloads.make_case(9, 700, model="gtr_i_g4", seed=21, compress=True)
    ref, _ = oracle.log_likelihood_flat(host.ft, host.model, host.patterns, host.mult)
    with capi.Context(library=emu_lib) as ctx:
        workloads.load_into(ctx, case)
        n, s2p, mult = ctx.compress_columns(rows)
        assert n == len(host.patterns) and np.array_equal(mult, host.mult)
        tot, per = ctx.log_likelihood(per_pattern=True)
        assert abs(tot - ref) <= 1e-9 * abs(ref)
        tot_a, states, _ = ctx.marginal_asr()
        assert states.shape == (case.ft["n_nodes"], n)


def test_asr_up_pass_sweep_equals_level_kernels(emu_lib):
    """The marginal-ASR up pass runs as the storing sweep by default; PHB_FLAG_LEVEL_KERNELS selects the level
    kernels.  Same states, same posteriors (to round-off), both against the oracle."""
    case = workloads.make_case(13, 700, model="gtr_i_g4", seed=17, ambiguity=0.1)
    assert _cases.check_asr(emu_lib, case, want_post=True) == 0
    assert _cases.check_asr(emu_lib, case, want_post=True, flags=capi.FLAG_LEVEL_KERNELS) == 0
    a = _cases.run_asr(emu_lib, case, want_post=True)
    b = _cases.run_asr(emu_lib, case, want_post=True, flags=capi.FLAG_LEVEL_KERNELS)
    assert np.array_equal(a[2], b[2]) and np.max(np.abs(a[3] - b[3])) < 1e-12 and abs(a[0] - b[0]) < 1e-9 * abs(b[0])
    with capi.Context(library=emu_lib) as ctx:
        workloads.load_into(ctx, case)
        ctx.marginal_asr()
        assert ctx.stats()["sweep_ops"] > 0
    deep = workloads.make_case(500, 40, model="gtr_g4", seed=18)  # exponents: stored CLVs are rescaled mid-tree
    assert _cases.check_asr(emu_lib, deep) == 0


def test_simulated_alignment_roundtrip(emu_lib):
    case = workloads.make_case(7, 10, model="gtr_g4", seed=13)
    with capi.Context(library=emu_lib) as ctx:
        workloads.load_into(ctx, case)
        ctx.simulate_alignment(700, seed=99)
        pats = ctx.get_alignment()
        assert pats.shape == (700, 7) and set(np.unique(pats)) <= set(b"ACGT")
        tot, per = ctx.log_likelihood(per_pattern=True)
        # shards of one logical alignment generated independently agree with the whole
        ctx.simulate_alignment(300, global_offset=400, seed=99)
        assert np.array_equal(ctx.get_alignment(), pats[400:])
    ft = dict(case.ft)
    leaves = np.where(ft["child_off"][1:] == ft["child_off"][:-1])[0]
    lr = np.full(ft["n_nodes"], -1, np.int32)
    lr[leaves] = np.arange(len(leaves))
    ft["leaf_row"] = lr
    ref, ref_per = oracle.log_likelihood_flat(ft, case.model, pats)
    _cases.assert_lnl_close(tot, per, ref, ref_per)


def test_fused_host_call_matches_two_call_path(emu_lib):
    """phb_log_likelihood_host (chunked copy -> pack -> sweep pipeline) == set_alignment + log_likelihood, bit for bit,
    over several chunks, with multiplicities, on one and on three virtual devices, and for the row-major fallback."""
    from phylo_rs_b200 import capi

    case = workloads.make_case(9, 7000, model="gtr_i_g4", seed=21, ambiguity=0.03)
    mult = (np.arange(7000) % 5 + 1).astype(np.uint64)
    case.mult = mult
    ref_tot, ref_per, _ = _cases.check_lnl(emu_lib, case)
    for devices in (None, [0, 0, 0]):
        with capi.Context(devices=devices, library=emu_lib) as ctx:
            m, ft = case.model, case.ft
            ctx.set_model(m.eigenvalues, m.eigenvectors, m.pi, m.sqrt_pi, m.inv_sqrt_pi, m.rates, m.weights)
            ctx.set_tree(ft["postorder"], ft["child_off"], ft["children"], ft["blen"], ft["leaf_row"], ft["root"])
            tot, per = ctx.log_likelihood_host(case.patterns, multiplicity=mult, per_pattern=True)
            assert np.array_equal(per, ref_per)
            if devices is None:
                assert tot == ref_tot
            else:
                assert abs(tot - ref_tot) <= 1e-12 * abs(ref_tot)
            # the alignment stays resident: the plain entry point gives the same number again (its site blocks may be
            # cut differently from the streamed chunks', so the block partial sums add up in another order)
            assert abs(ctx.log_likelihood() - tot) <= 1e-13 * abs(tot)
            tot_rm = ctx.log_likelihood_host(np.ascontiguousarray(case.patterns.T), layout=capi.LAYOUT_ROW_MAJOR,
                                             multiplicity=mult)
            assert abs(tot_rm - tot) <= 1e-13 * abs(tot)
            # errors: an invalid byte in a leaf row, and a leaf without a row
            bad = case.patterns.copy()
            bad[6500, 3] = ord("?")
            with pytest.raises(capi.PhbError) as e:
                ctx.log_likelihood_host(bad)
            assert e.value.variant == "AlphabetMismatch"
            with pytest.raises(capi.PhbError) as e:
                ctx.log_likelihood_host(case.patterns[:, :-1].copy())
            assert e.value.variant == "InvalidAlignment"
            assert ctx.log_likelihood_host(case.patterns, multiplicity=mult) == tot


def test_nonbinary_trees_all_engines(emu_lib):
    """polytomies of arity 3..6, unifurcations, nested polytomies: lnL (sweep and level kernels), marginal and joint ASR"""
    _cases.check_nonbinary(emu_lib)


def test_score_nni_cache_variants_and_repeated_calls(emu_lib):
    case = workloads.make_case(40, 300, model="gtr_i_g4", seed=77, ambiguity=0.03)
    _cases.check_nni_cache_variants(emu_lib, case)


def test_random_trees_all_engines(emu_lib):
    """random shapes (arities 1..5), zero / missing / tiny / long branches, random models, ambiguity, multiplicities"""
    _cases.check_random_trees(emu_lib, range(100, 112))


def test_host_call_from_pageable_memory_bounce_path():
    """phb_log_likelihood_host / phb_set_alignment handed pageable memory: slices go through two pinned bounce buffers filled
    by the copy pool.  A fresh process with tiny bounce buffers (many slices per chunk) and three copy threads must give the
    resident path's per-pattern values."""
    import os
    import subprocess
    import sys
    import textwrap

    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = textwrap.dedent("""
        import sys, numpy as np
        sys.path.insert(0, %r); sys.path.insert(0, %r)
        import __graft_entry__ as g
        from oracle import workloads
        from phylo_rs_b200 import capi
        lib = capi.Library(g.build_emu())
        case = workloads.make_case(23, 5000, model="gtr_i_g4", seed=3, ambiguity=0.05)
        m, ft = case.model, case.ft
        with capi.Context(devices=[0, 0], library=lib) as ctx:
            ctx.set_model(m.eigenvalues, m.eigenvectors, m.pi, m.sqrt_pi, m.inv_sqrt_pi, m.rates, m.weights)
            ctx.set_tree(ft["postorder"], ft["child_off"], ft["children"], ft["blen"], ft["leaf_row"], ft["root"])
            tot_h, per_h = ctx.log_likelihood_host(case.patterns, multiplicity=case.mult, per_pattern=True)
            ctx.set_alignment(case.patterns, multiplicity=case.mult)
            tot_r, per_r = ctx.log_likelihood(per_pattern=True)
            back = ctx.get_alignment()
        assert np.array_equal(per_h, per_r) and abs(tot_h - tot_r) <= 1e-12 * abs(tot_r)
        canon = case.patterns.copy()
        canon[canon == ord("U")] = ord("T")  # get_alignment returns the canonical letter of a code
        assert np.array_equal(back, canon)
        print("bounce ok")
    """) % (ROOT, os.path.join(ROOT, "tests"))
    env = dict(os.environ, PHB_FORCE_BOUNCE="1", PHB_BOUNCE_BYTES="8192", PHB_COPY_THREADS="3")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "bounce ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
