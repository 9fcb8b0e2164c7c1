"""Parity of the real sm_100a library (through its C ABI) against the CPU oracle.  `-m gpu` only.

Bars (BASELINE.json): total lnL within 1e-9 relative, per-pattern lnL within 1e-8 absolute (f64),
marginal-ASR argmax states bit-exact."""
import numpy as np
import pytest

import oracle
from oracle import RefAlignment, RefModel, parse_newick, workloads
from phylo_rs_b200 import capi

import _cases

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("model", ["jc69", "gtr", "gtr_g4", "gtr_i_g4", "gtr_i", "hky_g4"])
def test_lnl_models(gpu_lib, model):
    case = workloads.make_case(32, 2000, model=model, seed=21, ambiguity=0.05)
    _cases.check_lnl(gpu_lib, case)


def test_config1_jc69_64x10k(gpu_lib):
    """BASELINE.json configs[0]: JC69, 64-taxon Yule tree x 10k sites (compressed patterns + multiplicities)."""
    case = workloads.make_case(64, 10_000, model="jc69", seed=0x5EED0001, compress=True)
    _cases.check_lnl(gpu_lib, case)


def test_gtr_g4_1000_taxa_sample(gpu_lib):
    """configs[1] shape (1,000 taxa, GTR+G4) on a pattern sample the oracle finishes in seconds."""
    case = workloads.make_case(1000, 1536, model="gtr_g4", seed=0x5EED0002)
    _cases.check_lnl(gpu_lib, case)


def test_gtr_i_g4_deep_tree_scaling(gpu_lib):
    """2,500 taxa: per-site |lnL| ~ 1e4 and CLVs underflow without the exact power-of-two rescaling."""
    case = workloads.make_case(2500, 384, model="gtr_i_g4", seed=0x5EED0003)
    tot, per, _ = _cases.check_lnl(gpu_lib, case)
    assert per.min() < -2000


@pytest.mark.parametrize("model", ["jc69", "gtr_g4", "gtr_i_g4"])
def test_level_kernels_flag(gpu_lib, model):
    """The HBM-resident level-batched kernels (kept for the ASR up pass) stay parity-green for lnL too."""
    case = workloads.make_case(200, 3000, model=model, seed=22, ambiguity=0.05)
    _cases.check_lnl(gpu_lib, case, flags=capi.FLAG_LEVEL_KERNELS)


def test_sweep_shapes(gpu_lib):
    """Balanced tree (deepest stack), caterpillar (no stack), polytomies/unifurcations (generic ops)."""
    def balanced(lo, hi):
        if hi - lo == 1:
            return f"t{lo}:0.{1 + lo % 7}"
        mid = (lo + hi) // 2
        return f"({balanced(lo, mid)},{balanced(mid, hi)}):0.0{1 + (lo + hi) % 9}"

    def cater(n):
        s = "(t0:0.1,t1:0.2)"
        for i in range(2, n):
            s = f"({s}:0.0{1 + i % 9},t{i}:0.1{i % 7})"
        return s

    poly = "(((a:0.1,b:0.2,c:0.1)x:0.05,(d:0.3,e:0.2)y:0.01,((f:0.1,g:0.2)z:0.1)w:0.2,h:0.3,i:0.1)r:0.1,(j:0.2)u:0.3);"
    rng = np.random.default_rng(5)
    # balanced 1024: half the cherries are folded into their parents' ops, the cherry-level pending sibling stays in
    # the register accumulator, the other 7 pending siblings of the deepest path are stack levels
    for nwk, levels in ((balanced(0, 1024) + ";", 7), (cater(300) + ";", 0), (poly, None)):
        tree = parse_newick(nwk)
        names = [tree.nodes[i].taxa for i in tree.get_node_ids() if tree.is_leaf(i)]
        ft = workloads.flatten_fast(tree, names)
        for model in ("jc69", "gtr_i_g4"):
            m = workloads.make_model(model)
            pats = workloads.simulate_patterns(ft, m, 1500, rng)
            pats[rng.random(pats.shape) < 0.1] = ord("N")
            case = workloads.Case(tree, ft, m, model, pats, np.ones(1500, np.uint64), names)
            st = _cases.check_lnl(gpu_lib, case)[2]
            if levels is not None:
                assert st["sweep_stack_levels"] == levels


@pytest.mark.parametrize("k,depth", [(4, 110), (6, 110), (8, 110), (16, 110), (300, 1), (800, 1)])
def test_high_arity_with_deep_subtrees(gpu_lib, k, depth):
    """A root with k children, each a 110-leaf caterpillar (branches 0.9, random ACGT), and stars of 300 / 800 tips
    (likelihood.rs:113-131 multiplies any number of max-normalised child contributions, profile.rs:20-39).  The
    default fused sweep, the level kernels and both marginal-ASR up passes against the oracle; the 800-star
    underflows in the reference itself (-inf) and must do so here on every path."""
    case = workloads.polytomy_case(k, depth, 48, model="gtr_g4", seed=100 + k)
    tot, per, _ = _cases.check_lnl(gpu_lib, case)
    tot_l, per_l, _ = _cases.check_lnl(gpu_lib, case, flags=capi.FLAG_LEVEL_KERNELS)
    if (k, depth) == (800, 1):
        assert tot == -np.inf and tot_l == -np.inf
    else:
        assert np.isfinite(tot) and np.max(np.abs(per - per_l)) < 1e-9
    assert _cases.check_asr(gpu_lib, case) == 0
    assert _cases.check_asr(gpu_lib, case, flags=capi.FLAG_LEVEL_KERNELS) == 0


def test_literal_invariant_flag(gpu_lib):
    case = workloads.make_case(40, 3000, model="gtr_i_g4", seed=23, ambiguity=0.1)
    a = _cases.check_lnl(gpu_lib, case)[0]
    b = _cases.check_lnl(gpu_lib, case, flags=capi.FLAG_LITERAL_INVARIANT)[0]
    assert abs(a - b) <= 1e-10 * abs(a)


def test_all_17_codes(gpu_lib):
    case = workloads.make_case(24, 4000, model="gtr_g4", seed=24, ambiguity=0.5)
    assert set(np.unique(case.patterns)) >= set(b"ACGTURYSWKMBDHVN-")
    _cases.check_lnl(gpu_lib, case)
    assert _cases.check_asr(gpu_lib, case) == 0


def test_tiling_and_sharding_invariance(gpu_lib):
    case = workloads.make_case(50, 20_000, model="gtr_g4", seed=25)
    full = _cases.check_lnl(gpu_lib, case, flags=capi.FLAG_LEVEL_KERNELS)
    tiled = _cases.check_lnl(gpu_lib, case, budget=49 * 4 * 4 * 8 * 4096, flags=capi.FLAG_LEVEL_KERNELS)
    assert tiled[2]["n_tiles"] >= 4
    assert np.array_equal(full[1], tiled[1])
    sweep = _cases.check_lnl(gpu_lib, case)
    assert 25 <= sweep[2]["sweep_ops"] < 49 and np.max(np.abs(sweep[1] - full[1])) < 1e-10  # 49 nodes, cherries folded
    split = _cases.check_lnl(gpu_lib, case, devices=[0, 0, 0])  # three virtual ranks on one device
    assert np.array_equal(sweep[1], split[1])


@pytest.mark.parametrize("model", ["jc69", "gtr_g4", "gtr_i_g4"])
def test_asr_models(gpu_lib, model):
    case = workloads.make_case(40, 1500, model=model, seed=26, ambiguity=0.05)
    mism = _cases.check_asr(gpu_lib, case, want_post=True)
    # JC69 is exactly symmetric: ambiguity codes create EXACT posterior ties whose winner is decided by
    # the last bit of round-off (SURVEY.md §7 hard part 2); check_asr has already asserted that every
    # mismatch sits on such a tie (oracle top-2 gap < 1e-10).  Asymmetric models must match bit for bit.
    if model != "jc69":
        assert mism == 0


def test_asr_500_taxa_sample(gpu_lib):
    case = workloads.make_case(500, 512, model="gtr_g4", seed=0x5EED0004)
    assert _cases.check_asr(gpu_lib, case) == 0


def test_asr_up_pass_sweep_equals_level_kernels(gpu_lib):
    case = workloads.make_case(60, 5000, model="gtr_i_g4", seed=37, ambiguity=0.05)
    a = _cases.run_asr(gpu_lib, case, want_post=True)
    b = _cases.run_asr(gpu_lib, case, want_post=True, flags=capi.FLAG_LEVEL_KERNELS)
    assert np.array_equal(a[2], b[2]) and np.max(np.abs(a[3] - b[3])) < 1e-12 and abs(a[0] - b[0]) < 1e-9 * abs(b[0])
    assert _cases.check_asr(gpu_lib, case, want_post=True, flags=capi.FLAG_LEVEL_KERNELS) == 0
    deep = workloads.make_case(1500, 700, model="gtr_g4", seed=38)
    assert _cases.check_asr(gpu_lib, deep) == 0


def test_asr_tiled(gpu_lib):
    case = workloads.make_case(30, 6000, model="gtr_g4", seed=27)
    assert _cases.check_asr(gpu_lib, case, budget=29 * 4 * 4 * 8 * 1024) == 0


def test_polytomy_unifurcation_internal_row(gpu_lib):
    tree = parse_newick("((A:0.1,B:0.2,C:0.3)X:0.05,(D:0.1)Y:0.2,E:0.3)R;")
    aln = RefAlignment.from_fasta_bytes(b">A\nACGTN\n>B\nACGAR\n>C\nAGGT-\n>D\nTCGTA\n>E\nACCTA\n>X\n?????\n")
    m = RefModel.gtr([.3, .2, .2, .3], [1, 2.5, 1, 1, 2.5, 1]).with_gamma(0.5, 4)
    pats, _s2p, mult, lo = aln.compress_columns()
    case = workloads.Case(tree, oracle.flatten(tree, lo), m, "gtr_g4", pats, mult, lo)
    _cases.check_lnl(gpu_lib, case)
    assert _cases.check_asr(gpu_lib, case, want_post=True) == 0


def test_reference_fixture_kats(gpu_lib):
    """SURVEY.md §8c known answers through the CUDA path."""
    tree = parse_newick("((A:0.601,B:0.301)C:0.100,D:0.200)E:0.001;")
    aln = RefAlignment.from_fasta_bytes(
        b">A\n" + b"ACGT" * 4 + b"\n>B\n" + b"AGGT" * 4 + b"\n>C\n" + b"ACGT" * 4 + b"\n>D\n" + b"T" * 16 + b"\n")
    pats, _s2p, mult, lo = aln.compress_columns()
    ft = oracle.flatten(tree, lo)
    g4 = RefModel.gtr([.3, .2, .2, .3], [1, 2.5, 1, 1, 2.5, 1]).with_gamma(0.5, 4)
    for m, want in ((RefModel.jukes_cantor(), -66.3080496020529), (RefModel.hky85([.3, .2, .2, .3], 2.5), -68.11049472687087),
                    (g4, -69.66427000171137), (g4.with_invariant(0.2), -71.7982617885503)):
        case = workloads.Case(tree, ft, m, "", pats, mult, lo)
        tot, _per, _ = _cases.check_lnl(gpu_lib, case)
        assert tot == pytest.approx(want, abs=1e-10)


def test_zero_branch_conflict_and_all_gap(gpu_lib):
    tree = parse_newick("((A:0,B:0)C:0,D:0)E:0;")
    aln = RefAlignment.from_fasta_bytes(b">A\nAC-\n>B\nCC-\n>D\nGC-\n")
    pats, _s2p, mult, lo = aln.compress_columns()
    case = workloads.Case(tree, oracle.flatten(tree, lo), RefModel.jukes_cantor(), "jc69", pats, mult, lo)
    tot, per, _ = _cases.check_lnl(gpu_lib, case)
    assert tot == -np.inf and abs(per[2]) < 1e-13
    assert _cases.check_asr(gpu_lib, case) == 0


def test_errors_and_empty(gpu_lib):
    case = workloads.make_case(6, 20, model="jc69", seed=2)
    with capi.Context(library=gpu_lib) as ctx:
        with pytest.raises(capi.PhbError):
            ctx.log_likelihood()
        workloads.load_into(ctx, case)
        ft = case.ft
        bad = case.patterns.copy()
        leaf = int(np.where(ft["leaf_row"] >= 0)[0][0])
        bad[3, ft["leaf_row"][leaf]] = ord("X")
        ctx.set_alignment(bad)
        with pytest.raises(capi.PhbError) as e:
            ctx.log_likelihood()
        assert e.value.variant == "AlphabetMismatch"
        ctx.set_alignment(case.patterns[:0])
        assert ctx.log_likelihood() == 0.0


@pytest.mark.parametrize("model", ["jc69", "hky_g4", "gtr_g4", "gtr_i_g4"])
def test_joint_asr_models(gpu_lib, model):
    case = workloads.make_case(24, 4000, model=model, seed=51, ambiguity=0.08)
    _cases.check_joint(gpu_lib, case)


def test_joint_asr_500_taxa_tiled(gpu_lib):
    case = workloads.make_case(500, 1200, model="gtr_g4", seed=52)
    _cases.check_joint(gpu_lib, case)
    _, st = _cases.check_joint(gpu_lib, case, budget=499 * 4 * 32 * 512 + 1)
    assert st["n_tiles"] == 3


def test_joint_asr_edge_cases(gpu_lib):
    tree = parse_newick("((A:0.1,B:0.2,C:0.3)X:0.05,(D:0.1)Y:0.2,E:0.3)R;")
    aln = RefAlignment.from_fasta_bytes(b">A\nACGTN\n>B\nACGAR\n>C\nAGGT-\n>D\nTCGTA\n>E\nACCTA\n")
    m = RefModel.gtr([.3, .2, .2, .3], [1, 2.5, 1, 1, 2.5, 1]).with_gamma(0.5, 4)
    pats, _s2p, mult, lo = aln.compress_columns()
    _cases.check_joint(gpu_lib, workloads.Case(tree, oracle.flatten(tree, lo), m, "gtr_g4", pats, mult, lo))
    zero = parse_newick("((A:0,B:0)C:0,D:0)E:0;")
    zaln = RefAlignment.from_fasta_bytes(b">A\nACGT\n>B\nAGGT\n>D\nTTTT\n")
    pats, _s2p, mult, lo = zaln.compress_columns()
    zc = workloads.Case(zero, oracle.flatten(zero, lo), RefModel.jukes_cantor(), "jc69", pats, mult, lo)
    assert _cases.check_joint(gpu_lib, zc)[0] == 0
    gap = RefAlignment.from_fasta_bytes(b">A\n----\n>B\n----\n>D\n----\n")  # integration_test.rs:530-540
    pats, _s2p, mult, lo = gap.compress_columns()
    t3 = parse_newick("((A:0.601,B:0.301)C:0.100,D:0.200)E:0.001;")
    _cases.check_joint(gpu_lib, workloads.Case(t3, oracle.flatten(t3, lo), RefModel.jukes_cantor(), "jc69", pats, mult, lo))


def test_compress_columns(gpu_lib):
    rng = np.random.default_rng(78)
    base = rng.choice(np.frombuffer(b"ACGTUN-R?", np.uint8), size=(11, 300))
    _cases.check_compress(gpu_lib, base[:, rng.integers(0, 300, size=50_000)])
    _cases.check_compress(gpu_lib, rng.choice(np.frombuffer(b"ACGT", np.uint8), size=(64, 10_000)))   # config-1 shape
    _cases.check_compress(gpu_lib, rng.choice(np.frombuffer(b"ACGT", np.uint8), size=(4, 20_001)))    # <= 256 distinct columns
    _cases.check_compress(gpu_lib, np.full((3, 17), ord("A"), np.uint8))
    case = workloads.make_case(40, 6000, model="gtr_g4", seed=22)
    host = workloads.make_case(40, 6000, model="gtr_g4", seed=22, compress=True)
    ref, _ = oracle.log_likelihood_flat(host.ft, host.model, host.patterns, host.mult, mode=2, n_threads=8)
    with capi.Context(library=gpu_lib) as ctx:
        workloads.load_into(ctx, case)
        n, s2p, mult = ctx.compress_columns(np.ascontiguousarray(case.patterns.T))
        assert n == len(host.patterns) and np.array_equal(mult, host.mult)
        tot = ctx.log_likelihood()
    assert abs(tot - ref) <= 1e-9 * abs(ref)


def test_compress_columns_full_size_properties(gpu_lib):
    """1,000 rows x 1M sites drawn from 50,000 distinct columns: multiplicities sum to the width, site_to_pattern is
    a first-appearance numbering (each new maximum exceeds the previous by exactly one), sites mapped to the same
    pattern hold identical columns (sampled), and the pattern count equals numpy's unique count."""
    rng = np.random.default_rng(79)
    base = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=(1000, 50_000))
    pick = rng.integers(0, 50_000, size=1_000_000)
    rows = np.ascontiguousarray(base[:, pick])
    with capi.Context(library=gpu_lib) as ctx:
        n, s2p, mult = ctx.compress_columns(rows)
    s2p = s2p.astype(np.int64)
    assert int(mult.sum()) == rows.shape[1] and np.array_equal(np.bincount(s2p, minlength=n).astype(np.uint64), mult)
    run_max = np.maximum.accumulate(s2p)
    new = np.nonzero(np.diff(np.concatenate([[-1], run_max])) > 0)[0]
    assert np.array_equal(s2p[new], np.arange(n))           # first appearances are numbered 0,1,2,... in site order
    # distinct columns of `base` may coincide by chance only with negligible probability at 1,000 rows
    assert n == len(np.unique(pick))
    for s in rng.integers(0, rows.shape[1], size=2000):
        assert np.array_equal(rows[:, s], rows[:, new[s2p[s]]])


def test_score_nni(gpu_lib):
    """Path-only NNI rescoring (csrc/nni.cuh) vs the oracle on every rebuilt neighbour; vs full rescoring."""
    case = workloads.make_case(16, 500, model="gtr_g4", seed=28)
    got, st = _cases.check_nni(gpu_lib, case)
    assert st["nni_path_steps"] > 0
    full, _ = _cases.check_nni(gpu_lib, case, flags=capi.FLAG_NNI_FULL, expect_path=False)
    walk, _ = _cases.check_nni(gpu_lib, case, flags=capi.FLAG_NNI_WALK)
    ok = np.isfinite(got)
    assert np.array_equal(ok, np.isfinite(full)) and np.array_equal(ok, np.isfinite(walk))
    assert np.max(np.abs(got[ok] - full[ok]) / np.abs(full[ok])) < 1e-12
    assert np.max(np.abs(got[ok] - walk[ok]) / np.abs(walk[ok])) < 1e-12


@pytest.mark.parametrize("model,amb", [("jc69", 0.0), ("gtr_i_g4", 0.1)])
def test_score_nni_models_tiling(gpu_lib, model, amb):
    case = workloads.make_case(20, 3000, model=model, seed=41, ambiguity=amb)
    n_clv = 1 if model == "jc69" else 4
    got, st = _cases.check_nni(gpu_lib, case, budget=19 * n_clv * 72 * 1024 + 1)
    assert st["nni_tiles"] == 3
    one, _ = _cases.check_nni(gpu_lib, case)
    ok = np.isfinite(got)
    assert np.max(np.abs(got[ok] - one[ok]) / np.abs(one[ok])) < 1e-13


def test_score_nni_deep_tree(gpu_lib):
    """800 taxa (subtree likelihoods below 2^-256: cached exponents + the walk's own rescales)."""
    case = workloads.make_case(800, 300, model="gtr_i_g4", seed=43)
    ft = case.ft
    rng = np.random.default_rng(6)
    internal = [v for v in range(ft["n_nodes"]) if ft["child_off"][v + 1] > ft["child_off"][v]]
    moves = [(int(v), int(lc)) for v in rng.choice(internal, 48, replace=False) for lc in (0, 1)]
    _cases.check_nni(gpu_lib, case, moves=moves)
    _cases.check_nni(gpu_lib, case, moves=moves, flags=capi.FLAG_NNI_WALK)


def test_score_nni_full_size_config5_properties(gpu_lib):
    """configs[4] at FULL size (2,000 taxa x 200k sites, GTR+G4; 1,024 neighbours) through size-independent
    properties: path-only scores of a neighbour sample == full rescoring of the rebuilt tree (<= 1e-11 relative),
    a forced 3-way tiling gives the same totals, and nni(v,0)/nni(v,1) of the same v differ from the base."""
    from phylo_rs_b200 import synthetic
    tree = synthetic.yule_flat(2000, 0x5EED0005)
    model = synthetic.gtr_model(alpha=0.5, k=4)
    co = tree["child_off"]
    internal = [v for v in range(tree["n_nodes"]) if co[v + 1] > co[v] and v != tree["root"]]
    moves = [(v, lc) for v in internal for lc in (0, 1)][:1024]
    ids, left = [m[0] for m in moves], [m[1] for m in moves]
    with capi.Context(library=gpu_lib) as ctx:
        synthetic.load_into(ctx, tree, model)
        ctx.simulate_alignment(200_000, seed=0x5EED0005 ^ 0xA11C)
        base = ctx.log_likelihood()
        got = ctx.score_nni(ids, left)
        st = ctx.stats()
        assert st["nni_path_moves"] == 1024 and st["nni_tiles"] == 1
        ctx.set_clv_budget(int(st["nni_cache_bytes"] / 2.5))
        tiled = ctx.score_nni(ids, left)
        assert ctx.stats()["nni_tiles"] == 3
        assert np.max(np.abs(tiled - got) / np.abs(got)) < 1e-13
        ctx.set_clv_budget(0)
        assert ctx.log_likelihood() == base
    assert np.all(np.isfinite(got)) and np.all(got != base)
    sample = list(range(0, 1024, 128))
    with capi.Context(library=gpu_lib, flags=capi.FLAG_NNI_FULL) as ctx:
        synthetic.load_into(ctx, tree, model)
        ctx.simulate_alignment(200_000, seed=0x5EED0005 ^ 0xA11C)
        full = ctx.score_nni([ids[i] for i in sample], [left[i] for i in sample])
        assert ctx.stats()["nni_path_moves"] == 0
    assert np.max(np.abs(full - got[sample]) / np.abs(full)) < 1e-11


def test_full_size_properties_config2(gpu_lib):
    """configs[1] at FULL size (1,000 taxa x 100k sites, GTR+G4) through size-independent properties:
    the device-simulated alignment is scored whole, in 3 virtual ranks and under a forced 5-way tiling —
    per-pattern values must be identical; a 1,024-pattern sample is checked against the oracle; and
    duplicating every pattern (multiplicity 2) must double the total (linearity in the weights)."""
    case = workloads.make_case(1000, 8, model="gtr_g4", seed=0x5EED0002)
    P = 100_000
    with capi.Context(library=gpu_lib) as ctx:
        workloads.load_into(ctx, case)
        ctx.simulate_alignment(P, seed=0xA11C)
        tot, per = ctx.log_likelihood(per_pattern=True)
        assert 500 <= ctx.stats()["sweep_ops"] < 999  # cherries folded into their parents' ops
        pats = ctx.get_alignment(0, 1024)
        whole = ctx.get_alignment()
    with capi.Context(library=gpu_lib, flags=capi.FLAG_LEVEL_KERNELS) as ctx:  # HBM-resident CLVs, forced 5-way tiling
        workloads.load_into(ctx, case)
        ctx.set_alignment(whole)
        lr0 = np.full(case.ft["n_nodes"], -1, np.int32)
        lv = np.where(case.ft["child_off"][1:] == case.ft["child_off"][:-1])[0]
        lr0[lv] = np.arange(len(lv))
        ctx.set_tree(case.ft["postorder"], case.ft["child_off"], case.ft["children"], case.ft["blen"], lr0, case.ft["root"])
        ctx.set_clv_budget(999 * 4 * 4 * 8 * 20480)
        tot_t, per_t = ctx.log_likelihood(per_pattern=True)
        assert ctx.stats()["n_tiles"] >= 5
    assert np.max(np.abs(per - per_t)) < 1e-9 and abs(tot - tot_t) < 1e-11 * abs(tot)
    ft = dict(case.ft)
    leaves = np.where(ft["child_off"][1:] == ft["child_off"][:-1])[0]
    lr = np.full(ft["n_nodes"], -1, np.int32)
    lr[leaves] = np.arange(len(leaves))
    ft["leaf_row"] = lr
    ref, ref_per = oracle.log_likelihood_flat(ft, case.model, pats, n_threads=8)
    assert np.max(np.abs(per[:1024] - ref_per)) <= 1e-8
    with capi.Context(devices=[0, 0, 0], library=gpu_lib) as ctx:
        ctx.set_alignment(whole, multiplicity=np.full(P, 2, np.uint64))
        m = case.model
        ctx.set_model(m.eigenvalues, m.eigenvectors, m.pi, m.sqrt_pi, m.inv_sqrt_pi, m.rates, m.weights)
        ctx.set_tree(ft["postorder"], ft["child_off"], ft["children"], ft["blen"], lr, ft["root"])
        tot2, per2 = ctx.log_likelihood(per_pattern=True)
    assert np.array_equal(per2, per)
    assert abs(tot2 - 2 * tot) <= 1e-12 * abs(tot2)


def test_full_size_properties_config3(gpu_lib):
    """configs[2] at FULL size (10,000 taxa x 1M site patterns, GTR+I+G4 — the bench workload) through
    size-independent properties: the device-simulated alignment scored whole, in 4 virtual ranks (the multi-GPU
    shard arithmetic on one device) and with the HBM-resident level kernels under their tiling must give the same
    per-pattern values; the total is the sum of the per-pattern values; the first 256 patterns are checked against
    the oracle."""
    case = workloads.make_case(10_000, 8, model="gtr_i_g4", seed=0x5EED0003)
    P = 1_000_000
    with capi.Context(library=gpu_lib) as ctx:
        workloads.load_into(ctx, case)
        ctx.simulate_alignment(P, seed=0xB200)
        tot, per = ctx.log_likelihood(per_pattern=True)
        assert 5000 <= ctx.stats()["sweep_ops"] < 9999  # cherries folded into their parents' ops
        assert ctx.log_likelihood() == tot  # re-evaluation is bit-reproducible
        pats = ctx.get_alignment(0, 256)
    assert np.all(np.isfinite(per)) and abs(tot - float(np.sum(per))) <= 1e-11 * abs(tot)
    ft = dict(case.ft)
    leaves = np.where(ft["child_off"][1:] == ft["child_off"][:-1])[0]
    lr = np.full(ft["n_nodes"], -1, np.int32)
    lr[leaves] = np.arange(len(leaves))
    ft["leaf_row"] = lr
    _ref, ref_per = oracle.log_likelihood_flat(ft, case.model, pats, n_threads=8)
    assert np.max(np.abs(per[:256] - ref_per)) <= 1e-8
    with capi.Context(devices=[0, 0, 0, 0], library=gpu_lib) as ctx:
        workloads.load_into(ctx, case)
        ctx.simulate_alignment(P, seed=0xB200)
        tot4, per4 = ctx.log_likelihood(per_pattern=True)
    assert np.array_equal(per4, per) and abs(tot4 - tot) <= 1e-12 * abs(tot)
    with capi.Context(library=gpu_lib, flags=capi.FLAG_LEVEL_KERNELS) as ctx:
        workloads.load_into(ctx, case)
        ctx.simulate_alignment(P, seed=0xB200)
        tot_l, per_l = ctx.log_likelihood(per_pattern=True)
        assert ctx.stats()["n_tiles"] >= 2
    assert np.max(np.abs(per_l - per)) < 1e-8 and abs(tot_l - tot) <= 1e-11 * abs(tot)


def test_full_size_properties_config4(gpu_lib):
    """configs[3] at FULL size (marginal ASR, GTR+G4, 5,000 taxa x 500k site patterns) through size-independent
    properties: the device-simulated alignment reconstructed whole (tiled by the CLV budget) and in 3 virtual
    ranks gives identical states and per-pattern values; leaves keep their observed state; the joint
    reconstruction agrees with it at the leaves too; the first 128 patterns match the oracle state for state."""
    case = workloads.make_case(5_000, 8, model="gtr_g4", seed=0x5EED0004)
    P = 500_000
    code = np.full(256, 255, np.uint8)
    code[[ord(c) for c in "ACGT"]] = np.arange(4)
    with capi.Context(library=gpu_lib) as ctx:
        workloads.load_into(ctx, case)
        ctx.simulate_alignment(P, seed=0xA5A5)
        tot, states, _ = ctx.marginal_asr(False)
        per = ctx.per_pattern()
        assert ctx.stats()["n_tiles"] >= 2
        assert abs(ctx.log_likelihood() - tot) <= 1e-11 * abs(tot)
        pats = ctx.get_alignment(0, 128)
        sample = ctx.get_alignment(250_000, 4096)
        jtot, jstates = ctx.joint_asr()
    assert states.shape == (case.ft["n_nodes"], P) and states.max() <= 3
    ft = dict(case.ft)
    leaves = np.where(ft["child_off"][1:] == ft["child_off"][:-1])[0]
    lr = np.full(ft["n_nodes"], -1, np.int32)
    lr[leaves] = np.arange(len(leaves))
    ft["leaf_row"] = lr
    obs = code[sample].T  # [row][pattern]; the simulator emits A, C, G, T only
    assert obs.max() <= 3
    assert np.array_equal(states[leaves, 250_000:254_096], obs)
    assert np.array_equal(jstates[leaves, 250_000:254_096], obs) and jtot < tot
    ref_tot, ref_per, ref_states, _ = oracle.marginal_asr_flat(ft, case.model, pats, None, want_posteriors=False, n_threads=8)
    assert np.max(np.abs(per[:128] - ref_per)) <= 1e-8
    assert np.array_equal(states[:, :128], ref_states)
    del jstates
    with capi.Context(devices=[0, 0, 0], library=gpu_lib) as ctx:
        workloads.load_into(ctx, case)
        ctx.simulate_alignment(P, seed=0xA5A5)
        tot3, states3, _ = ctx.marginal_asr(False)
        per3 = ctx.per_pattern()
    assert np.array_equal(per3, per) and abs(tot3 - tot) <= 1e-12 * abs(tot)
    assert np.array_equal(states3, states)


def test_fused_host_call(gpu_lib):
    """phb_log_likelihood_host: chunked H2D copy overlapped with pack + sweep == the two-call path (several chunks)."""
    case = workloads.make_case(40, 230_000, model="gtr_i_g4", seed=31)
    with capi.Context(library=gpu_lib) as ctx:
        workloads.load_into(ctx, case)
        ref_tot, ref_per = ctx.log_likelihood(per_pattern=True)
        tot, per = ctx.log_likelihood_host(case.patterns, per_pattern=True)
        assert np.array_equal(per, ref_per)
        assert abs(tot - ref_tot) <= 1e-12 * abs(ref_tot)
        assert ctx.log_likelihood() == ref_tot  # alignment left resident
    small = workloads.make_case(40, 3000, model="gtr_i_g4", seed=31, ambiguity=0.05)
    o_tot, o_per = oracle.log_likelihood_flat(small.ft, small.model, small.patterns, small.mult, mode=0, n_threads=4)
    with capi.Context(library=gpu_lib) as ctx:
        m, ft = small.model, small.ft
        ctx.set_model(m.eigenvalues, m.eigenvectors, m.pi, m.sqrt_pi, m.inv_sqrt_pi, m.rates, m.weights)
        ctx.set_tree(ft["postorder"], ft["child_off"], ft["children"], ft["blen"], ft["leaf_row"], ft["root"])
        tot, per = ctx.log_likelihood_host(small.patterns, multiplicity=small.mult, per_pattern=True)
        _cases.assert_lnl_close(tot, per, o_tot, o_per)
        bad = small.patterns.copy()
        bad[2999, 0] = ord("!")
        with pytest.raises(capi.PhbError) as e:
            ctx.log_likelihood_host(bad)
        assert e.value.variant == "AlphabetMismatch"


def test_nonbinary_trees_all_engines(gpu_lib):
    """polytomies of arity 3..6, unifurcations, nested polytomies: lnL (sweep and level kernels), marginal and joint ASR"""
    _cases.check_nonbinary(gpu_lib)


def test_score_nni_cache_variants_and_repeated_calls(gpu_lib):
    case = workloads.make_case(300, 3000, model="gtr_i_g4", seed=77, ambiguity=0.03)
    _cases.check_nni_cache_variants(gpu_lib, case)


def test_random_trees_all_engines(gpu_lib):
    """random shapes (arities 1..5), zero / missing / tiny / long branches, random models, ambiguity, multiplicities"""
    _cases.check_random_trees(gpu_lib, range(100, 160))
