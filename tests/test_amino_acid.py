"""the 20-state AminoAcid instance of the engines (PHB_FLAG_AMINO_ACID contexts, csrc/aa.cuh).

CPU part: the oracle's 20-state build (oracle/libphylo_oracle_aa.so: the same C restatement with -DNS=20) is pinned on
the reference's own AminoAcid fixtures (src/tree/likelihood/integration_test.rs:75-109, 612-615), on brute-force
enumeration and on an independent numpy pruning + scipy expm; the product sources run on the emulation build.
GPU part (`-m gpu`): the real sm_100a library through its C ABI against that oracle, at BASELINE.json's bars
(total lnL 1e-9 relative, per-pattern 1e-8 absolute, ASR states bit-exact, posteriors 1e-9)."""
import os

import numpy as np
import pytest

import oracle
from oracle import AaModel, amino_acid_profile, workloads
from phylo_rs_b200 import capi

import _cases

AA = capi.FLAG_AMINO_ACID
CANONICAL = b"ACDEFGHIKLMNPQRSTVWY"


# ---- oracle, pinned on the reference's fixtures ------------------------------------------------------------------
def test_oracle_canonical_and_n_states():
    """integration_test.rs:75-83 (char_of / index_of round trip over CANONICAL), :612-615 (N_STATES == 20)."""
    assert oracle.lib_aa().orc_n_states() == 20 and len(CANONICAL) == 20
    for i, c in enumerate(CANONICAL):
        p = amino_acid_profile(c)
        assert p[i] == 1.0 and p.sum() == 1.0


def test_oracle_ambiguity_profiles():
    """integration_test.rs:86-109: B -> 0.5 at indices 3 and 11, Z -> 4 and 14, J -> 8 and 9, X all ones; alphabet.rs:113-114:
    the gap as X, anything else None."""
    for c, (a, b) in ((ord("B"), (3, 11)), (ord("Z"), (4, 14)), (ord("J"), (8, 9))):
        p = amino_acid_profile(c)
        assert abs(p[a] - 0.5) < 1e-10 and abs(p[b] - 0.5) < 1e-10 and np.count_nonzero(p) == 2
    assert np.all(amino_acid_profile(ord("X")) == 1.0) and np.all(amino_acid_profile(ord("-")) == 1.0)
    for c in b"OU*acgt?.":
        assert amino_acid_profile(c) is None


def test_oracle_rate_matrix_and_transition():
    scipy_linalg = pytest.importorskip("scipy.linalg")
    m = workloads.make_aa_model("aa")
    pi = m.pi
    assert abs(pi.sum() - 1) < 1e-14 and np.all(m.eigenvalues <= 1e-12) and np.sum(np.abs(m.eigenvalues) < 1e-10) == 1
    V = m.eigenvectors
    assert np.max(np.abs(V.T @ V - np.eye(20))) < 1e-12
    # rebuild Q the way RateMatrix::new does (rate_matrix.rs:34-97) and compare P(t) with expm(Q t)
    rng = np.random.default_rng(0xAA20)
    pi0 = rng.dirichlet(np.full(20, 8.0))
    w = rng.gamma(1.5, 1.0, (20, 20)) + 0.05
    w = (w + w.T) / 2
    np.fill_diagonal(w, 0.0)
    q = w * pi0[None, :]
    np.fill_diagonal(q, -q.sum(axis=1))
    q /= -(pi0 * np.diag(q)).sum()
    for t in (0.0, 1e-6, 0.05, 0.7, 5.0):
        P = m.transition(t)
        assert np.max(np.abs(P - scipy_linalg.expm(q * t))) < 1e-12
        assert np.max(np.abs(P.sum(axis=1) - 1)) < 1e-12
    assert np.array_equal(m.transition(0.0), np.eye(20))  # rate_matrix.rs:115-117
    with pytest.raises(oracle.OracleError):
        AaModel([0.25] * 4, np.ones((4, 4)))  # AlphabetMismatch, gtr.rs:39-41


def _numpy_site_lnl(ft, m, pat, rate=1.0):
    clv = [None] * ft["n_nodes"]
    for v in ft["postorder"]:
        ks = ft["children"][ft["child_off"][v]:ft["child_off"][v + 1]]
        if len(ks) == 0:
            clv[v] = amino_acid_profile(pat[ft["leaf_row"][v]])
        else:
            x = np.ones(20)
            for k in ks:
                x = x * (m.transition(rate * ft["blen"][k]) @ clv[k])
            clv[v] = x
    with np.errstate(divide="ignore"):
        return np.log(m.pi @ clv[ft["root"]])


def test_oracle_lnl_against_bruteforce_and_numpy():
    c = workloads.make_case(5, 12, model="aa", seed=4, ambiguity=0.25)
    tot, per = oracle.log_likelihood_flat(c.ft, c.model, c.patterns, c.mult)
    for i in range(len(per)):
        bf = oracle.bruteforce_site_likelihood(c.ft, c.model, c.patterns[i])  # 20^4 internal-state assignments
        assert abs(per[i] - np.log(bf)) < 1e-11
        assert abs(per[i] - _numpy_site_lnl(c.ft, c.model, c.patterns[i])) < 1e-11
    assert abs(tot - per.sum()) < 1e-10
    c = workloads.make_case(40, 30, model="aa_i_g4", seed=9, ambiguity=0.1)
    tot, per = oracle.log_likelihood_flat(c.ft, c.model, c.patterns, c.mult)
    m = c.model
    for i in range(0, 30, 7):
        cats = [np.log(w) + _numpy_site_lnl(c.ft, m, c.patterns[i], r) for r, w in zip(m.rates, m.weights)]
        assert abs(per[i] - np.logaddexp.reduce(cats)) < 1e-10


def test_oracle_marginal_posteriors_sum_to_one():
    c = workloads.make_case(12, 60, model="aa_g4", seed=2, ambiguity=0.1)
    tot, per, st, post = oracle.marginal_asr_flat(c.ft, c.model, c.patterns, c.mult, want_posteriors=True)
    assert np.allclose(post.sum(axis=2), 1.0, atol=1e-12) and st.max() < 20
    assert np.array_equal(post.shape, (c.ft["n_nodes"], 60, 20))
    # a leaf with a canonical letter is reconstructed as that letter
    for v in range(c.ft["n_nodes"]):
        r = c.ft["leaf_row"][v]
        if r >= 0 and c.ft["child_off"][v + 1] == c.ft["child_off"][v]:
            for p in range(60):
                b = c.patterns[p, r]
                if b in CANONICAL:
                    assert st[v, p] == CANONICAL.index(b)


# ---- product sources: emulation build (CPU) and the real library (GPU) -----------------------------------------------
def _parity(lib, n_taxa, n_sites, model, seed, ambiguity=0.05, budget=None, want_post=True):
    case = workloads.make_case(n_taxa, n_sites, model=model, seed=seed, ambiguity=ambiguity)
    _cases.check_lnl(lib, case, flags=AA, budget=budget)
    assert _cases.check_asr(lib, case, want_post=want_post, flags=AA, budget=budget) == 0
    _cases.check_joint(lib, case, flags=AA, budget=budget)
    return case


def _star_and_polytomy(lib):
    """any arity (likelihood.rs:113 iterates all children): a root with 5 children, each a 30-leaf caterpillar, and a
    unifurcation chain"""
    from oracle import parse_newick

    rng = np.random.default_rng(5)
    names, parts = [], []
    for j in range(5):
        s = f"(a{j}_0:0.3,a{j}_1:0.2)"
        names += [f"a{j}_0", f"a{j}_1"]
        for i in range(2, 30):
            names.append(f"a{j}_{i}")
            s = f"({s}:0.4,a{j}_{i}:0.25)"
        parts.append(f"(({s}:0.1):0.05):0.2")  # two unifurcations above each caterpillar
    tree = parse_newick("(" + ",".join(parts) + ");")
    ft = workloads.flatten_fast(tree, names)
    pats = np.frombuffer(workloads.AA_LETTERS + b"BZJX-", np.uint8)[rng.integers(25, size=(64, len(names)))]
    case = workloads.Case(tree, ft, workloads.make_aa_model("aa_g4"), "aa_g4", np.ascontiguousarray(pats), np.ones(64, np.uint64), names)
    _cases.check_lnl(lib, case, flags=AA)
    assert _cases.check_asr(lib, case, want_post=True, flags=AA) == 0
    _cases.check_joint(lib, case, flags=AA)


class _path_mode:
    """PHB_AA_NNI_PATH=1: aa_score_nni without upper partials (every neighbour re-evaluates its whole path v -> root);
    PHB_AA_NNI_NO_PAIR=1: upper partials, but v goes through a scratch slot instead of staying in registers"""

    def __init__(self, var="PHB_AA_NNI_PATH"):
        self.var = var

    def __enter__(self):
        os.environ[self.var] = "1"

    def __exit__(self, *exc):
        del os.environ[self.var]


def _rearrangements(lib):
    """phb_score_nni / phb_score_trees in a 20-state context (the reference's caller loop: clone, NNI::nni
    src/tree.rs:1213-1242 or SPR::spr src/tree/ops.rs:34-41, log_likelihood) against the oracle on each rebuilt tree.
    NNI neighbours are scored against the base tree's resident CLVs and upper partials (aa_score_nni, any arity); its
    cross-check mode without upper partials gives the same values, bit for bit, as re-scoring every rebuilt tree in full
    (PHB_FLAG_NNI_FULL).  Also with tiled patterns and more neighbours than a batch holds."""
    case = workloads.make_case(9, 40, model="aa_i_g4", seed=17, ambiguity=0.1)
    got, st = _cases.check_nni(lib, case, flags=AA)
    with _path_mode():
        walked, st_w = _cases.check_nni(lib, case, flags=AA)
    full, st_full = _cases.check_nni(lib, case, flags=AA | capi.FLAG_NNI_FULL, expect_path=False)
    assert np.array_equal(walked, full, equal_nan=True) and st_full["nni_path_moves"] == 0
    assert np.allclose(got, full, rtol=1e-12, atol=0, equal_nan=True) and 0 < st["nni_path_steps"] < st_w["nni_path_steps"]
    case4 = workloads.make_case(13, 70, model="aa_g4", seed=16, ambiguity=0.1)  # 4 categories: one launch per batch
    got4, _ = _cases.check_nni(lib, case4, flags=AA)
    with _path_mode("PHB_AA_NNI_NO_PAIR"):
        two4, _ = _cases.check_nni(lib, case4, flags=AA)
    full4, _ = _cases.check_nni(lib, case4, flags=AA | capi.FLAG_NNI_FULL, expect_path=False)
    assert np.allclose(got4, full4, rtol=1e-12, atol=0, equal_nan=True) and np.allclose(two4, full4, rtol=1e-12, atol=0, equal_nan=True)
    # 3 tiles of 256 patterns (29 CLVs + pi + 29 upper partials), every move ten times: 3 batches
    case = workloads.make_case(30, 600, model="aa_g4", seed=18, ambiguity=0.05)
    n = case.ft["n_nodes"]
    moves = [(v, lc) for v in range(n) for lc in (0, 1)]
    got, st = _cases.check_nni(lib, case, moves=moves, flags=AA, budget=59 * 4 * 164 * 256)
    assert st["nni_tiles"] == 3
    with capi.Context(flags=AA, library=lib) as ctx:
        workloads.load_into(ctx, case)
        many = ctx.score_nni([m[0] for m in moves] * 10, [m[1] for m in moves] * 10)
        assert ctx.stats()["nni_path_moves"] == 10 * int(np.isfinite(got).sum()) > 256
    assert np.array_equal(many, np.tile(got, 10), equal_nan=True)
    # polytomies and unifurcations: the moved child / sibling is appended last, as set_child does
    for nwk in _cases.NONBINARY_NEWICKS[:3]:
        c4 = _cases.nonbinary_case(nwk, "gtr", n_sites=8)
        rng = np.random.default_rng(len(nwk))
        pats = np.frombuffer(workloads.AA_LETTERS + b"BZJX-", np.uint8)[rng.integers(25, size=(50, len(c4.leaf_order)))]
        c20 = workloads.Case(c4.tree, c4.ft, workloads.make_aa_model("aa_g4"), "aa_g4", np.ascontiguousarray(pats), np.ones(50, np.uint64),
                             c4.leaf_order)
        a, _ = _cases.check_nni(lib, c20, flags=AA)
        with _path_mode():
            w, _ = _cases.check_nni(lib, c20, flags=AA)
        with _path_mode("PHB_AA_NNI_NO_PAIR"):
            t2, _ = _cases.check_nni(lib, c20, flags=AA)
        assert np.allclose(a, t2, rtol=1e-12, atol=0, equal_nan=True)
        b, _ = _cases.check_nni(lib, c20, flags=AA | capi.FLAG_NNI_FULL, expect_path=False)
        assert np.array_equal(w, b, equal_nan=True) and np.isfinite(a).sum() >= 4
        assert np.allclose(a, b, rtol=1e-12, atol=0, equal_nan=True)
    _cases.check_score_trees(lib, workloads.make_case(11, 30, model="aa_g4", seed=19, ambiguity=0.05), flags=AA)


def _compression(lib):
    """phb_compress_columns / phb_compress_rows in a 20-state context (Alignment::compress_columns, alignment.rs:93-139, is
    alphabet-agnostic): pattern numbering bit-exact against the oracle's restatement, and the evaluation of the adopted
    patterns equals the oracle on the host-compressed alignment"""
    from oracle import RefAlignment

    case = workloads.make_case(12, 90, model="aa_g4", seed=23, ambiguity=0.1)
    rng = np.random.default_rng(24)
    pick = rng.integers(90, size=700)
    rows = np.ascontiguousarray(case.patterns[pick].T)  # [row][site], 700 sites over <= 90 distinct columns
    n_rows, width = rows.shape
    pats, s2p, mult, _ = RefAlignment({f"r{i:02d}": bytes(rows[i]) for i in range(n_rows)}, width).compress_columns()
    ref_tot, ref_per = oracle.log_likelihood_flat(case.ft, case.model, pats, mult, mode=0, n_threads=4)
    m, ft = case.model, case.ft
    for by_rows in (False, True):
        with capi.Context(flags=AA, library=lib) as ctx:
            if by_rows:
                n, got_s2p, got_mult = ctx.compress_rows([rows[r].tobytes() for r in range(n_rows)])
            else:
                n, got_s2p, got_mult = ctx.compress_columns(rows)
            assert n == len(pats)
            assert np.array_equal(got_s2p.astype(np.int64), s2p) and np.array_equal(got_mult, mult)
            # the adopted patterns, read back in both layouts (all 25 accepted bytes are their own canonical byte)
            assert np.array_equal(ctx.get_alignment(), pats)
            assert np.array_equal(ctx.get_alignment(first=3, count=11, layout=capi.LAYOUT_ROW_MAJOR), pats[3:14].T)
            ctx.set_model(m.eigenvalues, m.eigenvectors, m.pi, m.sqrt_pi, m.inv_sqrt_pi, m.rates, m.weights)
            ctx.set_tree(ft["postorder"], ft["child_off"], ft["children"], ft["blen"], ft["leaf_row"], ft["root"])
            tot, per = ctx.log_likelihood(per_pattern=True)
            _cases.assert_lnl_close(tot, per, ref_tot, ref_per)
            tot_a, states, _ = ctx.marginal_asr()
            assert abs(tot_a - ref_tot) <= 1e-9 * abs(ref_tot) and states.shape == (ft["n_nodes"], n)
            # a byte AminoAcid::profile rejects, in a leaf row: AlphabetMismatch at evaluation, not at compression
            bad = rows.copy()
            bad[3, 5] = ord("O")
            ctx.compress_columns(bad)
            assert int((ctx.get_alignment() == ord("?")).sum()) == 1  # the rejected byte, and nothing else
            with pytest.raises(capi.PhbError) as ei:
                ctx.log_likelihood()
            assert ei.value.variant == "AlphabetMismatch"
            # no sites: no patterns
            n0, s0, m0 = ctx.compress_columns(np.zeros((n_rows, 0), np.uint8))
            assert n0 == 0 and len(s0) == 0 and len(m0) == 0
            assert ctx.log_likelihood() == 0.0


def _abi_behaviour(lib):
    case = workloads.make_case(6, 20, model="aa_g4", seed=3, ambiguity=0.2)
    m, ft = case.model, case.ft
    with capi.Context(flags=AA, library=lib) as ctx:
        with pytest.raises(capi.PhbError) as ei:
            ctx.log_likelihood()
        assert ei.value.variant == "State"
        with pytest.raises(ValueError):
            ctx.set_model(np.zeros(4), np.eye(4), [0.25] * 4, [0.5] * 4, [2.0] * 4, [1.0], [1.0])
        workloads.load_into(ctx, case)
        tot = ctx.log_likelihood()
        # multiplicities weight the per-pattern values; row-major input is the same alignment
        mult = np.arange(1, 21, dtype=np.uint64)
        ctx.set_alignment(np.ascontiguousarray(case.patterns.T), multiplicity=mult, layout=capi.LAYOUT_ROW_MAJOR)
        tot_m, per = ctx.log_likelihood(per_pattern=True)
        assert abs(tot_m - float((per * mult).sum())) <= 1e-12 * abs(tot_m)
        ctx.set_alignment(case.patterns)
        assert ctx.log_likelihood() == tot
        assert np.array_equal(ctx.get_alignment(), case.patterns)
        # host bytes in, lnL out: the same values as the two calls
        tot_h, per_h = ctx.log_likelihood_host(case.patterns, per_pattern=True)
        assert tot_h == tot and np.array_equal(per_h, ctx.log_likelihood(per_pattern=True)[1])
        # P(t) at 20 states, exact identity at t == 0
        Pm = ctx.transition_matrices(np.array([0.0, 0.3]), len(m.rates))
        assert np.array_equal(Pm[0, 0], np.eye(20)) and Pm.shape == (2, len(m.rates), 20, 20)
        for c in range(len(m.rates)):
            assert np.max(np.abs(Pm[1, c] - m.category_transition(c, 0.3))) <= 1e-13
        # entry points without a 20-state instance say so
        with pytest.raises(capi.PhbError) as ei:
            ctx.transition_matrices_noalloc([0.1], 4)
        assert ei.value.variant == "State" and "20-state" in str(ei.value)
        # a nucleotide-only byte in a LEAF row: AlphabetMismatch at evaluation (likelihood.rs:105-107)
        bad = case.patterns.copy()
        bad[7, 2] = ord("O")
        ctx.set_alignment(bad)
        with pytest.raises(capi.PhbError) as ei:
            ctx.log_likelihood()
        assert ei.value.variant == "AlphabetMismatch"
        # a leaf without a row: InvalidAlignment (likelihood.rs:99-103)
        ctx.set_alignment(case.patterns)
        lr = ft["leaf_row"].copy()
        lr[np.where(lr >= 0)[0][0]] = -1
        ctx.set_tree(ft["postorder"], ft["child_off"], ft["children"], ft["blen"], lr, ft["root"])
        with pytest.raises(capi.PhbError) as ei:
            ctx.marginal_asr()
        assert ei.value.variant == "InvalidAlignment"
    with pytest.raises(capi.PhbError):
        capi.Context(devices=[0, 0], flags=AA, library=lib)  # one device per 20-state context


@pytest.mark.parametrize("model", workloads.AA_MODELS)
def test_emu_lnl_and_asr(emu_lib, model):
    _parity(emu_lib, 14, 150, model, seed=11)


def test_emu_compression(emu_lib):
    _compression(emu_lib)


@pytest.mark.gpu
def test_gpu_compression(gpu_lib):
    _compression(gpu_lib)


def test_emu_rearrangements(emu_lib):
    _rearrangements(emu_lib)


@pytest.mark.gpu
def test_gpu_rearrangements(gpu_lib):
    _rearrangements(gpu_lib)


def test_emu_tiling_polytomy_abi(emu_lib):
    _parity(emu_lib, 9, 700, "aa_g4", seed=12, budget=9 * 4 * 164 * 256)  # 3 tiles of 256 patterns
    _star_and_polytomy(emu_lib)
    _abi_behaviour(emu_lib)


@pytest.mark.gpu
@pytest.mark.parametrize("model", workloads.AA_MODELS)
def test_gpu_lnl_and_asr(gpu_lib, model):
    _parity(gpu_lib, 48, 1200, model, seed=21)


@pytest.mark.gpu
def test_gpu_tiling_polytomy_abi(gpu_lib):
    _parity(gpu_lib, 9, 700, "aa_g4", seed=12, budget=9 * 4 * 164 * 256)
    _star_and_polytomy(gpu_lib)
    _abi_behaviour(gpu_lib)


@pytest.mark.gpu
def test_gpu_deep_tree_scaling(gpu_lib):
    """600 taxa: per-site |lnL| in the thousands, CLVs leave the f64 range without the power-of-two rescaling."""
    case = workloads.make_case(600, 96, model="aa_i_g4", seed=23, ambiguity=0.02)
    tot, per, _ = _cases.check_lnl(gpu_lib, case, flags=AA)
    assert per.min() < -800  # exp(-800) is below the smallest f64
    assert _cases.check_asr(gpu_lib, case, flags=AA) == 0


def test_emu_random_trees(emu_lib):
    _cases.check_random_trees_amino_acid(emu_lib, range(200, 210))


@pytest.mark.gpu
def test_gpu_random_trees(gpu_lib):
    _cases.check_random_trees_amino_acid(gpu_lib, range(200, 240))
