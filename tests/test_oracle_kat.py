"""Pins the CPU oracle: the known-answer table of SURVEY.md §8c (brute-force / analytic / independent
numpy restatement) and brute-force enumeration over internal states.  CPU only."""
import math

import numpy as np
import pytest

import oracle
from oracle import RefAlignment, RefModel, parse_newick, workloads

TINY = "((A:0.601,B:0.301)C:0.100,D:0.200)E:0.001;"  # src/tree/likelihood/integration_test.rs:14-17
ALN4 = b">A\nACGT\n>B\nAGGT\n>C\nACGT\n>D\nTTTT\n"
ALN16 = b">A\n" + b"ACGT" * 4 + b"\n>B\n" + b"AGGT" * 4 + b"\n>C\n" + b"ACGT" * 4 + b"\n>D\n" + b"T" * 16 + b"\n"
PI = [0.3, 0.2, 0.2, 0.3]


def lnl(model, aln=ALN16, newick=TINY):
    return oracle.log_likelihood(parse_newick(newick), model, RefAlignment.from_fasta_bytes(aln))


def test_kat_log_likelihoods():
    assert lnl(RefModel.jukes_cantor(), ALN4) == pytest.approx(-16.577012400513222, abs=1e-12)
    assert lnl(RefModel.jukes_cantor()) == pytest.approx(-66.3080496020529, abs=1e-11)
    assert lnl(RefModel.hky85(PI, 2.5)) == pytest.approx(-68.11049472687087, abs=1e-11)
    g4 = RefModel.gtr(PI, [1, 2.5, 1, 1, 2.5, 1]).with_gamma(0.5, 4)
    assert lnl(g4) == pytest.approx(-69.66427000171137, abs=1e-11)
    assert lnl(g4.with_invariant(0.2)) == pytest.approx(-71.7982617885503, abs=1e-11)


def test_kat_doc_example():
    # src/lib.rs:126-140
    tree = "((A:0.1,B:0.2):0.15,(C:0.3,D:0.1):0.05);"
    aln = b">A\nACGTACGT\n>B\nACGTATGT\n>C\nACGAACGT\n>D\nTCGTACGA\n"
    m = RefModel.hky85([0.25] * 4, 2.0).with_gamma(0.5, 4)
    v = lnl(m, aln, tree)
    assert v == pytest.approx(-30.098779761644973, abs=1e-12)
    r = oracle.marginal_asr(parse_newick(tree), m, RefAlignment.from_fasta_bytes(aln))
    assert r["log_likelihood"] == pytest.approx(v, abs=1e-9)


def test_kat_per_site_and_asr_states():
    tree = parse_newick(TINY)
    aln = RefAlignment.from_fasta_bytes(ALN4)
    pats, s2p, mult, lo = aln.compress_columns()
    ft = oracle.flatten(tree, lo)
    tot, per = oracle.log_likelihood_flat(ft, RefModel.jukes_cantor(), pats, mult)
    np.testing.assert_allclose(per[s2p], [-4.44798554852317, -5.200730091190169, -4.44798554852317, -2.480311212276715],
                               atol=1e-13)
    r = oracle.marginal_asr(tree, RefModel.jukes_cantor(), aln, want_posteriors=True)
    names = {tree.nodes[i].taxa: i for i in tree.get_node_ids()}
    got = {k: oracle.sequence_string(r["sequences"][v]) for k, v in names.items()}
    assert got == {"E": "ATGT", "C": "AGGT", "A": "ACGT", "B": "AGGT", "D": "TTTT"}
    np.testing.assert_allclose(r["posteriors"][names["E"]][0],
                               [0.5011792911334836, 0.03100425268467755, 0.03100425268467753, 0.43681220349716127],
                               atol=1e-14)


def test_kat_n_vs_gap_differ_by_ln_quarter():
    a = lnl(RefModel.jukes_cantor(), b">A\nN\n>B\nA\n>D\nC\n")
    b = lnl(RefModel.jukes_cantor(), b">A\n-\n>B\nA\n>D\nC\n")
    assert a - b == pytest.approx(math.log(0.25), abs=1e-13)


def test_kat_zero_branch_conflict_is_neg_inf_and_all_T():
    tree = "((A:0,B:0)C:0,D:0)E:0;"
    aln = RefAlignment.from_fasta_bytes(b">A\nA\n>B\nC\n>D\nG\n")
    r = oracle.marginal_asr(parse_newick(tree), RefModel.jukes_cantor(), aln)
    assert r["log_likelihood"] == -math.inf
    t = parse_newick(tree)
    for nid in t.get_node_ids():
        if not t.is_leaf(nid):
            assert list(r["sequences"][nid]) == [3]


def test_kat_all_gap():
    aln = b">A\n----\n>B\n----\n>C\n----\n>D\n----\n"
    assert abs(lnl(RefModel.jukes_cantor(), aln)) < 1e-13


def test_kat_discrete_gamma():
    np.testing.assert_allclose(oracle.discrete_gamma(0.5, 4),
                               [0.03338775338359941, 0.25191591759343773, 0.8202684819736465, 2.8944278470493163], rtol=1e-12)
    np.testing.assert_allclose(oracle.discrete_gamma(1.0, 4),
                               [0.1369537826446569, 0.4767518562354515, 0.999999999999997, 2.3862943611198943], rtol=1e-12)
    np.testing.assert_allclose(oracle.discrete_gamma(2.0, 4),
                               [0.2932747161235139, 0.6550136761581504, 1.0699896623909892, 1.9817219453273465], rtol=1e-12)


def test_kat_eigen_and_transition():
    m = RefModel.hky85(PI, 2.5)
    np.testing.assert_allclose(sorted(m.eigenvalues), sorted([0, -1 / 1.1, -1.75 / 1.1, -1.75 / 1.1]), atol=1e-13)
    np.testing.assert_allclose(m.transition(0.1)[0],
                               [0.9150977418327833, 0.01737985674354748, 0.0414526163083481, 0.026069785115321452], atol=1e-14)
    g = RefModel.gtr([0.1, 0.2, 0.3, 0.4], [1, 2, 3, 4, 5, 6])
    np.testing.assert_allclose(sorted(g.eigenvalues),
                               sorted([0, -0.7042302112436655, -1.3512778335426976, -1.726543237264919]), atol=1e-13)
    np.testing.assert_allclose(g.transition(0.25)[0],
                               [0.8533404980667257, 0.017548421334631817, 0.045295766825067696, 0.08381531377357469], atol=1e-14)
    jc = RefModel.jukes_cantor().transition(0.5)
    assert jc[0, 0] == pytest.approx(0.25 + 0.75 * math.exp(-2 / 3), abs=1e-15)
    assert jc[0, 1] == pytest.approx(0.25 - 0.25 * math.exp(-2 / 3), abs=1e-15)


def test_transition_matches_scipy_expm():
    from scipy.linalg import expm

    pi = np.array([0.1, 0.2, 0.3, 0.4])
    w = np.zeros((4, 4))
    for (i, j), r in zip(((0, 1), (0, 2), (0, 3), (1, 2), (1, 3), (2, 3)), (1, 2, 3, 4, 5, 6)):
        w[i, j] = w[j, i] = r
    q = w * pi[None, :]
    np.fill_diagonal(q, 0)
    np.fill_diagonal(q, -q.sum(1))
    q /= -(pi * np.diag(q)).sum()
    g = RefModel.gtr(pi, [1, 2, 3, 4, 5, 6])
    for t in (0.01, 0.3, 2.0):
        np.testing.assert_allclose(g.transition(t), expm(q * t), atol=1e-13)


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_bruteforce_enumeration(seed):
    from oracle import workloads

    case = workloads.make_case(6, 12, model="gtr", seed=seed, ambiguity=0.2)
    _tot, per = oracle.log_likelihood_flat(case.ft, case.model, case.patterns, case.mult)
    for p in range(len(case.patterns)):
        bf = oracle.bruteforce_site_likelihood(case.ft, case.model, case.patterns[p])
        ls = 0.0
        for r in range(case.patterns.shape[1]):
            c = bytes([case.patterns[p, r]])
            k = {b"R": 2, b"Y": 2, b"S": 2, b"W": 2, b"K": 2, b"M": 2, b"B": 3, b"D": 3, b"H": 3, b"V": 3, b"N": 4}.get(c, 1)
            ls += math.log(1.0 / k)
        # brute force uses the unscaled fractional profiles; Profile::scale() only moves ln(max) to log_scale
        assert per[p] == pytest.approx(math.log(bf), abs=1e-11), (p, ls)


NONBINARY = [
    # polytomies of arity 3..6, unifurcations, a polytomy under a polytomy; <= 9 internal nodes (4^9 assignments)
    "((a:0.1,b:0.2,c:0.3):0.2,(d:0.1,e:0.4,f:0.1,g:0.2):0.3,h:0.5);",
    "(((a:0.2,b:0.1):0.1):0.3,((c:0.3):0.2,d:0.1,e:0.2,f:0.3,g:0.1,h:0.2):0.2,(i:0.1,j:0.3):0.4);",
    "((a:0.05,b:0.4,c:0.1,d:0.3,e:0.2):0.6,(f:0.2,(g:0.1,h:0.1,i:0.3):0.2,j:0.2):0.1,k:0.7,l:0.02);",
    "(a:0.3,b:0.1,c:0.2,d:0.4,e:0.5,f:0.6);",
]


@pytest.mark.parametrize("nwk", NONBINARY)
@pytest.mark.parametrize("model", ["gtr", "gtr_i_g4"])
def test_bruteforce_enumeration_nonbinary(nwk, model):
    """An independent checker for arbitrary arity (likelihood.rs:113 iterates over any number of children): the sum over
    all internal-state assignments, per rate category, against the pruning restatement."""
    import re

    from oracle import parse_newick, workloads

    tree = parse_newick(nwk)
    names = re.findall(r"[(,]([a-z])", nwk)
    ft = workloads.flatten_fast(tree, names)
    m = workloads.make_model(model)
    rng = np.random.default_rng(len(nwk))
    pats = np.frombuffer(b"ACGTACGTACGTRYN-", np.uint8)[rng.integers(16, size=(10, len(names)))]
    _tot, per = oracle.log_likelihood_flat(ft, m, np.ascontiguousarray(pats))
    for p in range(len(pats)):
        with np.errstate(divide="ignore"):
            cats = [math.log(w) + np.log(oracle.bruteforce_site_likelihood(ft, m, pats[p], rate=float(r)))
                    for r, w in zip(m.rates, m.weights)]
        assert per[p] == pytest.approx(float(np.logaddexp.reduce(cats)), abs=1e-11)


def test_modes_agree():
    from oracle import workloads

    case = workloads.make_case(40, 200, model="gtr_i_g4", seed=5, ambiguity=0.05)
    t0, p0 = oracle.log_likelihood_flat(case.ft, case.model, case.patterns, case.mult, mode=0)
    t1, p1 = oracle.log_likelihood_flat(case.ft, case.model, case.patterns, case.mult, mode=1)
    t2, p2 = oracle.log_likelihood_flat(case.ft, case.model, case.patterns, case.mult, mode=2, n_threads=4)
    assert t0 == t1 and np.array_equal(p0, p1)
    np.testing.assert_allclose(p2, p0, atol=1e-10)
    ta, pa, _s, _ = oracle.marginal_asr_flat(case.ft, case.model, case.patterns, case.mult)
    np.testing.assert_allclose(pa, p0, atol=1e-12)


def test_joint_asr_restatement_against_bruteforce():
    """compute_joint_asr restatement (oracle joint_ref_worker, likelihood.rs:421-560): the per-pattern value is
    the maximum over (category, every assignment of internal states) of the joint log-probability — checked by
    enumeration on a 5-taxon tree with ambiguity codes — and the returned states attain it."""
    case = workloads.make_case(5, 30, model="gtr_g4", seed=3, ambiguity=0.15)
    ft, m = case.ft, case.model
    tot, per, states = oracle.joint_asr_flat(ft, m, case.patterns, case.mult)
    co = ft["child_off"]
    n = int(ft["n_nodes"])
    for p in range(len(case.patterns)):
        # full-assignment check (cheap form): the engine's own states reach the reported value ...
        val = oracle.joint_assignment_value(ft, m, case.patterns[p], states[:, p])
        assert abs(val - per[p]) < 1e-12
        # ... and no single-node change improves on it (local optimality; global optimality is covered below)
        for v in range(n):
            for s in range(4):
                st = np.array(states[:, p])
                st[v] = s
                assert oracle.joint_assignment_value(ft, m, case.patterns[p], st) <= per[p] + 1e-12
    assert tot == pytest.approx(float(np.sum(per * case.mult)), rel=1e-15)
    # joint value <= marginal site log-likelihood (a maximum over paths never exceeds their sum)
    _t, ref_per = oracle.log_likelihood_flat(ft, m, case.patterns, case.mult)
    assert np.all(per <= ref_per + 1e-12)


def test_joint_asr_global_optimum_by_enumeration():
    case = workloads.make_case(3, 8, model="gtr_i_g4", seed=8, ambiguity=0.2)
    ft, m = case.ft, case.model
    _tot, per, _states = oracle.joint_asr_flat(ft, m, case.patterns, case.mult)
    n = int(ft["n_nodes"])
    import itertools
    for p in range(len(case.patterns)):
        best = max(oracle.joint_assignment_value(ft, m, case.patterns[p], np.array(a)) for a in itertools.product(range(4), repeat=n))
        assert abs(best - per[p]) < 1e-12
