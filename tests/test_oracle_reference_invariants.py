"""The reference's own hot-path tests (src/tree/likelihood/integration_test.rs, src/models/*.rs tests,
src/tree/likelihood/profile.rs:52-78) restated against the oracle — every invariant they assert."""
import math

import numpy as np
import pytest

import oracle
from oracle import OracleError, RefAlignment, RefModel, parse_newick

TINY = "((A:0.601,B:0.301)C:0.100,D:0.200)E:0.001;"
PI = [0.3, 0.2, 0.2, 0.3]


def tiny():
    return parse_newick(TINY)


def aln(txt):
    return RefAlignment.from_fasta_bytes(txt)


# --- GTR tests, integration_test.rs:115-237 ---------------------------------------------------
def test_p0_is_identity_exactly():
    for m in (RefModel.jukes_cantor(), RefModel.gtr([.1, .2, .3, .4], [1, 2, 3, 4, 5, 6])):
        assert np.array_equal(m.transition(0.0), np.eye(4))


def test_rows_sum_to_one_and_nonnegative():
    m = RefModel.gtr([.1, .2, .3, .4], [1, 2, 3, 4, 5, 6])
    for t in (0.1, 0.5, 1.0, 2.0, 10.0):
        P = m.transition(t)
        np.testing.assert_allclose(P.sum(1), 1.0, atol=1e-10)
        assert P.min() >= -1e-15


def test_long_branch_converges_to_pi_and_stationarity():
    pi = np.array([.1, .2, .3, .4])
    m = RefModel.gtr(pi, [1, 2, 3, 4, 5, 6])
    np.testing.assert_allclose(m.transition(100.0), np.tile(pi, (4, 1)), atol=0.01)
    for t in (0.1, 1.0):
        np.testing.assert_allclose(pi @ m.transition(t), pi, atol=1e-8)


def test_custom_pi_is_normalised():
    m = RefModel.gtr([1, 2, 3, 4], [1, 1, 1, 1, 1, 1])
    np.testing.assert_allclose(m.pi, [.1, .2, .3, .4], atol=1e-15)


# --- models/gtr.rs tests :314-417 ----------------------------------------------------------------
def test_default_single_unit_category():
    assert RefModel.jukes_cantor().categories == [(1.0, 1.0)]


def test_hky_kappa1_equals_jc():
    np.testing.assert_allclose(RefModel.hky85([.25] * 4, 1.0).transition(0.5), RefModel.jukes_cantor().transition(0.5), atol=1e-10)


def test_gamma_and_invariant_categories():
    g = RefModel.jukes_cantor().with_gamma(0.5, 4)
    assert g.n_categories() == 4 and sum(g.weights) == pytest.approx(1.0)
    i = RefModel.jukes_cantor().with_invariant(0.3)
    assert i.categories[0] == (0.0, 0.3) and sum(i.weights) == pytest.approx(1.0)
    gi = RefModel.jukes_cantor().with_gamma(0.5, 4).with_invariant(0.2)
    assert gi.n_categories() == 5 and gi.categories[0] == (0.0, 0.2)
    assert all(w == pytest.approx(0.2) for w in gi.weights)
    for bad in (-0.1, 1.0, 1.5):
        with pytest.raises(OracleError):
            RefModel.jukes_cantor().with_invariant(bad)
    for bad in (0.0, -1.0):
        with pytest.raises(OracleError):
            RefModel.k80(bad)


# --- models/gamma.rs tests :231-276 --------------------------------------------------------------
def test_discrete_gamma_properties():
    assert list(oracle.discrete_gamma(1.0, 1)) == [1.0]
    prev_spread = None
    for a in (0.1, 0.5, 1.0, 2.0, 5.0):
        r = oracle.discrete_gamma(a, 4)
        assert np.mean(r) == pytest.approx(1.0, abs=1e-3)
        assert np.all(np.diff(r) > 0)
        spread = r[-1] - r[0]
        if prev_spread is not None:
            assert spread < prev_spread
        prev_spread = spread
    for a, k in ((0.0, 4), (-1.0, 4), (1.0, 0)):
        with pytest.raises(OracleError):
            oracle.discrete_gamma(a, k)


# --- Profile tests profile.rs:52-78 ----------------------------------------------------------------
def test_profile_scale_via_tip_tables():
    import ctypes as C

    p = np.zeros(4)
    assert oracle.lib().orc_nucleotide_profile(ord("R"), p.ctypes.data_as(C.POINTER(C.c_double))) == 1
    assert list(p) == [0.5, 0.0, 0.5, 0.0]
    assert oracle.lib().orc_nucleotide_profile(ord("N"), p.ctypes.data_as(C.POINTER(C.c_double))) == 1
    assert list(p) == [0.25] * 4
    assert oracle.lib().orc_nucleotide_profile(ord("-"), p.ctypes.data_as(C.POINTER(C.c_double))) == 1
    assert list(p) == [1.0] * 4
    assert oracle.lib().orc_nucleotide_profile(ord("U"), p.ctypes.data_as(C.POINTER(C.c_double))) == 1
    assert list(p) == [0, 0, 0, 1.0]
    assert oracle.lib().orc_nucleotide_profile(ord("X"), p.ctypes.data_as(C.POINTER(C.c_double))) == 0


# --- likelihood tests, integration_test.rs:344-706 -------------------------------------------------
ALN16 = b">A\n" + b"ACGT" * 4 + b"\n>B\n" + b"AGGT" * 4 + b"\n>C\n" + b"ACGT" * 4 + b"\n>D\n" + b"T" * 16 + b"\n"


@pytest.mark.parametrize("model", [
    RefModel.jukes_cantor(), RefModel.hky85(PI, 2.5), RefModel.gtr(PI, [1, 2.5, 1, 1, 2.5, 1]).with_gamma(0.5, 4)])
def test_log_likelihood_matches_marginal_asr(model):  # :659-684
    a = aln(ALN16)
    assert oracle.log_likelihood(tiny(), model, a) == pytest.approx(oracle.marginal_asr(tiny(), model, a)["log_likelihood"], abs=1e-9)


def test_log_likelihood_finite():  # :686-695
    v = oracle.log_likelihood(tiny(), RefModel.jukes_cantor(), aln(b">A\nACGT\n>B\nAGGT\n>C\nACGT\n>D\nTTTT\n"))
    assert math.isfinite(v) and v < 0


def test_missing_taxon_errors():  # :697-706, :499-510
    with pytest.raises(ValueError, match="AlphabetMismatch"):
        oracle.log_likelihood(tiny(), RefModel.jukes_cantor(), aln(b">A\nACGT\n>B\nAGGT\n>X\nACGT\n"))
    with pytest.raises(OracleError) as e:  # leaf D has no row
        oracle.log_likelihood(tiny(), RefModel.jukes_cantor(), aln(b">A\nACGT\n>B\nAGGT\n"))
    assert e.value.variant == "InvalidAlignment"
    with pytest.raises(OracleError) as e:
        oracle.log_likelihood(tiny(), RefModel.jukes_cantor(), aln(b">A\nACGT\n>B\nAGXT\n>D\nAAAA\n"))
    assert e.value.variant == "AlphabetMismatch"


def test_marginal_asr_posteriors():  # :360-390, :451-493
    m = RefModel.gtr([.9, .06, .02, .02], [1] * 6)
    a = aln(b">A\n" + b"ACGT" * 16 + b"\n>B\n" + b"AGGT" * 16 + b"\n>C\n" + b"ACGT" * 16 + b"\n>D\n" + b"T" * 64 + b"\n")
    r = oracle.marginal_asr(tiny(), m, a, want_posteriors=True)
    assert math.isfinite(r["log_likelihood"])
    for nid, post in r["posteriors"].items():
        np.testing.assert_allclose(post.sum(1), 1.0, atol=1e-3)
    assert set(r["sequences"]) == set(tiny().get_node_ids())
    assert all(len(s) == 64 for s in r["sequences"].values())


def test_all_gap_and_zero_branch_ok():  # :516-527, :546-559
    r = oracle.marginal_asr(tiny(), RefModel.jukes_cantor(), aln(b">A\n----\n>B\n----\n>C\n----\n>D\n----\n"))
    assert math.isfinite(r["log_likelihood"])
    z = parse_newick("((A:0,B:0)C:0,D:0)E:0;")
    r = oracle.marginal_asr(z, RefModel.jukes_cantor(), aln(b">A\nACGT\n>B\nCCGT\n>C\nACGT\n>D\nTTTT\n"))
    assert r["log_likelihood"] == -math.inf


# --- alignment tests :243-318, :565-605; alignment.rs:140-167 --------------------------------------
def test_fasta_parsing_rules():
    a = aln(b">s1 some description\r\nac\r\ngt\r\n>s2\nACGT\n")
    assert a.seqs == {"s1": b"ACGT", "s2": b"ACGT"} and a.width == 4
    for bad in (b"", b">\nACGT\n", b">a\nAC\n>b\nA\n", b">a\nAC\n>a\nAC\n"):
        with pytest.raises(ValueError):
            aln(bad)


def test_compress_columns_roundtrip():
    a = aln(b">x\nAACCA\n>y\nGGTTG\n")
    pats, s2p, mult, lo = a.compress_columns()
    assert pats.shape == (2, 2) and list(mult) == [3, 2] and list(s2p) == [0, 0, 1, 1, 0]
    rows = np.stack([np.frombuffer(a.seqs[n], np.uint8) for n in lo])
    assert np.array_equal(pats[s2p].T, rows)


def test_sequence_string():  # :324-338
    assert oracle.sequence_string([0, 1, 2, 3]) == "ACGT"


# --- NNI example tests/tree-tests.rs:377-389 (config-5 neighbour generator) ---------------------------
def test_nni_example():
    t = parse_newick("(A,(B,(C,D)));")
    assert t.nni(4, True) is None
    assert t.to_newick() == "(A,((C,B),D));"
    t = parse_newick("(A,(B,(C,D)));")
    assert t.nni(4, False) is None
    assert t.to_newick() == "(A,((D,B),C));"
    assert parse_newick("(A,(B,(C,D)));").nni(0, True) == "NoParent"
    assert parse_newick("(A,(B,(C,D)));").nni(1, True) == "NoParent"
