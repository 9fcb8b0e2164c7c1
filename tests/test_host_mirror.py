"""The reference's own hot-path tests, re-stated against the C++ host mirror (tests/hostshim/hostapi ->
host/libphylo_host.so -> C ABI -> CUDA kernels).  Names follow `src/tree/likelihood/integration_test.rs`
and the unit tests in `src/models/*.rs`, `src/alignment.rs`, `tests/tree-tests.rs`.

CPU part (`-m "not gpu"`): host logic (Newick ids, traversals, NNI, FASTA, compression, model categories,
discrete gamma, eigen-system) checked against the KAT table of SURVEY.md §8c and the oracle's host
restatement, plus the full engine path over the emulation build of the same .cu sources.
GPU part (`-m gpu`): the same engine tests through the real sm_100a library, parity against the oracle.
"""
import math

import numpy as np
import pytest

import oracle
from oracle import RefAlignment, RefModel, parse_newick
import hostapi as api
from hostapi import Alignment, AsrError, GtrModel, Nucleotide, PhyloTree

TINY = b"((A:0.601,B:0.301)C:0.100,D:0.200)E:0.001;"  # integration_test.rs:14-17
ALN4 = b">A\nACGT\n>B\nAGGT\n>C\nACGT\n>D\nTTTT\n"
ALN16 = b">A\nACGTACGTACGTACGT\n>B\nAGGTAGGTAGGTAGGT\n>C\nACGTACGTACGTACGT\n>D\nTTTTTTTTTTTTTTTT\n"
PI = [0.3, 0.2, 0.2, 0.3]


def build_tiny_tree():
    return PhyloTree.from_newick(TINY)


# ======================================================================================================
# host logic (no device)
# ======================================================================================================
def test_newick_ids_and_traversals_match_reference_rules():
    t = build_tiny_tree()
    ref = parse_newick(TINY)
    assert t.get_node_ids() == ref.get_node_ids()
    assert t.postord_ids() == ref.postord_ids()
    assert t.preord_ids() == ref.preord_ids()
    for v in t.get_node_ids():
        assert t.get_node_children_ids(v) == ref.nodes[v].children
        assert t.get_node_taxa(v) == ref.nodes[v].taxa
        w = t.get_edge_weight(t.get_node_parent_id(v), v)
        assert w == pytest.approx(float(ref.nodes[v].weight), abs=0)  # f32, bit-identical
    assert t.get_taxa_node_id("C") == ref.get_taxa_node_id("C")
    assert t.get_taxa_node_id("nope") is None
    assert np.float32(0.601) == t.get_edge_weight(None, t.get_taxa_node_id("A"))  # Appendix A-1: f32 parse


def test_newick_quotes_comments_and_errors():
    t = PhyloTree.from_newick("('a b':1,[c]B:2)'it''s';")
    assert sorted(x for x in (t.get_node_taxa(v) for v in t.get_node_ids()) if x) == ["B", "a b", "it's"]
    for bad in ("((A,B);", "A,B);", "", "(A:xyz,B);"):
        with pytest.raises(AsrError) as e:
            PhyloTree.from_newick(bad)
        assert e.value.variant == "Newick"


def test_nni_example():
    # tests/tree-tests.rs:377-389
    t = PhyloTree.from_newick("(A,(B,(C,D)));")
    a = t.clone()
    a.nni(4, True)
    assert a.to_newick() == "(A,((C,B),D));"
    b = t.clone()
    b.nni(4, False)
    assert b.to_newick() == "(A,((D,B),C));"
    for node, variant in ((0, "NoParent"), (1, "NoParent"), (99, "UnknownNode")):
        with pytest.raises(AsrError) as e:
            t.clone().nni(node, True)
        assert e.value.variant == variant


def test_yule_shape_and_labels():
    t = PhyloTree.yule(50, seed=0x5EED0001)
    ids = t.get_node_ids()
    assert len(ids) == 99 and ids == list(range(99))
    leaves = [v for v in ids if t.is_leaf(v)]
    assert sorted(int(t.get_node_taxa(v)) for v in leaves) == list(range(50))
    assert all(len(t.get_node_children_ids(v)) == 2 for v in ids if not t.is_leaf(v))
    assert all(t.get_edge_weight(None, v) is None for v in ids)  # the reference assigns no lengths
    assert PhyloTree.yule(50, seed=0x5EED0001).to_newick() == t.to_newick()
    assert PhyloTree.yule(50, seed=2).to_newick() != t.to_newick()
    assert PhyloTree.yule(2, 0).num_nodes() == 1 and PhyloTree.yule(3, 0).num_nodes() == 3  # tree.rs:507,526
    t.assign_exponential_lengths(7)
    w = [t.get_edge_weight(None, v) for v in ids if v != t.get_root_id()]
    assert all(1e-4 <= x <= 2.0 for x in w) and 0.05 < np.mean(w) < 0.2


def test_fasta_parsing_rules():
    a = Alignment.from_fasta_bytes(b">s1 desc\nac\ngt\n\n>s2\nAC-N\n")
    assert a.width == 4 and a.n_seqs() == 2
    c = a.compress_columns()
    assert c.leaf_order == ["s1", "s2"] and bytes(c.patterns[:, 0]) == b"ACGT"  # upper-cased, lines joined
    for bad, frag in ((b">a\nAC\n>a\nAC\n", "Duplicate"), (b">a\nAC\n>b\nA\n", "Ragged"), (b"\n\n", "Empty"),
                      (b"> x\nAC\n", "Empty taxon")):
        with pytest.raises(AsrError) as e:
            Alignment.from_fasta_bytes(bad)
        assert e.value.variant == "InvalidAlignment" and frag in str(e.value)


def test_compression_with_duplicate_patterns():
    # integration_test.rs:554-569
    aln = Alignment.from_seqs({"S1": b"ACGTACGT", "S2": b"ACGTACGT"}, width=8)
    comp = aln.compress_columns()
    assert len(comp.patterns) == 4
    assert all(m == 2 for m in comp.multiplicity)


def test_compression_re_expansion():
    # integration_test.rs:571-593
    seqs = {"S1": b"AATTGGCC", "S2": b"ATGCATGC"}
    comp = Alignment.from_seqs(seqs, width=8).compress_columns()
    for site in range(8):
        expected = bytes(seqs[name][site] for name in comp.leaf_order)
        assert bytes(comp.patterns[comp.site_to_pattern[site]]) == expected
    assert int(comp.multiplicity.sum()) == 8


def test_compression_matches_oracle_host_restatement():
    rng = np.random.default_rng(3)
    seqs = {f"t{i}": bytes(rng.choice(list(b"ACGT-N"), 300, p=[.3, .3, .2, .1, .05, .05]).astype(np.uint8)) for i in range(3)}
    comp = Alignment.from_seqs(seqs).compress_columns()
    pats, s2p, mult, lo = RefAlignment(dict(seqs), 300).compress_columns()
    assert comp.leaf_order == lo
    assert np.array_equal(comp.patterns, pats) and np.array_equal(comp.site_to_pattern, s2p)
    assert np.array_equal(comp.multiplicity, mult)


def test_nucleotide_profiles():
    # integration_test.rs:23-55
    assert Nucleotide.profile("A").tolist() == [1, 0, 0, 0] and Nucleotide.profile("U").tolist() == [0, 0, 0, 1]
    assert Nucleotide.profile("R").tolist() == [0.5, 0, 0.5, 0] and Nucleotide.profile("Y").tolist() == [0, 0.5, 0, 0.5]
    assert Nucleotide.profile("N").tolist() == [0.25] * 4 and Nucleotide.profile("-").tolist() == [1.0] * 4
    assert Nucleotide.profile("B").tolist() == [0, 1 / 3, 1 / 3, 1 / 3]
    assert Nucleotide.profile("X") is None and Nucleotide.profile("a") is None


def test_discrete_gamma_kat():
    # SURVEY.md §8c table
    np.testing.assert_allclose(api.discrete_gamma(0.5, 4), [0.03338775338359941, 0.25191591759343773, 0.8202684819736465,
                                                            2.8944278470493163], rtol=1e-13)
    np.testing.assert_allclose(api.discrete_gamma(1.0, 4), [0.1369537826446569, 0.4767518562354515, 0.999999999999997,
                                                            2.3862943611198943], rtol=1e-13)
    np.testing.assert_allclose(api.discrete_gamma(2.0, 4), [0.2932747161235139, 0.6550136761581504, 1.0699896623909892,
                                                            1.9817219453273465], rtol=1e-13)
    assert api.discrete_gamma(3.0, 1).tolist() == [1.0]
    for k in (2, 3, 8):
        for alpha in (0.1, 0.7, 5.0):
            np.testing.assert_allclose(api.discrete_gamma(alpha, k), oracle.discrete_gamma(alpha, k), rtol=1e-12)
            assert np.mean(api.discrete_gamma(alpha, k)) == pytest.approx(1.0, abs=1e-9)
    for alpha, k in ((0.0, 4), (-1.0, 4), (1.0, 0)):
        with pytest.raises(AsrError) as e:
            api.discrete_gamma(alpha, k)
        assert e.value.variant == "InvalidModelParameter"


def test_model_categories_gamma_and_invariant():
    # models/gtr.rs tests :361-417
    m = GtrModel.jukes_cantor()
    assert m.categories() == [(1.0, 1.0)]
    g = m.with_gamma(0.5, 4)
    assert g.n_categories() == 4 and all(w == 0.25 for _, w in g.categories())
    gi = g.with_invariant(0.2)
    cats = gi.categories()
    assert len(cats) == 5 and cats[0] == (0.0, 0.2)
    assert sum(w for _, w in cats) == pytest.approx(1.0, abs=1e-15)
    assert [r for r, _ in cats[1:]] == [r for r, _ in g.categories()]
    i_only = m.with_invariant(0.3)
    assert i_only.categories() == [(0.0, 0.3), (1.0, 0.7)]
    ig = i_only.with_gamma(1.0, 2)  # +G after +I keeps p_inv
    assert ig.categories()[0] == (0.0, 0.3) and len(ig.categories()) == 3
    assert gi.with_invariant(0.0).n_categories() == 4
    for bad in (1.0, -0.1, float("nan")):
        with pytest.raises(AsrError) as e:
            m.with_invariant(bad)
        assert e.value.variant == "InvalidModelParameter"
    for ctor in (lambda: GtrModel.k80(0.0), lambda: GtrModel.hky85(PI, -1), lambda: GtrModel.gtr(PI, [1, 1, 1, 1, 1, math.inf]),
                 lambda: GtrModel.new([0, 0, 0, 0], np.ones((4, 4)))):
        with pytest.raises(AsrError) as e:
            ctor()
        assert e.value.variant == "InvalidModelParameter"


def test_eigen_system_kat_and_oracle():
    lam, _ = GtrModel.hky85(PI, 2.5).eigen()
    np.testing.assert_allclose(np.sort(lam)[::-1], [0, -1 / 1.1, -1.75 / 1.1, -1.75 / 1.1], atol=1e-14)
    m = GtrModel.gtr([0.1, 0.2, 0.3, 0.4], [1, 2, 3, 4, 5, 6])
    lam, V = m.eigen()
    np.testing.assert_allclose(np.sort(lam)[::-1], [0, -0.7042302112436655, -1.3512778335426976, -1.726543237264919], atol=1e-13)
    np.testing.assert_allclose(V.T @ V, np.eye(4), atol=1e-13)
    np.testing.assert_allclose(m.equilibrium(), [0.1, 0.2, 0.3, 0.4], atol=1e-16)
    # custom pi is normalised (rate_matrix.rs:52-59)
    np.testing.assert_allclose(GtrModel.new([9, 0.6, 0.2, 0.2], np.ones((4, 4))).equilibrium(), [0.9, 0.06, 0.02, 0.02], atol=1e-16)
    # every named constructor agrees with the oracle's restatement of the same constructor
    pairs = [
        (GtrModel.k80(2.0), RefModel.k80(2.0)), (GtrModel.f81(PI), RefModel.f81(PI)),
        (GtrModel.tn93(PI, 2.0, 3.0), RefModel.tn93(PI, 2.0, 3.0)), (GtrModel.k81(1.5, 0.7, 1.2), RefModel.k81(1.5, 0.7, 1.2)),
        (GtrModel.tim(PI, 2.0, 3.0, 0.5, 1.5), RefModel.tim(PI, 2.0, 3.0, 0.5, 1.5)),
        (GtrModel.tvm(PI, 2.0, 0.5, 1.5, 0.8), RefModel.tvm(PI, 2.0, 0.5, 1.5, 0.8)),
        (GtrModel.sym([1, 2, 3, 4, 5, 6]), RefModel.sym([1, 2, 3, 4, 5, 6])),
    ]
    for mine, ref in pairs:
        lam, V = mine.eigen()
        rl = np.asarray(ref.eigenvalues)
        np.testing.assert_allclose(np.sort(lam), np.sort(rl), atol=1e-13)
        # same operator: V diag(lam) V^T
        S = V @ np.diag(lam) @ V.T
        rV = np.asarray(ref.eigenvectors).reshape(4, 4)
        np.testing.assert_allclose(S, rV @ np.diag(rl) @ rV.T, atol=1e-13)


def test_missing_taxon_is_raised_before_any_device_work():
    # integration_test.rs:476-489, :693-706 — raised while resolving leaf_order, no GPU needed
    tree = build_tiny_tree()
    aln = Alignment.from_fasta_bytes(b">A\nACGT\n>B\nAGGT\n>C\nAAAA\n>X\nTTTT\n")
    for call in (lambda: tree.log_likelihood(GtrModel.jukes_cantor(), aln),
                 lambda: tree.marginal_asr(GtrModel.jukes_cantor(), aln, False)):
        with pytest.raises(AsrError) as e:
            call()
        assert e.value.variant == "AlphabetMismatch" and "Taxon X" in str(e.value)


# ======================================================================================================
# engine tests: emulation build on CPU, real library on the GPU
# ======================================================================================================
@pytest.fixture
def emu_backend(emu_lib):
    api.set_library(emu_lib.path)
    yield
    api.set_library(None)


@pytest.fixture
def gpu_backend(gpu_lib):
    api.set_library(None)
    assert "sm_100a" in api.backend_version()
    yield
    api.release_backend()


def _engine_suite():
    """The reference's engine tests (integration_test.rs:116-222, :344-706) + the KAT table."""
    tree = build_tiny_tree()
    jc = GtrModel.jukes_cantor()
    hky = GtrModel.hky85(PI, 2.5)
    gtr_g = GtrModel.gtr(PI, [1.0, 2.5, 1.0, 1.0, 2.5, 1.0]).with_gamma(0.5, 4)
    aln4, aln16 = Alignment.from_fasta_bytes(ALN4), Alignment.from_fasta_bytes(ALN16)

    # --- transition matrices: test_p0_is_identity / rows sum / non-negative / convergence / stationarity
    for m in (jc, hky, gtr_g.with_invariant(0.2)):
        pi = m.equilibrium()
        for c in range(m.n_categories()):
            assert np.array_equal(m.category_transition(c, 0.0), np.eye(4))
            for t in (0.01, 0.1, 1.0, 10.0):
                p = m.category_transition(c, t)
                np.testing.assert_allclose(p.sum(1), 1.0, atol=1e-10)
                assert (p >= -1e-15).all()
                np.testing.assert_allclose(pi @ p, pi, atol=1e-8)
            if m.categories()[c][0] > 0:
                np.testing.assert_allclose(m.category_transition(c, 1000.0 / m.categories()[c][0]), np.tile(pi, (4, 1)), atol=1e-6)
            else:
                assert np.array_equal(m.category_transition(c, 5.0), np.eye(4))  # rate 0 => identity
    np.testing.assert_allclose(hky.category_transition(0, 0.1)[0],
                               [0.9150977418327833, 0.01737985674354748, 0.0414526163083481, 0.026069785115321452], atol=1e-14)
    np.testing.assert_allclose(GtrModel.gtr([.1, .2, .3, .4], [1, 2, 3, 4, 5, 6]).category_transition(0, 0.25)[0],
                               [0.8533404980667257, 0.017548421334631817, 0.045295766825067696, 0.08381531377357469], atol=1e-14)
    np.testing.assert_allclose(jc.category_transition(0, 0.5)[0], [0.635062839274444] + [0.121645720241852] * 3, atol=1e-14)
    np.testing.assert_allclose(GtrModel.hky85([0.25] * 4, 1.0).category_transition(0, 0.3), jc.category_transition(0, 0.3), atol=1e-10)

    # --- test_log_likelihood_matches_marginal_asr + KAT values
    kat = {id(jc): -66.3080496020529, id(hky): -68.11049472687087, id(gtr_g): -69.66427000171137}
    for model in (jc, hky, gtr_g):
        direct = tree.log_likelihood(model, aln16)
        via_asr = tree.marginal_asr(model, aln16, False).log_likelihood
        assert abs(direct - via_asr) < 1e-9
        assert direct == pytest.approx(kat[id(model)], rel=1e-12)
    assert tree.log_likelihood(gtr_g.with_invariant(0.2), aln16) == pytest.approx(-71.7982617885503, rel=1e-12)
    # test_log_likelihood_finite / test_marginal_asr_log_likelihood_positive
    assert tree.log_likelihood(jc, aln4) == pytest.approx(-16.577012400513222, rel=1e-13)
    assert math.isfinite(tree.marginal_asr(jc, aln4, False).log_likelihood)
    # doc example, src/lib.rs:123-147
    doc_tree = PhyloTree.from_newick("((A:0.1,B:0.2):0.15,(C:0.3,D:0.1):0.05);")
    doc_aln = Alignment.from_fasta_bytes(b">A\nACGTACGT\n>B\nACGTATGT\n>C\nACGAACGT\n>D\nTCGTACGA\n")
    doc_model = GtrModel.hky85([0.25] * 4, 2.0).with_gamma(0.5, 4)
    assert doc_tree.log_likelihood(doc_model, doc_aln) == pytest.approx(-30.098779761644973, rel=1e-12)

    # --- test_marginal_asr_basic / _with_posteriors: entries for EVERY node, posteriors sum to 1
    recon = tree.marginal_asr(jc, aln16, True)
    assert recon.n_nodes() == 5 and recon.has_posteriors()
    for v in tree.get_node_ids():
        assert len(recon.sequence(v)) == 16
        np.testing.assert_allclose(recon.posteriors(v).sum(1), 1.0, atol=1e-12)
    assert not tree.marginal_asr(jc, aln16, False).has_posteriors()
    # KAT: states and root posterior
    r4 = tree.marginal_asr(jc, aln4, True)
    names = {tree.get_node_taxa(v): v for v in tree.get_node_ids()}
    assert {k: r4.sequence_string(v) for k, v in names.items()} == {"E": "ATGT", "C": "AGGT", "A": "ACGT", "B": "AGGT", "D": "TTTT"}
    np.testing.assert_allclose(r4.posteriors(names["E"])[0],
                               [0.5011792911334836, 0.03100425268467755, 0.03100425268467753, 0.43681220349716127], atol=1e-13)
    rg = tree.marginal_asr(gtr_g, aln16, False)
    assert {k: rg.sequence_string(v) for k, v in names.items()} == {
        "E": "T" * 16, "C": "AGGT" * 4, "A": "ACGT" * 4, "B": "AGGT" * 4, "D": "T" * 16}
    assert r4.sequence_string(12345) is None

    # --- test_marginal_asr_likelihood_normalization (custom pi, root posteriors sum to 1)
    model = GtrModel.new([0.9, 0.06, 0.02, 0.02], np.ones((4, 4)), True)
    big = Alignment.from_fasta_bytes(
        b">A\n" + b"A" * 16 + b"C" * 16 + b"G" * 16 + b"T" * 16 + b"\n>B\n" + (b"AAAACCCCGGGGTTTT" * 4) +
        b"\n>C\n" + b"ACGT" * 16 + b"\n>D\n" + b"A" * 64 + b"\n")
    rec = tree.marginal_asr(model, big, True)
    assert math.isfinite(rec.log_likelihood)
    np.testing.assert_allclose(rec.posteriors(tree.get_root_id()).sum(1), 1.0, atol=1e-3)
    o = oracle.marginal_asr(parse_newick(TINY), RefModel([0.9, 0.06, 0.02, 0.02], np.ones((4, 4)), True),
                            RefAlignment.from_fasta_bytes(big_bytes(big)), want_posteriors=True)
    assert rec.log_likelihood == pytest.approx(o["log_likelihood"], rel=1e-12)
    for v in tree.get_node_ids():
        assert np.array_equal(rec.sequence(v), o["sequences"][v])
        np.testing.assert_allclose(rec.posteriors(v), o["posteriors"][v], atol=1e-12)

    # --- JointAsr: test_joint_asr_basic, test_joint_vs_marginal_same_sequences_simple, test_joint_asr_all_gap_alignment
    jr = tree.joint_asr(jc, aln16)
    assert not jr.has_posteriors() and jr.n_nodes() == 5
    oj = oracle.joint_asr(parse_newick(TINY), RefModel.jukes_cantor(), RefAlignment.from_fasta_bytes(big_bytes(aln16)))
    assert jr.log_likelihood == pytest.approx(oj["log_likelihood"], rel=1e-12)
    for v in tree.get_node_ids():
        assert np.array_equal(jr.sequence(v), oj["sequences"][v])
    assert jr.log_likelihood <= tree.log_likelihood(jc, aln16)  # a maximum over paths never exceeds their sum
    sim = Alignment.from_fasta_bytes(b">A\nAAAAAAAA\n>B\nAAAAAAAC\n>C\nAAAAAAAA\n>D\nCCCCCCCC\n")
    assert tree.joint_asr(jc, sim).n_nodes() == tree.marginal_asr(jc, sim, False).n_nodes()
    jg = tree.joint_asr(gtr_g.with_invariant(0.2), aln16)
    ojg = oracle.joint_asr(parse_newick(TINY), RefModel.gtr([.3, .2, .2, .3], [1, 2.5, 1, 1, 2.5, 1]).with_gamma(0.5, 4).with_invariant(0.2),
                           RefAlignment.from_fasta_bytes(big_bytes(aln16)))
    assert jg.log_likelihood == pytest.approx(ojg["log_likelihood"], rel=1e-12)
    for v in tree.get_node_ids():
        assert np.array_equal(jg.sequence(v), ojg["sequences"][v])

    # --- all-gap, zero branch lengths, N vs gap
    gap = Alignment.from_fasta_bytes(b">A\n----\n>B\n----\n>C\n----\n>D\n----\n")
    assert abs(tree.marginal_asr(jc, gap, False).log_likelihood) < 1e-12
    assert math.isfinite(tree.joint_asr(jc, gap).log_likelihood)
    zero = PhyloTree.from_newick(b"((A:0.0,B:0.0)C:0.0,D:0.0)E:0.0;")
    zaln = Alignment.from_fasta_bytes(b">A\nACGT\n>B\nAGGT\n>C\nAAAA\n>D\nTTTT\n")
    zr = zero.marginal_asr(jc, zaln, False)
    assert zr.log_likelihood == -math.inf  # conflicting leaves at distance 0: Ok(-inf), not an error
    assert zr.sequence_string(zero.get_root_id()) == "TTTT"  # last-max over all-zero posteriors
    zj = zero.joint_asr(jc, zaln)  # columns 0-2 conflict: no category has a value > -inf => state 0 everywhere
    assert zj.log_likelihood == -math.inf and zj.sequence_string(zero.get_root_id()) == "AAAT"
    assert zj.sequence_string([v for v in zero.get_node_ids() if zero.get_node_taxa(v) == "D"][0]) == "AAAT"
    three = PhyloTree.from_newick("((A:0.1,B:0.2):0.1,D:0.3);")
    a = three.log_likelihood(jc, Alignment.from_fasta_bytes(b">A\nN\n>B\nA\n>D\nC\n"))
    b = three.log_likelihood(jc, Alignment.from_fasta_bytes(b">A\n-\n>B\nA\n>D\nC\n"))
    assert a - b == pytest.approx(math.log(0.25), abs=1e-12)

    # --- errors raised by the device library keep their AsrError variant
    with pytest.raises(AsrError) as e:  # D missing: every row names a node, but leaf D has no row
        tree.log_likelihood(jc, Alignment.from_fasta_bytes(b">A\nACGT\n>B\nAGGT\n>C\nAAAA\n"))
    assert e.value.variant == "InvalidAlignment"
    with pytest.raises(AsrError) as e:
        tree.log_likelihood(jc, Alignment.from_fasta_bytes(b">A\nACGT\n>B\nAG?T\n>C\nAAAA\n>D\nTTTT\n"))
    assert e.value.variant == "AlphabetMismatch"
    # invalid bytes in a row naming an INTERNAL node are never read (Appendix A-4)
    assert math.isfinite(tree.log_likelihood(jc, Alignment.from_fasta_bytes(b">A\nACGT\n>B\nAGGT\n>C\n????\n>D\nTTTT\n")))

    # --- config-5 caller pattern: batch NNI scoring == clone + nni + log_likelihood
    t8 = PhyloTree.yule(8, seed=11)
    t8.assign_exponential_lengths(5)
    leaves = [v for v in t8.get_node_ids() if t8.is_leaf(v)]
    rng = np.random.default_rng(0)
    aln8 = Alignment.from_seqs({t8.get_node_taxa(v): bytes(rng.choice(list(b"ACGT"), 40).astype(np.uint8)) for v in leaves})
    moves = [(v, left) for v in t8.get_node_ids() for left in (False, True)]
    got = t8.score_nni_neighbours(gtr_g, aln8, moves)
    for (v, left), val in zip(moves, got):
        if t8.is_leaf(v) or v == t8.get_root_id():
            assert math.isnan(val)
        else:
            nb = t8.clone()
            nb.nni(v, left)
            assert val == pytest.approx(nb.log_likelihood(gtr_g, aln8), rel=1e-12)


def big_bytes(aln):
    c = aln.compress_columns()
    cols = c.patterns[c.site_to_pattern]  # [site][row]
    return b"".join(b">" + n.encode() + b"\n" + bytes(cols[:, r]) + b"\n" for r, n in enumerate(c.leaf_order))


def test_engine_suite_on_emulation_build(emu_backend):
    _engine_suite()


@pytest.mark.gpu
def test_engine_suite_on_b200(gpu_backend):
    _engine_suite()


@pytest.mark.gpu
def test_mirror_parity_yule_gtr_i_g4(gpu_backend):
    """A 200-taxon Yule tree through the mirror end to end (compress -> flatten -> device) vs the oracle."""
    t = PhyloTree.yule(200, seed=0x5EED0002)
    t.assign_exponential_lengths(0x5EED0002 ^ 0xA11C)
    ref = parse_newick(t.to_newick())  # the parser renumbers the nodes; post-order positions correspond
    to_ref = dict(zip(t.postord_ids(), ref.postord_ids()))
    assert len(to_ref) == t.num_nodes()
    rng = np.random.default_rng(5)
    leaves = [v for v in t.get_node_ids() if t.is_leaf(v)]
    seqs = {t.get_node_taxa(v): bytes(rng.choice(list(b"ACGTRYN-"), 2000, p=[.24, .24, .24, .24, .01, .01, .01, .01]).astype(np.uint8))
            for v in leaves}
    model = GtrModel.gtr(PI, [1.0, 2.5, 1.0, 1.0, 2.5, 1.0]).with_gamma(0.5, 4).with_invariant(0.2)
    rmodel = RefModel.gtr(PI, [1.0, 2.5, 1.0, 1.0, 2.5, 1.0]).with_gamma(0.5, 4).with_invariant(0.2)
    want = oracle.log_likelihood(ref, rmodel, RefAlignment(dict(seqs), 2000))
    got = t.log_likelihood(model, Alignment.from_seqs(seqs))
    assert abs(got - want) <= 1e-9 * abs(want)
    rec = t.marginal_asr(model, Alignment.from_seqs(seqs), False)
    o = oracle.marginal_asr(ref, rmodel, RefAlignment(dict(seqs), 2000), want_posteriors=True)
    assert abs(rec.log_likelihood - want) <= 1e-9 * abs(want)
    mism = 0
    for v in t.get_node_ids():
        d = rec.sequence(v) != np.asarray(o["sequences"][to_ref[v]])
        if d.any():
            srt = np.sort(np.asarray(o["posteriors"][to_ref[v]]), axis=1)
            assert np.all((srt[:, 3] - srt[:, 2])[d] < 1e-10)
            mism += int(d.sum())
    assert mism <= 2
