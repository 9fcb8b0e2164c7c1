"""Golden fixtures (tests/golden/*.json, written by tests/golden/make_golden.py).

CPU: the oracle still reproduces every fixture (kat.json pins the oracle; cases.json are its regression
vectors) and so does the full product path over the emulation build.
GPU: the CUDA path reproduces the fixtures through the reference-shaped interface with NO oracle in the loop.
Bars: total lnL 1e-9 relative, per-site lnL 1e-8 absolute, ASR strings bit-exact (all fixtures have a top-2
posterior gap > 1e-3, recorded in the fixture, so there are no ties to audit)."""
import json
import math
import os

import numpy as np
import pytest

import hostapi as api
from hostapi import Alignment, GtrModel, PhyloTree

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
KAT = json.load(open(os.path.join(HERE, "kat.json")))
CASES = json.load(open(os.path.join(HERE, "cases.json")))["cases"]
PI, RATES = [0.3, 0.2, 0.2, 0.3], [1.0, 2.5, 1.0, 1.0, 2.5, 1.0]


def mirror_model(name):
    return {
        "jc69": lambda: GtrModel.jukes_cantor(),
        "hky85": lambda: GtrModel.hky85(PI, 2.5),
        "hky_g4": lambda: GtrModel.hky85([0.25] * 4, 2.0).with_gamma(0.5, 4),
        "gtr": lambda: GtrModel.gtr(PI, RATES),
        "gtr_g4": lambda: GtrModel.gtr(PI, RATES).with_gamma(0.5, 4),
        "gtr_i_g4": lambda: GtrModel.gtr(PI, RATES).with_gamma(0.5, 4).with_invariant(0.2),
        "gtr_i": lambda: GtrModel.gtr(PI, RATES).with_invariant(0.3),
    }[name]()


def oracle_model(name):
    from oracle import RefModel, workloads

    return RefModel.hky85(PI, 2.5) if name == "hky85" else workloads.make_model(name)


# ---- CPU: the oracle against the fixtures ------------------------------------------------------------------
def test_oracle_reproduces_kat_table():
    import oracle
    from oracle import RefAlignment, parse_newick

    tiny = parse_newick(KAT["tiny_tree"])
    for key, model, aln in (("jc69_aln4", "jc69", "aln4"), ("jc69_aln16", "jc69", "aln16"), ("hky85_aln16", "hky85", "aln16"),
                            ("gtr_g4_aln16", "gtr_g4", "aln16"), ("gtr_i_g4_aln16", "gtr_i_g4", "aln16"),
                            ("jc69_aln8", "jc69", "aln8")):
        got = oracle.log_likelihood(tiny, oracle_model(model), RefAlignment.from_fasta_bytes(KAT[aln].encode()))
        assert got == pytest.approx(KAT["lnl"][key], abs=1e-11), key
    for alpha, want in KAT["discrete_gamma"].items():
        np.testing.assert_allclose(oracle.discrete_gamma(float(alpha), 4), want, rtol=1e-13)


def test_oracle_reproduces_cases():
    import oracle
    from oracle import RefAlignment, parse_newick

    for c in CASES:
        tree = parse_newick(c["newick"])
        aln = RefAlignment.from_fasta_bytes(c["fasta"].encode())
        m = oracle_model(c["model"])
        assert oracle.log_likelihood(tree, m, aln) == pytest.approx(c["log_likelihood"], rel=1e-13)
        r = oracle.marginal_asr(tree, m, aln)
        for nid, s in c["asr_sequences"].items():
            assert oracle.sequence_string(r["sequences"][int(nid)]) == s
        j = oracle.joint_asr(tree, m, aln)
        assert j["log_likelihood"] == pytest.approx(c["joint_log_likelihood"], rel=1e-13)
        for nid, s in c["joint_sequences"].items():
            assert oracle.sequence_string(j["sequences"][int(nid)]) == s
        pats, s2p, mult, _lo = aln.compress_columns()
        assert len(pats) == c["n_patterns"] and list(s2p) == c["site_to_pattern"] and list(mult) == c["multiplicity"]


# ---- product path against the fixtures ----------------------------------------------------------------------
def _product_vs_golden():
    tiny = PhyloTree.from_newick(KAT["tiny_tree"])
    names = {tiny.get_node_taxa(v): v for v in tiny.get_node_ids()}
    for key, model, aln in (("jc69_aln4", "jc69", "aln4"), ("jc69_aln16", "jc69", "aln16"), ("hky85_aln16", "hky85", "aln16"),
                            ("gtr_g4_aln16", "gtr_g4", "aln16"), ("gtr_i_g4_aln16", "gtr_i_g4", "aln16"),
                            ("jc69_aln8", "jc69", "aln8")):
        got = tiny.log_likelihood(mirror_model(model), Alignment.from_fasta_bytes(KAT[aln].encode()))
        assert abs(got - KAT["lnl"][key]) <= 1e-9 * abs(KAT["lnl"][key]), key
    doc = KAT["doc_example"]
    got = PhyloTree.from_newick(doc["tree"]).log_likelihood(mirror_model("hky_g4"), Alignment.from_fasta_bytes(doc["aln"].encode()))
    assert abs(got - KAT["lnl"]["doc_example"]) <= 1e-9 * abs(got)
    r = tiny.marginal_asr(mirror_model("jc69"), Alignment.from_fasta_bytes(KAT["aln4"].encode()), True)
    assert {k: r.sequence_string(v) for k, v in names.items()} == KAT["asr"]["jc69_aln4_states"]
    np.testing.assert_allclose(r.posteriors(names["E"])[0], KAT["asr"]["jc69_aln4_root_posterior_site0"], atol=1e-12)
    r = tiny.marginal_asr(mirror_model("gtr_g4"), Alignment.from_fasta_bytes(KAT["aln16"].encode()), False)
    assert {k: r.sequence_string(v) for k, v in names.items()} == KAT["asr"]["gtr_g4_aln16_states"]
    for key, want in KAT["transition_row_A"].items():
        m, t = {"hky85_pi.3.2.2.3_k2.5_t0.1": (GtrModel.hky85(PI, 2.5), 0.1),
                "gtr_pi.1.2.3.4_r123456_t0.25": (GtrModel.gtr([.1, .2, .3, .4], [1, 2, 3, 4, 5, 6]), 0.25),
                "jc69_t0.5": (GtrModel.jukes_cantor(), 0.5)}[key]
        np.testing.assert_allclose(m.category_transition(0, t)[0], want, atol=1e-14)

    for c in CASES:
        tree = PhyloTree.from_newick(c["newick"])
        aln = Alignment.from_fasta_bytes(c["fasta"].encode())
        model = mirror_model(c["model"])
        got = tree.log_likelihood(model, aln)
        assert abs(got - c["log_likelihood"]) <= 1e-9 * abs(c["log_likelihood"]), c["name"]
        rec = tree.marginal_asr(model, aln, False)
        assert abs(rec.log_likelihood - c["log_likelihood"]) <= 1e-9 * abs(c["log_likelihood"])
        assert c["asr_min_top2_gap"] > 1e-6
        for nid, s in c["asr_sequences"].items():
            assert rec.sequence_string(int(nid)) == s, (c["name"], nid)
        # joint ASR and the NNI neighbourhood through the reference-shaped interface
        jr = tree.joint_asr(model, aln)
        assert abs(jr.log_likelihood - c["joint_log_likelihood"]) <= 1e-9 * abs(c["joint_log_likelihood"])
        for nid, s in c["joint_sequences"].items():
            assert jr.sequence_string(int(nid)) == s, (c["name"], "joint", nid)
        moves = [(int(v), bool(lc)) for v, lc, _ in c["nni_scores"]]
        got_nni = tree.score_nni_neighbours(model, aln, moves)
        want_nni = np.asarray([x[2] for x in c["nni_scores"]])
        assert np.max(np.abs(got_nni - want_nni) / np.abs(want_nni)) <= 1e-9, c["name"]
        # per-site values through the raw C ABI (the reference API does not expose them)
        from phylo_rs_b200 import capi

        comp = aln.compress_columns()
        lib = capi.Library(_current_lib_path())
        with capi.Context(library=lib) as ctx:
            _load_raw(ctx, tree, model, comp)
            tot, per = ctx.log_likelihood(per_pattern=True)
        assert math.isclose(tot, c["log_likelihood"], rel_tol=1e-9)
        assert np.max(np.abs(per[comp.site_to_pattern] - np.asarray(c["per_site"]))) <= 1e-8
        # device-side compress_columns against the fixture (bit-exact)
        fa = dict(zip(c["fasta"].split()[0::2], c["fasta"].split()[1::2]))
        rows = [fa[">" + nm].encode() for nm in comp.leaf_order]
        with capi.Context(library=lib) as ctx:
            n, s2p, mult = ctx.compress_rows(rows)
        assert n == c["n_patterns"] and list(s2p) == c["site_to_pattern"] and list(mult) == c["multiplicity"]


_LIB_PATH = [None]


def _current_lib_path():
    return _LIB_PATH[0]


def _load_raw(ctx, tree, model, comp):
    """The flattening a binding does, here for the raw-ABI per-pattern check."""
    ids = tree.get_node_ids()
    dense = {v: k for k, v in enumerate(ids)}
    leaf_row = np.full(len(ids), -1, np.int32)
    for r, name in enumerate(comp.leaf_order):
        v = tree.get_taxa_node_id(name)
        if tree.is_leaf(v) and leaf_row[dense[v]] < 0:
            leaf_row[dense[v]] = r
    off, kids, blen = [0], [], []
    for v in ids:
        kids += [dense[c] for c in tree.get_node_children_ids(v)]
        off.append(len(kids))
        w = tree.get_edge_weight(None, v)
        blen.append(0.0 if w is None else float(w))
    lam, V = model.eigen()
    pi = model.equilibrium()
    cats = model.categories()
    ctx.set_alignment(comp.patterns, multiplicity=comp.multiplicity)
    ctx.set_model(lam, V, pi, np.sqrt(pi), 1.0 / np.sqrt(pi), [c[0] for c in cats], [c[1] for c in cats])
    ctx.set_tree([dense[v] for v in tree.postord_ids()], off, kids, blen, leaf_row, dense[tree.get_root_id()])


def test_product_path_reproduces_golden_on_emulation_build(emu_lib):
    api.set_library(emu_lib.path)
    _LIB_PATH[0] = emu_lib.path
    try:
        _product_vs_golden()
    finally:
        api.set_library(None)


@pytest.mark.gpu
def test_cuda_path_reproduces_golden(gpu_lib):
    api.set_library(None)
    _LIB_PATH[0] = gpu_lib.path
    _product_vs_golden()
    api.release_backend()
