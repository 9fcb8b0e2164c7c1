"""CPU ORACLE — seeded synthetic inputs for the parity tests (test infrastructure, NOT product code).

Trees come from the restated Yule simulator (oracle/refhost.py: src/tree.rs:505-553) with i.i.d.
Exp(mean 0.1) branch lengths clamped to [1e-4, 2.0] and rounded to f32 (the reference stores f32,
src/tree.rs:45); alignments are simulated down the tree under the model with numpy.  SURVEY.md §8d.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from . import refhost

MODELS = ("jc69", "hky_g4", "gtr", "gtr_g4", "gtr_i_g4", "gtr_i")
PI = (0.30, 0.20, 0.20, 0.30)
GTR_RATES = (1.0, 2.5, 1.0, 1.0, 2.5, 1.0)
AMBIG = b"RYSWKMBDHVN-U"
AA_LETTERS = b"ACDEFGHIKLMNPQRSTVWY"  # AminoAcid::CANONICAL, src/alphabet.rs:73
AA_AMBIG = b"BZJX-"
AA_MODELS = ("aa", "aa_g4", "aa_i_g4")


def make_model(name):
    from . import RefModel

    if name == "jc69":
        return RefModel.jukes_cantor()
    if name == "hky_g4":
        return RefModel.hky85([0.25] * 4, 2.0).with_gamma(0.5, 4)
    if name == "gtr":
        return RefModel.gtr(PI, GTR_RATES)
    if name == "gtr_g4":
        return RefModel.gtr(PI, GTR_RATES).with_gamma(0.5, 4)
    if name == "gtr_i_g4":
        return RefModel.gtr(PI, GTR_RATES).with_gamma(0.5, 4).with_invariant(0.2)
    if name == "gtr_i":
        return RefModel.gtr(PI, GTR_RATES).with_invariant(0.3)
    if name in AA_MODELS:
        return make_aa_model(name)
    raise ValueError(name)


def make_aa_model(name, seed=0xAA20):
    """A seeded general 20-state GTR (GtrModel::<AminoAcid>::new(pi, W, true), src/models/gtr.rs:39-50): the reference
    ships no named protein matrices, the caller brings pi and the symmetric exchangeabilities."""
    from . import AaModel

    rng = np.random.default_rng(seed)
    pi = rng.dirichlet(np.full(20, 8.0))
    w = rng.gamma(1.5, 1.0, (20, 20)) + 0.05
    w = (w + w.T) / 2
    np.fill_diagonal(w, 0.0)
    m = AaModel(pi, w, True)
    if name == "aa_g4":
        m = m.with_gamma(0.6, 4)
    elif name == "aa_i_g4":
        m = m.with_gamma(0.6, 4).with_invariant(0.15)
    return m


def assign_branch_lengths(tree, rng, mean=0.1):
    for nid in tree.get_node_ids():
        if nid != tree.root:
            t = min(max(rng.exponential(mean), 1e-4), 2.0)
            tree.set_edge_weight(nid, np.float32(t))


def simulate_patterns(ft, model, n_sites, rng):
    """[n_sites][rows] ASCII bytes for the leaves of the flattened tree (rows = leaf_row order)."""
    n = int(ft["n_nodes"])
    rates, weights = model.rates, model.weights
    cat = rng.choice(len(rates), size=n_sites, p=weights / weights.sum())
    state = np.zeros((n, n_sites), np.int8)
    pi = model.pi
    n_rows = int(ft["leaf_row"].max()) + 1
    out = np.full((n_sites, n_rows), ord("-"), np.uint8)
    ns = len(pi)
    letters = np.frombuffer(b"ACGT" if ns == 4 else AA_LETTERS, np.uint8)
    for v in ft["preorder"]:
        if v == ft["root"]:
            state[v] = rng.choice(ns, size=n_sites, p=pi / pi.sum())
        else:
            par = state[ft["parent"][v]]
            u = rng.random(n_sites)
            s = np.zeros(n_sites, np.int8)
            for c in range(len(rates)):
                sel = cat == c
                if not sel.any():
                    continue
                P = model.transition(rates[c] * ft["blen"][v])
                cum = np.cumsum(np.clip(P, 0, None), axis=1)
                cum /= cum[:, -1:]
                s[sel] = (u[sel, None] > cum[par[sel]]).sum(axis=1).clip(0, ns - 1)
            state[v] = s
        r = ft["leaf_row"][v]
        if r >= 0:
            out[:, r] = letters[state[v]]
    return out


@dataclass
class Case:
    tree: object
    ft: dict
    model: object
    model_name: str
    patterns: np.ndarray  # [P][rows] ASCII
    mult: np.ndarray      # [P] uint64
    leaf_order: list


def make_case(n_taxa, n_sites, model="gtr_g4", seed=1, ambiguity=0.0, compress=False):
    rng = np.random.default_rng(seed)
    tree = refhost.yule(n_taxa, rng)
    assign_branch_lengths(tree, rng)
    leaf_order = [tree.nodes[i].taxa for i in tree.get_node_ids() if tree.nodes[i].taxa is not None]
    ft = flatten_fast(tree, leaf_order)
    m = make_model(model)
    pats = simulate_patterns(ft, m, n_sites, rng)
    if ambiguity > 0:
        mask = rng.random(pats.shape) < ambiguity
        codes = np.frombuffer(AA_AMBIG if model in AA_MODELS else AMBIG, np.uint8)
        pats[mask] = codes[rng.integers(len(codes), size=int(mask.sum()))]
    if compress:
        uniq, inv, counts = np.unique(pats, axis=0, return_inverse=True, return_counts=True)
        first = np.full(len(uniq), len(pats), np.int64)
        np.minimum.at(first, inv.ravel(), np.arange(len(pats)))
        order = np.argsort(first, kind="stable")  # first-appearance order (src/alignment.rs:104-120)
        pats, mult = uniq[order], counts[order].astype(np.uint64)
    else:
        mult = np.ones(len(pats), np.uint64)
    return Case(tree, ft, m, model, np.ascontiguousarray(pats), mult, leaf_order)


def polytomy_case(k, depth, n_sites, model="gtr_g4", seed=1, blen=0.9):
    """A root with k children.  depth >= 2: every child is a `depth`-leaf caterpillar (deep subtrees under a
    high-arity node: each child's CLV carries a large scale); depth == 1: a star of k tips
    (the reference's own product of k contributions underflows for k >~ 460: likelihood.rs:123-125).
    All branches `blen`, i.i.d. uniform ACGT columns."""
    from . import parse_newick

    names = []

    def cat(j):
        if depth == 1:
            names.append(f"s{j}")
            return f"s{j}:{blen}"
        names.append(f"c{j}_0")
        names.append(f"c{j}_1")
        s = f"(c{j}_0:{blen},c{j}_1:{blen})"
        for i in range(2, depth):
            names.append(f"c{j}_{i}")
            s = f"({s}:{blen},c{j}_{i}:{blen})"
        return s + f":{blen}"

    nwk = "(" + ",".join(cat(j) for j in range(k)) + ");"
    tree = parse_newick(nwk)
    rng = np.random.default_rng(seed)
    ft = flatten_fast(tree, names)
    pats = np.frombuffer(b"ACGT", np.uint8)[rng.integers(4, size=(n_sites, len(names)))]
    return Case(tree, ft, make_model(model), model, np.ascontiguousarray(pats), np.ones(n_sites, np.uint64), names)


def flatten_fast(tree, leaf_order):
    """refhost.flatten with a dict for the name lookup (same result when labels are unique)."""
    by_name = {}
    for n in tree.nodes:
        if n is not None and n.taxa is not None and n.taxa not in by_name:
            by_name[n.taxa] = n.id
    saved = tree.get_taxa_node_id
    tree.get_taxa_node_id = lambda name: by_name.get(name)
    try:
        return refhost.flatten(tree, leaf_order)
    finally:
        tree.get_taxa_node_id = saved


def load_into(ctx, case):
    """Hand a Case to a phylo_rs_b200.capi.Context (the C-ABI product library)."""
    m, ft = case.model, case.ft
    ctx.set_alignment(case.patterns, multiplicity=case.mult)
    ctx.set_model(m.eigenvalues, m.eigenvectors, m.pi, m.sqrt_pi, m.inv_sqrt_pi, m.rates, m.weights)
    ctx.set_tree(ft["postorder"], ft["child_off"], ft["children"], ft["blen"], ft["leaf_row"], ft["root"])
