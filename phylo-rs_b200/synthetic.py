"""Synthetic benchmark inputs: seeded Yule trees with branch lengths and GTR+I+G models,
flattened into the arrays the C ABI takes.  Host-side harness code for bench.py — no likelihood
arithmetic happens here; the alignment itself is simulated on the device (phb_simulate_alignment).

Yule process as the reference's simulator (src/tree.rs:505-553): root 0 with leaves 1, 2; then for
i in 2..n pick a uniformly random current leaf, split its parent edge with a new internal node
(appended LAST among the parent's children, src/tree/simple_rtree.rs:143-153) and hang the new leaf
under it.  The reference assigns no branch lengths and uses an unseeded RNG; here the RNG is seeded and
lengths are i.i.d. Exp(mean 0.1) clamped to [1e-4, 2] and rounded to f32 (PhyloTree stores f32).
"""
from __future__ import annotations

import numpy as np


def yule_flat(n_taxa, seed):
    rng = np.random.default_rng(seed)
    n_nodes = 2 * n_taxa - 1
    parent = np.full(n_nodes, -1, np.int64)
    kids = [[] for _ in range(n_nodes)]
    kids[0] = [1, 2]
    parent[1] = parent[2] = 0
    leaves = [1, 2]
    nxt = 3
    picks = rng.random(n_taxa)
    for i in range(2, n_taxa):
        leaf = leaves[int(picks[i] * len(leaves))]
        p = parent[leaf]
        s, new = nxt, nxt + 1
        nxt += 2
        kids[p].remove(leaf)
        kids[p].append(s)
        parent[s] = p
        kids[s] = [leaf, new]
        parent[leaf] = s
        parent[new] = s
        leaves.append(new)
    blen = np.clip(rng.exponential(0.1, n_nodes), 1e-4, 2.0).astype(np.float32).astype(np.float64)
    blen[0] = 0.0
    child_off = np.zeros(n_nodes + 1, np.int32)
    children = []
    for v in range(n_nodes):
        children.extend(kids[v])
        child_off[v + 1] = len(children)
    # DFSPostOrderIterator order (src/iter.rs:112-138): left-to-right post-order
    post = []
    stack = [(0, 0)]
    while stack:
        v, k = stack.pop()
        if k < len(kids[v]):
            stack.append((v, k + 1))
            stack.append((kids[v][k], 0))
        else:
            post.append(v)
    is_leaf = child_off[1:] == child_off[:-1]
    leaf_row = np.full(n_nodes, -1, np.int32)
    leaf_row[is_leaf] = np.arange(int(is_leaf.sum()), dtype=np.int32)  # rows = leaves in ascending id
    return dict(n_nodes=n_nodes, n_taxa=n_taxa, postorder=np.asarray(post, np.int32), child_off=child_off,
                children=np.asarray(children, np.int32), parent=parent.astype(np.int32), blen=blen,
                leaf_row=leaf_row, root=0)


def _discrete_gamma(alpha, k):
    """Mean-of-bin discrete gamma (Yang 1994), the scheme of src/models/gamma.rs:15-55."""
    from scipy.special import gammainc, gammaincinv

    if k == 1:
        return np.ones(1)
    q = gammaincinv(alpha, np.arange(1, k) / k) / alpha       # bin boundaries of Gamma(alpha, rate alpha)
    cdf = np.concatenate(([0.0], gammainc(alpha + 1.0, q * alpha), [1.0]))
    return np.diff(cdf) * k


def gtr_model(pi=(0.30, 0.20, 0.20, 0.30), rates=(1.0, 2.5, 1.0, 1.0, 2.5, 1.0), alpha=None, k=4, p_inv=0.0):
    """Eigen-system + rate categories as GtrModel exposes them (src/models/rate_matrix.rs:34-110,
    src/models/gtr.rs:85-165): Q_ij = pi_j W_ij normalised to mean rate 1, S = D Q D^-1 symmetric,
    S = V diag(lambda) V^T.  Category order: [invariant, gamma_1..gamma_k]."""
    pi = np.asarray(pi, np.float64)
    pi = pi / pi.sum()
    w = np.zeros((4, 4))
    for (i, j), r in zip(((0, 1), (0, 2), (0, 3), (1, 2), (1, 3), (2, 3)), rates):
        w[i, j] = w[j, i] = r
    q = w * pi[None, :]
    np.fill_diagonal(q, 0.0)
    np.fill_diagonal(q, -q.sum(1))
    q /= -(pi * np.diag(q)).sum()
    sp, isp = np.sqrt(pi), 1.0 / np.sqrt(pi)
    s = sp[:, None] * q * isp[None, :]
    lam, vec = np.linalg.eigh(0.5 * (s + s.T))
    cats = []
    if p_inv > 0:
        cats.append((0.0, p_inv))
    if alpha is None:
        cats.append((1.0, 1.0 - p_inv))
    else:
        for r in _discrete_gamma(alpha, k):
            cats.append((float(r), (1.0 - p_inv) / k))
    return dict(eigenvalues=lam, eigenvectors=vec, pi=pi, sqrt_pi=sp, inv_sqrt_pi=isp,
                rates=np.array([c[0] for c in cats]), weights=np.array([c[1] for c in cats]))


def load_into(ctx, tree, model):
    ctx.set_model(model["eigenvalues"], model["eigenvectors"], model["pi"], model["sqrt_pi"], model["inv_sqrt_pi"],
                  model["rates"], model["weights"])
    ctx.set_tree(tree["postorder"], tree["child_off"], tree["children"], tree["blen"], tree["leaf_row"], tree["root"])
