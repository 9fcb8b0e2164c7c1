"""CPU ORACLE package (test infrastructure, NOT product code).

ctypes front-end of oracle/phylo_oracle.c plus the model constructors of the reference
(src/models/gtr.rs) restated on top of it.  Only tests/, __graft_entry__.smoke() and bench.py's
CPU-baseline legs may import this package; nothing under phylo-rs_b200/ does.

Parity status: see the header of phylo_oracle.c ("parity unpinned" by reference-blessed absolute
values; pinned on the reference's own test invariants + brute force + SURVEY.md §8c KATs).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from . import refhost
from .refhost import RefAlignment, RefTree, parse_newick, yule, flatten  # noqa: F401

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    """Compile the C restatement (gcc).  Called by __graft_entry__.build()."""
    src = os.path.join(_HERE, "phylo_oracle.c")
    outs = [os.path.join(_HERE, "libphylo_oracle.so"), os.path.join(_HERE, "libphylo_oracle_v3.so"),
            os.path.join(_HERE, "libphylo_oracle_aa.so")]
    stale = force or any((not os.path.exists(o)) or os.path.getmtime(o) < os.path.getmtime(src) for o in outs)
    if stale:
        subprocess.run(["make", "-C", _HERE, "-B", "all"], check=True, capture_output=True)
    return outs


def _has_avx2():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return " avx2 " in line + " " and " fma " in line + " "
    except OSError:
        pass
    return False


def _rate_matrix_struct(ns):
    class RM(C.Structure):
        _fields_ = [("pi", C.c_double * ns), ("lam", C.c_double * ns), ("V", C.c_double * (ns * ns)),
                    ("sqrt_pi", C.c_double * ns), ("inv_sqrt_pi", C.c_double * ns)]
    return RM


_RateMatrix = _rate_matrix_struct(4)
_RateMatrixAA = _rate_matrix_struct(20)  # libphylo_oracle_aa.so (-DNS=20)
_LIB_AA = None


def _bind(L, RM, profile_fn):
    dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
    L.orc_discrete_gamma.argtypes = [C.c_double, C.c_int, dp]
    L.orc_discrete_gamma.restype = C.c_int
    L.orc_rate_matrix_new.argtypes = [dp, dp, C.c_int, C.POINTER(RM)]
    L.orc_rate_matrix_new.restype = C.c_int
    L.orc_transition.argtypes = [C.POINTER(RM), C.c_double, dp]
    L.orc_transition.restype = None
    getattr(L, profile_fn).argtypes = [C.c_uint8, dp]
    getattr(L, profile_fn).restype = C.c_int
    for nm in ("orc_test_ln_gamma", "orc_test_inverse_normal_cdf"):
        getattr(L, nm).argtypes = [C.c_double]
        getattr(L, nm).restype = C.c_double
    L.orc_test_incomplete_gamma.argtypes = [C.c_double, C.c_double]
    L.orc_test_incomplete_gamma.restype = C.c_double
    tree_args = [C.c_int, ip, ip, ip, ip, ip, dp, ip, C.c_int]
    model_args = [C.POINTER(RM), C.c_int, dp, dp]
    aln_args = [C.c_int, C.c_int64, C.POINTER(C.c_uint8), C.POINTER(C.c_uint64)]
    L.orc_log_likelihood.argtypes = tree_args + model_args + aln_args + [C.c_int, C.c_int, dp, dp]
    L.orc_log_likelihood.restype = C.c_int
    L.orc_marginal_asr.argtypes = tree_args + model_args + aln_args + [C.c_int, C.POINTER(C.c_uint8), dp, dp, dp]
    L.orc_marginal_asr.restype = C.c_int
    L.orc_joint_asr.argtypes = tree_args + model_args + aln_args + [C.c_int, C.POINTER(C.c_uint8), dp, dp]
    L.orc_joint_asr.restype = C.c_int
    L.orc_bruteforce_site_likelihood.argtypes = [C.c_int, ip, ip, ip, dp, ip, C.c_int,
                                                 C.POINTER(RM), C.c_double, C.POINTER(C.c_uint8)]
    L.orc_bruteforce_site_likelihood.restype = C.c_double
    L.orc_n_states.argtypes = []
    L.orc_n_states.restype = C.c_int


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    base = os.path.join(_HERE, "libphylo_oracle.so")
    v3 = os.path.join(_HERE, "libphylo_oracle_v3.so")
    if not os.path.exists(base):
        build()
    path = v3 if (_has_avx2() and os.path.exists(v3)) else base
    L = C.CDLL(path)
    _bind(L, _RateMatrix, "orc_nucleotide_profile")
    _LIB = L
    return L


def lib_aa():
    """The 20-state (AminoAcid, src/alphabet.rs:66-116) build of the same C restatement."""
    global _LIB_AA
    if _LIB_AA is None:
        path = os.path.join(_HERE, "libphylo_oracle_aa.so")
        if not os.path.exists(path):
            build(force=True)
        L = C.CDLL(path)
        _bind(L, _RateMatrixAA, "orc_amino_acid_profile")
        assert L.orc_n_states() == 20
        _LIB_AA = L
    return _LIB_AA


ERRORS = {1: "MissingBranchLength", 2: "AlphabetMismatch", 3: "InvalidAlignment",
          4: "EigendecompFailure", 5: "NumericalInstability", 6: "InvalidModelParameter"}


class OracleError(ValueError):
    def __init__(self, code, msg=""):
        self.code = code
        self.variant = ERRORS.get(code, f"code{code}")
        super().__init__(f"{self.variant}: {msg}" if msg else self.variant)


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int))


def discrete_gamma(alpha, k):
    """discrete_gamma, src/models/gamma.rs:15-55."""
    out = np.zeros(max(k, 1), np.float64)
    rc = lib().orc_discrete_gamma(float(alpha), int(k), _dp(out))
    if rc:
        raise OracleError(rc, "discrete_gamma")
    return out[:k]


def _symmetric_matrix(ac, ag, at, cg, ct, gt):
    """symmetric_matrix, src/models/gtr.rs:291-306 (state order A,C,G,T)."""
    w = np.zeros((4, 4))
    for i, j, r in ((0, 1, ac), (0, 2, ag), (0, 3, at), (1, 2, cg), (1, 3, ct), (2, 3, gt)):
        w[i, j] = w[j, i] = r
    return w


def _validate_positive(v, name):
    """validate_positive, src/models/gtr.rs:280-289."""
    if v <= 0.0 or not np.isfinite(v):
        raise OracleError(6, f"{name} must be positive and finite")


class RefModel:
    """GtrModel<Nucleotide>, src/models/gtr.rs:30-34 (+ RateMatrix, src/models/rate_matrix.rs)."""

    NS = 4
    _RM = _RateMatrix
    _lib = staticmethod(lambda: lib())

    def __init__(self, pi, w, normalize=True, _rm=None, _cats=None):
        if _rm is not None:
            self.rm, self.categories = _rm, list(_cats)
            return
        pi = np.ascontiguousarray(pi, np.float64)
        w = np.ascontiguousarray(w, np.float64)
        if pi.shape != (self.NS,):
            raise OracleError(2, "pi length does not match alphabet states")
        if w.shape != (self.NS, self.NS):
            raise OracleError(2, "W matrix dimensions do not match alphabet states")
        self.rm = self._RM()
        rc = self._lib().orc_rate_matrix_new(_dp(pi), _dp(w), int(bool(normalize)), C.byref(self.rm))
        if rc:
            raise OracleError(rc, "RateMatrix::new")
        self.categories = [(1.0, 1.0)]  # (rate, weight), src/models/gtr.rs:43-46

    # accessors ------------------------------------------------------------------------------
    @property
    def pi(self):
        return np.array(self.rm.pi[:])

    @property
    def eigenvalues(self):
        return np.array(self.rm.lam[:])

    @property
    def eigenvectors(self):
        """4x4, [i, k] = component i of eigenvector k."""
        return np.array(self.rm.V[:]).reshape(self.NS, self.NS).T.copy()

    @property
    def sqrt_pi(self):
        return np.array(self.rm.sqrt_pi[:])

    @property
    def inv_sqrt_pi(self):
        return np.array(self.rm.inv_sqrt_pi[:])

    @property
    def rates(self):
        return np.array([c[0] for c in self.categories], np.float64)

    @property
    def weights(self):
        return np.array([c[1] for c in self.categories], np.float64)

    def n_categories(self):
        return len(self.categories)

    def transition(self, t):
        """RateMatrix::transition, src/models/rate_matrix.rs:113-135 -> 4x4 row-major."""
        P = np.zeros(self.NS * self.NS)
        self._lib().orc_transition(C.byref(self.rm), float(t), _dp(P))
        return P.reshape(self.NS, self.NS)

    def category_transition(self, cat, t):
        """src/models/gtr.rs:61-64."""
        return self.transition(self.categories[cat][0] * float(t))

    # named constructors, src/models/gtr.rs:52-57,182-278 --------------------------------------
    @staticmethod
    def jukes_cantor():
        return RefModel([0.25] * 4, np.ones((4, 4)), True)

    @staticmethod
    def k80(kappa):
        _validate_positive(kappa, "kappa")
        return RefModel([0.25] * 4, _symmetric_matrix(1.0, kappa, 1.0, 1.0, kappa, 1.0), True)

    @staticmethod
    def f81(pi):
        return RefModel(pi, np.ones((4, 4)), True)

    @staticmethod
    def hky85(pi, kappa):
        _validate_positive(kappa, "kappa")
        return RefModel(pi, _symmetric_matrix(1.0, kappa, 1.0, 1.0, kappa, 1.0), True)

    @staticmethod
    def tn93(pi, kappa1, kappa2):
        _validate_positive(kappa1, "kappa1")
        _validate_positive(kappa2, "kappa2")
        return RefModel(pi, _symmetric_matrix(1.0, kappa1, 1.0, 1.0, kappa2, 1.0), True)

    @staticmethod
    def k81(a, b, c):
        """src/models/gtr.rs:207-216: AC=b, AG=a, AT=c, CG=c, CT=a, GT=b."""
        for nm, v in (("a", a), ("b", b), ("c", c)):
            _validate_positive(v, nm)
        return RefModel([0.25] * 4, _symmetric_matrix(b, a, c, c, a, b), True)

    @staticmethod
    def tim(pi, kappa1, kappa2, a, b):
        """src/models/gtr.rs:221-230: AC=a, AG=kappa1, AT=b, CG=b, CT=kappa2, GT=a."""
        for nm, v in (("kappa1", kappa1), ("kappa2", kappa2), ("a", a), ("b", b)):
            _validate_positive(v, nm)
        return RefModel(pi, _symmetric_matrix(a, kappa1, b, b, kappa2, a), True)

    @staticmethod
    def tvm(pi, kappa, a, b, c):
        """src/models/gtr.rs:234-244: GT is the geometric mean (a*b*c)^(1/3)."""
        for nm, v in (("kappa", kappa), ("a", a), ("b", b), ("c", c)):
            _validate_positive(v, nm)
        return RefModel(pi, _symmetric_matrix(a, kappa, b, c, kappa, float(np.cbrt(a * b * c))), True)

    @staticmethod
    def sym(rates):
        """src/models/gtr.rs:248-259."""
        for nm, r in zip(("ac", "ag", "at", "cg", "ct", "gt"), rates):
            _validate_positive(r, nm)
        return RefModel([0.25] * 4, _symmetric_matrix(*rates), True)

    @staticmethod
    def gtr(pi, rates):
        for nm, r in zip(("ac", "ag", "at", "cg", "ct", "gt"), rates):
            _validate_positive(r, nm)
        return RefModel(pi, _symmetric_matrix(*rates), True)

    def _proportion_invariant(self):
        return sum(w for r, w in self.categories if r == 0.0)

    def with_gamma(self, alpha, k):
        """src/models/gtr.rs:85-112."""
        p_inv = self._proportion_invariant()
        rates = discrete_gamma(alpha, k)
        vw = 1.0 - p_inv
        cats = []
        if p_inv > 0.0:
            cats.append((0.0, p_inv))
        for r in rates:
            cats.append((float(r), vw / float(k)))
        return type(self)(None, None, _rm=self.rm, _cats=cats)

    def with_invariant(self, p_inv):
        """src/models/gtr.rs:117-165."""
        if not (0.0 <= p_inv < 1.0):
            raise OracleError(6, "proportion of invariant sites must be in [0, 1)")
        old = self._proportion_invariant()
        var = [(r, w) for r, w in self.categories if r != 0.0]
        old_vw, new_vw = 1.0 - old, 1.0 - p_inv
        cats = []
        if p_inv > 0.0:
            cats.append((0.0, p_inv))
        for r, w in var:
            scale = new_vw / old_vw if old_vw > 0.0 else 0.0
            cats.append((r, w * scale))
        if all(r == 0.0 for r, _ in cats):
            cats.append((1.0, new_vw))
        return type(self)(None, None, _rm=self.rm, _cats=cats)


class AaModel(RefModel):
    """GtrModel<AminoAcid> (src/models/gtr.rs:30-58 with A::N_STATES = 20): GtrModel::new(pi, W, normalize) and the
    rate-heterogeneity builders; the named constructors of gtr.rs:180-278 are Nucleotide-only in the reference too."""
    NS = 20
    _RM = _RateMatrixAA
    _lib = staticmethod(lambda: lib_aa())
    CANONICAL = b"ACDEFGHIKLMNPQRSTVWY"  # src/alphabet.rs:73


def amino_acid_profile(c):
    """AminoAcid::profile, src/alphabet.rs:76-116 -> 20 values or None."""
    p = np.zeros(20)
    return p if lib_aa().orc_amino_acid_profile(int(c), _dp(p)) else None


def _tree_args(ft):
    return [int(ft["n_nodes"]), _ip(ft["postorder"]), _ip(ft["preorder"]), _ip(ft["child_off"]),
            _ip(ft["children"]), _ip(ft["parent"]), _dp(ft["blen"]), _ip(ft["leaf_row"]), int(ft["root"])]


def log_likelihood_flat(ft, model, patterns, mult=None, mode=0, n_threads=1):
    """compute_log_likelihood (src/tree/likelihood.rs:156-209) on a flattened tree.
    patterns: uint8 [P][rows] ASCII.  Returns (total, per_pattern)."""
    patterns = np.ascontiguousarray(patterns, np.uint8)
    P, rows = patterns.shape
    rates, weights = np.ascontiguousarray(model.rates), np.ascontiguousarray(model.weights)
    per = np.zeros(max(P, 1), np.float64)
    tot = C.c_double(0.0)
    m = None if mult is None else np.ascontiguousarray(mult, np.uint64)
    rc = model._lib().orc_log_likelihood(
        *_tree_args(ft), C.byref(model.rm), len(rates), _dp(rates), _dp(weights), rows, P,
        patterns.ctypes.data_as(C.POINTER(C.c_uint8)),
        None if m is None else m.ctypes.data_as(C.POINTER(C.c_uint64)),
        int(mode), int(n_threads), C.byref(tot), _dp(per))
    if rc:
        raise OracleError(rc)
    return tot.value, per[:P]


def marginal_asr_flat(ft, model, patterns, mult=None, want_posteriors=False, n_threads=1):
    """compute_marginal_asr (src/tree/likelihood.rs:224-408) on a flattened tree.
    Returns (total, per_pattern, states [node][P] uint8, posteriors [node][P][4] or None)."""
    patterns = np.ascontiguousarray(patterns, np.uint8)
    P, rows = patterns.shape
    n = int(ft["n_nodes"])
    rates, weights = np.ascontiguousarray(model.rates), np.ascontiguousarray(model.weights)
    per = np.zeros(max(P, 1), np.float64)
    states = np.zeros((n, max(P, 1)), np.uint8)
    post = np.zeros((n, max(P, 1), model.NS), np.float64) if want_posteriors else None
    tot = C.c_double(0.0)
    m = None if mult is None else np.ascontiguousarray(mult, np.uint64)
    rc = model._lib().orc_marginal_asr(
        *_tree_args(ft), C.byref(model.rm), len(rates), _dp(rates), _dp(weights), rows, P,
        patterns.ctypes.data_as(C.POINTER(C.c_uint8)),
        None if m is None else m.ctypes.data_as(C.POINTER(C.c_uint64)),
        int(n_threads), states.ctypes.data_as(C.POINTER(C.c_uint8)),
        None if post is None else _dp(post), C.byref(tot), _dp(per))
    if rc:
        raise OracleError(rc)
    return tot.value, per[:P], states[:, :P], (None if post is None else post[:, :P])


def joint_asr_flat(ft, model, patterns, mult=None, n_threads=1):
    """compute_joint_asr (src/tree/likelihood.rs:421-560) on a flattened tree.
    Returns (total, per_pattern best ln(weight)+Viterbi value, states [node][P] uint8)."""
    patterns = np.ascontiguousarray(patterns, np.uint8)
    P, rows = patterns.shape
    n = int(ft["n_nodes"])
    rates, weights = np.ascontiguousarray(model.rates), np.ascontiguousarray(model.weights)
    per = np.zeros(max(P, 1), np.float64)
    states = np.zeros((n, max(P, 1)), np.uint8)
    tot = C.c_double(0.0)
    m = None if mult is None else np.ascontiguousarray(mult, np.uint64)
    rc = model._lib().orc_joint_asr(
        *_tree_args(ft), C.byref(model.rm), len(rates), _dp(rates), _dp(weights), rows, P,
        patterns.ctypes.data_as(C.POINTER(C.c_uint8)),
        None if m is None else m.ctypes.data_as(C.POINTER(C.c_uint64)),
        int(n_threads), states.ctypes.data_as(C.POINTER(C.c_uint8)), C.byref(tot), _dp(per))
    if rc:
        raise OracleError(rc)
    return tot.value, per[:P], states[:, :P]


def joint_asr(tree, model, aln):
    """JointAsr::joint_asr for PhyloTree (src/tree.rs:58-66) -> dict like Reconstruction (posteriors None)."""
    patterns, s2p, mult, leaf_order = aln.compress_columns()
    ft = flatten(tree, leaf_order)
    tot, _per, states = joint_asr_flat(ft, model, patterns, mult)
    seqs = {}
    for k, nid in enumerate(ft["ids"]):
        seqs[int(nid)] = states[k][s2p].astype(np.int64) if len(s2p) else np.zeros(0, np.int64)
    return dict(sequences=seqs, posteriors=None, log_likelihood=tot)


def joint_assignment_value(ft, model, pattern, states):
    """max over categories of ln w_c + ln pi[root state] + sum over edges ln P_c[s_parent][s_child] + sum over
    leaves ln profile[s_leaf]: the joint value compute_joint_asr maximises, evaluated for ONE given assignment
    (pure Python; used to audit ties)."""
    n = int(ft["n_nodes"])
    co, ch, blen, leaf_row, root = ft["child_off"], ft["children"], ft["blen"], ft["leaf_row"], int(ft["root"])
    best = -np.inf
    with np.errstate(divide="ignore", invalid="ignore"):
        for r, w in zip(model.rates, model.weights):
            val = np.log(w) + np.log(model.rm.pi[int(states[root])])
            for v in range(n):
                for k in range(co[v], co[v + 1]):
                    c = int(ch[k])
                    ns = model.NS
                    P = np.zeros(ns * ns)
                    model._lib().orc_transition(C.byref(model.rm), float(r * blen[c]), _dp(P))
                    val += np.log(P[ns * int(states[v]) + int(states[c])])
                if co[v + 1] == co[v]:
                    prof = np.zeros(model.NS)
                    L = model._lib()
                    (L.orc_nucleotide_profile if model.NS == 4 else L.orc_amino_acid_profile)(int(pattern[leaf_row[v]]), _dp(prof))
                    val += np.log(prof[int(states[v])])
            if val > best:
                best = val
    return float(best)


def log_likelihood(tree, model, aln):
    """TreeLikelihood::log_likelihood for PhyloTree (src/tree.rs:68-76)."""
    patterns, _s2p, mult, leaf_order = aln.compress_columns()
    ft = flatten(tree, leaf_order)
    tot, _ = log_likelihood_flat(ft, model, patterns, mult)
    return tot


def marginal_asr(tree, model, aln, want_posteriors=False):
    """MarginalAsr::marginal_asr for PhyloTree (src/tree.rs:47-56) -> dict like Reconstruction:
    sequences {NodeID: [state per site]}, posteriors {NodeID: [site][4]} | None, log_likelihood."""
    patterns, s2p, mult, leaf_order = aln.compress_columns()
    ft = flatten(tree, leaf_order)
    tot, _per, states, post = marginal_asr_flat(ft, model, patterns, mult, want_posteriors)
    seqs, posts = {}, ({} if want_posteriors else None)
    for k, nid in enumerate(ft["ids"]):
        seqs[int(nid)] = states[k][s2p].astype(np.int64) if len(s2p) else np.zeros(0, np.int64)
        if want_posteriors:
            posts[int(nid)] = post[k][s2p]
    return dict(sequences=seqs, posteriors=posts, log_likelihood=tot)


def bruteforce_site_likelihood(ft, model, pattern, rate=1.0):
    pattern = np.ascontiguousarray(pattern, np.uint8)
    return model._lib().orc_bruteforce_site_likelihood(
        int(ft["n_nodes"]), _ip(ft["child_off"]), _ip(ft["children"]), _ip(ft["parent"]), _dp(ft["blen"]),
        _ip(ft["leaf_row"]), int(ft["root"]), C.byref(model.rm), float(rate),
        pattern.ctypes.data_as(C.POINTER(C.c_uint8)))


def sequence_string(seq):
    """Reconstruction::sequence_string, src/tree/likelihood/reconstruction.rs:19-26."""
    return "".join("ACGT"[int(i)] for i in seq)
