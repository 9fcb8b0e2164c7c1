"""TEST HARNESS: Python binding of the C++ host mirror (phylo-rs_b200/host/, the reference-shaped consumer of the C ABI)
through the `phh_*` C facade in tests/hostshim/phylo_capi.cpp, so that pytest can drive the C++ classes.  Neither the
facade nor this module is part of the product: the product is libphylo_b200.so (C ABI) + the C++ mirror above it.

The classes carry the reference's names and call shapes so that tests read like the reference's own
(`src/tree/likelihood/integration_test.rs`):

    tree  = PhyloTree.from_newick(b"((A:0.601,B:0.301)C:0.100,D:0.200)E:0.001;")
    model = GtrModel.gtr([.3,.2,.2,.3], [1,2.5,1,1,2.5,1]).with_gamma(0.5, 4)
    aln   = Alignment.from_fasta_bytes(b">A\\nACGT\\n...")
    ll    = tree.log_likelihood(model, aln)                       # TreeLikelihood::log_likelihood
    recon = tree.marginal_asr(model, aln, want_posteriors=True)    # MarginalAsr::marginal_asr

`Result::Err(AsrError::X(msg))` becomes `raise AsrError` with `.variant == "X"`.  All likelihood arithmetic
runs in libphylo_b200.so on the GPU; this module holds no numerical code and has no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
HOST_LIB = os.path.join(_HERE, "libphylo_hostshim.so")


class AsrError(RuntimeError):
    def __init__(self, kind, variant, msg):
        self.kind = kind
        self.variant = variant
        super().__init__(f"{variant}: {msg}")


_L = None


def _lib():
    global _L
    if _L is None:
        if not os.path.exists(HOST_LIB):
            raise RuntimeError(f"{HOST_LIB} not found: run __graft_entry__.build() first (build_hostshim)")
        L = C.CDLL(HOST_LIB)
        L.phh_last_error.restype = C.c_char_p
        L.phh_error_kind_name.restype = C.c_char_p
        L.phh_error_kind_name.argtypes = [C.c_int]
        for nm in ("phh_tree_num_nodes", "phh_tree_root", "phh_aln_width", "phh_aln_n_seqs", "phh_comp_n_patterns",
                   "phh_comp_n_rows", "phh_comp_n_sites", "phh_model_n_categories", "phh_recon_n_nodes"):
            getattr(L, nm).restype = C.c_size_t
            getattr(L, nm).argtypes = [C.c_void_p]
        L.phh_recon_log_likelihood.restype = C.c_double
        L.phh_recon_log_likelihood.argtypes = [C.c_void_p]
        L.phh_recon_has_posteriors.argtypes = [C.c_void_p]
        for nm in ("phh_tree_free", "phh_aln_free", "phh_comp_free", "phh_model_free", "phh_recon_free"):
            getattr(L, nm).restype = None
            getattr(L, nm).argtypes = [C.c_void_p]
        vp, sz, vpp = C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)
        szp, ip, dp = C.POINTER(C.c_size_t), C.POINTER(C.c_int), C.POINTER(C.c_double)
        sig = {
            "phh_backend_version": [C.c_char_p, sz], "phh_set_library": [C.c_char_p], "phh_set_devices": [ip, C.c_int, C.c_uint64],
            "phh_tree_from_newick": [C.c_char_p, sz, vpp], "phh_tree_yule": [sz, C.c_uint64, vpp],
            "phh_tree_clone": [vp, vpp], "phh_tree_node_ids": [vp, szp], "phh_tree_postord_ids": [vp, szp],
            "phh_tree_preord_ids": [vp, szp], "phh_tree_children": [vp, sz, szp, sz, szp],
            "phh_tree_parent": [vp, sz, szp, ip], "phh_tree_is_leaf": [vp, sz, ip],
            "phh_tree_taxa": [vp, sz, C.c_char_p, sz, ip], "phh_tree_taxa_node_id": [vp, C.c_char_p, szp, ip],
            "phh_tree_edge_weight": [vp, sz, C.POINTER(C.c_float), ip],
            "phh_tree_set_edge_weight": [vp, sz, C.c_float, C.c_int],
            "phh_tree_assign_exponential_lengths": [vp, C.c_uint64, C.c_double, C.c_double, C.c_double],
            "phh_tree_nni": [vp, sz, C.c_int], "phh_tree_to_newick": [vp, C.c_char_p, sz, szp],
            "phh_aln_from_fasta": [C.c_char_p, sz, vpp], "phh_aln_new": [sz, vpp],
            "phh_aln_insert": [vp, C.c_char_p, C.c_char_p, sz], "phh_aln_compress": [vp, vpp],
            "phh_comp_export": [vp, vp, vp, vp], "phh_comp_leaf_name": [vp, sz, C.c_char_p, sz],
            "phh_model_new": [dp, dp, C.c_int, vpp], "phh_model_named": [C.c_char_p, dp, C.c_int, vpp],
            "phh_model_with_gamma": [vp, C.c_double, sz, vpp], "phh_model_with_invariant": [vp, C.c_double, vpp],
            "phh_model_categories": [vp, dp, dp], "phh_model_equilibrium": [vp, dp], "phh_model_eigen": [vp, dp, dp],
            "phh_model_category_transition": [vp, sz, C.c_double, dp], "phh_discrete_gamma": [C.c_double, sz, dp],
            "phh_nucleotide_profile": [C.c_uint8, dp, ip],
            "phh_log_likelihood": [vp, vp, vp, dp], "phh_marginal_asr": [vp, vp, vp, C.c_int, vpp],
            "phh_joint_asr": [vp, vp, vp, vpp],
            "phh_recon_sequence": [vp, sz, vp, sz, szp, ip], "phh_recon_posteriors": [vp, sz, vp, sz, ip],
            "phh_score_nni_neighbours": [vp, vp, vp, sz, vp, vp, dp],
        }
        for nm, args in sig.items():
            getattr(L, nm).argtypes = args
            getattr(L, nm).restype = C.c_int
        _L = L
    return _L


def _ck(rc):
    if rc:
        L = _lib()
        raise AsrError(rc, L.phh_error_kind_name(rc).decode(), L.phh_last_error().decode(errors="replace"))


def _d(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def backend_version():
    buf = C.create_string_buffer(128)
    _ck(_lib().phh_backend_version(buf, 128))
    return buf.value.decode()


def set_devices(devices, flags=0):
    d = (C.c_int * len(devices))(*devices)
    _ck(_lib().phh_set_devices(d, len(devices), flags))


def set_library(path=None):
    """Point the mirror at another build of the C ABI (None = the sm_100a product library)."""
    _ck(_lib().phh_set_library(None if path is None else os.fsencode(path)))


def release_backend():
    _lib().phh_release_backend()


def discrete_gamma(alpha, k):
    out = np.zeros(max(k, 1))
    _ck(_lib().phh_discrete_gamma(alpha, k, _d(out)))
    return out[:k]


class Nucleotide:
    N_STATES = 4
    CANONICAL = b"ACGT"
    GAP = ord("-")

    @staticmethod
    def profile(c):
        out, has = np.zeros(4), C.c_int(0)
        _ck(_lib().phh_nucleotide_profile(c if isinstance(c, int) else ord(c), _d(out), C.byref(has)))
        return out if has.value else None

    @staticmethod
    def char_of(i):
        return Nucleotide.CANONICAL[i]


class _Handle:
    _free = None

    def __init__(self, h):
        self._h = h

    def __del__(self):
        try:
            if self._h:
                getattr(_lib(), self._free)(self._h)
                self._h = None
        except Exception:
            pass


class PhyloTree(_Handle):
    _free = "phh_tree_free"

    @staticmethod
    def from_newick(src):
        if isinstance(src, str):
            src = src.encode()
        h = C.c_void_p()
        _ck(_lib().phh_tree_from_newick(src, len(src), C.byref(h)))
        return PhyloTree(h)

    @staticmethod
    def yule(num_taxa, seed=0):
        h = C.c_void_p()
        _ck(_lib().phh_tree_yule(num_taxa, seed, C.byref(h)))
        return PhyloTree(h)

    def clone(self):
        h = C.c_void_p()
        _ck(_lib().phh_tree_clone(self._h, C.byref(h)))
        return PhyloTree(h)

    def num_nodes(self):
        return _lib().phh_tree_num_nodes(self._h)

    def get_root_id(self):
        return _lib().phh_tree_root(self._h)

    def _ids(self, fn):
        out = (C.c_size_t * max(self.num_nodes(), 1))()
        _ck(getattr(_lib(), fn)(self._h, out))
        return list(out[: self.num_nodes()])

    def get_node_ids(self):
        return self._ids("phh_tree_node_ids")

    def postord_ids(self):
        return self._ids("phh_tree_postord_ids")

    def preord_ids(self):
        return self._ids("phh_tree_preord_ids")

    def get_node_children_ids(self, node_id):
        n = C.c_size_t(0)
        cap = 8
        while True:
            out = (C.c_size_t * cap)()
            _ck(_lib().phh_tree_children(self._h, node_id, out, cap, C.byref(n)))
            if n.value <= cap:
                return list(out[: n.value])
            cap = n.value

    def get_node_parent_id(self, node_id):
        p, has = C.c_size_t(0), C.c_int(0)
        _ck(_lib().phh_tree_parent(self._h, node_id, C.byref(p), C.byref(has)))
        return p.value if has.value else None

    def is_leaf(self, node_id):
        out = C.c_int(0)
        _ck(_lib().phh_tree_is_leaf(self._h, node_id, C.byref(out)))
        return bool(out.value)

    def get_node_taxa(self, node_id):
        buf, has = C.create_string_buffer(4096), C.c_int(0)
        _ck(_lib().phh_tree_taxa(self._h, node_id, buf, 4096, C.byref(has)))
        return buf.value.decode() if has.value else None

    def get_taxa_node_id(self, name):
        i, has = C.c_size_t(0), C.c_int(0)
        _ck(_lib().phh_tree_taxa_node_id(self._h, name.encode(), C.byref(i), C.byref(has)))
        return i.value if has.value else None

    def get_edge_weight(self, parent_id, child_id):
        w, has = C.c_float(0), C.c_int(0)
        _ck(_lib().phh_tree_edge_weight(self._h, child_id, C.byref(w), C.byref(has)))
        return w.value if has.value else None

    def set_edge_weight(self, child_id, w):
        _ck(_lib().phh_tree_set_edge_weight(self._h, child_id, 0.0 if w is None else w, w is not None))

    def assign_exponential_lengths(self, seed, mean=0.1, lo=1e-4, hi=2.0):
        _ck(_lib().phh_tree_assign_exponential_lengths(self._h, seed, mean, lo, hi))

    def nni(self, node_id, left_ch):
        _ck(_lib().phh_tree_nni(self._h, node_id, int(bool(left_ch))))

    def to_newick(self):
        n = C.c_size_t(0)
        _ck(_lib().phh_tree_to_newick(self._h, None, 0, C.byref(n)))
        buf = C.create_string_buffer(n.value + 1)
        _ck(_lib().phh_tree_to_newick(self._h, buf, n.value + 1, C.byref(n)))
        return buf.value.decode()

    # -- the two traits ------------------------------------------------------------------------------
    def log_likelihood(self, model, aln):
        out = C.c_double(0.0)
        _ck(_lib().phh_log_likelihood(self._h, model._h, aln._h, C.byref(out)))
        return out.value

    def marginal_asr(self, model, aln, want_posteriors=False):
        h = C.c_void_p()
        _ck(_lib().phh_marginal_asr(self._h, model._h, aln._h, int(want_posteriors), C.byref(h)))
        return Reconstruction(h, aln.width)

    def joint_asr(self, model, aln):
        """JointAsr::joint_asr (src/tree/asr.rs:30-37): Reconstruction with posteriors None."""
        h = C.c_void_p()
        _ck(_lib().phh_joint_asr(self._h, model._h, aln._h, C.byref(h)))
        return Reconstruction(h, aln.width)

    def score_nni_neighbours(self, model, aln, moves):
        n = len(moves)
        nodes = (C.c_size_t * max(n, 1))(*[m[0] for m in moves])
        left = (C.c_uint8 * max(n, 1))(*[int(bool(m[1])) for m in moves])
        out = np.zeros(max(n, 1))
        _ck(_lib().phh_score_nni_neighbours(self._h, model._h, aln._h, n, nodes, left, _d(out)))
        return out[:n]


class Alignment(_Handle):
    _free = "phh_aln_free"

    @staticmethod
    def from_fasta_bytes(data):
        h = C.c_void_p()
        _ck(_lib().phh_aln_from_fasta(data, len(data), C.byref(h)))
        return Alignment(h)

    @staticmethod
    def from_seqs(seqs, width=None):
        """`Alignment { seqs, width }` literal: seqs = {name: bytes} (insertion order kept)."""
        seqs = dict(seqs)
        if width is None:
            width = len(next(iter(seqs.values()))) if seqs else 0
        h = C.c_void_p()
        _ck(_lib().phh_aln_new(width, C.byref(h)))
        a = Alignment(h)
        for name, s in seqs.items():
            s = bytes(s)
            _ck(_lib().phh_aln_insert(h, name.encode(), s, len(s)))
        return a

    @property
    def width(self):
        return _lib().phh_aln_width(self._h)

    def n_seqs(self):
        return _lib().phh_aln_n_seqs(self._h)

    def compress_columns(self):
        L = _lib()
        h = C.c_void_p()
        _ck(L.phh_aln_compress(self._h, C.byref(h)))
        try:
            P, rows, sites = L.phh_comp_n_patterns(h), L.phh_comp_n_rows(h), L.phh_comp_n_sites(h)
            pat = np.zeros((max(P, 1), max(rows, 1)), np.uint8)
            mult, s2p = np.zeros(max(P, 1), np.uint64), np.zeros(max(sites, 1), np.uint64)
            _ck(L.phh_comp_export(h, pat.ctypes.data, mult.ctypes.data, s2p.ctypes.data))
            names = []
            for r in range(rows):
                buf = C.create_string_buffer(4096)
                _ck(L.phh_comp_leaf_name(h, r, buf, 4096))
                names.append(buf.value.decode())
        finally:
            L.phh_comp_free(h)
        return CompressedColumns(pat[:P, :rows], s2p[:sites].astype(np.int64), mult[:P], names)


class CompressedColumns:
    def __init__(self, patterns, site_to_pattern, multiplicity, leaf_order):
        self.patterns, self.site_to_pattern = patterns, site_to_pattern
        self.multiplicity, self.leaf_order = multiplicity, leaf_order


class GtrModel(_Handle):
    _free = "phh_model_free"

    @staticmethod
    def new(pi, w, normalize=True):
        pi = np.ascontiguousarray(pi, np.float64)
        w = np.ascontiguousarray(w, np.float64).reshape(16)
        h = C.c_void_p()
        _ck(_lib().phh_model_new(_d(pi), _d(w), int(normalize), C.byref(h)))
        return GtrModel(h)

    @staticmethod
    def _named(name, *params):
        p = np.ascontiguousarray(np.concatenate([np.atleast_1d(np.asarray(x, np.float64)) for x in params])
                                 if params else np.zeros(0), np.float64)
        h = C.c_void_p()
        _ck(_lib().phh_model_named(name.encode(), _d(p) if p.size else None, int(p.size), C.byref(h)))
        return GtrModel(h)

    jukes_cantor = staticmethod(lambda: GtrModel._named("jc69"))
    k80 = staticmethod(lambda kappa: GtrModel._named("k80", kappa))
    f81 = staticmethod(lambda pi: GtrModel._named("f81", pi))
    hky85 = staticmethod(lambda pi, kappa: GtrModel._named("hky85", pi, kappa))
    tn93 = staticmethod(lambda pi, k1, k2: GtrModel._named("tn93", pi, k1, k2))
    k81 = staticmethod(lambda a, b, c: GtrModel._named("k81", a, b, c))
    tim = staticmethod(lambda pi, k1, k2, a, b: GtrModel._named("tim", pi, k1, k2, a, b))
    tvm = staticmethod(lambda pi, kappa, a, b, c: GtrModel._named("tvm", pi, kappa, a, b, c))
    sym = staticmethod(lambda rates: GtrModel._named("sym", rates))
    gtr = staticmethod(lambda pi, rates: GtrModel._named("gtr", pi, rates))

    def with_gamma(self, alpha, k):
        h = C.c_void_p()
        _ck(_lib().phh_model_with_gamma(self._h, alpha, k, C.byref(h)))
        return GtrModel(h)

    def with_invariant(self, p_inv):
        h = C.c_void_p()
        _ck(_lib().phh_model_with_invariant(self._h, p_inv, C.byref(h)))
        return GtrModel(h)

    def n_categories(self):
        return _lib().phh_model_n_categories(self._h)

    def categories(self):
        k = self.n_categories()
        r, w = np.zeros(k), np.zeros(k)
        _ck(_lib().phh_model_categories(self._h, _d(r), _d(w)))
        return list(zip(r.tolist(), w.tolist()))

    def equilibrium(self):
        pi = np.zeros(4)
        _ck(_lib().phh_model_equilibrium(self._h, _d(pi)))
        return pi

    def eigen(self):
        """(eigenvalues[4], eigenvectors[4,4] with [i, k] = component i of eigenvector k)."""
        lam, v = np.zeros(4), np.zeros(16)
        _ck(_lib().phh_model_eigen(self._h, _d(lam), _d(v)))
        return lam, v.reshape(4, 4).T.copy()

    def category_transition(self, category, t):
        p = np.zeros(16)
        _ck(_lib().phh_model_category_transition(self._h, category, t, _d(p)))
        return p.reshape(4, 4)


class Reconstruction(_Handle):
    _free = "phh_recon_free"

    def __init__(self, h, width):
        super().__init__(h)
        self.width = width

    @property
    def log_likelihood(self):
        return _lib().phh_recon_log_likelihood(self._h)

    def n_nodes(self):
        return _lib().phh_recon_n_nodes(self._h)

    def has_posteriors(self):
        return bool(_lib().phh_recon_has_posteriors(self._h))

    def sequence(self, node_id):
        out = np.zeros(max(self.width, 1), np.uint64)
        w, has = C.c_size_t(0), C.c_int(0)
        _ck(_lib().phh_recon_sequence(self._h, node_id, out.ctypes.data, len(out), C.byref(w), C.byref(has)))
        return out[: w.value].astype(np.int64) if has.value else None

    def sequence_string(self, node_id):
        s = self.sequence(node_id)
        return None if s is None else bytes(Nucleotide.CANONICAL[i] for i in s).decode()

    def posteriors(self, node_id):
        out, has = np.zeros((max(self.width, 1), 4)), C.c_int(0)
        _ck(_lib().phh_recon_posteriors(self._h, node_id, out.ctypes.data, len(out), C.byref(has)))
        return out[: self.width] if has.value else None
