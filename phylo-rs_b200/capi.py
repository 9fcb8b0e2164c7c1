"""ctypes binding of the C ABI in include/phylo_b200.h (libphylo_b200.so).

This is the Python spelling of the `extern "C"` block a Rust shim would declare;
argument meaning, ownership and status codes are exactly those of the header.  `Library()` loads the
sm_100a product library built in-tree under csrc/ and raises if it is missing — there is no CPU
path.  (tests/ may point `Library(path=...)` at the host-emulation harness build to exercise the
scheduler without a GPU; nothing in this package ever does.)
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_LIB = os.environ.get("PHYLO_B200_LIB") or os.path.join(_HERE, "csrc", "libphylo_b200.so")

PHB_OK = 0
STATUS_NAMES = {
    0: "OK", 1: "MissingBranchLength", 2: "AlphabetMismatch", 3: "InvalidAlignment",
    4: "EigendecompFailure", 5: "NumericalInstability", 6: "InvalidModelParameter",
    16: "InvalidArgument", 17: "Cuda", 18: "Nccl", 19: "OutOfMemory", 20: "State",
}
FLAG_LITERAL_INVARIANT = 1 << 0
FLAG_TIMING = 1 << 1
FLAG_LEVEL_KERNELS = 1 << 2
FLAG_NNI_FULL = 1 << 3
FLAG_NNI_WALK = 1 << 4
FLAG_AMINO_ACID = 1 << 5  # 20-state context (AminoAcid, src/alphabet.rs:66-116)
LAYOUT_PATTERN_MAJOR = 0
LAYOUT_ROW_MAJOR = 1

# every symbol include/phylo_b200.h declares (checked by tests/test_abi.py)
EXPORTS = [
    "phb_ctx_create", "phb_ctx_destroy", "phb_set_clv_budget", "phb_set_alignment", "phb_compress_columns", "phb_compress_rows", "phb_set_model",
    "phb_set_tree", "phb_log_likelihood", "phb_marginal_asr", "phb_joint_asr", "phb_get_per_pattern", "phb_score_nni", "phb_score_trees",
    "phb_simulate_alignment", "phb_get_alignment", "phb_get_stats", "phb_measure_fp64_tflops",
    "phb_timer_start", "phb_timer_stop", "phb_transition_matrices", "phb_transition_matrices_noalloc", "phb_log_likelihood_host",
    "phb_last_error", "phb_version",
]


class PhbError(RuntimeError):
    """A non-zero status from the library; `.variant` names the AsrError variant (src/error.rs:106-125)."""

    def __init__(self, code, msg):
        self.code = code
        self.variant = STATUS_NAMES.get(code, f"code{code}")
        super().__init__(f"{self.variant}: {msg}")


class Stats(C.Structure):
    _fields_ = [
        ("n_patterns", C.c_int64), ("tile_patterns", C.c_int64), ("n_tiles", C.c_int32),
        ("n_levels", C.c_int32), ("n_internal", C.c_int32), ("n_clv_categories", C.c_int32),
        ("kernel_launches", C.c_int64), ("clv_updates", C.c_int64), ("algorithmic_bytes", C.c_double),
        ("ms_transition", C.c_double), ("ms_clv", C.c_double), ("ms_root", C.c_double),
        ("ms_down", C.c_double), ("ms_total", C.c_double), ("clv_launches", C.c_int32),
        ("sweep_ops", C.c_int32), ("sweep_stack_levels", C.c_int32), ("sweep_smem_levels", C.c_int32),
        ("sweep_stages", C.c_int32), ("sweep_block_sites", C.c_int32), ("sweep_stream_bytes", C.c_int64),
        ("nni_path_moves", C.c_int32), ("nni_tiles", C.c_int32), ("nni_path_steps", C.c_int64),
        ("nni_cache_bytes", C.c_double), ("nni_read_bytes", C.c_double),
        ("sweep_guards", C.c_int32), ("sweep_normouts", C.c_int32),
        ("sweep_fp64_instr", C.c_double), ("sweep_flops", C.c_double),
        ("ms_nni_cache", C.c_double), ("ms_nni_upper", C.c_double), ("ms_nni_edge", C.c_double),
    ]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "reserved"}


class TreeDesc(C.Structure):
    """phb_tree_desc: a flattened candidate tree (the arguments of phb_set_tree)."""
    _fields_ = [("n_nodes", C.c_int32), ("postorder", C.POINTER(C.c_int32)), ("child_offsets", C.POINTER(C.c_int32)),
                ("children", C.POINTER(C.c_int32)), ("branch_len", C.POINTER(C.c_double)), ("leaf_row", C.POINTER(C.c_int32)),
                ("root", C.c_int32)]


class Library:
    def __init__(self, path=None):
        path = path or DEFAULT_LIB
        if not os.path.exists(path):
            raise RuntimeError(
                f"{path} not found: build the sm_100a library first (python -c 'import __graft_entry__ as g; g.build()'). "
                "phylo-rs_b200 has no CPU fallback.")
        self.path = path
        L = self.L = C.CDLL(path)
        vp, dp, ip = C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int32)
        u8p, u64p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint64)
        L.phb_ctx_create.argtypes = [ip, C.c_int32, C.c_uint64, C.POINTER(vp)]
        L.phb_ctx_destroy.argtypes = [vp]
        L.phb_ctx_destroy.restype = None
        L.phb_set_clv_budget.argtypes = [vp, C.c_uint64]
        L.phb_set_alignment.argtypes = [vp, C.c_int32, C.c_int64, vp, C.c_int32, u64p]
        L.phb_compress_columns.argtypes = [vp, C.c_int32, C.c_int64, vp, C.POINTER(C.c_int64), u64p, u64p]
        L.phb_compress_rows.argtypes = [vp, C.c_int32, C.c_int64, C.POINTER(vp), C.POINTER(C.c_int64), u64p, u64p]
        L.phb_set_model.argtypes = [vp, dp, dp, dp, dp, dp, C.c_int32, dp, dp]
        L.phb_set_tree.argtypes = [vp, C.c_int32, ip, ip, ip, dp, ip, C.c_int32]
        L.phb_log_likelihood.argtypes = [vp, dp, dp]
        L.phb_log_likelihood_host.argtypes = [vp, C.c_int32, C.c_int64, vp, C.c_int32, u64p, dp, dp]
        L.phb_marginal_asr.argtypes = [vp, u8p, dp, dp]
        L.phb_joint_asr.argtypes = [vp, u8p, dp]
        L.phb_get_per_pattern.argtypes = [vp, dp]
        L.phb_score_nni.argtypes = [vp, C.c_int32, ip, u8p, dp]
        L.phb_score_trees.argtypes = [vp, C.c_int32, C.POINTER(TreeDesc), dp]
        L.phb_simulate_alignment.argtypes = [vp, C.c_int64, C.c_int64, C.c_uint64]
        L.phb_get_alignment.argtypes = [vp, vp, C.c_int32, C.c_int64, C.c_int64]
        L.phb_get_stats.argtypes = [vp, C.POINTER(Stats)]
        L.phb_measure_fp64_tflops.argtypes = [vp, dp]
        L.phb_transition_matrices.argtypes = [vp, C.c_int32, dp, dp]
        L.phb_timer_start.argtypes = [vp]
        L.phb_timer_stop.argtypes = [vp, dp]
        L.phb_last_error.argtypes = [vp]
        L.phb_last_error.restype = C.c_char_p
        L.phb_version.argtypes = []
        L.phb_version.restype = C.c_char_p
        for nm in EXPORTS:
            if nm not in ("phb_ctx_destroy", "phb_last_error", "phb_version"):
                getattr(L, nm).restype = C.c_int32

    def version(self):
        return self.L.phb_version().decode()


_DEFAULT = None


def default_library():
    global _DEFAULT
    if _DEFAULT is None:
        _DEFAULT = Library(os.environ.get("PHB_LIB"))  # PHB_LIB: another build of the same sources (kernel experiments)
    return _DEFAULT


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


class Context:
    """Owns one phb_ctx.  Inputs are numpy arrays (copied by the library before each call returns)."""

    def __init__(self, devices=None, flags=0, library=None):
        self.lib = library or default_library()
        self._h = C.c_void_p()
        if devices is None:
            rc = self.lib.L.phb_ctx_create(None, 0, flags, C.byref(self._h))
        else:
            d = np.ascontiguousarray(devices, np.int32)
            rc = self.lib.L.phb_ctx_create(_ip(d), len(d), flags, C.byref(self._h))
        if rc:
            self._h = C.c_void_p()
            raise PhbError(rc, "phb_ctx_create failed (no usable B200 / CUDA device?)")
        self.n_states = 20 if (flags & FLAG_AMINO_ACID) else 4
        self.n_rows = 0
        self.n_patterns = 0
        self.n_nodes = 0

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self.lib.L.phb_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc):
        if rc:
            raise PhbError(rc, self.lib.L.phb_last_error(self._h).decode(errors="replace"))

    # -- inputs ------------------------------------------------------------------------------
    def set_clv_budget(self, nbytes):
        self._check(self.lib.L.phb_set_clv_budget(self._h, int(nbytes)))

    def compress_columns(self, rows):
        """phb_compress_columns: rows = uint8 [n_rows][width] raw alignment bytes (one sequence per row).
        Returns (n_patterns, site_to_pattern uint64[width], multiplicity uint64[n_patterns]); the compressed
        patterns become the context's alignment."""
        rows = np.ascontiguousarray(rows, np.uint8)
        if rows.ndim != 2:
            raise ValueError("rows must be 2-D")
        n_rows, width = rows.shape
        s2p = np.zeros(max(width, 1), np.uint64)
        mult = np.zeros(max(width, 1), np.uint64)
        n = C.c_int64(0)
        self._check(self.lib.L.phb_compress_columns(
            self._h, n_rows, width, rows.ctypes.data_as(C.c_void_p), C.byref(n),
            s2p.ctypes.data_as(C.POINTER(C.c_uint64)), mult.ctypes.data_as(C.POINTER(C.c_uint64))))
        self.n_rows, self.n_patterns = n_rows, int(n.value)
        return int(n.value), s2p[:width], mult[: int(n.value)]

    def compress_rows(self, seqs):
        """phb_compress_rows: seqs = list of equal-length bytes objects / uint8 arrays, one per row."""
        arrs = [np.ascontiguousarray(np.frombuffer(s, np.uint8) if isinstance(s, (bytes, bytearray)) else s, np.uint8) for s in seqs]
        width = len(arrs[0]) if arrs else 0
        ptrs = (C.c_void_p * max(len(arrs), 1))(*[a.ctypes.data for a in arrs])
        s2p = np.zeros(max(width, 1), np.uint64)
        mult = np.zeros(max(width, 1), np.uint64)
        n = C.c_int64(0)
        self._check(self.lib.L.phb_compress_rows(
            self._h, len(arrs), width, ptrs, C.byref(n), s2p.ctypes.data_as(C.POINTER(C.c_uint64)),
            mult.ctypes.data_as(C.POINTER(C.c_uint64))))
        self.n_rows, self.n_patterns = len(arrs), int(n.value)
        return int(n.value), s2p[:width], mult[: int(n.value)]

    def set_alignment(self, data, layout=LAYOUT_PATTERN_MAJOR, multiplicity=None):
        """data: uint8 array, [P][rows] (pattern-major) or [rows][P] (row-major) ASCII bytes."""
        data = np.ascontiguousarray(data, np.uint8)
        if data.ndim != 2:
            raise ValueError("alignment must be 2-D")
        if layout == LAYOUT_PATTERN_MAJOR:
            P, rows = data.shape
        else:
            rows, P = data.shape
        m = None
        if multiplicity is not None:
            m = np.ascontiguousarray(multiplicity, np.uint64)
            if m.shape != (P,):
                raise ValueError("multiplicity length must equal the number of patterns")
        self._check(self.lib.L.phb_set_alignment(
            self._h, rows, P, data.ctypes.data, layout,
            None if m is None else m.ctypes.data_as(C.POINTER(C.c_uint64))))
        self.n_rows, self.n_patterns = rows, P

    def set_alignment_ptr(self, ptr, n_rows, n_patterns, layout=LAYOUT_PATTERN_MAJOR):
        """Raw host pointer variant (e.g. pinned memory); multiplicity 1."""
        self._check(self.lib.L.phb_set_alignment(self._h, n_rows, n_patterns, ptr, layout, None))
        self.n_rows, self.n_patterns = n_rows, n_patterns

    def set_model(self, eigenvalues, eigenvectors, pi, sqrt_pi, inv_sqrt_pi, rates, weights):
        """eigenvectors: 4x4 with [i, k] = component i of eigenvector k (passed column-major)."""
        lam = np.ascontiguousarray(eigenvalues, np.float64)
        V = np.ascontiguousarray(np.asarray(eigenvectors, np.float64).T).ravel()  # V[i + ns * k]
        if lam.shape != (self.n_states,) or V.shape != (self.n_states * self.n_states,):
            raise ValueError(f"this context evaluates {self.n_states} states")
        pi, sp, isp = (np.ascontiguousarray(x, np.float64) for x in (pi, sqrt_pi, inv_sqrt_pi))
        r, w = np.ascontiguousarray(rates, np.float64), np.ascontiguousarray(weights, np.float64)
        self._check(self.lib.L.phb_set_model(self._h, _dp(lam), _dp(V), _dp(pi), _dp(sp), _dp(isp), len(r), _dp(r), _dp(w)))

    def set_tree(self, postorder, child_off, children, blen, leaf_row, root):
        po = np.ascontiguousarray(postorder, np.int32)
        co = np.ascontiguousarray(child_off, np.int32)
        ch = np.ascontiguousarray(children, np.int32)
        if ch.size == 0:
            ch = np.zeros(1, np.int32)
        bl = np.ascontiguousarray(blen, np.float64)
        lr = np.ascontiguousarray(leaf_row, np.int32)
        self._check(self.lib.L.phb_set_tree(self._h, len(po), _ip(po), _ip(co), _ip(ch), _dp(bl), _ip(lr), int(root)))
        self.n_nodes = len(po)
        self.n_leaves = int(np.sum(co[1:] == co[:-1]))

    # -- engines -----------------------------------------------------------------------------
    def log_likelihood(self, per_pattern=False):
        tot = C.c_double(0.0)
        pp = np.zeros(max(self.n_patterns, 1), np.float64) if per_pattern else None
        self._check(self.lib.L.phb_log_likelihood(self._h, C.byref(tot), None if pp is None else _dp(pp)))
        return (tot.value, pp[: self.n_patterns]) if per_pattern else tot.value

    def log_likelihood_host(self, data, layout=LAYOUT_PATTERN_MAJOR, multiplicity=None, per_pattern=False):
        """set_alignment + log_likelihood in one call; the H2D copy is pipelined with the packing and the sweep.
        data: uint8 array ([P][rows] / [rows][P]) or a (ptr, n_rows, n_patterns) tuple for raw (pinned) memory."""
        if isinstance(data, tuple):
            ptr, rows, P = data
        else:
            data = np.ascontiguousarray(data, np.uint8)
            P, rows = data.shape if layout == LAYOUT_PATTERN_MAJOR else data.shape[::-1]
            ptr = data.ctypes.data
        m = None if multiplicity is None else np.ascontiguousarray(multiplicity, np.uint64)
        tot = C.c_double(0.0)
        pp = np.zeros(max(P, 1), np.float64) if per_pattern else None
        self._check(self.lib.L.phb_log_likelihood_host(
            self._h, rows, P, ptr, layout, None if m is None else m.ctypes.data_as(C.POINTER(C.c_uint64)),
            C.byref(tot), None if pp is None else _dp(pp)))
        self.n_rows, self.n_patterns = rows, P
        return (tot.value, pp[:P]) if per_pattern else tot.value

    def marginal_asr(self, want_posteriors=False):
        n, P = self.n_nodes, self.n_patterns
        states = np.zeros((n, max(P, 1)), np.uint8)
        post = np.zeros((n, max(P, 1), self.n_states), np.float64) if want_posteriors else None
        tot = C.c_double(0.0)
        self._check(self.lib.L.phb_marginal_asr(
            self._h, states.ctypes.data_as(C.POINTER(C.c_uint8)), None if post is None else _dp(post), C.byref(tot)))
        if P == 0:
            return tot.value, states[:, :0], (None if post is None else post[:, :0])
        return tot.value, states, post

    def joint_asr(self):
        """phb_joint_asr: (Reconstruction.log_likelihood, states [node][pattern] uint8)."""
        n, P = self.n_nodes, self.n_patterns
        states = np.zeros((n, max(P, 1)), np.uint8)
        tot = C.c_double(0.0)
        self._check(self.lib.L.phb_joint_asr(self._h, states.ctypes.data_as(C.POINTER(C.c_uint8)), C.byref(tot)))
        return tot.value, (states[:, :0] if P == 0 else states)

    def transition_matrices(self, t, n_cat):
        """P(rate_c * t_k) for every t_k and model category: [len(t)][n_cat][ns][ns] (device-evaluated)."""
        t = np.ascontiguousarray(np.atleast_1d(t), np.float64)
        out = np.zeros((len(t), n_cat, self.n_states, self.n_states), np.float64)
        self._check(self.lib.L.phb_transition_matrices(self._h, len(t), _dp(t), _dp(out)))
        return out

    def transition_matrices_noalloc(self, t, n_cat):
        """As transition_matrices, through the allocation-free entry point (fixed device staging array; fails with
        InvalidArgument beyond PHB_NOALLOC_SCRATCH_BYTES)."""
        t = np.ascontiguousarray(np.atleast_1d(t), np.float64)
        out = np.zeros((len(t), n_cat, 4, 4), np.float64)
        self._check(self.lib.L.phb_transition_matrices_noalloc(self._h, len(t), _dp(t), _dp(out)))
        return out

    def per_pattern(self):
        pp = np.zeros(max(self.n_patterns, 1), np.float64)
        self._check(self.lib.L.phb_get_per_pattern(self._h, _dp(pp)))
        return pp[: self.n_patterns]

    def score_nni(self, node_ids, left_ch):
        ids = np.ascontiguousarray(node_ids, np.int32)
        lc = np.ascontiguousarray(left_ch, np.uint8)
        out = np.zeros(max(len(ids), 1), np.float64)
        self._check(self.lib.L.phb_score_nni(self._h, len(ids), _ip(ids), lc.ctypes.data_as(C.POINTER(C.c_uint8)), _dp(out)))
        return out[: len(ids)]

    def score_trees(self, trees):
        """phb_score_trees: trees = list of (postorder, child_off, children, branch_len, leaf_row, root) flattened
        candidates; returns their log-likelihoods against the resident alignment and model (NaN: unschedulable)."""
        keep, descs = [], (TreeDesc * max(len(trees), 1))()
        for i, (post, co, ch, bl, lr, root) in enumerate(trees):
            a = [np.ascontiguousarray(post, np.int32), np.ascontiguousarray(co, np.int32), np.ascontiguousarray(ch, np.int32),
                 np.ascontiguousarray(bl, np.float64), np.ascontiguousarray(lr, np.int32)]
            keep.append(a)
            descs[i] = TreeDesc(len(a[0]), _ip(a[0]), _ip(a[1]), _ip(a[2]), _dp(a[3]), _ip(a[4]), int(root))
        out = np.zeros(max(len(trees), 1), np.float64)
        self._check(self.lib.L.phb_score_trees(self._h, len(trees), descs, _dp(out)))
        return out[: len(trees)]

    # -- harness -----------------------------------------------------------------------------
    def simulate_alignment(self, n_patterns, global_offset=0, seed=0):
        self._check(self.lib.L.phb_simulate_alignment(self._h, int(n_patterns), int(global_offset), int(seed)))
        self.n_patterns = int(n_patterns)
        self.n_rows = self.n_leaves

    def get_alignment(self, first=0, count=None, layout=LAYOUT_PATTERN_MAJOR, n_rows=None):
        n_rows = self.n_rows if n_rows is None else n_rows
        count = self.n_patterns - first if count is None else count
        shape = (count, n_rows) if layout == LAYOUT_PATTERN_MAJOR else (n_rows, count)
        out = np.zeros(shape, np.uint8)
        self._check(self.lib.L.phb_get_alignment(self._h, out.ctypes.data, layout, int(first), int(count)))
        return out

    def stats(self):
        s = Stats()
        self._check(self.lib.L.phb_get_stats(self._h, C.byref(s)))
        return s.as_dict()

    def timer_start(self):
        self._check(self.lib.L.phb_timer_start(self._h))

    def timer_stop(self):
        v = C.c_double(0.0)
        self._check(self.lib.L.phb_timer_stop(self._h, C.byref(v)))
        return v.value

    def measure_fp64_tflops(self):
        v = C.c_double(0.0)
        self._check(self.lib.L.phb_measure_fp64_tflops(self._h, C.byref(v)))
        return v.value
