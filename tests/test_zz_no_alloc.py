"""phb_transition_matrices_noalloc: the allocation-free variant of phb_transition_matrices (staging memory in a fixed
file-scope device array, include/phylo_b200.h) against the existing entry point, through the C ABI.

Collected last on purpose; passes alone in a fresh process and after the other tests.
The kernel is the same, so the two entry points must agree bit for bit; the oracle bounds both at the tests' f64 bar."""
import numpy as np
import pytest

from oracle import workloads
from phylo_rs_b200 import capi

SCRATCH_BYTES = 2 << 20  # PHB_NOALLOC_SCRATCH_BYTES


def _model_ctx(lib, model):
    case = workloads.make_case(8, 16, model=model, seed=0x5EED0A11)
    m = case.model
    ctx = capi.Context(library=lib)
    ctx.set_model(m.eigenvalues, m.eigenvectors, m.pi, m.sqrt_pi, m.inv_sqrt_pi, m.rates, m.weights)
    return ctx, m


def _check(lib, model):
    ctx, m = _model_ctx(lib, model)
    n_cat = len(m.rates)
    rng = np.random.default_rng(7)
    with ctx:
        # two calls in a row on the shared array: the second is SHORTER than the first and has other branch lengths,
        # so anything the first call left behind would show up as a wrong or shifted matrix
        t1 = np.concatenate([[0.0, 1e-9, 5.0], rng.exponential(0.1, 1500)])
        t2 = np.concatenate([rng.exponential(0.3, 97), [0.0]])
        for t in (t1, t2, t1):
            want = ctx.transition_matrices(t, n_cat)
            got = ctx.transition_matrices_noalloc(t, n_cat)
            assert got.shape == want.shape == (len(t), n_cat, 4, 4)
            assert np.array_equal(got, want)
            for k in (0, 1, 2, len(t) - 1):
                for c in range(n_cat):
                    ref = m.category_transition(c, t[k])
                    assert np.max(np.abs(got[k, c] - ref)) <= 1e-12
            assert np.allclose(got.sum(axis=3), 1.0, atol=1e-12)
        # empty input: nothing to do, nothing touched
        assert ctx.transition_matrices_noalloc(np.zeros(0), n_cat).shape == (0, n_cat, 4, 4)
        # the largest request that fits, and the first one that does not
        n_max = SCRATCH_BYTES // (8 * (1 + 16 * n_cat))
        t = rng.exponential(0.2, n_max)
        assert np.array_equal(ctx.transition_matrices_noalloc(t, n_cat), ctx.transition_matrices(t, n_cat))
        with pytest.raises(capi.PhbError) as ei:
            ctx.transition_matrices_noalloc(rng.exponential(0.2, n_max + 1), n_cat)
        assert ei.value.variant == "InvalidArgument" and "fixed device array" in str(ei.value)
        # the context is still usable after the rejection
        assert np.array_equal(ctx.transition_matrices_noalloc(t2, n_cat), ctx.transition_matrices(t2, n_cat))


@pytest.mark.gpu
@pytest.mark.parametrize("model", ["gtr_i_g4", "jc69"])
def test_transition_matrices_noalloc_gpu(gpu_lib, model):
    _check(gpu_lib, model)


@pytest.mark.gpu
def test_transition_matrices_noalloc_needs_model(gpu_lib):
    with capi.Context(library=gpu_lib) as ctx:
        with pytest.raises(capi.PhbError):
            ctx.transition_matrices_noalloc(np.array([0.1]), 1)


def test_transition_matrices_noalloc_emu(emu_lib):
    """Host logic (limit arithmetic, argument checks, symbol-address plumbing) on the emulation build."""
    _check(emu_lib, "gtr_g4")
