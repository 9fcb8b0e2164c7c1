"""phylo-rs_b200 — B200 (sm_100a) implementation of phylo-rs's Felsenstein-pruning likelihood /
marginal-ASR hot path behind the reference's own interface.

Layers:
  csrc/   hand-written CUDA kernels + the C-ABI shared library libphylo_b200.so (include/phylo_b200.h)
  capi    ctypes binding of that C ABI (what a Rust `extern "C"` block would bind)
  host/   C++ mirror of the reference's host-side interface (PhyloTree / GtrModel / Alignment /
          TreeLikelihood / MarginalAsr) above the C ABI (driven from pytest through tests/hostshim/)
There is no CPU implementation of the path in this package: without the CUDA library and a B200
the entry points raise.
"""
from . import capi  # noqa: F401

__all__ = ["capi"]
