#!/usr/bin/env python
"""Regenerates bindings/phylo-b200-shim.patch from a checkout of the reference crate (default /root/reference):
the `b200` feature, the four eigen-system accessors of RateMatrix (+ GtrModel::rate_matrix), the module
declaration, the three trait impls of src/tree.rs:47-76 switched to the backend, and src/tree/likelihood/b200.rs
(= bindings/shim/b200.rs).  The reference tree is only read; the edited copy lives under a temporary directory.

    python bindings/make_patch.py [/path/to/phylo-rs]
"""
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
FILES = ["Cargo.toml", "src/tree.rs", "src/tree/likelihood.rs", "src/models/rate_matrix.rs", "src/models/gtr.rs"]


def edit(path, old, new):
    s = open(path).read()
    assert s.count(old) == 1, f"{path}: anchor not found exactly once: {old[:60]!r}"
    open(path, "w").write(s.replace(old, new))


def main():
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    tmp = tempfile.mkdtemp(prefix="phb_patch_")
    a, b = os.path.join(tmp, "a"), os.path.join(tmp, "b")
    for root in (a, b):
        for f in FILES:
            os.makedirs(os.path.dirname(os.path.join(root, f)), exist_ok=True)
            shutil.copy(os.path.join(ref, f), os.path.join(root, f))
    os.makedirs(os.path.join(b, "src/tree/likelihood"), exist_ok=True)
    shutil.copy(os.path.join(HERE, "shim", "b200.rs"), os.path.join(b, "src/tree/likelihood/b200.rs"))
    edit(os.path.join(b, "Cargo.toml"), 'parallel = ["dep:rayon"]\n',
         'parallel = ["dep:rayon"]\n# Likelihood / ASR engines on an NVIDIA B200 through libphylo_b200.so (no host fallback behind it)\n'
         'b200 = ["simple_rooted_tree", "dep:phylo-b200-sys"]\n')
    edit(os.path.join(b, "Cargo.toml"), 'rayon = { version = "1.10.0", optional = true }\n',
         'rayon = { version = "1.10.0", optional = true }\n'
         'phylo-b200-sys = { path = "../phylo-b200/bindings/phylo-b200-sys", optional = true }\n')
    edit(os.path.join(b, "src/models/rate_matrix.rs"),
         "    pub fn equilibrium(&self) -> &DVector<f64> {\n        &self.pi\n    }\n",
         "    pub fn equilibrium(&self) -> &DVector<f64> {\n        &self.pi\n    }\n\n"
         "    /// Eigenvalues of the symmetrized generator `S = V diag(lambda) V^T`.\n"
         "    pub fn eigenvalues(&self) -> &DVector<f64> {\n        &self.eigenvalues\n    }\n\n"
         "    /// Eigenvectors of the symmetrized generator (columns of `V`; nalgebra stores them column-major).\n"
         "    pub fn eigenvectors(&self) -> &DMatrix<f64> {\n        &self.eigenvectors\n    }\n\n"
         "    /// `sqrt(pi)`, the right scaling of `P(t) = diag(1/sqrt(pi)) exp(S t) diag(sqrt(pi))`.\n"
         "    pub fn sqrt_pi(&self) -> &DVector<f64> {\n        &self.sqrt_pi\n    }\n\n"
         "    /// `1 / sqrt(pi)`, the left scaling.\n"
         "    pub fn inv_sqrt_pi(&self) -> &DVector<f64> {\n        &self.inv_sqrt_pi\n    }\n")
    edit(os.path.join(b, "src/models/gtr.rs"),
         "    pub fn equilibrium(&self) -> &nalgebra::DVector<f64> {\n        self.matrix.equilibrium()\n    }\n",
         "    pub fn equilibrium(&self) -> &nalgebra::DVector<f64> {\n        self.matrix.equilibrium()\n    }\n\n"
         "    /// The eigendecomposed rate matrix behind this model (what an accelerator backend uploads).\n"
         "    pub fn rate_matrix(&self) -> &RateMatrix {\n        &self.matrix\n    }\n")
    edit(os.path.join(b, "src/tree/likelihood.rs"),
         "/// Result type carrying reconstructed states and the tree log-likelihood.\npub mod reconstruction;\n",
         "/// Result type carrying reconstructed states and the tree log-likelihood.\npub mod reconstruction;\n\n"
         "/// NVIDIA B200 backend of the engines below (`libphylo_b200.so` through `phylo-b200-sys`).\n"
         "#[cfg(feature = \"b200\")]\npub mod b200;\n")
    for engine, call in (("compute_marginal_asr(self, model, aln, want_posteriors)", "marginal_asr(self, model, aln, want_posteriors)"),
                         ("compute_joint_asr(self, model, aln)", "joint_asr(self, model, aln)"),
                         ("compute_log_likelihood(self, model, aln)", "log_likelihood(self, model, aln)")):
        edit(os.path.join(b, "src/tree.rs"), f"            crate::tree::likelihood::{engine}\n",
             f"            #[cfg(feature = \"b200\")]\n            return crate::tree::likelihood::b200::{call};\n"
             f"            #[cfg(not(feature = \"b200\"))]\n            crate::tree::likelihood::{engine}\n")
    out = subprocess.run(["diff", "-ruN", "a", "b"], cwd=tmp, capture_output=True, text=True).stdout
    lines = []
    for ln in out.splitlines(keepends=True):
        if ln.startswith("diff -ruN a/"):
            f = ln.split()[2][2:]
            ln = f"diff --git a/{f} b/{f}\n"
        elif ln.startswith("--- a/") or ln.startswith("+++ b/"):
            ln = ln.split("\t")[0].rstrip("\n") + "\n"
        lines.append(ln)
    # a file that exists only in b/: the header git wants for a new file
    for i, ln in enumerate(lines):
        if ln.startswith("@@ -0,0 +") and lines[i - 2].startswith("--- a/"):
            lines[i - 2] = "--- /dev/null\n"
            lines[i - 3] = lines[i - 3] + "new file mode 100644\n"
    dst = os.path.join(HERE, "phylo-b200-shim.patch")
    open(dst, "w").write("".join(lines))
    shutil.rmtree(tmp)
    print(f"wrote {dst} ({len(lines)} lines)")


if __name__ == "__main__":
    main()
