"""The reference-side binding as files (bindings/): the raw FFI crate declares exactly the header's entry points, the
Rust sources balance, the patch applies to the reference, and the same call sequence runs from plain C
(tests/test_abi_c.c) against the emulation build (CPU) and the product library (`-m gpu`)."""
import os
import re
import shutil
import subprocess
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SYS = os.path.join(ROOT, "bindings", "phylo-b200-sys", "src", "lib.rs")
SHIM = os.path.join(ROOT, "bindings", "shim", "b200.rs")
PATCH = os.path.join(ROOT, "bindings", "phylo-b200-shim.patch")


def strip_rust(src):
    """drop comments, string / char literals and lifetimes so that only structure is left"""
    src = re.sub(r"//[^\n]*", "", src)
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r'"(?:\\.|[^"\\])*"', '""', src)
    src = re.sub(r"'(?:\\.|[^'\\])'", "' '", src)
    return src


def balanced(src):
    pairs = {")": "(", "]": "[", "}": "{"}
    stack = []
    for ch in strip_rust(src):
        if ch in "([{":
            stack.append(ch)
        elif ch in pairs:
            if not stack or stack.pop() != pairs[ch]:
                return False
    return not stack


def header_decls():
    txt = open(os.path.join(ROOT, "include", "phylo_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return txt


def test_rust_sources_balance():
    for path in (SYS, SHIM, os.path.join(ROOT, "bindings", "phylo-b200-sys", "build.rs")):
        assert balanced(open(path).read()), path


def test_sys_crate_declares_the_header():
    h = header_decls()
    declared = sorted(set(re.findall(r"\b(phb_[a-z0-9_]+)\s*\(", h)))
    rs = strip_rust(open(SYS).read())
    bound = sorted(set(re.findall(r"pub fn (phb_[a-z0-9_]+)\s*\(", rs)))
    assert bound == declared
    # same number of parameters per entry point
    for name in declared:
        c_args = re.search(rf"\b{name}\s*\(([^;]*?)\)\s*;", h, flags=re.S).group(1)
        r_args = re.search(rf"pub fn {name}\s*\(([^;]*?)\)\s*(?:->[^;]*)?;", rs, flags=re.S).group(1)
        n_c = 0 if c_args.strip() in ("", "void") else c_args.count(",") + 1
        n_r = 0 if not r_args.strip() else len([a for a in r_args.split(",") if a.strip()])
        assert n_c == n_r, (name, n_c, n_r)
    # status codes and flags carry the header's values
    for m in re.finditer(r"#define (PHB_(?:OK|ERR_[A-Z_]+|LAYOUT_[A-Z_]+)) (\d+)", h):
        assert re.search(rf"pub const {m.group(1)}: i32 = {m.group(2)};", rs), m.group(1)
    for m in re.finditer(r"#define (PHB_FLAG_[A-Z_]+) \(1ull << (\d+)\)", h):
        assert re.search(rf"pub const {m.group(1)}: u64 = 1 << {m.group(2)};", rs), m.group(1)
    # phb_stats: field for field
    c_fields = re.findall(r"\b(?:int32_t|int64_t|double)\s+([a-z0-9_]+);", re.search(r"typedef struct phb_stats \{(.*?)\} phb_stats;", h, flags=re.S).group(1))
    r_fields = re.findall(r"pub ([a-z0-9_]+):", re.search(r"pub struct phb_stats \{(.*?)\}", rs, flags=re.S).group(1))
    assert c_fields == r_fields


def test_shim_uses_only_bound_entry_points_and_real_fields():
    rs = strip_rust(open(SYS).read())
    bound = set(re.findall(r"pub fn (phb_[a-z0-9_]+)\s*\(", rs))
    shim = strip_rust(open(SHIM).read())
    used = set(re.findall(r"sys::(phb_[a-z0-9_]+)\s*\(", shim))
    assert used and used <= bound
    # Reconstruction's field is `alphabet` (src/tree/likelihood/reconstruction.rs:14), AsrError's unit variants take no payload
    assert "alphabet: PhantomData" in shim and "_marker" not in shim
    for unit in ("MissingBranchLength", "EigendecompFailure", "NumericalInstability"):
        assert not re.search(rf"AsrError::{unit}\s*\(", shim), unit


@pytest.mark.skipif(not os.path.isdir("/root/reference/src"), reason="the reference checkout is only present in the build container")
def test_patch_applies_to_the_reference():
    tmp = tempfile.mkdtemp(prefix="phb_ref_")
    try:
        dst = os.path.join(tmp, "ref")
        shutil.copytree("/root/reference", dst, ignore=shutil.ignore_patterns(".git", "comparison", "target"))
        r = subprocess.run(["git", "apply", "--check", PATCH], cwd=dst, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        subprocess.run(["git", "apply", PATCH], cwd=dst, check=True)
        assert open(os.path.join(dst, "src/tree/likelihood/b200.rs")).read() == open(SHIM).read()
        for f in ("src/tree.rs", "src/tree/likelihood.rs", "src/models/rate_matrix.rs", "src/models/gtr.rs"):
            assert balanced(open(os.path.join(dst, f)).read()), f
        assert 'b200 = ["simple_rooted_tree", "dep:phylo-b200-sys"]' in open(os.path.join(dst, "Cargo.toml")).read()
    finally:
        shutil.rmtree(tmp)


def _run_c_test(binary, env=None):
    r = subprocess.run([binary], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    return r.stdout


def test_c_abi_program_on_the_emulation_build():
    import __graft_entry__ as g

    g.build_emu()
    out = _run_c_test(g.build_abi_c_test(emu=True))
    assert "-> ok" in out


@pytest.mark.gpu
def test_c_abi_program_on_the_product_library():
    import __graft_entry__ as g

    out = _run_c_test(g.build_abi_c_test(emu=False))
    assert "sm_100a" in out and "-> ok" in out
