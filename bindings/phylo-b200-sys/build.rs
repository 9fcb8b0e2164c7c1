// Links libphylo_b200.so.  PHYLO_B200_LIB_DIR names the directory that holds it (default: the in-tree build
// location phylo-rs_b200/csrc of the repository this crate is vendored from); the directory is also put on the
// binary's rpath so that `cargo test` finds the library without LD_LIBRARY_PATH.
use std::env;
use std::path::PathBuf;

fn main() {
    let dir = env::var("PHYLO_B200_LIB_DIR").map(PathBuf::from).unwrap_or_else(|_| {
        PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../../phylo-rs_b200/csrc")
    });
    println!("cargo:rerun-if-env-changed=PHYLO_B200_LIB_DIR");
    println!("cargo:rustc-link-search=native={}", dir.display());
    println!("cargo:rustc-link-lib=dylib=phylo_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", dir.display());
}
