//! Raw bindings of `libphylo_b200.so`, one `extern "C"` item per declaration of `include/phylo_b200.h`.
//!
//! The safe layer (flattening a `PhyloTree`, mapping status codes to `AsrError`) lives in the `phylo`
//! crate behind its `b200` feature (`src/tree/likelihood/b200.rs`, see `bindings/phylo-b200-shim.patch`).
#![allow(non_camel_case_types)]

use std::os::raw::c_char;

/// Opaque context (`struct phb_ctx`).
#[repr(C)]
pub struct phb_ctx {
    _private: [u8; 0],
}

// ---- status codes (1..=6 are `AsrError`'s variants in declaration order, src/error.rs:106-125) ----
pub const PHB_OK: i32 = 0;
pub const PHB_ERR_MISSING_BRANCH_LENGTH: i32 = 1;
pub const PHB_ERR_ALPHABET_MISMATCH: i32 = 2;
pub const PHB_ERR_INVALID_ALIGNMENT: i32 = 3;
pub const PHB_ERR_EIGENDECOMP_FAILURE: i32 = 4;
pub const PHB_ERR_NUMERICAL_INSTABILITY: i32 = 5;
pub const PHB_ERR_INVALID_MODEL_PARAMETER: i32 = 6;
pub const PHB_ERR_INVALID_ARGUMENT: i32 = 16;
pub const PHB_ERR_CUDA: i32 = 17;
pub const PHB_ERR_NCCL: i32 = 18;
pub const PHB_ERR_OOM: i32 = 19;
pub const PHB_ERR_STATE: i32 = 20;

// ---- context flags ----
pub const PHB_FLAG_LITERAL_INVARIANT: u64 = 1 << 0;
pub const PHB_FLAG_TIMING: u64 = 1 << 1;
pub const PHB_FLAG_LEVEL_KERNELS: u64 = 1 << 2;
pub const PHB_FLAG_NNI_FULL: u64 = 1 << 3;
pub const PHB_FLAG_NNI_WALK: u64 = 1 << 4;
/// 20-state context: the AminoAcid alphabet (`src/alphabet.rs:66-116`) instead of Nucleotide.
pub const PHB_FLAG_AMINO_ACID: u64 = 1 << 5;

// ---- alignment layouts ----
pub const PHB_LAYOUT_PATTERN_MAJOR: i32 = 0;
pub const PHB_LAYOUT_ROW_MAJOR: i32 = 1;

/// `phb_tree_desc`: a flattened candidate tree (the arguments of `phb_set_tree`).
#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct phb_tree_desc {
    pub n_nodes: i32,
    pub postorder: *const i32,
    pub child_offsets: *const i32,
    pub children: *const i32,
    pub branch_len: *const f64,
    pub leaf_row: *const i32,
    pub root: i32,
}

/// `phb_stats`: counters and timings of the last evaluation.
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct phb_stats {
    pub n_patterns: i64,
    pub tile_patterns: i64,
    pub n_tiles: i32,
    pub n_levels: i32,
    pub n_internal: i32,
    pub n_clv_categories: i32,
    pub kernel_launches: i64,
    pub clv_updates: i64,
    pub algorithmic_bytes: f64,
    pub ms_transition: f64,
    pub ms_clv: f64,
    pub ms_root: f64,
    pub ms_down: f64,
    pub ms_total: f64,
    pub clv_launches: i32,
    pub sweep_ops: i32,
    pub sweep_stack_levels: i32,
    pub sweep_smem_levels: i32,
    pub sweep_stages: i32,
    pub sweep_block_sites: i32,
    pub sweep_stream_bytes: i64,
    pub nni_path_moves: i32,
    pub nni_tiles: i32,
    pub nni_path_steps: i64,
    pub nni_cache_bytes: f64,
    pub nni_read_bytes: f64,
    pub sweep_guards: i32,
    pub sweep_normouts: i32,
    pub sweep_fp64_instr: f64,
    pub sweep_flops: f64,
    pub ms_nni_cache: f64,
    pub ms_nni_upper: f64,
    pub ms_nni_edge: f64,
}

/// Size of the fixed device staging array behind `phb_transition_matrices_noalloc`.
pub const PHB_NOALLOC_SCRATCH_BYTES: usize = 2 << 20;

extern "C" {
    pub fn phb_version() -> *const c_char;
    pub fn phb_last_error(ctx: *mut phb_ctx) -> *const c_char;
    pub fn phb_ctx_create(devices: *const i32, n_dev: i32, flags: u64, out: *mut *mut phb_ctx) -> i32;
    pub fn phb_ctx_destroy(ctx: *mut phb_ctx);
    pub fn phb_set_clv_budget(ctx: *mut phb_ctx, bytes: u64) -> i32;
    pub fn phb_set_alignment(
        ctx: *mut phb_ctx,
        n_rows: i32,
        n_patterns: i64,
        bytes: *const u8,
        layout: i32,
        multiplicity: *const u64,
    ) -> i32;
    pub fn phb_compress_columns(
        ctx: *mut phb_ctx,
        n_rows: i32,
        width: i64,
        bytes: *const u8,
        n_patterns: *mut i64,
        site_to_pattern: *mut u64,
        multiplicity: *mut u64,
    ) -> i32;
    pub fn phb_compress_rows(
        ctx: *mut phb_ctx,
        n_rows: i32,
        width: i64,
        rows: *const *const u8,
        n_patterns: *mut i64,
        site_to_pattern: *mut u64,
        multiplicity: *mut u64,
    ) -> i32;
    pub fn phb_set_model(
        ctx: *mut phb_ctx,
        lambda: *const f64,
        v: *const f64,
        pi: *const f64,
        sqrt_pi: *const f64,
        inv_sqrt_pi: *const f64,
        n_cat: i32,
        rates: *const f64,
        weights: *const f64,
    ) -> i32;
    pub fn phb_set_tree(
        ctx: *mut phb_ctx,
        n_nodes: i32,
        postorder: *const i32,
        child_offsets: *const i32,
        children: *const i32,
        branch_len: *const f64,
        leaf_row: *const i32,
        root: i32,
    ) -> i32;
    pub fn phb_log_likelihood(ctx: *mut phb_ctx, total: *mut f64, per_pattern: *mut f64) -> i32;
    pub fn phb_log_likelihood_host(
        ctx: *mut phb_ctx,
        n_rows: i32,
        n_patterns: i64,
        bytes: *const u8,
        layout: i32,
        multiplicity: *const u64,
        total: *mut f64,
        per_pattern: *mut f64,
    ) -> i32;
    pub fn phb_get_per_pattern(ctx: *mut phb_ctx, per_pattern: *mut f64) -> i32;
    pub fn phb_marginal_asr(ctx: *mut phb_ctx, states: *mut u8, posteriors: *mut f64, total: *mut f64) -> i32;
    pub fn phb_joint_asr(ctx: *mut phb_ctx, states: *mut u8, total: *mut f64) -> i32;
    pub fn phb_score_nni(
        ctx: *mut phb_ctx,
        n_moves: i32,
        node_id: *const i32,
        left_ch: *const u8,
        totals: *mut f64,
    ) -> i32;
    pub fn phb_score_trees(ctx: *mut phb_ctx, n_trees: i32, trees: *const phb_tree_desc, totals: *mut f64) -> i32;
    pub fn phb_simulate_alignment(ctx: *mut phb_ctx, n_patterns: i64, global_offset: i64, seed: u64) -> i32;
    pub fn phb_get_alignment(ctx: *mut phb_ctx, bytes: *mut u8, layout: i32, first_pattern: i64, n_patterns: i64) -> i32;
    pub fn phb_get_stats(ctx: *mut phb_ctx, out: *mut phb_stats) -> i32;
    pub fn phb_timer_start(ctx: *mut phb_ctx) -> i32;
    pub fn phb_timer_stop(ctx: *mut phb_ctx, ms: *mut f64) -> i32;
    pub fn phb_transition_matrices(ctx: *mut phb_ctx, n_t: i32, t: *const f64, out: *mut f64) -> i32;
    /// Allocation-free variant (fixed device staging array of `PHB_NOALLOC_SCRATCH_BYTES`; no reference counterpart).
    pub fn phb_transition_matrices_noalloc(ctx: *mut phb_ctx, n_t: i32, t: *const f64, out: *mut f64) -> i32;
    pub fn phb_measure_fp64_tflops(ctx: *mut phb_ctx, tflops: *mut f64) -> i32;
}

#[cfg(test)]
mod tests {
    use super::*;

    /// Needs a B200 at run time; on any other machine `phb_ctx_create` must fail loudly (PHB_ERR_CUDA).
    #[test]
    fn context_or_loud_failure() {
        let mut ctx: *mut phb_ctx = std::ptr::null_mut();
        let rc = unsafe { phb_ctx_create(std::ptr::null(), 0, 0, &mut ctx) };
        if rc == PHB_OK {
            unsafe { phb_ctx_destroy(ctx) };
        } else {
            assert_eq!(rc, PHB_ERR_CUDA);
        }
    }
}
