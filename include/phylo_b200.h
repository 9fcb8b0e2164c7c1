/*
 * phylo_b200.h — C ABI of libphylo_b200.so: the B200 (sm_100a) replacement for the reference's
 * Felsenstein-pruning likelihood / marginal-ASR engines.
 *
 * The reference (sriram98v/phylo-rs, crate `phylo` 6.0.0) has NO FFI layer today; its engines are
 * the Rust functions
 *     compute_log_likelihood   src/tree/likelihood.rs:156-209   (called from src/tree.rs:68-76)
 *     compute_marginal_asr     src/tree/likelihood.rs:224-408   (called from src/tree.rs:47-56)
 *     prune_pattern_category   src/tree/likelihood.rs:83-143    (shared pruning core)
 *     RateMatrix::transition   src/models/rate_matrix.rs:113-135 (per-branch P(t))
 * A Rust shim keeps the `TreeLikelihood` / `MarginalAsr` trait surface, flattens
 * the arena-allocated `PhyloTree`, the `GtrModel` eigen-system and the `CompressedColumns` into the
 * plain arrays below and binds exactly these entry points.
 *
 * Conventions: all functions are extern "C", return an int32 status (0 = OK), never throw or abort
 * across the ABI.  The caller owns every input and output buffer; inputs are copied before a
 * `phb_set_*` call returns; outputs are written into caller-allocated memory.  The context owns all
 * device memory.  A context is NOT thread-safe (one in-flight call); use one per thread or a mutex.
 * -inf / NaN log-likelihoods pass through as values (zero-likelihood data is not an error in the
 * reference: src/tree/likelihood.rs:140, :61-68).
 */
#ifndef PHYLO_B200_H
#define PHYLO_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- status codes: 1..6 map 1:1 onto AsrError (src/error.rs:106-125) ------------------------- */
#define PHB_OK 0
#define PHB_ERR_MISSING_BRANCH_LENGTH 1   /* AsrError::MissingBranchLength (defined, never raised) */
#define PHB_ERR_ALPHABET_MISMATCH 2       /* invalid byte in a LEAF row (likelihood.rs:105-107)     */
#define PHB_ERR_INVALID_ALIGNMENT 3       /* a leaf has no alignment row (likelihood.rs:99-103)     */
#define PHB_ERR_EIGENDECOMP_FAILURE 4     /* AsrError::EigendecompFailure (defined, never raised)   */
#define PHB_ERR_NUMERICAL_INSTABILITY 5   /* malformed tree (child missing: likelihood.rs:114,135)  */
#define PHB_ERR_INVALID_MODEL_PARAMETER 6 /* AsrError::InvalidModelParameter                        */
#define PHB_ERR_INVALID_ARGUMENT 16       /* null pointer, bad size, bad layout flag                */
#define PHB_ERR_CUDA 17                   /* CUDA runtime error or no usable sm_100 device          */
#define PHB_ERR_NCCL 18                   /* reserved for the in-library collective                 */
#define PHB_ERR_OOM 19                    /* device allocation failed                               */
#define PHB_ERR_STATE 20                  /* call order (e.g. log_likelihood before set_tree)       */

/* ---- context flags ------------------------------------------------------------------------- */
/* Treat rate-0 (+I) categories as ordinary CLV-carrying categories (the reference's literal
 * K+1-category evaluation, src/models/gtr.rs:94-99) instead of the closed form over tip masks. */
#define PHB_FLAG_LITERAL_INVARIANT (1ull << 0)
/* Record CUDA events around every phase of an evaluation (see phb_get_stats). */
#define PHB_FLAG_TIMING (1ull << 1)
/* Level-batched kernels that keep every CLV in HBM, for the log-likelihood AND for the up pass of
 * phb_marginal_asr, instead of the fused on-chip sweep (default log-likelihood path) and its storing
 * variant (default marginal-ASR up pass).  Same results; kept for cross-checking. */
#define PHB_FLAG_LEVEL_KERNELS (1ull << 2)
/* phb_score_nni rescores every neighbour from scratch (the reference's caller loop) instead of walking
 * v -> root against the cached CLVs of the base tree; for cross-checking the two. */
#define PHB_FLAG_NNI_FULL (1ull << 3)
/* phb_score_nni walks v -> root per neighbour (one cached CLV read per ancestor) instead of using the
 * pre-order "upper" partials that make a neighbour a constant-cost edge-local evaluation.  The walk kernel is a
 * cross-check only: it is compiled with -DPHB_WITH_WALK (the CPU emulation harness does); the product build
 * ignores the flag. */
#define PHB_FLAG_NNI_WALK (1ull << 4)

/* The context evaluates the 20-state AminoAcid alphabet (src/alphabet.rs:66-116; the reference's engines are generic
 * over A: Alphabet, likelihood.rs:156-166, 224-235, with GtrModel<A>::new(pi, W, normalize), gtr.rs:39-50) instead of
 * Nucleotide.  In such a context
 *   phb_set_alignment        accepts AminoAcid::profile's bytes: ACDEFGHIKLMNPQRSTVWY, B, Z, J (two states at 0.5 --
 *                            the state INDICES the reference sets: 3|11, 4|14, 8|9), X and '-' (all ones);
 *   phb_set_model            takes lambda[20], V[400] (V[i + 20*k]), pi / sqrt_pi / inv_sqrt_pi[20];
 *   phb_set_tree, phb_log_likelihood, phb_get_per_pattern, phb_set_clv_budget   as documented below;
 *   phb_marginal_asr         states are indices 0..19 into CANONICAL, posteriors [n_nodes][n_patterns][20];
 *   phb_joint_asr            as documented below, states 0..19;
 *   phb_transition_matrices  out[(k * n_cat + c) * 400 + 20 * i + j];
 *   phb_compress_columns, phb_compress_rows   as documented below (the compressor compares bytes and does not look at
 *                            the alphabet, as Alignment::compress_columns does not); the unique columns are encoded to
 *                            20-state tip codes on the device;
 *   phb_get_alignment        the 25 accepted bytes read back as themselves, any other byte as '?';
 *   phb_log_likelihood_host  phb_set_alignment + phb_log_likelihood in one call (not overlapped);
 *   phb_score_nni            every neighbour is two node updates and a dot product against the base tree's resident
 *                            partials and upper partials (any arity);
 *   phb_score_trees          every candidate is re-scheduled and scored in full against the resident alignment;
 * rate-0 categories are evaluated literally (identity P).  One device per context.  phb_simulate_alignment and
 * phb_transition_matrices_noalloc have no 20-state form and return PHB_ERR_STATE. */
#define PHB_FLAG_AMINO_ACID (1ull << 5)

/* ---- alignment layouts --------------------------------------------------------------------- */
#define PHB_LAYOUT_PATTERN_MAJOR 0 /* bytes[pattern][row] — CompressedColumns.patterns, alignment.rs:129-138 */
#define PHB_LAYOUT_ROW_MAJOR 1     /* bytes[row][pattern] */

typedef struct phb_ctx phb_ctx;

/* Create a context driving `n_dev` CUDA devices (device ordinals in `devices`; n_dev == 0 or
 * devices == NULL means device 0 only).  Site patterns are sharded contiguously across the devices;
 * every device holds the full tree and model.  Partial log-likelihoods are combined by a
 * rank-ordered host sum (deterministic). */
int32_t phb_ctx_create(const int32_t *devices, int32_t n_dev, uint64_t flags, phb_ctx **out);
void phb_ctx_destroy(phb_ctx *ctx);

/* Cap on device bytes used for CLV tiles per device (0 = 85 % of free memory).  Smaller caps force
 * more site tiles; results are identical for any tiling. */
int32_t phb_set_clv_budget(phb_ctx *ctx, uint64_t bytes);

/* Site patterns.  Replaces `Alignment::compress_columns()` output (src/alignment.rs:93-125) as the
 * engines consume it at src/tree/likelihood.rs:164,186-187.
 *   n_rows       rows per pattern (= CompressedColumns.leaf_order.len(); rows naming internal nodes
 *                are legal and ignored by pruning, likelihood.rs:98-108)
 *   n_patterns   number of distinct patterns P
 *   bytes        ASCII alignment bytes (the 17 codes of Nucleotide::profile, src/alphabet.rs:44-64),
 *                layout per `layout`
 *   multiplicity CompressedColumns.multiplicity (nullable => all 1)
 * Bytes are validated lazily and only for rows that turn out to be leaves (as the reference does). */
int32_t phb_set_alignment(phb_ctx *ctx, int32_t n_rows, int64_t n_patterns, const uint8_t *bytes,
                          int32_t layout, const uint64_t *multiplicity);

/* Substitution model.  Replaces GtrModel::{equilibrium,categories,category_transition}
 * (src/models/gtr.rs:61-74) and the RateMatrix eigen-system (src/models/rate_matrix.rs:15-25,100):
 *   lambda[4]      eigenvalues of the symmetrised rate matrix
 *   V[16]          eigenvectors, COLUMN-major as nalgebra stores them: V[i + 4*k] = component i of
 *                  eigenvector k
 *   pi, sqrt_pi, inv_sqrt_pi   normalised equilibrium frequencies and their roots
 *   n_cat, rates, weights      RateCategory list in the reference's order (+I first, gtr.rs:94-99)
 * P(t) is evaluated on the device per (edge, category) as
 *   diag(inv_sqrt_pi) * V * diag(exp(lambda * rate * t)) * V^T * diag(sqrt_pi),
 * and is the exact identity when rate * t == 0.0 (rate_matrix.rs:115-117). */
/* Replaces Alignment::compress_columns (src/alignment.rs:93-125) on the device and leaves the result as the
 * context's alignment (as if phb_set_alignment had been called with CompressedColumns.patterns and
 * .multiplicity, src/alignment.rs:129-138).  bytes: the alignment as the reference holds it, one sequence per
 * row, [n_rows][width] raw bytes (already upper-cased by from_fasta_bytes); the caller's row order is
 * CompressedColumns.leaf_order.  Columns are compared over ALL rows byte for byte ('T' != 'U'; rows naming
 * internal nodes take part), and patterns are numbered in order of first appearance, exactly as the
 * reference's HashMap walk does.  *n_patterns receives the pattern count; site_to_pattern (nullable) receives
 * `width` indices; multiplicity (nullable, capacity `width`) receives *n_patterns counts.
 * The canonical pattern bytes can be read back with phb_get_alignment. */
int32_t phb_compress_columns(phb_ctx *ctx, int32_t n_rows, int64_t width, const uint8_t *bytes, int64_t *n_patterns,
                             uint64_t *site_to_pattern, uint64_t *multiplicity);

/* Same, with one pointer per row: rows[r] -> the `width` bytes of sequence r (the reference keeps each sequence
 * in its own Vec<u8>, Alignment.seqs src/alignment.rs:5-10, so the Rust shim passes the Vec pointers as they are). */
int32_t phb_compress_rows(phb_ctx *ctx, int32_t n_rows, int64_t width, const uint8_t *const *rows, int64_t *n_patterns,
                          uint64_t *site_to_pattern, uint64_t *multiplicity);

int32_t phb_set_model(phb_ctx *ctx, const double *lambda, const double *V, const double *pi,
                      const double *sqrt_pi, const double *inv_sqrt_pi, int32_t n_cat,
                      const double *rates, const double *weights);

/* Tree schedule.  Replaces the accessor calls of src/tree/likelihood.rs:94-127,180-183,266-273:
 *   n_nodes        dense node count (the shim re-indexes the sparse arena, src/tree.rs:118)
 *   postorder      n_nodes dense ids in DFSPostOrderIterator order (src/iter.rs:112-138)
 *   child_offsets  CSR offsets, n_nodes + 1 entries
 *   children       CSR child ids in arena (left-to-right) order; any arity >= 0
 *   branch_len     per node: weight of the edge (parent -> node) widened f32 -> f64, None -> 0.0
 *                  (likelihood.rs:115-118); the root's entry is ignored
 *   leaf_row       per node: row inside a pattern for leaves (first row naming the node), -1 if none
 *   root           dense id of the root
 * A leaf with leaf_row == -1 yields PHB_ERR_INVALID_ALIGNMENT at evaluation time. */
int32_t phb_set_tree(phb_ctx *ctx, int32_t n_nodes, const int32_t *postorder,
                     const int32_t *child_offsets, const int32_t *children, const double *branch_len,
                     const int32_t *leaf_row, int32_t root);

/* compute_log_likelihood (src/tree/likelihood.rs:156-209).
 *   total        sum_p multiplicity[p] * logsumexp_c(cat_ll[p][c])
 *   per_pattern  nullable; n_patterns per-pattern log-likelihoods (site_ll before the multiplicity) */
int32_t phb_log_likelihood(phb_ctx *ctx, double *total, double *per_pattern);

/* The reference's call shape in ONE entry: `tree.log_likelihood(&model, &aln)` hands over the alignment on
 * every call (src/tree/likelihood.rs:156-164), so the binding's natural unit is "host bytes in, lnL out".
 * Equivalent to phb_set_alignment + phb_log_likelihood (same results bit for bit), but for the pattern-major
 * layout the host->device copy of chunk k+1 overlaps the packing and the sweep of chunk k, so the call
 * costs about max(PCIe copy, compute) instead of their sum.  Requires set_model + set_tree; leaves the
 * alignment resident for later phb_log_likelihood / phb_marginal_asr / phb_score_nni calls. */
int32_t phb_log_likelihood_host(phb_ctx *ctx, int32_t n_rows, int64_t n_patterns, const uint8_t *bytes,
                                int32_t layout, const uint64_t *multiplicity, double *total,
                                double *per_pattern);

/* compute_marginal_asr (src/tree/likelihood.rs:224-408).
 *   states       [n_nodes][n_patterns] argmax state index 0..3 (A,C,G,T) for EVERY node incl. leaves
 *                (likelihood.rs:262-264), last-maximal-index tie rule (:385-390)
 *   posteriors   nullable; [n_nodes][n_patterns][4] mixed, renormalised posteriors (:365-382)
 *   total        as phb_log_likelihood
 * The shim expands patterns to sites through CompressedColumns.site_to_pattern (:391-398). */
int32_t phb_marginal_asr(phb_ctx *ctx, uint8_t *states, double *posteriors, double *total);

/* GtrModel::category_transition (src/models/gtr.rs:61-64) -> RateMatrix::transition
 * (src/models/rate_matrix.rs:113-135), evaluated by the same device kernel the engines use:
 *   out[(k * n_cat + c) * 16 + 4 * i + j] = P(rates[c] * t[k])[i][j]      (row-major 4x4)
 * The exact identity when rates[c] * t[k] == 0.0.  Requires set_model. */
int32_t phb_transition_matrices(phb_ctx *ctx, int32_t n_t, const double *t, double *out);

/* Allocation-free variant of phb_transition_matrices.  It has no counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments, same result (bit for bit: the same kernel), but the
 * device staging memory is a fixed file-scope device array of PHB_NOALLOC_SCRATCH_BYTES that every GPU loading the
 * library holds until exit, and the call performs no cudaMalloc / cudaFree / allocating library call.
 *   limit    8 * n_t * (1 + 16 * n_cat) <= PHB_NOALLOC_SCRATCH_BYTES (n_cat = categories incl. the invariant one:
 *            4,032 branch lengths for GTR+G4, 3,236 for GTR+I+G4); beyond it PHB_ERR_INVALID_ARGUMENT with a message
 *            naming the limit, nothing launched
 *   sharing  every call on a device uses the same array, which keeps the previous call's contents; nothing in it has
 *            to start at zero (each call overwrites all it reads back).  Calls that overlap on one GPU -- two contexts,
 *            two threads -- race on it: serialise them.  The call returns after its stream has drained. */
#define PHB_NOALLOC_SCRATCH_BYTES (2u << 20)
int32_t phb_transition_matrices_noalloc(phb_ctx *ctx, int32_t n_t, const double *t, double *out);

/* Replaces compute_joint_asr (src/tree/likelihood.rs:421-560; trait JointAsr::joint_asr src/tree/asr.rs:30-37,
 * impl src/tree.rs:58-66): Viterbi reconstruction per rate category in log space, the category with the
 * greatest ln(weight) + root maximum wins (first on ties), states by traceback.  states: [n_nodes][n_patterns]
 * state indices 0..3 for EVERY node; a pattern for which no category has a value > -inf keeps state 0
 * everywhere, as in the reference.  *total = sum_p multiplicity_p * best value_p (Reconstruction.log_likelihood);
 * the per-pattern best values are available through phb_get_per_pattern. */
int32_t phb_joint_asr(phb_ctx *ctx, uint8_t *states, double *total);

/* Per-pattern log-likelihoods of the most recent evaluation (either entry point). */
int32_t phb_get_per_pattern(phb_ctx *ctx, double *per_pattern);

/* Batch scoring of NNI neighbours of the tree set by phb_set_tree (config 5; caller loop
 * `t.nni(node, left_ch); t.log_likelihood(..)`, NNI::nni src/tree.rs:1213-1242).  For each i the
 * tree is the base tree with nni(node_id[i], left_ch[i]) applied (dense ids); totals[i] receives
 * its log-likelihood.  Invalid moves yield NaN in totals[i]: an unknown node, a leaf or the root (the reference's
 * Err), and the two moves its indexing panics on (the second child of a unifurcation, a node without a sibling). */
int32_t phb_score_nni(phb_ctx *ctx, int32_t n_moves, const int32_t *node_id, const uint8_t *left_ch,
                      double *totals);

/* A flattened candidate tree, with the arguments of phb_set_tree. */
typedef struct phb_tree_desc {
    int32_t n_nodes;
    const int32_t *postorder;
    const int32_t *child_offsets;
    const int32_t *children;
    const double *branch_len;
    const int32_t *leaf_row;
    int32_t root;
} phb_tree_desc;

/* Replaces the caller loop of any rearrangement other than NNI -- SPR::spr (src/tree/ops.rs:34-41: prune + graft),
 * rerooting, ... --  `let mut t = base.clone(); t.spr(e1, e2)?; t.log_likelihood(&model, &aln)?`  with one call:
 * totals[i] = log-likelihood of trees[i] against the context's resident alignment and model (no per-candidate
 * compress_columns, no re-upload).  The context's own tree is untouched.  A candidate that cannot be scheduled
 * (malformed child lists, a leaf without an alignment row, an unaccepted byte in one of its rows) scores NaN. */
int32_t phb_score_trees(phb_ctx *ctx, int32_t n_trees, const phb_tree_desc *trees, double *totals);

/* ---- synthetic inputs (benchmark harness) ------------------------------------ */
/* Simulate `n_patterns` site patterns down the current tree under the current model directly into
 * the context's device-resident alignment (rows = leaves in ascending dense id, multiplicity 1).
 * Pattern p of this call is global site `global_offset + p`, so shards of one logical alignment
 * can be generated independently per rank.  Requires set_model + set_tree.  Rewrites leaf_row. */
int32_t phb_simulate_alignment(phb_ctx *ctx, int64_t n_patterns, int64_t global_offset, uint64_t seed);
/* Copy the context's alignment back to the host as ASCII bytes in the given layout. */
int32_t phb_get_alignment(phb_ctx *ctx, uint8_t *bytes, int32_t layout, int64_t first_pattern,
                          int64_t n_patterns);

/* ---- introspection ------------------------------------------------------------------------- */
typedef struct phb_stats {
    int64_t n_patterns;        /* patterns held by this context (all devices)                     */
    int64_t tile_patterns;     /* patterns per CLV tile on device 0                               */
    int32_t n_tiles;           /* tiles per evaluation on device 0                                */
    int32_t n_levels;          /* post-order levels (kernel batches per tile)                     */
    int32_t n_internal;        /* internal nodes m                                                */
    int32_t n_clv_categories;  /* categories that carry CLVs (rate > 0 unless LITERAL_INVARIANT)  */
    int64_t kernel_launches;   /* kernels launched by this context since creation                 */
    int64_t clv_updates;       /* CLV site-node updates performed by the last evaluation          */
    double algorithmic_bytes;  /* SURVEY.md §8d bytes of the last evaluation's CLV kernels        */
    double ms_transition;      /* PHB_FLAG_TIMING: device ms in the P(t) kernel (last evaluation)  */
    double ms_clv;             /* ... in the level-batched CLV update kernels (sum over launches)  */
    double ms_root;            /* ... in the tip-scale, root and reduce kernels                    */
    double ms_down;            /* ... in the ASR pre-order kernels                                 */
    double ms_total;           /* ... first launch to last launch of the evaluation                */
    int32_t clv_launches;      /* number of CLV kernel launches in the last evaluation             */
    int32_t sweep_ops;         /* fused sweep: ops per site block (0 = the level kernels ran)       */
    int32_t sweep_stack_levels;/* ... stack levels the schedule needs (Sethi-Ullman number)         */
    int32_t sweep_smem_levels; /* ... of which held in shared memory (the rest: L2-resident stack)  */
    int32_t sweep_stages;      /* ... TMA ring depth                                                */
    int32_t sweep_block_sites; /* ... site patterns per CTA pass                                    */
    int64_t sweep_stream_bytes;/* ... bytes of the op-ordered P(t)/tip-table stream                 */
    int32_t nni_path_moves;    /* last phb_score_nni: neighbours scored by the path-only kernel     */
    int32_t nni_tiles;         /* ... CLV-cache tiles per device                                    */
    int64_t nni_path_steps;    /* ... walk steps summed over those neighbours                       */
    double nni_cache_bytes;    /* ... bytes of cached partials (lower rows, upper rows, exponents) RESIDENT on device 0: one tile at a time */
    double nni_read_bytes;     /* ... algorithmic bytes the walks read (sibling CLVs, exponents, tips)*/
    int32_t sweep_guards;      /* fused sweep: nodes followed by the lazy underflow check (2^-256 threshold) */
    int32_t sweep_normouts;    /* ... nodes normalised to max in [1,2) on completion (extreme branch x rate)  */
    double sweep_fp64_instr;   /* ... FP64 instructions (DMUL/DFMA) one sweep executes per (pattern, CLV category) */
    double sweep_flops;        /* ... the same in flops (DFMA = 2)                                           */
    double ms_nni_cache;       /* PHB_FLAG_TIMING, last phb_score_nni: device ms caching the base tree's CLVs  */
    double ms_nni_upper;       /* ... in the pre-order pass that forms the upper partials                      */
    double ms_nni_edge;        /* ... in the edge kernel that scores the neighbours (+ its row reduction)      */
} phb_stats;
int32_t phb_get_stats(phb_ctx *ctx, phb_stats *out);

/* Device-side stopwatch for harnesses: phb_timer_start records a CUDA event on every device stream of
 * the context, phb_timer_stop records a second one, waits for it, and returns the elapsed device time
 * in ms (max over the context's devices).  Everything the context enqueued in between is covered. */
int32_t phb_timer_start(phb_ctx *ctx);
int32_t phb_timer_stop(phb_ctx *ctx, double *ms);

/* FP64 FMA throughput microbenchmark on device 0 (TFLOP/s); confirms the path is HBM-bound. */
int32_t phb_measure_fp64_tflops(phb_ctx *ctx, double *tflops);

const char *phb_last_error(phb_ctx *ctx);
const char *phb_version(void);

#ifdef __cplusplus
}
#endif
#endif /* PHYLO_B200_H */
