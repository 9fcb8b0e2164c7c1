//! NVIDIA B200 backend for the likelihood / ancestral-reconstruction engines (feature `b200`).
//!
//! Same traits, same call sites: with the feature on, [`TreeLikelihood`](super::TreeLikelihood),
//! [`MarginalAsr`](crate::tree::asr::MarginalAsr) and [`JointAsr`](crate::tree::asr::JointAsr) for
//! `PhyloTree` hand the flattened tree, the model's eigen-system and the compressed alignment to
//! `libphylo_b200.so` (crate `phylo-b200-sys`) instead of running `prune_pattern_category` on the host.
//! Nothing here does arithmetic: it is the marshalling the C ABI needs, in the order the host engines
//! themselves prepare their inputs (`compute_log_likelihood`, lines 156-183 of `likelihood.rs`).

use std::cell::RefCell;
use std::collections::HashMap;
use std::ffi::CStr;
use std::marker::PhantomData;
use std::ptr;

use num_traits::NumCast;
use phylo_b200_sys as sys;

use crate::alignment::{Alignment, CompressedColumns};
use crate::alphabet::Alphabet;
use crate::error::AsrError;
use crate::models::GtrModel;
use crate::node::NodeID;
use crate::prelude::*;
use crate::tree::likelihood::Reconstruction;
use crate::tree::PhyloTree;

/// One context per thread: a `phb_ctx` is not thread-safe, and keeping it alive keeps the packed
/// alignment, the schedules and the device buffers alive between calls.
struct Ctx(*mut sys::phb_ctx);

impl Drop for Ctx {
    fn drop(&mut self) {
        if !self.0.is_null() {
            unsafe { sys::phb_ctx_destroy(self.0) };
        }
    }
}

thread_local! {
    static CTX: RefCell<Ctx> = RefCell::new(Ctx(ptr::null_mut()));      // Nucleotide (4 states)
    static CTX_AA: RefCell<Ctx> = RefCell::new(Ctx(ptr::null_mut()));   // AminoAcid (20 states)
}

/// The library has one instance of the engines per alphabet of the crate: 4 states, and 20 behind
/// `PHB_FLAG_AMINO_ACID`.
fn ctx<A: Alphabet>() -> Result<*mut sys::phb_ctx, AsrError> {
    let (key, flags) = match A::N_STATES {
        4 => (&CTX, 0u64),
        20 => (&CTX_AA, sys::PHB_FLAG_AMINO_ACID),
        n => {
            return Err(AsrError::AlphabetMismatch(format!(
                "the B200 backend has no {n}-state instance"
            )))
        }
    };
    key.with(|c| {
        let mut c = c.borrow_mut();
        if c.0.is_null() {
            let rc = unsafe { sys::phb_ctx_create(ptr::null(), 0, flags, &mut c.0) };
            if rc != sys::PHB_OK {
                // no usable sm_100 device / CUDA failure: there is no host fallback behind this feature
                return Err(AsrError::NumericalInstability);
            }
        }
        Ok(c.0)
    })
}

/// Status code -> `AsrError` (codes 1..=6 are the variants in declaration order; 16..=20 have no
/// counterpart and surface as `NumericalInstability`).
fn err(c: *mut sys::phb_ctx, rc: i32) -> AsrError {
    let msg = unsafe { CStr::from_ptr(sys::phb_last_error(c)) }
        .to_string_lossy()
        .into_owned();
    match rc {
        sys::PHB_ERR_MISSING_BRANCH_LENGTH => AsrError::MissingBranchLength,
        sys::PHB_ERR_ALPHABET_MISMATCH => AsrError::AlphabetMismatch(msg),
        sys::PHB_ERR_INVALID_ALIGNMENT => AsrError::InvalidAlignment(msg),
        sys::PHB_ERR_EIGENDECOMP_FAILURE => AsrError::EigendecompFailure,
        sys::PHB_ERR_INVALID_MODEL_PARAMETER => AsrError::InvalidModelParameter(msg),
        _ => AsrError::NumericalInstability,
    }
}

fn check(c: *mut sys::phb_ctx, rc: i32) -> Result<(), AsrError> {
    if rc == sys::PHB_OK {
        Ok(())
    } else {
        Err(err(c, rc))
    }
}

/// The tree as `phb_set_tree` wants it: dense node ids, CSR child lists, one branch length per node
/// (the edge above it), the alignment row of every leaf.
struct Flat {
    ids: Vec<NodeID>,
    post: Vec<i32>,
    off: Vec<i32>,
    kids: Vec<i32>,
    blen: Vec<f64>,
    leaf_row: Vec<i32>,
    root: i32,
}

impl Flat {
    fn desc(&self) -> sys::phb_tree_desc {
        sys::phb_tree_desc {
            n_nodes: self.ids.len() as i32,
            postorder: self.post.as_ptr(),
            child_offsets: self.off.as_ptr(),
            children: self.kids.as_ptr(),
            branch_len: self.blen.as_ptr(),
            leaf_row: self.leaf_row.as_ptr(),
            root: self.root,
        }
    }
}

/// Everything `compute_log_likelihood` derives from the tree before its pattern loop.
fn flatten(tree: &PhyloTree, leaf_order: &[String]) -> Result<Flat, AsrError> {
    // only the nodes the traversal from the root reaches: `SPR::prune` leaves detached copies behind
    let root_id = tree.get_root_id();
    let reached: Vec<NodeID> = tree
        .postord_ids(root_id)
        .expect("invariant: the root id always names a node")
        .collect();
    let mut ids = reached.clone();
    ids.sort_unstable();
    let dense: HashMap<NodeID, i32> = ids.iter().enumerate().map(|(k, &v)| (v, k as i32)).collect();
    let mut leaf_row = vec![-1i32; ids.len()];
    for (r, name) in leaf_order.iter().enumerate() {
        let id = tree.get_taxa_node_id(name).ok_or_else(|| {
            AsrError::AlphabetMismatch(format!("Taxon {} in alignment not found in tree", name))
        })?;
        // the host engine finds a leaf's row with `position`, i.e. the first row naming it
        if let Some(&k) = dense.get(&id) {
            if tree.is_leaf(id) && leaf_row[k as usize] < 0 {
                leaf_row[k as usize] = r as i32;
            }
        }
    }
    let mut off = vec![0i32];
    let mut kids: Vec<i32> = Vec::new();
    let mut blen = vec![0f64; ids.len()];
    for &v in &ids {
        for c in tree.get_node_children_ids(v) {
            kids.push(dense[&c]);
            // f32 -> f64; a missing weight reads as 0.0, as in `prune_pattern_category`
            let w: f64 = tree.get_edge_weight(v, c).and_then(NumCast::from).unwrap_or(0.0);
            blen[dense[&c] as usize] = w;
        }
        off.push(kids.len() as i32);
    }
    let post: Vec<i32> = reached.iter().map(|v| dense[v]).collect();
    Ok(Flat {
        root: dense[&root_id],
        ids,
        post,
        off,
        kids,
        blen,
        leaf_row,
    })
}

fn set_model<A: Alphabet>(c: *mut sys::phb_ctx, model: &GtrModel<A>) -> Result<(), AsrError> {
    debug_assert_eq!(model.rate_matrix().eigenvalues().len(), A::N_STATES);
    let rm = model.rate_matrix();
    let cats = model.categories();
    let rates: Vec<f64> = cats.iter().map(|c| c.rate).collect();
    let weights: Vec<f64> = cats.iter().map(|c| c.weight).collect();
    // nalgebra stores matrices column-major, which is the layout `phb_set_model` documents for V
    let rc = unsafe {
        sys::phb_set_model(
            c,
            rm.eigenvalues().as_ptr(),
            rm.eigenvectors().as_ptr(),
            model.equilibrium().as_ptr(),
            rm.sqrt_pi().as_ptr(),
            rm.inv_sqrt_pi().as_ptr(),
            cats.len() as i32,
            rates.as_ptr(),
            weights.as_ptr(),
        )
    };
    check(c, rc)
}

fn set_tree(c: *mut sys::phb_ctx, f: &Flat) -> Result<(), AsrError> {
    let rc = unsafe {
        sys::phb_set_tree(
            c,
            f.ids.len() as i32,
            f.post.as_ptr(),
            f.off.as_ptr(),
            f.kids.as_ptr(),
            f.blen.as_ptr(),
            f.leaf_row.as_ptr(),
            f.root,
        )
    };
    check(c, rc)
}

/// Host-side `compress_columns` (exactly the reference's), then alignment, model and tree across the ABI.
fn load<A: Alphabet>(
    tree: &PhyloTree,
    model: &GtrModel<A>,
    aln: &Alignment,
) -> Result<(*mut sys::phb_ctx, Flat, CompressedColumns), AsrError> {
    let c = ctx::<A>()?;
    let comp = aln.compress_columns();
    let f = flatten(tree, &comp.leaf_order)?;
    let rows = comp.leaf_order.len();
    let bytes: Vec<u8> = comp.patterns.iter().flatten().copied().collect(); // [pattern][row]
    let mult: Vec<u64> = comp.multiplicity.iter().map(|&m| m as u64).collect();
    let rc = unsafe {
        sys::phb_set_alignment(
            c,
            rows as i32,
            comp.patterns.len() as i64,
            bytes.as_ptr(),
            sys::PHB_LAYOUT_PATTERN_MAJOR,
            mult.as_ptr(),
        )
    };
    check(c, rc)?;
    set_model(c, model)?;
    set_tree(c, &f)?;
    Ok((c, f, comp))
}

/// The faster load: the sequences go across as they lie in `Alignment::seqs` and the device performs
/// `compress_columns` (patterns numbered by first appearance, as on the host), for either alphabet.
fn load_dev<A: Alphabet>(
    tree: &PhyloTree,
    model: &GtrModel<A>,
    aln: &Alignment,
) -> Result<(*mut sys::phb_ctx, Flat, Vec<u64>, usize), AsrError> {
    let c = ctx::<A>()?;
    let leaf_order: Vec<String> = aln.seqs.keys().cloned().collect();
    let f = flatten(tree, &leaf_order)?; // errors before the device is touched
    let rows: Vec<*const u8> = leaf_order.iter().map(|n| aln.seqs[n].as_ptr()).collect();
    set_model(c, model)?;
    set_tree(c, &f)?;
    let mut n_patterns = 0i64;
    let mut s2p = vec![0u64; aln.width.max(1)];
    let rc = unsafe {
        sys::phb_compress_rows(
            c,
            rows.len() as i32,
            aln.width as i64,
            rows.as_ptr(),
            &mut n_patterns,
            s2p.as_mut_ptr(),
            ptr::null_mut(),
        )
    };
    check(c, rc)?;
    s2p.truncate(aln.width);
    Ok((c, f, s2p, n_patterns as usize))
}

/// `TreeLikelihood::log_likelihood`.
pub fn log_likelihood<A: Alphabet>(
    tree: &PhyloTree,
    model: &GtrModel<A>,
    aln: &Alignment,
) -> Result<f64, AsrError> {
    let (c, _f, _s2p, _p) = load_dev(tree, model, aln)?;
    let mut total = 0f64;
    let rc = unsafe { sys::phb_log_likelihood(c, &mut total, ptr::null_mut()) };
    check(c, rc)?;
    Ok(total)
}

/// States (and posteriors) per pattern -> per site, for every node the library reconstructed.  Nodes of
/// the arena that the traversal does not reach keep the all-zero sequence `compute_marginal_asr` seeds.
fn expand<A: Alphabet>(
    tree: &PhyloTree,
    f: &Flat,
    site_to_pattern: &[usize],
    n_patterns: usize,
    states: &[u8],
    post: Option<&[f64]>,
    total: f64,
) -> Reconstruction<A> {
    let width = site_to_pattern.len();
    let mut sequences: HashMap<NodeID, Vec<usize>> = HashMap::new();
    let mut posteriors: Option<HashMap<NodeID, Vec<Vec<f64>>>> = post.map(|_| HashMap::new());
    for id in tree.get_node_ids() {
        sequences.insert(id, vec![0; width]);
        if let Some(map) = posteriors.as_mut() {
            map.insert(id, vec![vec![0.0; A::N_STATES]; width]);
        }
    }
    for (k, &id) in f.ids.iter().enumerate() {
        let seq: Vec<usize> = site_to_pattern
            .iter()
            .map(|&s| states[k * n_patterns + s] as usize)
            .collect();
        sequences.insert(id, seq);
        if let (Some(map), Some(p)) = (posteriors.as_mut(), post) {
            let rows: Vec<Vec<f64>> = site_to_pattern
                .iter()
                .map(|&s| {
                    let at = (k * n_patterns + s) * A::N_STATES;
                    p[at..at + A::N_STATES].to_vec()
                })
                .collect();
            map.insert(id, rows);
        }
    }
    Reconstruction {
        sequences,
        posteriors,
        log_likelihood: total,
        alphabet: PhantomData,
    }
}

/// `MarginalAsr::marginal_asr`.
pub fn marginal_asr<A: Alphabet>(
    tree: &PhyloTree,
    model: &GtrModel<A>,
    aln: &Alignment,
    want_posteriors: bool,
) -> Result<Reconstruction<A>, AsrError> {
    let (c, f, comp) = load(tree, model, aln)?;
    let (n, p) = (f.ids.len(), comp.patterns.len());
    let mut states = vec![0u8; n * p.max(1)];
    let mut post = if want_posteriors {
        vec![0f64; n * p.max(1) * A::N_STATES]
    } else {
        Vec::new()
    };
    let mut total = 0f64;
    let post_ptr = if want_posteriors {
        post.as_mut_ptr()
    } else {
        ptr::null_mut()
    };
    let rc = unsafe { sys::phb_marginal_asr(c, states.as_mut_ptr(), post_ptr, &mut total) };
    check(c, rc)?;
    let post_ref = if want_posteriors { Some(post.as_slice()) } else { None };
    Ok(expand(tree, &f, &comp.site_to_pattern, p, &states, post_ref, total))
}

/// `JointAsr::joint_asr`.
pub fn joint_asr<A: Alphabet>(
    tree: &PhyloTree,
    model: &GtrModel<A>,
    aln: &Alignment,
) -> Result<Reconstruction<A>, AsrError> {
    let (c, f, s2p, p) = load_dev(tree, model, aln)?;
    let mut states = vec![0u8; f.ids.len() * p.max(1)];
    let mut total = 0f64;
    let rc = unsafe { sys::phb_joint_asr(c, states.as_mut_ptr(), &mut total) };
    check(c, rc)?;
    let s2p: Vec<usize> = s2p.iter().map(|&s| s as usize).collect();
    Ok(expand(tree, &f, &s2p, p, &states, None, total))
}

/// The topology-search caller loop `t.nni(v, left); t.log_likelihood(..)` over many `(v, left)` as one
/// call (`NNI::nni`): log-likelihood of every neighbour, `NaN` where `nni` itself would return `Err`.
pub fn score_nni<A: Alphabet>(
    tree: &PhyloTree,
    model: &GtrModel<A>,
    aln: &Alignment,
    moves: &[(NodeID, bool)],
) -> Result<Vec<f64>, AsrError> {
    let (c, f, _s2p, _p) = load_dev(tree, model, aln)?;
    let dense: HashMap<NodeID, i32> = f.ids.iter().enumerate().map(|(k, &v)| (v, k as i32)).collect();
    let ids: Vec<i32> = moves.iter().map(|(v, _)| *dense.get(v).unwrap_or(&-1)).collect();
    let left: Vec<u8> = moves.iter().map(|(_, l)| *l as u8).collect();
    let mut totals = vec![0f64; moves.len()];
    let rc = unsafe {
        sys::phb_score_nni(c, moves.len() as i32, ids.as_ptr(), left.as_ptr(), totals.as_mut_ptr())
    };
    check(c, rc)?;
    Ok(totals)
}

/// The same for arbitrary rearrangements (`SPR::spr`, rerooting, ...): every candidate tree is flattened
/// and scored against the alignment and model loaded for `base`.
pub fn score_trees<A: Alphabet>(
    base: &PhyloTree,
    model: &GtrModel<A>,
    aln: &Alignment,
    candidates: &[PhyloTree],
) -> Result<Vec<f64>, AsrError> {
    let leaf_order: Vec<String> = aln.seqs.keys().cloned().collect();
    let (c, _f, _s2p, _p) = load_dev(base, model, aln)?;
    let flats: Vec<Flat> = candidates
        .iter()
        .map(|t| flatten(t, &leaf_order))
        .collect::<Result<_, _>>()?;
    let descs: Vec<sys::phb_tree_desc> = flats.iter().map(Flat::desc).collect();
    let mut totals = vec![0f64; candidates.len()];
    let rc = unsafe {
        sys::phb_score_trees(c, descs.len() as i32, descs.as_ptr(), totals.as_mut_ptr())
    };
    check(c, rc)?;
    Ok(totals)
}
