// phylo_host.hpp — C++ mirror of the reference's host-side interface for the likelihood path, above the
// C ABI of libphylo_b200.so (include/phylo_b200.h).
//
// The reference (sriram98v/phylo-rs, crate `phylo` 6.0.0) is Rust and there is no Rust toolchain in the
// build image, so the host side that would stay in Rust (INTEGRATION.md §3 gives that shim as source) is
// mirrored here with the same names, argument meaning and error behaviour:
//
//   reference                                               here
//   ------------------------------------------------------  -------------------------------------------
//   PhyloTree = SimpleRootedTree<String,f32,f32> tree.rs:45  phylo::PhyloTree (arena of optional nodes)
//   Newick::from_newick            tree/newick.rs:168-284    PhyloTree::from_newick
//   Yule::yule                     tree.rs:505-553           PhyloTree::yule (seeded; reference: thread_rng)
//   NNI::nni                       tree.rs:1213-1242         PhyloTree::nni
//   Alignment, compress_columns    alignment.rs:5-138        phylo::Alignment, CompressedColumns
//   Alphabet for Nucleotide        alphabet.rs:23-65         phylo::Nucleotide
//   RateMatrix::new                models/rate_matrix.rs:34  phylo::RateMatrix (host: Q, symmetrise, Jacobi)
//   GtrModel ctors, +G, +I         models/gtr.rs:39-306      phylo::GtrModel
//   discrete_gamma                 models/gamma.rs:15-229    phylo::discrete_gamma
//   TreeLikelihood::log_likelihood tree/likelihood.rs:36-44  PhyloTree::log_likelihood
//   MarginalAsr::marginal_asr      tree/asr.rs:19-27         PhyloTree::marginal_asr
//   Reconstruction                 likelihood/reconstruction.rs:5-27  phylo::Reconstruction
//   AsrError                       error.rs:106-125          phylo::AsrError (thrown; Result::Err in Rust)
//
// Everything numerical on the path — P(t), the pruning recursion, the root mix, the ASR down pass — runs
// in the CUDA library; this layer only prepares its inputs (what compute_log_likelihood does before its
// pattern loop, likelihood.rs:164-183) and rebuilds the reference's result containers.  There is no CPU
// implementation of the path here: without libphylo_b200.so and a B200 every engine call throws.
#pragma once

#include <array>
#include <cstddef>
#include <cstdint>
#include <map>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <string_view>
#include <unordered_map>
#include <utility>
#include <vector>

namespace phylo {

using NodeID = std::size_t;

// ---- errors (src/error.rs) -----------------------------------------------------------------------
enum class ErrorKind : int {
    // AsrError variants, numbered as the C ABI status codes (include/phylo_b200.h)
    MissingBranchLength = 1,
    AlphabetMismatch = 2,
    InvalidAlignment = 3,
    EigendecompFailure = 4,
    NumericalInstability = 5,
    InvalidModelParameter = 6,
    // tree-edit / parser errors used by the callers of the path (src/error.rs: TreeEditError, NewickError)
    UnknownNode = 32,
    NoParent = 33,
    Newick = 34,
    // no reference counterpart: the device library is missing / failed (status 16..20)
    Backend = 48,
};
const char *error_kind_name(ErrorKind k);

class AsrError : public std::runtime_error {
  public:
    AsrError(ErrorKind kind, const std::string &msg) : std::runtime_error(msg), kind_(kind) {}
    ErrorKind kind() const { return kind_; }

  private:
    ErrorKind kind_;
};

// ---- alphabet (src/alphabet.rs:23-65) -------------------------------------------------------------
struct Nucleotide {
    static constexpr std::size_t N_STATES = 4;
    static constexpr const char *CANONICAL = "ACGT";
    static constexpr std::uint8_t GAP = '-';
    static std::optional<std::size_t> index_of(std::uint8_t c);
    static std::uint8_t char_of(std::size_t i);
    static std::optional<std::array<double, 4>> profile(std::uint8_t c);
};

// ---- tree (src/tree.rs, src/tree/simple_rtree.rs, src/node.rs) --------------------------------------
struct Node {
    NodeID id = 0;
    std::optional<NodeID> parent;
    std::vector<NodeID> children;      // ordered; set_child appends last
    std::optional<std::string> taxa;
    std::optional<float> weight;       // weight of the edge (parent -> this), f32 as in the reference
};

class GtrModel;
class Alignment;
struct Reconstruction;

class PhyloTree {
  public:
    explicit PhyloTree(NodeID root_id = 0);

    static PhyloTree from_newick(std::string_view src);
    // Yule process of tree.rs:505-553 with an explicit generator: splitmix64-seeded xoshiro256**,
    // uniform leaf pick by 64-bit multiply-shift.  No branch lengths, as the reference.
    static PhyloTree yule(std::size_t num_taxa, std::uint64_t seed);
    // Harness helper: i.i.d. Exp(mean) lengths clamped to [lo, hi], rounded to f32, on every non-root node.
    void assign_exponential_lengths(std::uint64_t seed, double mean = 0.1, double lo = 1e-4, double hi = 2.0);

    // accessors the engines read
    NodeID get_root_id() const { return root_; }
    std::vector<NodeID> get_node_ids() const;
    std::size_t num_nodes() const;
    bool contains(NodeID id) const { return id < nodes_.size() && nodes_[id].has_value(); }
    const Node &get_node(NodeID id) const;
    const std::vector<NodeID> &get_node_children_ids(NodeID id) const { return get_node(id).children; }
    std::optional<NodeID> get_node_parent_id(NodeID id) const { return get_node(id).parent; }
    bool is_leaf(NodeID id) const { return get_node(id).children.empty(); }
    std::optional<NodeID> get_taxa_node_id(const std::string &taxa) const;
    std::optional<float> get_edge_weight(NodeID parent, NodeID child) const;
    std::vector<NodeID> postord_ids(std::optional<NodeID> start = std::nullopt) const; // iter.rs:112-138
    std::vector<NodeID> preord_ids(std::optional<NodeID> start = std::nullopt) const;  // node_iter.rs:137-161

    // edits
    NodeID next_id() const;
    void set_node(Node node);
    void set_child(NodeID parent, NodeID child);
    void add_child(NodeID parent, Node node);
    void delete_edge(NodeID parent, NodeID child);
    void split_edge(std::pair<NodeID, NodeID> edge, Node node);
    void set_node_taxa(NodeID id, std::optional<std::string> taxa);
    void set_edge_weight(NodeID child, std::optional<float> w);
    void nni(NodeID node_id, bool left_ch); // throws UnknownNode / NoParent as NNI::nni
    std::string to_newick() const;

    // TreeLikelihood::log_likelihood (likelihood.rs:36-44; impl tree.rs:68-76)
    double log_likelihood(const GtrModel &model, const Alignment &aln) const;
    // MarginalAsr::marginal_asr (asr.rs:19-27; impl tree.rs:47-56)
    Reconstruction marginal_asr(const GtrModel &model, const Alignment &aln, bool want_posteriors) const;
    // JointAsr::joint_asr (asr.rs:30-37; impl tree.rs:58-66)
    Reconstruction joint_asr(const GtrModel &model, const Alignment &aln) const;
    // Config-5 caller loop `for (v, left) in moves { t = self.clone(); t.nni(v, left); t.log_likelihood(..) }`
    // in one device batch.  Invalid moves give NaN.
    std::vector<double> score_nni_neighbours(const GtrModel &model, const Alignment &aln,
                                             const std::vector<std::pair<NodeID, bool>> &moves) const;

  private:
    NodeID root_;
    std::vector<std::optional<Node>> nodes_;
    Node &node_mut(NodeID id);
};

// ---- alignment (src/alignment.rs) -----------------------------------------------------------------
struct CompressedColumns {
    std::vector<std::vector<std::uint8_t>> patterns; // [pattern][row]
    std::vector<std::size_t> site_to_pattern;
    std::vector<std::size_t> multiplicity;
    std::vector<std::string> leaf_order;
};

class Alignment {
  public:
    // The reference keeps a HashMap (iteration order random per process, alignment.rs:100); here rows keep
    // insertion order, which only fixes the row order inside a pattern and no arithmetic.
    std::vector<std::pair<std::string, std::vector<std::uint8_t>>> seqs;
    std::size_t width = 0;

    Alignment() = default;
    static Alignment from_fasta_bytes(std::string_view data);
    void insert(const std::string &name, std::vector<std::uint8_t> seq); // replaces an existing entry
    const std::vector<std::uint8_t> *get(const std::string &name) const;
    CompressedColumns compress_columns() const;
};

// ---- models (src/models/*.rs) ---------------------------------------------------------------------
struct RateCategory {
    double rate;
    double weight;
};

std::vector<double> discrete_gamma(double alpha, std::size_t k); // gamma.rs:15-55

class RateMatrix {
  public:
    // pi: n_states frequencies (normalised here); w: n_states x n_states row-major exchangeabilities
    RateMatrix(std::size_t n_states, std::vector<double> pi, const std::vector<double> &w, bool normalize);
    const std::array<double, 4> &equilibrium() const { return pi_; }
    const std::array<double, 4> &eigenvalues() const { return eigenvalues_; }
    const std::array<double, 16> &eigenvectors() const { return eigenvectors_; } // column-major (nalgebra)
    const std::array<double, 4> &sqrt_pi() const { return sqrt_pi_; }
    const std::array<double, 4> &inv_sqrt_pi() const { return inv_sqrt_pi_; }

  private:
    std::array<double, 4> pi_{}, eigenvalues_{}, sqrt_pi_{}, inv_sqrt_pi_{};
    std::array<double, 16> eigenvectors_{};
};

class GtrModel {
  public:
    GtrModel(std::vector<double> pi, const std::vector<double> &w /* 4x4 row-major */, bool normalize);
    static GtrModel jukes_cantor();
    static GtrModel k80(double kappa);
    static GtrModel f81(const std::array<double, 4> &pi);
    static GtrModel hky85(const std::array<double, 4> &pi, double kappa);
    static GtrModel tn93(const std::array<double, 4> &pi, double kappa1, double kappa2);
    static GtrModel k81(double a, double b, double c);
    static GtrModel tim(const std::array<double, 4> &pi, double kappa1, double kappa2, double a, double b);
    static GtrModel tvm(const std::array<double, 4> &pi, double kappa, double a, double b, double c);
    static GtrModel sym(const std::array<double, 6> &rates);
    static GtrModel gtr(const std::array<double, 4> &pi, const std::array<double, 6> &rates);

    GtrModel with_gamma(double alpha, std::size_t k) const;
    GtrModel with_invariant(double p_inv) const;

    const std::array<double, 4> &equilibrium() const { return matrix_.equilibrium(); }
    const std::vector<RateCategory> &categories() const { return categories_; }
    std::size_t n_categories() const { return categories_.size(); }
    const RateMatrix &rate_matrix() const { return matrix_; }
    // P(rate_c * t), row-major 4x4, evaluated by the device kernel the engines use (phb_transition_matrices).
    std::array<double, 16> category_transition(std::size_t category, double t) const;

  private:
    RateMatrix matrix_;
    std::vector<RateCategory> categories_;
    double proportion_invariant() const;
};

// ---- result container (src/tree/likelihood/reconstruction.rs) -------------------------------------
struct Reconstruction {
    std::unordered_map<NodeID, std::vector<std::size_t>> sequences;                          // [site] state index
    std::optional<std::unordered_map<NodeID, std::vector<std::array<double, 4>>>> posteriors; // [site][state]
    double log_likelihood = 0.0;
    std::optional<std::string> sequence_string(NodeID id) const;
};

// ---- device backend ----------------------------------------------------------------------------------
// One lazily created phb_ctx per thread (the ctx is not thread-safe; the reference's engines take every
// input by shared reference, so callers may evaluate from many threads).  The library is located through
// $PHYLO_B200_LIB, else next to this library (../csrc/libphylo_b200.so).
void set_devices(const std::vector<int> &devices, std::uint64_t flags = 0); // applies to contexts created afterwards (this thread: recreated)
void release_backend();                                                     // destroys this thread's context
std::string backend_version();
// Use another build of the C ABI (tests point this at the CPU emulation build of the same sources);
// an empty path restores the default.  Contexts of OTHER threads must be released first.
void set_library(const std::string &path);

} // namespace phylo
