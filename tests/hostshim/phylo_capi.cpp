// phylo_capi.cpp — flat C facade (`phh_*`) over the C++ host mirror so that other hosts (the Python
// binding phylo_rs_b200/api.py, via ctypes) can drive phylo::PhyloTree / GtrModel / Alignment exactly as
// a Rust caller drives the reference's types.  Handles are opaque heap objects; every function returns 0
// on success or the ErrorKind of the AsrError that was thrown (message via phh_last_error).
#include "phylo_host.hpp"

#include <cstring>
#include <new>

using namespace phylo;

namespace {
thread_local std::string t_err;
thread_local int t_kind = 0;

template <class F> int guard(F &&f)
{
    try {
        f();
        t_kind = 0;
        return 0;
    } catch (const AsrError &e) {
        t_err = e.what();
        t_kind = (int)e.kind();
        return t_kind;
    } catch (const std::exception &e) {
        t_err = e.what();
        t_kind = (int)ErrorKind::Backend;
        return t_kind;
    }
}
size_t copy_str(const std::string &s, char *buf, size_t cap)
{
    if (buf && cap) {
        const size_t n = std::min(cap - 1, s.size());
        std::memcpy(buf, s.data(), n);
        buf[n] = 0;
    }
    return s.size();
}
} // namespace

extern "C" {

const char *phh_last_error(void) { return t_err.c_str(); }
int phh_last_error_kind(void) { return t_kind; }
const char *phh_error_kind_name(int kind) { return error_kind_name(static_cast<ErrorKind>(kind)); }
int phh_backend_version(char *buf, size_t cap)
{
    return guard([&] { copy_str(backend_version(), buf, cap); });
}
int phh_set_devices(const int *devices, int n, uint64_t flags)
{
    return guard([&] { set_devices(std::vector<int>(devices, devices + (n > 0 ? n : 0)), flags); });
}
int phh_set_library(const char *path)
{
    return guard([&] { set_library(path ? path : ""); });
}
void phh_release_backend(void) { guard([] { release_backend(); }); }

// ---- tree ---------------------------------------------------------------------------------------------
int phh_tree_from_newick(const char *src, size_t len, void **out)
{
    return guard([&] { *out = new PhyloTree(PhyloTree::from_newick(std::string_view(src, len))); });
}
int phh_tree_yule(size_t num_taxa, uint64_t seed, void **out)
{
    return guard([&] { *out = new PhyloTree(PhyloTree::yule(num_taxa, seed)); });
}
int phh_tree_clone(const void *t, void **out)
{
    return guard([&] { *out = new PhyloTree(*static_cast<const PhyloTree *>(t)); });
}
void phh_tree_free(void *t) { delete static_cast<PhyloTree *>(t); }
size_t phh_tree_num_nodes(const void *t) { return static_cast<const PhyloTree *>(t)->num_nodes(); }
size_t phh_tree_root(const void *t) { return static_cast<const PhyloTree *>(t)->get_root_id(); }
int phh_tree_node_ids(const void *t, size_t *out)
{
    return guard([&] {
        const auto ids = static_cast<const PhyloTree *>(t)->get_node_ids();
        std::copy(ids.begin(), ids.end(), out);
    });
}
int phh_tree_postord_ids(const void *t, size_t *out)
{
    return guard([&] {
        const auto ids = static_cast<const PhyloTree *>(t)->postord_ids();
        std::copy(ids.begin(), ids.end(), out);
    });
}
int phh_tree_preord_ids(const void *t, size_t *out)
{
    return guard([&] {
        const auto ids = static_cast<const PhyloTree *>(t)->preord_ids();
        std::copy(ids.begin(), ids.end(), out);
    });
}
// children of `id` into out (capacity cap); *n receives the count
int phh_tree_children(const void *t, size_t id, size_t *out, size_t cap, size_t *n)
{
    return guard([&] {
        const auto &ch = static_cast<const PhyloTree *>(t)->get_node_children_ids(id);
        *n = ch.size();
        for (size_t k = 0; k < ch.size() && k < cap; ++k) out[k] = ch[k];
    });
}
// *parent receives the parent id; *has = 0 for the root
int phh_tree_parent(const void *t, size_t id, size_t *parent, int *has)
{
    return guard([&] {
        const auto p = static_cast<const PhyloTree *>(t)->get_node_parent_id(id);
        *has = p.has_value();
        *parent = p.value_or(0);
    });
}
int phh_tree_is_leaf(const void *t, size_t id, int *out)
{
    return guard([&] { *out = static_cast<const PhyloTree *>(t)->is_leaf(id); });
}
int phh_tree_taxa(const void *t, size_t id, char *buf, size_t cap, int *has)
{
    return guard([&] {
        const auto &tx = static_cast<const PhyloTree *>(t)->get_node(id).taxa;
        *has = tx.has_value();
        copy_str(tx.value_or(""), buf, cap);
    });
}
int phh_tree_taxa_node_id(const void *t, const char *name, size_t *id, int *has)
{
    return guard([&] {
        const auto r = static_cast<const PhyloTree *>(t)->get_taxa_node_id(name);
        *has = r.has_value();
        *id = r.value_or(0);
    });
}
int phh_tree_edge_weight(const void *t, size_t child, float *w, int *has)
{
    return guard([&] {
        const auto r = static_cast<const PhyloTree *>(t)->get_node(child).weight;
        *has = r.has_value();
        *w = r.value_or(0.0f);
    });
}
int phh_tree_set_edge_weight(void *t, size_t child, float w, int has)
{
    return guard([&] { static_cast<PhyloTree *>(t)->set_edge_weight(child, has ? std::optional<float>(w) : std::nullopt); });
}
int phh_tree_assign_exponential_lengths(void *t, uint64_t seed, double mean, double lo, double hi)
{
    return guard([&] { static_cast<PhyloTree *>(t)->assign_exponential_lengths(seed, mean, lo, hi); });
}
int phh_tree_nni(void *t, size_t node, int left_ch)
{
    return guard([&] { static_cast<PhyloTree *>(t)->nni(node, left_ch != 0); });
}
// returns the needed length (excluding NUL) through *len
int phh_tree_to_newick(const void *t, char *buf, size_t cap, size_t *len)
{
    return guard([&] { *len = copy_str(static_cast<const PhyloTree *>(t)->to_newick(), buf, cap); });
}

// ---- alignment ------------------------------------------------------------------------------------------
int phh_aln_from_fasta(const char *data, size_t len, void **out)
{
    return guard([&] { *out = new Alignment(Alignment::from_fasta_bytes(std::string_view(data, len))); });
}
int phh_aln_new(size_t width, void **out)
{
    return guard([&] {
        auto *a = new Alignment();
        a->width = width;
        *out = a;
    });
}
int phh_aln_insert(void *a, const char *name, const uint8_t *seq, size_t len)
{
    return guard([&] { static_cast<Alignment *>(a)->insert(name, std::vector<uint8_t>(seq, seq + len)); });
}
void phh_aln_free(void *a) { delete static_cast<Alignment *>(a); }
size_t phh_aln_width(const void *a) { return static_cast<const Alignment *>(a)->width; }
size_t phh_aln_n_seqs(const void *a) { return static_cast<const Alignment *>(a)->seqs.size(); }
int phh_aln_compress(const void *a, void **out)
{
    return guard([&] { *out = new CompressedColumns(static_cast<const Alignment *>(a)->compress_columns()); });
}
void phh_comp_free(void *c) { delete static_cast<CompressedColumns *>(c); }
size_t phh_comp_n_patterns(const void *c) { return static_cast<const CompressedColumns *>(c)->patterns.size(); }
size_t phh_comp_n_rows(const void *c) { return static_cast<const CompressedColumns *>(c)->leaf_order.size(); }
size_t phh_comp_n_sites(const void *c) { return static_cast<const CompressedColumns *>(c)->site_to_pattern.size(); }
// patterns [P][rows] u8, multiplicity [P] u64, site_to_pattern [sites] u64
int phh_comp_export(const void *c, uint8_t *patterns, uint64_t *mult, uint64_t *s2p)
{
    return guard([&] {
        const auto *cc = static_cast<const CompressedColumns *>(c);
        const size_t rows = cc->leaf_order.size();
        for (size_t p = 0; p < cc->patterns.size(); ++p) {
            if (rows) std::memcpy(patterns + p * rows, cc->patterns[p].data(), rows);
            mult[p] = cc->multiplicity[p];
        }
        for (size_t s = 0; s < cc->site_to_pattern.size(); ++s) s2p[s] = cc->site_to_pattern[s];
    });
}
int phh_comp_leaf_name(const void *c, size_t row, char *buf, size_t cap)
{
    return guard([&] { copy_str(static_cast<const CompressedColumns *>(c)->leaf_order.at(row), buf, cap); });
}

// ---- model ----------------------------------------------------------------------------------------------
int phh_model_new(const double *pi, const double *w16, int normalize, void **out)
{
    return guard([&] { *out = new GtrModel(std::vector<double>(pi, pi + 4), std::vector<double>(w16, w16 + 16), normalize != 0); });
}
// named constructors: "jc69" (0 params), "k80" (kappa), "f81" (pi4), "hky85" (pi4,kappa), "tn93" (pi4,k1,k2),
// "k81" (a,b,c), "tim" (pi4,k1,k2,a,b), "tvm" (pi4,kappa,a,b,c), "sym" (6 rates), "gtr" (pi4, 6 rates)
int phh_model_named(const char *name, const double *p, int n, void **out)
{
    return guard([&] {
        const std::string nm(name);
        auto need = [&](int k) {
            if (n != k) throw AsrError(ErrorKind::InvalidModelParameter, nm + ": wrong number of parameters");
        };
        auto pi = [&] { return std::array<double, 4>{p[0], p[1], p[2], p[3]}; };
        if (nm == "jc69") { need(0); *out = new GtrModel(GtrModel::jukes_cantor()); }
        else if (nm == "k80") { need(1); *out = new GtrModel(GtrModel::k80(p[0])); }
        else if (nm == "f81") { need(4); *out = new GtrModel(GtrModel::f81(pi())); }
        else if (nm == "hky85") { need(5); *out = new GtrModel(GtrModel::hky85(pi(), p[4])); }
        else if (nm == "tn93") { need(6); *out = new GtrModel(GtrModel::tn93(pi(), p[4], p[5])); }
        else if (nm == "k81") { need(3); *out = new GtrModel(GtrModel::k81(p[0], p[1], p[2])); }
        else if (nm == "tim") { need(8); *out = new GtrModel(GtrModel::tim(pi(), p[4], p[5], p[6], p[7])); }
        else if (nm == "tvm") { need(8); *out = new GtrModel(GtrModel::tvm(pi(), p[4], p[5], p[6], p[7])); }
        else if (nm == "sym") { need(6); *out = new GtrModel(GtrModel::sym({p[0], p[1], p[2], p[3], p[4], p[5]})); }
        else if (nm == "gtr") { need(10); *out = new GtrModel(GtrModel::gtr(pi(), {p[4], p[5], p[6], p[7], p[8], p[9]})); }
        else throw AsrError(ErrorKind::InvalidModelParameter, "unknown model name " + nm);
    });
}
int phh_model_with_gamma(const void *m, double alpha, size_t k, void **out)
{
    return guard([&] { *out = new GtrModel(static_cast<const GtrModel *>(m)->with_gamma(alpha, k)); });
}
int phh_model_with_invariant(const void *m, double p_inv, void **out)
{
    return guard([&] { *out = new GtrModel(static_cast<const GtrModel *>(m)->with_invariant(p_inv)); });
}
void phh_model_free(void *m) { delete static_cast<GtrModel *>(m); }
size_t phh_model_n_categories(const void *m) { return static_cast<const GtrModel *>(m)->n_categories(); }
int phh_model_categories(const void *m, double *rates, double *weights)
{
    return guard([&] {
        const auto &c = static_cast<const GtrModel *>(m)->categories();
        for (size_t i = 0; i < c.size(); ++i) {
            rates[i] = c[i].rate;
            weights[i] = c[i].weight;
        }
    });
}
int phh_model_equilibrium(const void *m, double *pi4)
{
    return guard([&] { std::memcpy(pi4, static_cast<const GtrModel *>(m)->equilibrium().data(), 32); });
}
int phh_model_eigen(const void *m, double *lambda4, double *v16_colmajor)
{
    return guard([&] {
        const auto &rm = static_cast<const GtrModel *>(m)->rate_matrix();
        std::memcpy(lambda4, rm.eigenvalues().data(), 32);
        std::memcpy(v16_colmajor, rm.eigenvectors().data(), 128);
    });
}
int phh_model_category_transition(const void *m, size_t cat, double t, double *p16_rowmajor)
{
    return guard([&] {
        const auto p = static_cast<const GtrModel *>(m)->category_transition(cat, t);
        std::memcpy(p16_rowmajor, p.data(), 128);
    });
}
int phh_discrete_gamma(double alpha, size_t k, double *out)
{
    return guard([&] {
        const auto r = discrete_gamma(alpha, k);
        std::copy(r.begin(), r.end(), out);
    });
}
int phh_nucleotide_profile(uint8_t c, double *out4, int *has)
{
    return guard([&] {
        const auto p = Nucleotide::profile(c);
        *has = p.has_value();
        if (p) std::memcpy(out4, p->data(), 32);
    });
}

// ---- engines --------------------------------------------------------------------------------------------
int phh_log_likelihood(const void *t, const void *m, const void *a, double *out)
{
    return guard([&] {
        *out = static_cast<const PhyloTree *>(t)->log_likelihood(*static_cast<const GtrModel *>(m), *static_cast<const Alignment *>(a));
    });
}
int phh_marginal_asr(const void *t, const void *m, const void *a, int want_posteriors, void **out)
{
    return guard([&] {
        *out = new Reconstruction(static_cast<const PhyloTree *>(t)->marginal_asr(
            *static_cast<const GtrModel *>(m), *static_cast<const Alignment *>(a), want_posteriors != 0));
    });
}
int phh_joint_asr(const void *t, const void *m, const void *a, void **out)
{
    return guard([&] {
        *out = new Reconstruction(static_cast<const PhyloTree *>(t)->joint_asr(*static_cast<const GtrModel *>(m),
                                                                              *static_cast<const Alignment *>(a)));
    });
}
void phh_recon_free(void *r) { delete static_cast<Reconstruction *>(r); }
double phh_recon_log_likelihood(const void *r) { return static_cast<const Reconstruction *>(r)->log_likelihood; }
size_t phh_recon_n_nodes(const void *r) { return static_cast<const Reconstruction *>(r)->sequences.size(); }
int phh_recon_has_posteriors(const void *r) { return static_cast<const Reconstruction *>(r)->posteriors.has_value(); }
// *has = 0 when the node has no entry; out holds `width` state indices
int phh_recon_sequence(const void *r, size_t node, uint64_t *out, size_t cap, size_t *width, int *has)
{
    return guard([&] {
        const auto &seqs = static_cast<const Reconstruction *>(r)->sequences;
        auto it = seqs.find(node);
        *has = it != seqs.end();
        *width = *has ? it->second.size() : 0;
        if (*has)
            for (size_t s = 0; s < it->second.size() && s < cap; ++s) out[s] = it->second[s];
    });
}
int phh_recon_posteriors(const void *r, size_t node, double *out /* [width][4] */, size_t cap_sites, int *has)
{
    return guard([&] {
        const auto &po = static_cast<const Reconstruction *>(r)->posteriors;
        *has = 0;
        if (!po) return;
        auto it = po->find(node);
        if (it == po->end()) return;
        *has = 1;
        for (size_t s = 0; s < it->second.size() && s < cap_sites; ++s) std::memcpy(out + 4 * s, it->second[s].data(), 32);
    });
}
int phh_score_nni_neighbours(const void *t, const void *m, const void *a, size_t n, const size_t *nodes,
                             const uint8_t *left, double *out)
{
    return guard([&] {
        std::vector<std::pair<NodeID, bool>> moves;
        for (size_t i = 0; i < n; ++i) moves.emplace_back(nodes[i], left[i] != 0);
        const auto r = static_cast<const PhyloTree *>(t)->score_nni_neighbours(
            *static_cast<const GtrModel *>(m), *static_cast<const Alignment *>(a), moves);
        std::copy(r.begin(), r.end(), out);
    });
}

} // extern "C"
