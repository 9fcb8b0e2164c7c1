// phylo_engine.cpp — the engine calls of the C++ host mirror: what `compute_log_likelihood` /
// `compute_marginal_asr` do BEFORE their pattern loops (likelihood.rs:164-183, :232-273) and AFTER them
// (scatter to sites, :391-398), with the loops themselves replaced by calls into libphylo_b200.so.
// Same call sequence as the Rust shim of INTEGRATION.md §3.
#include "phylo_host.hpp"

#include "../../include/phylo_b200.h"

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <limits>
#include <mutex>

namespace phylo {

// =====================================================================================================
// backend: libphylo_b200.so through dlopen, one context per thread
// =====================================================================================================
namespace {

struct Api {
    void *handle = nullptr;
    decltype(&phb_ctx_create) ctx_create = nullptr;
    decltype(&phb_ctx_destroy) ctx_destroy = nullptr;
    decltype(&phb_set_alignment) set_alignment = nullptr;
    decltype(&phb_compress_rows) compress_rows = nullptr;
    decltype(&phb_set_model) set_model = nullptr;
    decltype(&phb_set_tree) set_tree = nullptr;
    decltype(&phb_log_likelihood) log_likelihood = nullptr;
    decltype(&phb_log_likelihood_host) log_likelihood_host = nullptr;
    decltype(&phb_marginal_asr) marginal_asr = nullptr;
    decltype(&phb_joint_asr) joint_asr = nullptr;
    decltype(&phb_score_nni) score_nni = nullptr;
    decltype(&phb_transition_matrices) transition_matrices = nullptr;
    decltype(&phb_last_error) last_error = nullptr;
    decltype(&phb_version) version = nullptr;
};

std::string default_library_path()
{
    if (const char *env = std::getenv("PHYLO_B200_LIB"))
        if (*env) return env;
    Dl_info info;
    if (dladdr(reinterpret_cast<void *>(&default_library_path), &info) && info.dli_fname) {
        std::string here(info.dli_fname);
        const auto slash = here.rfind('/');
        const std::string dir = slash == std::string::npos ? std::string(".") : here.substr(0, slash);
        return dir + "/../csrc/libphylo_b200.so";
    }
    return "libphylo_b200.so";
}

std::mutex g_api_mutex;
Api g_api;
std::string g_lib_override;

// Loads the library on first use (or after set_library).  Throws Backend when it cannot be loaded.
const Api &api()
{
    std::lock_guard<std::mutex> g(g_api_mutex);
    Api &a = g_api;
    if (a.handle) return a;
    const std::string path = g_lib_override.empty() ? default_library_path() : g_lib_override;
    void *h = dlopen(path.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (!h) {
        const char *why = dlerror();
        throw AsrError(ErrorKind::Backend, "cannot load " + path + ": " + (why ? why : "?") +
                                               " (build it with __graft_entry__.build(); there is no CPU fallback)");
    }
    std::string missing;
    auto sym = [&](const char *name) {
        void *p = dlsym(h, name);
        if (!p && missing.empty()) missing = name;
        return p;
    };
    a.ctx_create = reinterpret_cast<decltype(a.ctx_create)>(sym("phb_ctx_create"));
    a.ctx_destroy = reinterpret_cast<decltype(a.ctx_destroy)>(sym("phb_ctx_destroy"));
    a.set_alignment = reinterpret_cast<decltype(a.set_alignment)>(sym("phb_set_alignment"));
    a.compress_rows = reinterpret_cast<decltype(a.compress_rows)>(sym("phb_compress_rows"));
    a.set_model = reinterpret_cast<decltype(a.set_model)>(sym("phb_set_model"));
    a.set_tree = reinterpret_cast<decltype(a.set_tree)>(sym("phb_set_tree"));
    a.log_likelihood = reinterpret_cast<decltype(a.log_likelihood)>(sym("phb_log_likelihood"));
    a.log_likelihood_host = reinterpret_cast<decltype(a.log_likelihood_host)>(sym("phb_log_likelihood_host"));
    a.marginal_asr = reinterpret_cast<decltype(a.marginal_asr)>(sym("phb_marginal_asr"));
    a.joint_asr = reinterpret_cast<decltype(a.joint_asr)>(sym("phb_joint_asr"));
    a.score_nni = reinterpret_cast<decltype(a.score_nni)>(sym("phb_score_nni"));
    a.transition_matrices = reinterpret_cast<decltype(a.transition_matrices)>(sym("phb_transition_matrices"));
    a.last_error = reinterpret_cast<decltype(a.last_error)>(sym("phb_last_error"));
    a.version = reinterpret_cast<decltype(a.version)>(sym("phb_version"));
    if (!missing.empty()) {
        dlclose(h);
        a = Api();
        throw AsrError(ErrorKind::Backend, "symbol missing from " + path + ": " + missing);
    }
    a.handle = h;
    return a;
}

std::mutex g_cfg_mutex;
std::vector<int> g_devices;
std::uint64_t g_flags = 0;

struct ThreadCtx {
    phb_ctx *ctx = nullptr;
    ~ThreadCtx()
    {
        if (ctx && g_api.handle) g_api.ctx_destroy(ctx);
    }
};
thread_local ThreadCtx t_ctx;

phb_ctx *context()
{
    if (t_ctx.ctx) return t_ctx.ctx;
    const Api &a = api();
    std::vector<int32_t> devs;
    std::uint64_t flags;
    {
        std::lock_guard<std::mutex> g(g_cfg_mutex);
        devs.assign(g_devices.begin(), g_devices.end());
        flags = g_flags;
    }
    phb_ctx *c = nullptr;
    const int32_t rc = a.ctx_create(devs.empty() ? nullptr : devs.data(), (int32_t)devs.size(), flags, &c);
    if (rc != PHB_OK || !c)
        throw AsrError(ErrorKind::Backend, "phb_ctx_create failed with status " + std::to_string(rc) +
                                               " (no usable sm_100 device?)");
    t_ctx.ctx = c;
    return c;
}

// status -> the AsrError variant of the same number; 16..20 have no reference counterpart
[[noreturn]] void raise_status(phb_ctx *c, int32_t rc)
{
    const std::string msg = api().last_error(c);
    if (rc >= 1 && rc <= 6) throw AsrError(static_cast<ErrorKind>(rc), msg);
    throw AsrError(ErrorKind::Backend, "status " + std::to_string(rc) + ": " + msg);
}
void check(phb_ctx *c, int32_t rc)
{
    if (rc != PHB_OK) raise_status(c, rc);
}

// ---- flattening: sparse arena -> dense ids + CSR children + f32->f64 branch lengths -------------------
struct Flat {
    std::vector<NodeID> ids;              // dense -> arena id
    std::vector<int32_t> post, off, kids, leaf_row;
    std::vector<double> blen;
    int32_t root = 0;
    std::vector<int32_t> dense_of;        // arena id -> dense (-1 = vacant)
};

Flat flatten(const PhyloTree &tree, const std::vector<std::string> &leaf_order)
{
    Flat f;
    f.ids = tree.get_node_ids();
    NodeID max_id = 0;
    for (NodeID id : f.ids) max_id = std::max(max_id, id);
    f.dense_of.assign(max_id + 1, -1);
    for (std::size_t k = 0; k < f.ids.size(); ++k) f.dense_of[f.ids[k]] = (int32_t)k;
    const std::size_t n = f.ids.size();
    f.leaf_row.assign(n, -1);
    // likelihood.rs:172-178: every row must name SOME node; :99 uses the first row that resolves to the leaf
    for (std::size_t r = 0; r < leaf_order.size(); ++r) {
        const auto id = tree.get_taxa_node_id(leaf_order[r]);
        if (!id) throw AsrError(ErrorKind::AlphabetMismatch, "Taxon " + leaf_order[r] + " in alignment not found in tree");
        const int32_t d = f.dense_of[*id];
        if (tree.is_leaf(*id) && f.leaf_row[d] < 0) f.leaf_row[d] = (int32_t)r;
    }
    f.off.assign(1, 0);
    f.blen.assign(n, 0.0);
    for (std::size_t k = 0; k < n; ++k) {
        const Node &nd = tree.get_node(f.ids[k]);
        for (NodeID c : nd.children) f.kids.push_back(f.dense_of[c]);
        f.off.push_back((int32_t)f.kids.size());
        f.blen[k] = nd.weight ? (double)*nd.weight : 0.0; // NumCast f32 -> f64, None -> 0.0 (likelihood.rs:115-118)
    }
    for (NodeID v : tree.postord_ids()) f.post.push_back(f.dense_of[v]);
    f.root = f.dense_of[tree.get_root_id()];
    if (f.kids.empty()) f.kids.push_back(0);
    return f;
}

void load_model(phb_ctx *c, const GtrModel &m)
{
    const RateMatrix &rm = m.rate_matrix();
    std::vector<double> rates, weights;
    for (const auto &cat : m.categories()) {
        rates.push_back(cat.rate);
        weights.push_back(cat.weight);
    }
    check(c, api().set_model(c, rm.eigenvalues().data(), rm.eigenvectors().data(), rm.equilibrium().data(),
                             rm.sqrt_pi().data(), rm.inv_sqrt_pi().data(), (int32_t)rates.size(), rates.data(),
                             weights.data()));
}

// The engines' common prologue with the column compression done ON THE DEVICE (phb_compress_rows replaces
// Alignment::compress_columns, alignment.rs:93-125): rows resolve to nodes first (the reference's errors,
// likelihood.rs:172-178), then model, tree, and the raw sequences go over as they lie in Alignment.seqs.
// s2p (nullable) receives site_to_pattern; returns the pattern count through n_patterns.
phb_ctx *load_device_compressed(const PhyloTree &tree, const GtrModel &model, const Alignment &aln, Flat &flat,
                                std::vector<std::uint64_t> *s2p, std::size_t &n_patterns)
{
    std::vector<std::string> leaf_order;
    std::vector<const std::uint8_t *> rows;
    leaf_order.reserve(aln.seqs.size());
    for (const auto &kv : aln.seqs) {
        leaf_order.push_back(kv.first);
        rows.push_back(kv.second.data());
    }
    flat = flatten(tree, leaf_order);
    phb_ctx *c = context();
    load_model(c, model);
    check(c, api().set_tree(c, (int32_t)flat.ids.size(), flat.post.data(), flat.off.data(), flat.kids.data(),
                            flat.blen.data(), flat.leaf_row.data(), flat.root));
    if (s2p) s2p->assign(std::max<std::size_t>(aln.width, 1), 0);
    if (rows.empty()) rows.push_back(nullptr);
    int64_t P = 0;
    check(c, api().compress_rows(c, (int32_t)aln.seqs.size(), (int64_t)aln.width, rows.data(), &P, s2p ? s2p->data() : nullptr,
                                 nullptr));
    n_patterns = (std::size_t)P;
    return c;
}

} // namespace

void set_devices(const std::vector<int> &devices, std::uint64_t flags)
{
    {
        std::lock_guard<std::mutex> g(g_cfg_mutex);
        g_devices = devices;
        g_flags = flags;
    }
    release_backend();
}

void release_backend()
{
    if (t_ctx.ctx) {
        if (g_api.handle) g_api.ctx_destroy(t_ctx.ctx);
        t_ctx.ctx = nullptr;
    }
}

std::string backend_version() { return api().version(); }

void set_library(const std::string &path)
{
    release_backend();
    std::lock_guard<std::mutex> g(g_api_mutex);
    if (g_api.handle) dlclose(g_api.handle);
    g_api = Api();
    g_lib_override = path;
}

// =====================================================================================================
// GtrModel::category_transition — device-evaluated P(rate_c * t)
// =====================================================================================================
std::array<double, 16> GtrModel::category_transition(std::size_t category, double t) const
{
    if (category >= categories_.size()) throw AsrError(ErrorKind::InvalidModelParameter, "category index out of range");
    phb_ctx *c = context();
    load_model(c, *this);
    std::vector<double> out(16 * categories_.size());
    check(c, api().transition_matrices(c, 1, &t, out.data()));
    std::array<double, 16> p{};
    std::memcpy(p.data(), out.data() + 16 * category, sizeof(double) * 16);
    return p;
}

// =====================================================================================================
// TreeLikelihood::log_likelihood
// =====================================================================================================
double PhyloTree::log_likelihood(const GtrModel &model, const Alignment &aln) const
{
    // raw sequences in, lnL out: compression (alignment.rs:93-125) and pruning (likelihood.rs:156-209) on the device
    Flat flat;
    std::size_t P = 0;
    phb_ctx *c = load_device_compressed(*this, model, aln, flat, nullptr, P);
    double total = 0.0;
    check(c, api().log_likelihood(c, &total, nullptr));
    return total;
}

// =====================================================================================================
// MarginalAsr::marginal_asr
// =====================================================================================================
Reconstruction PhyloTree::marginal_asr(const GtrModel &model, const Alignment &aln, bool want_posteriors) const
{
    Flat flat;
    std::vector<std::uint64_t> s2p;
    std::size_t P = 0;
    phb_ctx *c = load_device_compressed(*this, model, aln, flat, &s2p, P);
    const std::size_t n = flat.ids.size(), width = aln.width;
    std::vector<std::uint8_t> states(std::max<std::size_t>(n * P, 1));
    std::vector<double> post;
    if (want_posteriors) post.resize(std::max<std::size_t>(n * P * 4, 1));
    Reconstruction rec;
    check(c, api().marginal_asr(c, states.data(), want_posteriors ? post.data() : nullptr, &rec.log_likelihood));
    // expand patterns to sites (likelihood.rs:391-398); every node id gets an entry, leaves included (:262-264)
    if (want_posteriors) rec.posteriors.emplace();
    for (std::size_t k = 0; k < n; ++k) {
        std::vector<std::size_t> seq(width, 0);
        const std::uint8_t *row = states.data() + k * P;
        for (std::size_t s = 0; s < width; ++s) seq[s] = row[s2p[s]];
        rec.sequences.emplace(flat.ids[k], std::move(seq));
        if (want_posteriors) {
            std::vector<std::array<double, 4>> pp(width, std::array<double, 4>{0.0, 0.0, 0.0, 0.0});
            const double *prow = post.data() + k * P * 4;
            for (std::size_t s = 0; s < width; ++s)
                std::memcpy(pp[s].data(), prow + s2p[s] * 4, sizeof(double) * 4);
            rec.posteriors->emplace(flat.ids[k], std::move(pp));
        }
    }
    return rec;
}

// =====================================================================================================
// JointAsr::joint_asr (compute_joint_asr, likelihood.rs:421-560): posteriors stay None (:553-558)
// =====================================================================================================
Reconstruction PhyloTree::joint_asr(const GtrModel &model, const Alignment &aln) const
{
    Flat flat;
    std::vector<std::uint64_t> s2p;
    std::size_t P = 0;
    phb_ctx *c = load_device_compressed(*this, model, aln, flat, &s2p, P);
    const std::size_t n = flat.ids.size(), width = aln.width;
    std::vector<std::uint8_t> states(std::max<std::size_t>(n * P, 1));
    Reconstruction rec;
    check(c, api().joint_asr(c, states.data(), &rec.log_likelihood));
    for (std::size_t k = 0; k < n; ++k) { // every node id gets an entry (:453-456), expanded to sites (:545-551)
        std::vector<std::size_t> seq(width, 0);
        const std::uint8_t *row = states.data() + k * P;
        for (std::size_t s = 0; s < width; ++s) seq[s] = row[s2p[s]];
        rec.sequences.emplace(flat.ids[k], std::move(seq));
    }
    return rec;
}

// =====================================================================================================
// config 5: NNI neighbourhood scoring in one device batch
// =====================================================================================================
std::vector<double> PhyloTree::score_nni_neighbours(const GtrModel &model, const Alignment &aln,
                                                    const std::vector<std::pair<NodeID, bool>> &moves) const
{
    Flat flat;
    std::size_t P = 0;
    phb_ctx *c = load_device_compressed(*this, model, aln, flat, nullptr, P);
    std::vector<int32_t> ids(std::max<std::size_t>(moves.size(), 1), -1);
    std::vector<std::uint8_t> left(std::max<std::size_t>(moves.size(), 1), 0);
    for (std::size_t i = 0; i < moves.size(); ++i) {
        const NodeID v = moves[i].first;
        ids[i] = (v < flat.dense_of.size()) ? flat.dense_of[v] : -1;
        left[i] = moves[i].second ? 1 : 0;
    }
    std::vector<double> out(std::max<std::size_t>(moves.size(), 1), std::numeric_limits<double>::quiet_NaN());
    check(c, api().score_nni(c, (int32_t)moves.size(), ids.data(), left.data(), out.data()));
    out.resize(moves.size());
    return out;
}

} // namespace phylo
