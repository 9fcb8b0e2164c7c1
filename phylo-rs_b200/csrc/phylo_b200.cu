// phylo_b200.cu — context, host scheduler (H1) and C ABI of libphylo_b200.so.
//
// Replaces the bodies of the reference's engines (paths relative to /root/reference):
//   compute_log_likelihood  src/tree/likelihood.rs:156-209
//   compute_marginal_asr    src/tree/likelihood.rs:224-408
// behind the C ABI declared in include/phylo_b200.h.  The host side turns the post-order into
// levels (a node's level = 1 + max level of its children), assigns CLV slots, tiles the site axis
// so that a tile's CLVs fit in HBM, and launches the kernels of kernels.cuh level by level.
#include "kernels.cuh"
#include "sweep.cuh"
#include "nni.cuh"
#include "joint.cuh"
#include "compress.cuh"
#include "sweep_sched.h"
#include "down_sweep.cuh"
#include "aa.cuh"
#include "../../include/phylo_b200.h"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>
#include <string>
#include <chrono>
#include <vector>

using namespace phb;

namespace {

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes ? bytes : 16);
        if (e == cudaSuccess) cap = bytes ? bytes : 16;
        return e;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <typename T> T *as() const { return reinterpret_cast<T *>(p); }
};

// Host-side schedule derived from the flattened tree (phb_set_tree).
struct TreeSched {
    int n = 0, root = -1;
    std::vector<int> child_off, children, parent, leaf_row, level, depth, slot;
    std::vector<double> blen;
    int n_internal = 0, n_levels = 0, max_depth = 0;
    std::vector<NodeOp> ops;        // binary updates, grouped by level
    std::vector<int> level_off;     // n_levels + 1 (level l in 1..n_levels -> index l-1)
    std::vector<GenOp> gops;        // generic-arity updates, grouped by level
    std::vector<GenChild> gkids;
    std::vector<int> glevel_off;
    std::vector<GenOp> uops;        // every internal node as an any-arity op, grouped by level (joint ASR)
    std::vector<GenChild> ukids;
    std::vector<int> ulevel_off;
    std::vector<DownOp> dops;       // ASR pre-order ops (one per internal node), grouped by depth
    std::vector<DownChild> dkids;
    std::vector<int> dlevel_off;    // max_depth + 1
    std::vector<int> leaf_nodes, leaf_rows;
    double sum_one_plus_i = 0;      // sum over internal nodes of (1 + #internal children)
    double sum_t = 0;               // total tip children
    bool valid = false;

    // Builds everything from child lists.  Returns a PHB status.
    int build(std::string &err)
    {
        valid = false;
        parent.assign(n, -1);
        for (int v = 0; v < n; ++v)
            for (int k = child_off[v]; k < child_off[v + 1]; ++k) {
                int c = children[k];
                if (c < 0 || c >= n || c == root || parent[c] != -1) {
                    err = "tree: child id out of range or node with two parents";
                    return PHB_ERR_NUMERICAL_INSTABILITY;
                }
                parent[c] = v;
            }
        // iterative left-to-right DFS from the root: post-order + depth
        std::vector<int> post;
        post.reserve(n);
        depth.assign(n, 0);
        {
            std::vector<std::pair<int, int>> st;
            st.emplace_back(root, child_off[root]);
            while (!st.empty()) {
                auto &top = st.back();
                int v = top.first;
                if (top.second < child_off[v + 1]) {
                    int c = children[top.second++];
                    depth[c] = depth[v] + 1;
                    st.emplace_back(c, child_off[c]);
                } else {
                    post.push_back(v);
                    st.pop_back();
                }
            }
        }
        if ((int)post.size() != n) {
            err = "tree: not every node is reachable from the root";
            return PHB_ERR_NUMERICAL_INSTABILITY;
        }
        level.assign(n, 0);
        slot.assign(n, -1);
        n_internal = 0;
        n_levels = 0;
        max_depth = 0;
        leaf_nodes.clear();
        leaf_rows.clear();
        for (int v : post) {
            max_depth = std::max(max_depth, depth[v]);
            int nch = child_off[v + 1] - child_off[v];
            if (nch == 0) {
                leaf_nodes.push_back(v);
                continue;
            }
            int lv = 0;
            for (int k = child_off[v]; k < child_off[v + 1]; ++k) lv = std::max(lv, level[children[k]]);
            level[v] = lv + 1;
            n_levels = std::max(n_levels, level[v]);
            slot[v] = n_internal++;
        }
        std::sort(leaf_nodes.begin(), leaf_nodes.end());
        for (int v : leaf_nodes) leaf_rows.push_back(leaf_row[v]);
        // group ops by level
        std::vector<std::vector<int>> by_level(n_levels + 1);
        for (int v : post)
            if (level[v] > 0) by_level[level[v]].push_back(v);
        ops.clear(); gops.clear(); gkids.clear(); uops.clear(); ukids.clear();
        level_off.assign(1, 0);
        glevel_off.assign(1, 0);
        ulevel_off.assign(1, 0);
        sum_one_plus_i = 0;
        sum_t = 0;
        for (int l = 1; l <= n_levels; ++l) {
            for (int v : by_level[l]) {
                int b = child_off[v], e = child_off[v + 1];
                int n_int = 0, n_tip = 0;
                for (int k = b; k < e; ++k) (slot[children[k]] >= 0 ? n_int : n_tip)++;
                sum_one_plus_i += 1 + n_int;
                sum_t += n_tip;
                {
                    GenOp u;
                    u.out_slot = slot[v];
                    u.child_begin = (int)ukids.size();
                    for (int k = b; k < e; ++k) {
                        int c = children[k];
                        GenChild uc;
                        uc.is_tip = slot[c] < 0;
                        uc.src = uc.is_tip ? leaf_row[c] : slot[c];
                        uc.edge = c;
                        uc.pad = 0;
                        ukids.push_back(uc);
                    }
                    u.child_end = (int)ukids.size();
                    u.pad = 0;
                    uops.push_back(u);
                }
                if (e - b == 2) {
                    int c0 = children[b], c1 = children[b + 1];
                    NodeOp op;
                    op.out_slot = slot[v];
                    op.e0 = c0; op.e1 = c1;
                    op.kind = (slot[c0] < 0 ? 1 : 0) | (slot[c1] < 0 ? 2 : 0);
                    op.c0 = slot[c0] < 0 ? leaf_row[c0] : slot[c0];
                    op.c1 = slot[c1] < 0 ? leaf_row[c1] : slot[c1];
                    ops.push_back(op);
                } else {
                    GenOp g;
                    g.out_slot = slot[v];
                    g.child_begin = (int)gkids.size();
                    for (int k = b; k < e; ++k) {
                        int c = children[k];
                        GenChild gc;
                        gc.is_tip = slot[c] < 0;
                        gc.src = gc.is_tip ? leaf_row[c] : slot[c];
                        gc.edge = c;
                        gc.pad = 0;
                        gkids.push_back(gc);
                    }
                    g.child_end = (int)gkids.size();
                    g.pad = 0;
                    gops.push_back(g);
                }
            }
            level_off.push_back((int)ops.size());
            glevel_off.push_back((int)gops.size());
            ulevel_off.push_back((int)uops.size());
        }
        // ASR pre-order ops grouped by the parent's depth
        std::vector<std::vector<int>> by_depth(max_depth + 1);
        for (int v = 0; v < n; ++v)
            if (slot[v] >= 0) by_depth[depth[v]].push_back(v);
        dops.clear(); dkids.clear();
        dlevel_off.assign(1, 0);
        for (int d = 0; d <= max_depth; ++d) {
            for (int v : by_depth[d]) {
                DownOp o;
                o.parent_slot = slot[v];
                o.child_begin = (int)dkids.size();
                for (int k = child_off[v]; k < child_off[v + 1]; ++k) {
                    int c = children[k];
                    DownChild dc;
                    dc.node = c;
                    dc.is_tip = slot[c] < 0;
                    dc.src = dc.is_tip ? leaf_row[c] : slot[c];
                    dc.pad = 0; // 1: internal child with an internal child of its own (the only upper partials ever read)
                    if (!dc.is_tip)
                        for (int k2 = child_off[c]; k2 < child_off[c + 1]; ++k2)
                            if (slot[children[k2]] >= 0) dc.pad = 1;
                    dkids.push_back(dc);
                }
                o.child_end = (int)dkids.size();
                o.pad = v; // dense id of the parent: row of the states output (joint traceback)
                dops.push_back(o);
            }
            dlevel_off.push_back((int)dops.size());
        }
        valid = true;
        return PHB_OK;
    }
};

struct EventPair {
    cudaEvent_t a = nullptr, b = nullptr;
};

// bytes per pinned bounce buffer (two per device); PHB_BOUNCE_BYTES shrinks it so that tests walk through many slices
inline size_t bounce_bytes()
{
    static const size_t v = [] {
        const char *e = getenv("PHB_BOUNCE_BYTES");
        const long long n = e ? atoll(e) : 0;
        return n >= 4096 ? (size_t)n / 4096 * 4096 : (size_t)32 << 20;
    }();
    return v;
}

// pageable copies below this size take the driver's path (0 in the tests' tiny-buffer mode)
inline size_t bounce_min_bytes() { return getenv("PHB_BOUNCE_BYTES") ? 0 : (size_t)16 << 20; }

// Host threads that copy slices of a caller's PAGEABLE buffer into pinned bounce buffers (h2d_from_host): the driver's own
// path for pageable memory stages through one thread at about 11 GB/s, a fifth of what the link takes from pinned memory.
struct CopyPool {
    std::vector<std::thread> workers;
    std::mutex m;
    std::condition_variable cv, done_cv;
    struct Task { const uint8_t *src; uint8_t *dst; size_t n; };
    std::vector<Task> tasks;
    size_t pending = 0;
    bool stop = false;
    explicit CopyPool(int n)
    {
        for (int i = 0; i < n; ++i)
            workers.emplace_back([this] {
                for (;;) {
                    Task t;
                    {
                        std::unique_lock<std::mutex> lk(m);
                        cv.wait(lk, [this] { return stop || !tasks.empty(); });
                        if (tasks.empty()) return; // stop
                        t = tasks.back();
                        tasks.pop_back();
                    }
                    memcpy(t.dst, t.src, t.n);
                    {
                        std::lock_guard<std::mutex> lk(m);
                        if (--pending == 0) done_cv.notify_all();
                    }
                }
            });
    }
    ~CopyPool()
    {
        {
            std::lock_guard<std::mutex> lk(m);
            stop = true;
        }
        cv.notify_all();
        for (std::thread &t : workers) t.join();
    }
    // every task of `list`, the last ones by the calling thread while it waits
    void run(std::vector<Task> &list)
    {
        if (workers.empty() || list.size() < 2) {
            for (const Task &t : list) memcpy(t.dst, t.src, t.n);
            return;
        }
        Task own = list.back();
        list.pop_back();
        {
            std::lock_guard<std::mutex> lk(m);
            for (const Task &t : list) {
                tasks.push_back(t);
                ++pending;
            }
        }
        cv.notify_all();
        memcpy(own.dst, own.src, own.n);
        std::unique_lock<std::mutex> lk(m);
        done_cv.wait(lk, [this] { return pending == 0; });
    }
    // dst[0..n) = src[0..n), split over the workers and the calling thread
    void copy(uint8_t *dst, const uint8_t *src, size_t n)
    {
        const size_t parts = workers.size() + 1;
        const size_t slice = (n / parts + 4095) / 4096 * 4096;
        if (workers.empty() || n < std::min<size_t>((size_t)1 << 20, bounce_bytes() / 2) || slice == 0) {
            memcpy(dst, src, n);
            return;
        }
        size_t own = std::min(slice, n);
        {
            std::lock_guard<std::mutex> lk(m);
            for (size_t off = own; off < n; off += slice) {
                tasks.push_back({src + off, dst + off, std::min(slice, n - off)});
                ++pending;
            }
        }
        cv.notify_all();
        memcpy(dst, src, own);
        std::unique_lock<std::mutex> lk(m);
        done_cv.wait(lk, [this] { return pending == 0; });
    }
};

struct Device {
    int ordinal = 0;
    cudaStream_t stream = nullptr;
    int64_t p_begin = 0, P = 0, P_pad = 0; // shard of the global pattern range
    DevBuf tips, mult, leaf_ls, inv_mask, per_pattern, row_invalid, Pmat, blen, clv, scale, cat_ll,
        partials, total, ops, gops, gkids, dops, dkids, leaf_rows, states, posteriors, staging,
        sim_pre, sim_parent, sim_depth, sim_leafrow, sim_stack, scratch, sw_descs, sw_opoff, sw_blocks, sw_stream, sw_gstack,
        sw_err, sw_ops, dn_ops, dn_descs, dn_blocks, dn_stream, dn_opoff, dn_gstack, uops, ukids, lnP, jbest, nni_upper, nni_uexps, nni_jobs, nni_rows, nni_exps, nni_partials, nni_totals, nni_moves, nni_steps, nni_need, nni_cnt, nni_pred, nni_dops, nni_dkids;
    bool sweep_uploaded = false;
    bool sweep_ops_stale = false; // the guard flags of the op table changed since the last upload
    bool down_uploaded = false;   // pre-order sweep schedule (down_sweep.cuh)
    int sm_count = 0;
    size_t smem_optin = 0;
    cudaStream_t copy_stream = nullptr;          // H2D of alignment chunks, overlapped with packing / sweeping
    DevBuf staging2;                             // second staging buffer (double buffering)
    cudaEvent_t ev_copied[2] = {nullptr, nullptr}, ev_free[2] = {nullptr, nullptr};
    double *h_total = nullptr; // pinned
    uint8_t *bounce[2] = {nullptr, nullptr};           // pinned, bounce_bytes() each: pageable host input goes through them
    cudaEvent_t ev_bounce[2] = {nullptr, nullptr};     // the DMA out of bounce[i] has finished
    uint64_t bounce_uses = 0;                          // slices sent so far: slice k uses bounce[k & 1]
    cudaEvent_t tm_a = nullptr, tm_b = nullptr; // phb_timer_start / phb_timer_stop
    int64_t tile_pad = 0;
    int n_tiles = 0;
    bool tip_constants_valid = false;
    std::vector<int> tip_constants_rows;
    // timing (device 0 only, PHB_FLAG_TIMING)
    std::vector<EventPair> ev_pool;
    size_t ev_used = 0;
    std::vector<std::pair<int, size_t>> ev_log; // (phase, pool index)
};

// 20-state (AminoAcid) engines of a PHB_FLAG_AMINO_ACID context: aa.cuh, aa_engine.cuh
struct AaState {
    AaModelDev M;
    DevBuf model, tips, mult, leaf_ls, per_pattern, Pmat, blen, clv, exps, cat_ll, partials, uops, ukids, down, leaf_rows, states, post, row_flags, lnP, jbest, parent, nni_ops, nni_kids, nni_roots;
    std::vector<AaDown> down_nodes;
    std::vector<int> down_off; // by depth, max_depth + 2 entries
    int64_t P_pad = 0;
    bool tip_const_valid = false;
};

enum Phase { PH_TRANSITION = 0, PH_CLV = 1, PH_ROOT = 2, PH_DOWN = 3, PH_UPPER = 4 };

} // namespace

struct phb_ctx {
    std::vector<Device> devs;
    uint64_t flags = 0;
    uint64_t clv_budget = 0;
    std::string err;
    // alignment
    bool have_aln = false;
    int n_rows = 0;
    int64_t n_patterns = 0;
    std::vector<int> row_invalid_host;   // per row: an unaccepted byte somewhere in the row
    std::vector<uint8_t> row_first_invalid; // per row: the byte of pattern 0 is unaccepted (error precedence, check_leaves)
    // model
    bool have_model = false;
    ModelDev model;      // as used by the lnL path (analytic +I unless LITERAL_INVARIANT)
    ModelDev model_lit;  // literal categories (ASR path)
    // tree
    TreeSched tree;
    bool tree_uploaded = false;
    SweepSched sweep;      // fused-sweep op stream for (tree, n_clv)
    SweepGeom sweep_geom;
    bool sweep_guards_valid = false; // guard flags planned for the current branch lengths and rates
    // the NNI cache pass runs a storing sweep with its own schedule (only the partials some scored edge reads are
    // written); it is swapped into `sweep` for the duration of the pass (score_nni_paths)
    SweepSched nni_sweep;
    SweepGeom nni_geom;
    bool nni_guards_valid = false;
    DownSched down;        // pre-order sweep schedule of the marginal-ASR down pass for (tree, n_clv)
    // marginal ASR: caller's output buffers, so that a finished tile's states go home while the next tile runs
    uint8_t *asr_states_out = nullptr;
    double *asr_post_out = nullptr;
    // stats
    phb_stats stats;
    int64_t launches = 0;
    std::unique_ptr<CopyPool> copy_pool; // created by the first call that is handed pageable memory
    // 20 states (PHB_FLAG_AMINO_ACID)
    bool is_aa = false;
    AaState aa;

    int fail(int code, const char *fmt, ...)
    {
        char buf[512];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof buf, fmt, ap);
        va_end(ap);
        err = buf;
        return code;
    }
};

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return ctx->fail(e_ == cudaErrorMemoryAllocation ? PHB_ERR_OOM : PHB_ERR_CUDA,         \
                             "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

#define CHECK_LAUNCH()                                                                             \
    do {                                                                                           \
        ctx->launches++;                                                                           \
        cudaError_t e_ = cudaGetLastError();                                                       \
        if (e_ != cudaSuccess)                                                                     \
            return ctx->fail(PHB_ERR_CUDA, "kernel launch failed: %s (%s:%d)", cudaGetErrorString(e_), \
                             __FILE__, __LINE__);                                                  \
    } while (0)

namespace {

inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }
constexpr size_t TIPS_SLACK = 4096; // the sweep's last site block may read past the final row

// timing helpers: bracket one launch with an event pair on device 0 when PHB_FLAG_TIMING is set
struct ScopedTimer {
    phb_ctx *ctx;
    Device *d;
    bool on;
    size_t idx = 0;
    ScopedTimer(phb_ctx *c, Device *dev, int phase) : ctx(c), d(dev)
    {
        on = (c->flags & PHB_FLAG_TIMING) && dev == &c->devs[0];
        if (!on) return;
        if (d->ev_used == d->ev_pool.size()) {
            EventPair ep;
            cudaEventCreate(&ep.a);
            cudaEventCreate(&ep.b);
            d->ev_pool.push_back(ep);
        }
        idx = d->ev_used++;
        d->ev_log.emplace_back(phase, idx);
        cudaEventRecord(d->ev_pool[idx].a, d->stream);
    }
    ~ScopedTimer()
    {
        if (on) cudaEventRecord(d->ev_pool[idx].b, d->stream);
    }
};

void build_model_dev(const ModelDev &base, bool literal, ModelDev &out)
{
    out = base;
    int n_clv = 0;
    for (int c = 0; c < out.n_cat; ++c) {
        if (!literal && out.rates[c] == 0.0) out.clv_idx[c] = -1;
        else {
            out.clv_idx[c] = n_clv;
            out.cat_of_clv[n_clv] = c;
            ++n_clv;
        }
    }
    out.n_clv = n_clv;
}

int upload_tree(phb_ctx *ctx)
{
    TreeSched &T = ctx->tree;
    for (Device &d : ctx->devs) {
        CU(cudaSetDevice(d.ordinal));
        auto up = [&](DevBuf &b, const void *src, size_t bytes) -> cudaError_t {
            cudaError_t e = b.ensure(bytes);
            if (e != cudaSuccess) return e;
            if (bytes) return cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, d.stream);
            return cudaSuccess;
        };
        CU(up(d.blen, T.blen.data(), sizeof(double) * T.n));
        CU(up(d.ops, T.ops.data(), sizeof(NodeOp) * T.ops.size()));
        CU(up(d.gops, T.gops.data(), sizeof(GenOp) * T.gops.size()));
        CU(up(d.gkids, T.gkids.data(), sizeof(GenChild) * T.gkids.size()));
        CU(up(d.dops, T.dops.data(), sizeof(DownOp) * T.dops.size()));
        CU(up(d.dkids, T.dkids.data(), sizeof(DownChild) * T.dkids.size()));
        CU(up(d.uops, T.uops.data(), sizeof(GenOp) * T.uops.size()));
        CU(up(d.ukids, T.ukids.data(), sizeof(GenChild) * T.ukids.size()));
        CU(up(d.leaf_rows, T.leaf_rows.data(), sizeof(int) * T.leaf_rows.size()));
        if (d.tip_constants_rows != T.leaf_rows) d.tip_constants_valid = false;
        // the host vectors stay alive in ctx->tree until the next set_tree, which first syncs
        CU(cudaStreamSynchronize(d.stream));
    }
    ctx->tree_uploaded = true;
    if (!ctx->sweep.same_topology(T.root, T.child_off, T.children, T.leaf_row, T.slot)) ctx->sweep.valid = false;
    if (!ctx->nni_sweep.same_topology(T.root, T.child_off, T.children, T.leaf_row, T.slot)) ctx->nni_sweep.valid = false;
    ctx->nni_guards_valid = false;
    ctx->sweep_guards_valid = false; // branch lengths may have changed: re-plan the guards (and re-check the folded cherries)
    return PHB_OK;
}

// Validation shared by both engines: every leaf needs a row (likelihood.rs:99-103) holding only
// accepted bytes (likelihood.rs:105-107).
// The reference meets these errors pattern by pattern, walking the leaves in post-order: on pattern 0 a leaf
// without a row (InvalidAlignment) or with an unaccepted byte IN PATTERN 0 (AlphabetMismatch), whichever leaf
// comes first; an unaccepted byte in a later pattern can only surface once pattern 0 went through, i.e. when
// every leaf has a row.  leaf_nodes is in that post-order (children in stored order, src/iter.rs:112-138).
int check_leaves(phb_ctx *ctx)
{
    const TreeSched &T = ctx->tree;
    const bool have_first = ctx->row_first_invalid.size() >= (size_t)ctx->n_rows;
    for (size_t i = 0; i < T.leaf_nodes.size(); ++i) {
        int r = T.leaf_rows[i];
        if (r < 0 || r >= ctx->n_rows)
            return ctx->fail(PHB_ERR_INVALID_ALIGNMENT, "Leaf in tree not found in alignment leaf order (node %d)",
                             T.leaf_nodes[i]);
        if (have_first && ctx->row_first_invalid[r])
            return ctx->fail(PHB_ERR_ALPHABET_MISMATCH, "Invalid char in alignment (row %d)", r);
    }
    for (size_t i = 0; i < T.leaf_nodes.size(); ++i)
        if (ctx->row_invalid_host[T.leaf_rows[i]])
            return ctx->fail(PHB_ERR_ALPHABET_MISMATCH, "Invalid char in alignment (row %d)", T.leaf_rows[i]);
    return PHB_OK;
}

// Plans the site tiling of one device and (re)allocates the per-evaluation buffers.
int plan_device(phb_ctx *ctx, Device &d, const ModelDev &M, bool asr, bool want_post, bool use_sweep, int64_t wave_sites = 0)
{
    const TreeSched &T = ctx->tree;
    CU(cudaSetDevice(d.ordinal));
    if (use_sweep) { // no CLV tiles in HBM: the whole shard is one pass
        CU(d.per_pattern.ensure(sizeof(double) * (size_t)d.P_pad));
        CU(d.leaf_ls.ensure(sizeof(double) * (size_t)d.P_pad));
        CU(d.inv_mask.ensure((size_t)d.P_pad));
        CU(d.total.ensure(2 * sizeof(double)));
        d.tile_pad = d.P_pad;
        d.n_tiles = d.P > 0 ? 1 : 0;
        return PHB_OK;
    }
    if (asr) {
        CU(d.states.ensure((size_t)T.n * (size_t)d.P_pad));
        if (want_post) CU(d.posteriors.ensure((size_t)T.n * (size_t)d.P_pad * 4 * sizeof(double)));
    }
    CU(d.Pmat.ensure(sizeof(double) * 16 * (size_t)T.n * M.n_cat));
    CU(d.per_pattern.ensure(sizeof(double) * (size_t)d.P_pad));
    CU(d.leaf_ls.ensure(sizeof(double) * (size_t)d.P_pad));
    CU(d.inv_mask.ensure((size_t)d.P_pad));
    CU(d.total.ensure(2 * sizeof(double)));
    const int n_slots = std::max(T.n_internal, 1);
    const size_t per_site = (size_t)n_slots * M.n_clv * 4 * sizeof(double);
    size_t need_full = per_site * (size_t)d.P_pad;
    size_t budget = ctx->clv_budget;
    if (!budget) {
        size_t free_b = 0, total_b = 0;
        CU(cudaMemGetInfo(&free_b, &total_b));
        budget = (size_t)((double)(free_b + d.clv.cap) * 0.85);
    }
    int64_t tile = d.P_pad;
    if (need_full > budget) {
        tile = (int64_t)(budget / per_site) / SITES_PER_BLOCK * SITES_PER_BLOCK;
        if (tile < SITES_PER_BLOCK)
            return ctx->fail(PHB_ERR_OOM, "CLV budget %zu B cannot hold one %d-site tile of %d slots", budget,
                             SITES_PER_BLOCK, n_slots);
    }
    if (asr && d.P_pad >= 8 * 32768) // at least 8 tiles: the D2H of a finished tile overlaps the next tile's kernels
        tile = std::min<int64_t>(tile, round_up((d.P_pad + 7) / 8, SITES_PER_BLOCK));
    // storing sweep: a tile that is a whole number of persistent-grid passes (SMs x patterns per CTA pass)
    if (wave_sites >= 16 && wave_sites % 16 == 0 && tile > wave_sites && tile < d.P_pad) tile = tile / wave_sites * wave_sites;
    d.tile_pad = tile;
    d.n_tiles = (int)((d.P + tile - 1) / tile);
    if (d.P == 0) d.n_tiles = 0;
    CU(d.clv.ensure(per_site * (size_t)(tile + 512))); // + slack: the blocked layout of the sweep path rounds the tile up to whole blocks
    CU(d.scale.ensure(sizeof(int) * (size_t)std::max(M.n_clv, 1) * (size_t)tile));
    if (asr) CU(d.cat_ll.ensure(sizeof(double) * (size_t)M.n_cat * (size_t)tile));
    if (asr) { // joint ASR: ln P table and the winning category per pattern
        CU(d.lnP.ensure(sizeof(double) * 16 * (size_t)T.n * M.n_cat));
        CU(d.jbest.ensure((size_t)tile));
    }
    size_t n_part = 0;
    for (int t = 0; t < d.n_tiles; ++t) {
        int64_t len = std::min<int64_t>(tile, d.P - (int64_t)t * tile);
        n_part += (size_t)((len + ROOT_THREADS - 1) / ROOT_THREADS);
    }
    CU(d.partials.ensure(sizeof(double) * std::max<size_t>(n_part, 1)));
    return PHB_OK;
}

// The up pass of one tile: all levels, then the root kernel.
int run_tile_up(phb_ctx *ctx, Device &d, const ModelDev &M, int64_t tile_off, int tile_len, size_t part_off,
                bool want_cat_ll)
{
    const TreeSched &T = ctx->tree;
    const size_t tile_pad = (size_t)d.tile_pad;
    const size_t slot_stride = (size_t)M.n_clv * 4 * tile_pad;
    const int bpr = (tile_len + SITES_PER_BLOCK - 1) / SITES_PER_BLOCK;
    if (M.n_clv > 0) CU(cudaMemsetAsync(d.scale.p, 0, sizeof(int) * (size_t)M.n_clv * tile_pad, d.stream));
    if (T.n_internal == 0) {
        // single-leaf tree: the root is a tip
        if (M.n_clv > 0) {
            ScopedTimer tm(ctx, &d, PH_CLV);
            PHB_LAUNCH(clv_from_tip_kernel, dim3((tile_len + 255) / 256), dim3(256), d.stream, d.clv.as<double>(),
                       slot_stride, 0, M.n_clv, tile_pad, tile_len, d.tips.as<uint8_t>(), d.P_pad, tile_off,
                       T.leaf_row[T.root]);
            CHECK_LAUNCH();
        }
    } else if (M.n_clv > 0) {
        for (int l = 1; l <= T.n_levels; ++l) {
            int nb = T.level_off[l] - T.level_off[l - 1];
            if (nb > 0) {
                ScopedTimer tm(ctx, &d, PH_CLV);
                int64_t grid = (int64_t)nb * M.n_clv * bpr;
                PHB_LAUNCH(clv_update_level_kernel, dim3((unsigned)grid), dim3(CLV_THREADS), d.stream,
                           d.ops.as<NodeOp>() + T.level_off[l - 1], nb, M, d.Pmat.as<double>(), d.clv.as<double>(),
                           slot_stride, tile_pad, tile_len, bpr, d.tips.as<uint8_t>(), d.P_pad, tile_off,
                           d.scale.as<int>());
                CHECK_LAUNCH();
                if (&d == &ctx->devs[0]) ctx->stats.clv_launches++;
            }
            int ng = T.glevel_off[l] - T.glevel_off[l - 1];
            if (ng > 0) {
                ScopedTimer tm(ctx, &d, PH_CLV);
                int64_t grid = (int64_t)ng * M.n_clv * bpr;
                PHB_LAUNCH(clv_update_generic_kernel, dim3((unsigned)grid), dim3(CLV_THREADS), d.stream,
                           d.gops.as<GenOp>() + T.glevel_off[l - 1], ng, d.gkids.as<GenChild>(), M,
                           d.Pmat.as<double>(), d.clv.as<double>(), slot_stride, tile_pad, tile_len, bpr,
                           d.tips.as<uint8_t>(), d.P_pad, tile_off, d.scale.as<int>());
                CHECK_LAUNCH();
                if (&d == &ctx->devs[0]) ctx->stats.clv_launches++;
            }
        }
    }
    {
        ScopedTimer tm(ctx, &d, PH_ROOT);
        int root_slot = T.n_internal == 0 ? 0 : T.slot[T.root];
        PHB_LAUNCH(root_kernel, dim3((tile_len + ROOT_THREADS - 1) / ROOT_THREADS), dim3(ROOT_THREADS), d.stream,
                   d.clv.as<double>() + (size_t)root_slot * slot_stride, tile_pad, tile_len, d.scale.as<int>(),
                   d.leaf_ls.as<double>() + tile_off, d.inv_mask.as<uint8_t>() + tile_off,
                   d.mult.as<double>() + tile_off, M, d.per_pattern.as<double>() + tile_off,
                   want_cat_ll ? d.cat_ll.as<double>() : (double *)nullptr, d.partials.as<double>() + part_off);
        CHECK_LAUNCH();
    }
    return PHB_OK;
}

int run_tile_down(phb_ctx *ctx, Device &d, const ModelDev &M, int64_t tile_off, int tile_len, bool want_post)
{
    const TreeSched &T = ctx->tree;
    const size_t tile_pad = (size_t)d.tile_pad;
    const size_t slot_stride = (size_t)M.n_clv * 4 * tile_pad;
    const int bpr = (tile_len + ROOT_THREADS - 1) / ROOT_THREADS;
    const int bpr2 = (tile_len + 2 * DOWN_THREADS - 1) / (2 * DOWN_THREADS); // down kernel: a site pair per thread
    uint8_t *states = d.states.as<uint8_t>();
    double *post = want_post ? d.posteriors.as<double>() : nullptr;
    {
        ScopedTimer tm(ctx, &d, PH_DOWN);
        int root_slot = T.n_internal == 0 ? 0 : T.slot[T.root];
        PHB_LAUNCH(asr_root_kernel, dim3(bpr), dim3(ROOT_THREADS), d.stream,
                   d.clv.as<double>() + (size_t)root_slot * slot_stride, tile_pad, tile_len, M, d.cat_ll.as<double>(),
                   d.per_pattern.as<double>() + tile_off, states + (int64_t)T.root * d.P_pad + tile_off,
                   post ? post + ((int64_t)T.root * d.P_pad + tile_off) * 4 : (double *)nullptr);
        CHECK_LAUNCH();
    }
    for (int dl = 0; dl <= T.max_depth; ++dl) {
        int nb = T.dlevel_off[dl + 1] - T.dlevel_off[dl];
        if (nb <= 0) continue;
        ScopedTimer tm(ctx, &d, PH_DOWN);
        int64_t grid = (int64_t)nb * bpr2;
        PHB_LAUNCH(asr_down_level_kernel, dim3((unsigned)grid), dim3(DOWN_THREADS), d.stream,
                   d.dops.as<DownOp>() + T.dlevel_off[dl], nb, d.dkids.as<DownChild>(), M, d.Pmat.as<double>(),
                   d.clv.as<double>(), slot_stride, tile_pad, tile_len, bpr2, d.tips.as<uint8_t>(), d.P_pad, tile_off,
                   d.cat_ll.as<double>(), states, post);
        CHECK_LAUNCH();
    }
    return PHB_OK;
}

// Joint (Viterbi) ASR of one tile: max-plus up pass over all levels, root kernel, traceback (joint.cuh).
int run_tile_joint(phb_ctx *ctx, Device &d, const ModelDev &M, int64_t tile_off, int tile_len, size_t part_off)
{
    const TreeSched &T = ctx->tree;
    const size_t tile_pad = (size_t)d.tile_pad;
    const size_t slot_stride = (size_t)M.n_clv * 4 * tile_pad; // literal categories: n_clv == n_cat
    const int bpr = (tile_len + SITES_PER_BLOCK - 1) / SITES_PER_BLOCK;
    uint8_t *states = d.states.as<uint8_t>();
    for (int l = 1; l <= T.n_levels; ++l) {
        const int nb = T.level_off[l] - T.level_off[l - 1]; // binary nodes: both children loaded up front
        if (nb > 0) {
            ScopedTimer tm(ctx, &d, PH_CLV);
            PHB_LAUNCH(joint_up_binary_kernel, dim3((unsigned)((int64_t)nb * M.n_cat * bpr)), dim3(CLV_THREADS), d.stream,
                       d.ops.as<NodeOp>() + T.level_off[l - 1], nb, M, d.lnP.as<double>(), d.clv.as<double>(), slot_stride,
                       tile_pad, tile_len, bpr, d.tips.as<uint8_t>(), d.P_pad, tile_off);
            CHECK_LAUNCH();
            if (&d == &ctx->devs[0]) ctx->stats.clv_launches++;
        }
        const int ng = T.glevel_off[l] - T.glevel_off[l - 1]; // any other arity
        if (ng > 0) {
            ScopedTimer tm(ctx, &d, PH_CLV);
            PHB_LAUNCH(joint_up_level_kernel, dim3((unsigned)((int64_t)ng * M.n_cat * bpr)), dim3(CLV_THREADS), d.stream,
                       d.gops.as<GenOp>() + T.glevel_off[l - 1], ng, d.gkids.as<GenChild>(), M, d.lnP.as<double>(),
                       d.clv.as<double>(), slot_stride, tile_pad, tile_len, bpr, d.tips.as<uint8_t>(), d.P_pad, tile_off);
            CHECK_LAUNCH();
            if (&d == &ctx->devs[0]) ctx->stats.clv_launches++;
        }
    }
    {
        ScopedTimer tm(ctx, &d, PH_ROOT);
        const bool leaf_root = T.n_internal == 0;
        const int root_slot = leaf_root ? 0 : T.slot[T.root];
        PHB_LAUNCH(joint_root_kernel, dim3((tile_len + ROOT_THREADS - 1) / ROOT_THREADS), dim3(ROOT_THREADS), d.stream,
                   d.clv.as<double>() + (size_t)root_slot * slot_stride, tile_pad, tile_len, M,
                   leaf_root ? d.tips.as<uint8_t>() + (int64_t)T.leaf_row[T.root] * d.P_pad + tile_off : (const uint8_t *)nullptr,
                   d.mult.as<double>() + tile_off, d.per_pattern.as<double>() + tile_off, d.jbest.as<uint8_t>(),
                   states + (int64_t)T.root * d.P_pad + tile_off, d.partials.as<double>() + part_off);
        CHECK_LAUNCH();
    }
    const int bpr1 = (tile_len + ROOT_THREADS - 1) / ROOT_THREADS;
    for (int dl = 0; dl <= T.max_depth; ++dl) {
        const int nb = T.dlevel_off[dl + 1] - T.dlevel_off[dl];
        if (nb <= 0) continue;
        ScopedTimer tm(ctx, &d, PH_DOWN);
        PHB_LAUNCH(joint_down_level_kernel, dim3((unsigned)((int64_t)nb * bpr1)), dim3(ROOT_THREADS), d.stream,
                   d.dops.as<DownOp>() + T.dlevel_off[dl], nb, d.dkids.as<DownChild>(), M, d.lnP.as<double>(),
                   d.clv.as<double>(), slot_stride, tile_pad, tile_len, bpr1, d.tips.as<uint8_t>(), d.P_pad, tile_off,
                   d.jbest.as<uint8_t>(), states);
        CHECK_LAUNCH();
    }
    return PHB_OK;
}

// ---- fused sweep (sweep.cuh) ---------------------------------------------------------------------
// (Re)builds the op stream for the current tree and category count and uploads it to every device.
// keep: nodes whose own CLV the caller wants written (NNI cache pass): their cherries are not folded away
int ensure_sweep(phb_ctx *ctx, const ModelDev &M, bool store, const std::vector<char> *keep = nullptr)
{
    const auto t_begin = std::chrono::steady_clock::now();
    struct Trace {
        std::chrono::steady_clock::time_point t0;
        ~Trace()
        {
            if (getenv("PHB_SWEEP_TRACE"))
                fprintf(stderr, "ensure_sweep: %.3f ms on the host\n",
                        std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
        }
    } trace{t_begin};
    const TreeSched &T = ctx->tree;
    SweepSched &S = ctx->sweep;
    int64_t shard_P = 0;
    for (Device &d : ctx->devs) shard_P = std::max(shard_P, d.P);
    size_t smem_max = ctx->devs[0].smem_optin;
    for (Device &d : ctx->devs) smem_max = std::min(smem_max, d.smem_optin);
    // the plain sweep folds cherries into their parents' ops (sweep_sched.h); the storing one needs every node's CLV
    const bool collapse = !store && !getenv("PHB_SWEEP_NO_COLLAPSE");
    std::vector<double> gb, gbd;
    std::vector<char> no_collapse;
    const bool replan = keep || !S.valid || S.n_clv != M.n_clv || S.collapsed_build != collapse || !ctx->sweep_guards_valid;
    if (replan) { // per-edge shrink bounds: which cherries may be folded, where the underflow guard runs
        sweep_guard_bounds(M, T.blen, gb, gbd);
        if (collapse) sweep_no_collapse(T.child_off, T.children, gb, gbd, no_collapse);
        if (collapse && keep) {
            no_collapse.resize(T.n, 0);
            for (int v = 0; v < T.n; ++v) no_collapse[v] = no_collapse[v] || (*keep)[v];
        }
    }
    if (!S.valid || S.n_clv != M.n_clv || S.collapsed_build != collapse || (replan && collapse && no_collapse != S.built_no_collapse)) {
        for (Device &d : ctx->devs) { // the old host image may still be in flight
            cudaSetDevice(d.ordinal);
            cudaStreamSynchronize(d.stream);
            d.sweep_uploaded = false;
        }
        S.build(T.n, T.root, T.child_off, T.children, T.leaf_row, M.n_clv, T.slot, collapse, no_collapse);
        S.built_no_collapse = no_collapse;
        ctx->sweep_guards_valid = false;
        ctx->sweep_geom.planned_P = -1;
        if (getenv("PHB_SWEEP_TRACE"))
            fprintf(stderr, "sweep schedule: ops %d (TT %d TI %d II %d TS %d SI %d APPLY %d MULTIP %d MULPOP %d LOADTIP %d) pushes %d "
                            "register siblings %d levels %d stream %zu B record cap %d B rows %d\n", S.n_ops, S.count_kind[SW_TT],
                    S.count_kind[SW_TI], S.count_kind[SW_II], S.count_kind[SW_TS], S.count_kind[SW_SI], S.count_kind[SW_APPLY],
                    S.count_kind[SW_MULTIP], S.count_kind[SW_MULPOP], S.count_kind[SW_LOADTIP], S.n_push, S.n_reg_sibling,
                    S.n_levels, S.stream_bytes, S.rec_cap, S.max_rows);
    }
    if (ctx->sweep_geom.planned_P != shard_P) {
        if (!ctx->sweep_geom.plan(S, smem_max, shard_P, std::max(ctx->devs[0].sm_count, 1), 0, true)) // run_sweep splits the tail off
            return ctx->fail(PHB_ERR_INVALID_MODEL_PARAMETER, "sweep: %d CLV categories do not fit one CTA", M.n_clv);
    }
    if (!ctx->sweep_guards_valid) { // branch lengths or rates changed: re-plan where the underflow guard runs
        std::vector<uint32_t> before(S.ops.size());
        for (size_t k = 0; k < S.ops.size(); ++k) before[k] = S.ops[k].w0;
        for (Device &d : ctx->devs) // the host image of the op table may still be in flight
            if (d.sweep_uploaded) {
                cudaSetDevice(d.ordinal);
                cudaStreamSynchronize(d.stream);
            }
        S.plan_guards(gb, gbd);
        bool changed = false;
        for (size_t k = 0; k < S.ops.size() && !changed; ++k) changed = before[k] != S.ops[k].w0;
        if (changed)
            for (Device &d : ctx->devs) d.sweep_ops_stale = true;
        ctx->sweep_guards_valid = true;
        if (getenv("PHB_SWEEP_TRACE"))
            fprintf(stderr, "sweep guards: %d lazy + %d full of %d nodes\n", S.n_guards, S.n_normout, T.n_internal);
    }
    for (Device &d : ctx->devs) {
        if (d.sweep_uploaded && !d.sweep_ops_stale) continue;
        CU(cudaSetDevice(d.ordinal));
        if (!d.sweep_uploaded) {
            CU(d.sw_descs.ensure(sizeof(SweepStage) * S.stages.size()));
            CU(d.sw_opoff.ensure(sizeof(uint64_t) * std::max<size_t>(S.op_off.size(), 1)));
            CU(d.sw_blocks.ensure(sizeof(SweepBlock) * S.blocks.size()));
            CU(d.sw_stream.ensure(S.stream_bytes + 16));
            CU(d.sw_err.ensure(sizeof(unsigned int)));
            CU(cudaMemcpyAsync(d.sw_descs.p, S.stages.data(), sizeof(SweepStage) * S.stages.size(), cudaMemcpyHostToDevice, d.stream));
            CU(cudaMemcpyAsync(d.sw_opoff.p, S.op_off.data(), sizeof(uint64_t) * S.op_off.size(), cudaMemcpyHostToDevice, d.stream));
            CU(cudaMemcpyAsync(d.sw_blocks.p, S.blocks.data(), sizeof(SweepBlock) * S.blocks.size(), cudaMemcpyHostToDevice, d.stream));
            CU(d.sw_ops.ensure(sizeof(SweepOp) * S.ops.size()));
            CU(cudaMemsetAsync(d.sw_err.p, 0, sizeof(unsigned int), d.stream));
        }
        CU(cudaMemcpyAsync(d.sw_ops.p, S.ops.data(), sizeof(SweepOp) * S.ops.size(), cudaMemcpyHostToDevice, d.stream));
        d.sweep_uploaded = true;
        d.sweep_ops_stale = false;
    }
    return PHB_OK;
}

// P(t) / tip tables of every op, written straight into the op-ordered stream.
int launch_sweep_transition(phb_ctx *ctx, Device &d, const ModelDev &M)
{
    const SweepSched &S = ctx->sweep;
    ScopedTimer tm(ctx, &d, PH_TRANSITION);
    int nthreads = (int)S.blocks.size() * M.n_clv + S.n_ops;
    PHB_LAUNCH(sweep_transition_kernel, dim3((nthreads + 127) / 128), dim3(128), d.stream, d.sw_blocks.as<SweepBlock>(),
               (int)S.blocks.size(), d.blen.as<double>(), M, d.sw_stream.as<unsigned char>(), d.sw_ops.as<SweepOp>(),
               d.sw_opoff.as<uint64_t>(), S.n_ops);
    CHECK_LAUNCH();
    return PHB_OK;
}

// One persistent sweep launch over patterns [p0, p0 + np) of the device's shard with geometry G; the per-block
// partial sums land at partials[part_off ...].  p0 must be a multiple of 16 (TMA source alignment).
// clv_out != null selects the storing variant (marginal-ASR up pass): CLVs of the launch's patterns go to
// clv_out[slot][cat][state][site - p0] with row stride tile_pad, per-category log-likelihoods to cat_ll_out.
int launch_sweep_range(phb_ctx *ctx, Device &d, const ModelDev &M, const SweepGeom &G, int64_t p0, int64_t np,
                       size_t part_off, double *clv_out = nullptr, double *cat_ll_out = nullptr, size_t tile_pad = 0,
                       bool blocked = false, int *exp_out = nullptr, uint32_t n_store = 0xffffffffu)
{
    const SweepSched &S = ctx->sweep;
    const int sites_per_block = 4 * G.Q;
    const int64_t n_blocks = (np + sites_per_block - 1) / sites_per_block;
    const int grid = (int)std::min<int64_t>(n_blocks, std::max(d.sm_count, 1));
    if (G.g_levels > 0) CU(d.sw_gstack.ensure((size_t)std::max(d.sm_count, 1) * G.g_levels * G.level_bytes));
    SweepParams sp;
    memset(&sp, 0, sizeof sp);
    sp.ops = d.sw_ops.as<SweepOp>();
    sp.stages = d.sw_descs.as<SweepStage>();
    sp.n_stage_descs = S.n_stage_descs;
    sp.stream = d.sw_stream.as<unsigned char>();
    sp.n_ops = S.n_ops;
    sp.Q = G.Q;
    sp.n_consumers = G.n_consumers;
    sp.n_stages = G.n_stages;
    sp.stage_bytes = G.stage_bytes;
    sp.rec_cap = (S.stage_rec_cap + 15) / 16 * 16;
    sp.smem_levels = G.smem_levels;
    sp.g_levels = G.g_levels;
    sp.gstack = d.sw_gstack.as<double>();
    sp.tips = d.tips.as<uint8_t>() + p0;
    sp.P_pad = d.P_pad;
    sp.P = np;
    sp.n_blocks = n_blocks;
    sp.leaf_ls = d.leaf_ls.as<double>() + p0;
    sp.inv_mask = d.inv_mask.as<uint8_t>() + p0;
    sp.mult = d.mult.as<double>() + p0;
    sp.per_pattern = d.per_pattern.as<double>() + p0;
    sp.partials = d.partials.as<double>() + part_off;
    sp.error_flag = d.sw_err.as<unsigned int>();
    sp.clv_out = clv_out;
    sp.cat_ll_out = cat_ll_out;
    sp.tile_pad = tile_pad;
    sp.slot_stride = blocked ? (size_t)n_blocks * M.n_clv * 4 * sites_per_block : (size_t)M.n_clv * 4 * tile_pad;
    sp.blocked = blocked ? 1 : 0;
    sp.exp_out = exp_out;
    sp.n_store = n_store;
    ScopedTimer tm(ctx, &d, PH_CLV);
#ifndef PHB_HOST_EMU
    static bool attr_set[64] = {false};
    if (!attr_set[d.ordinal & 63]) {
        CU(cudaFuncSetAttribute(sweep_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)d.smem_optin));
        CU(cudaFuncSetAttribute(sweep_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)d.smem_optin));
        attr_set[d.ordinal & 63] = true;
    }
#endif
    if (clv_out) PHB_LAUNCH_SMEM(sweep_kernel<true>, dim3(grid), dim3(G.threads), G.smem_bytes, d.stream, sp, M);
    else PHB_LAUNCH_SMEM(sweep_kernel<false>, dim3(grid), dim3(G.threads), G.smem_bytes, d.stream, sp, M);
    CHECK_LAUNCH();
    if (&d == &ctx->devs[0]) ctx->stats.clv_launches++;
    return PHB_OK;
}

int launch_sweep_reduce(phb_ctx *ctx, Device &d, size_t n_partials)
{
    ScopedTimer tm(ctx, &d, PH_ROOT);
    PHB_LAUNCH(final_reduce_kernel, dim3(1), dim3(1024), d.stream, d.partials.as<double>(), (int)n_partials,
               d.total.as<double>(), d.sw_err.as<unsigned int>());
    CHECK_LAUNCH();
    return PHB_OK;
}

// ---- pre-order sweep of the marginal-ASR down pass (down_sweep.cuh) ---------------------------------
constexpr int DN_MAX_K = 5; // categories the pre-order kernel is instantiated for (GTR+I+G4 carries 5 literal ones)

int ensure_down(phb_ctx *ctx, const ModelDev &M)
{
    const TreeSched &T = ctx->tree;
    DownSched &D = ctx->down;
    if (!D.same_topology(T.root, T.child_off, T.children, T.leaf_row, T.slot, M.n_clv)) {
        for (Device &d : ctx->devs) { // the old host image may still be in flight
            cudaSetDevice(d.ordinal);
            cudaStreamSynchronize(d.stream);
            d.down_uploaded = false;
        }
        D.build(T.n, T.root, T.child_off, T.children, T.leaf_row, T.slot, M.n_clv);
        if (getenv("PHB_SWEEP_TRACE"))
            fprintf(stderr, "pre-order schedule: ops %d (LEAF %d DESCEND %d DEFER %d RESUME %d) levels %d stream %zu B\n", D.n_ops,
                    D.count_kind[DN_LEAF], D.count_kind[DN_DESCEND], D.count_kind[DN_DEFER], D.count_kind[DN_RESUME], D.n_levels,
                    D.stream_bytes);
    }
    for (Device &d : ctx->devs) {
        if (d.down_uploaded) continue;
        CU(cudaSetDevice(d.ordinal));
        CU(d.dn_ops.ensure(sizeof(SweepOp) * D.ops.size()));
        CU(d.dn_descs.ensure(sizeof(DownDesc) * D.descs.size()));
        CU(d.dn_blocks.ensure(sizeof(SweepBlock) * std::max<size_t>(D.blocks.size(), 1)));
        CU(d.dn_opoff.ensure(sizeof(uint64_t) * std::max<size_t>(D.op_off.size(), 1)));
        CU(d.dn_stream.ensure(D.stream_bytes + 16));
        CU(cudaMemcpyAsync(d.dn_ops.p, D.ops.data(), sizeof(SweepOp) * D.ops.size(), cudaMemcpyHostToDevice, d.stream));
        CU(cudaMemcpyAsync(d.dn_descs.p, D.descs.data(), sizeof(DownDesc) * D.descs.size(), cudaMemcpyHostToDevice, d.stream));
        CU(cudaMemcpyAsync(d.dn_blocks.p, D.blocks.data(), sizeof(SweepBlock) * D.blocks.size(), cudaMemcpyHostToDevice, d.stream));
        CU(cudaMemcpyAsync(d.dn_opoff.p, D.op_off.data(), sizeof(uint64_t) * D.op_off.size(), cudaMemcpyHostToDevice, d.stream));
        d.down_uploaded = true;
    }
    return PHB_OK;
}

// transposed P(t) of every edge + the op headers, written into the pre-order stream (once per evaluation)
int launch_down_transition(phb_ctx *ctx, Device &d, const ModelDev &M)
{
    const DownSched &D = ctx->down;
    ScopedTimer tm(ctx, &d, PH_TRANSITION);
    const int nthreads = (int)D.blocks.size() * M.n_clv + D.n_ops;
    PHB_LAUNCH(sweep_transition_kernel, dim3((nthreads + 127) / 128), dim3(128), d.stream, d.dn_blocks.as<SweepBlock>(),
               (int)D.blocks.size(), d.blen.as<double>(), M, d.dn_stream.as<unsigned char>(), d.dn_ops.as<SweepOp>(),
               d.dn_opoff.as<uint64_t>(), D.n_ops);
    CHECK_LAUNCH();
    return PHB_OK;
}

// One persistent pre-order launch over patterns [p0, p0 + np) whose up-pass CLVs the storing sweep left in blocked
// layout with block width S.
int launch_down_sweep(phb_ctx *ctx, Device &d, const ModelDev &M, int S, int64_t p0, int64_t np, bool want_post)
{
    const DownSched &D = ctx->down;
    const TreeSched &T = ctx->tree;
    const int K = M.n_clv;
    const int64_t n_blocks = (np + S - 1) / S;
    const size_t tile_bytes = (size_t)K * 4 * S * 8;
    const int data_off = (DN_STAGE_OPS * D.rec_bytes + 127) / 128 * 128;
    const int rows_off = data_off + (int)((tile_bytes + 127) / 128 * 128);
    const int row_pitch = (S + 15) / 16 * 16;
    const int stage_bytes = (rows_off + DN_STAGE_ROWS * row_pitch + 127) / 128 * 128;
    // stages first (each holds one node's tile: they are what hides the HBM latency), the rest is the stack
    int n_stages = 3, smem_levels = 0;
    if (const char *e = getenv("PHB_DOWN_STAGES")) n_stages = std::max(2, std::min(6, atoi(e)));
    while (n_stages > 2 && SW_BAR_BYTES + (size_t)n_stages * stage_bytes > d.smem_optin) --n_stages;
    if (SW_BAR_BYTES + (size_t)n_stages * stage_bytes > d.smem_optin)
        return ctx->fail(PHB_ERR_INVALID_MODEL_PARAMETER, "pre-order sweep: %d categories do not fit one CTA", K);
    smem_levels = (int)std::min<size_t>((d.smem_optin - SW_BAR_BYTES - (size_t)n_stages * stage_bytes) / tile_bytes, (size_t)D.n_levels);
    const int g_levels = std::max(0, D.n_levels - smem_levels);
    const size_t smem_bytes = SW_BAR_BYTES + (size_t)n_stages * stage_bytes + (size_t)smem_levels * tile_bytes;
    const int grid = (int)std::min<int64_t>(n_blocks, std::max(d.sm_count, 1));
    if (g_levels > 0) CU(d.dn_gstack.ensure((size_t)std::max(d.sm_count, 1) * g_levels * tile_bytes));
    DownParams dp;
    memset(&dp, 0, sizeof dp);
    dp.ops = d.dn_ops.as<SweepOp>();
    dp.descs = d.dn_descs.as<DownDesc>();
    dp.stream = d.dn_stream.as<unsigned char>();
    dp.n_ops = D.n_ops;
    dp.n_stage_descs = D.n_stage_descs;
    dp.rows_off = rows_off;
    dp.row_pitch = row_pitch;
    dp.S = S;
    dp.n_stages = n_stages;
    dp.stage_bytes = stage_bytes;
    dp.data_off = data_off;
    dp.smem_levels = smem_levels;
    dp.g_levels = g_levels;
    dp.gstack = d.dn_gstack.as<double>();
    dp.clv = d.clv.as<double>();
    dp.slot_stride = (size_t)n_blocks * K * 4 * S;
    dp.tips = d.tips.as<uint8_t>() + p0;
    dp.P_pad = d.P_pad;
    dp.P = np;
    dp.n_blocks = n_blocks;
    dp.cat_ll = d.cat_ll.as<double>();
    dp.tile_pad = (size_t)d.tile_pad;
    dp.site_ll = d.per_pattern.as<double>() + p0;
    dp.states = d.states.as<uint8_t>() + p0;
    dp.posteriors = want_post ? d.posteriors.as<double>() + p0 * 4 : nullptr;
    (void)T;
    ScopedTimer tm(ctx, &d, PH_DOWN);
#ifndef PHB_HOST_EMU
    static bool attr_set[64] = {false};
    if (!attr_set[d.ordinal & 63]) {
        CU(cudaFuncSetAttribute(asr_down_sweep_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)d.smem_optin));
        CU(cudaFuncSetAttribute(asr_down_sweep_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)d.smem_optin));
        CU(cudaFuncSetAttribute(asr_down_sweep_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)d.smem_optin));
        CU(cudaFuncSetAttribute(asr_down_sweep_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)d.smem_optin));
        CU(cudaFuncSetAttribute(asr_down_sweep_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)d.smem_optin));
        attr_set[d.ordinal & 63] = true;
    }
#endif
    switch (K) {
    case 1: PHB_LAUNCH_SMEM(asr_down_sweep_kernel<1>, dim3(grid), dim3(S + 32), smem_bytes, d.stream, dp, M); break;
    case 2: PHB_LAUNCH_SMEM(asr_down_sweep_kernel<2>, dim3(grid), dim3(S + 32), smem_bytes, d.stream, dp, M); break;
    case 3: PHB_LAUNCH_SMEM(asr_down_sweep_kernel<3>, dim3(grid), dim3(S + 32), smem_bytes, d.stream, dp, M); break;
    case 4: PHB_LAUNCH_SMEM(asr_down_sweep_kernel<4>, dim3(grid), dim3(S + 32), smem_bytes, d.stream, dp, M); break;
    default: PHB_LAUNCH_SMEM(asr_down_sweep_kernel<5>, dim3(grid), dim3(S + 32), smem_bytes, d.stream, dp, M); break;
    }
    CHECK_LAUNCH();
    return PHB_OK;
}

// compute_log_likelihood's pattern loop on one device: P(t) stream, one persistent sweep launch.
int run_sweep(phb_ctx *ctx, Device &d, const ModelDev &M)
{
    const SweepGeom &G = ctx->sweep_geom;
    const int64_t n_blocks = (d.P + 4 * G.Q - 1) / (4 * G.Q);
    const int n_sm = std::max(d.sm_count, 1);
    // A pass of the persistent grid over the op list costs about the same for 9-11 consumer warps, 14 % less for 5-7
    // and 31 % less for <= 3 (one, two or three warps per scheduler: profiles/r2_pass_time_vs_width.txt), so the
    // patterns left over after the last FULL pass go out as a launch of their own, just wide enough to take them in one
    // pass (10k x 1M: the 20th pass holds 10,144 of the 1M patterns; a 125k shard's third pass 20,808).
    int64_t main_blocks = n_blocks / n_sm * n_sm;
    SweepGeom tail;
    int64_t tail_P = 0, tail_blocks = 0;
    static const bool no_tail = getenv("PHB_SWEEP_NO_TAIL") != nullptr;
    if (!no_tail && main_blocks > 0 && main_blocks < n_blocks) {
        tail_P = d.P - main_blocks * 4 * G.Q;
        const int step = sweep_q_step(M.n_clv);
        int q_tail = (int)((tail_P + (int64_t)4 * n_sm - 1) / ((int64_t)4 * n_sm));
        q_tail = (q_tail + step - 1) / step * step;
        auto warps_per_scheduler = [&](int q) { return (q * M.n_clv / 32 + 1 + 3) / 4; };
        if (q_tail < G.Q && warps_per_scheduler(q_tail) < warps_per_scheduler(G.Q) && tail.plan(ctx->sweep, d.smem_optin, tail_P, n_sm, q_tail))
            tail_blocks = (tail_P + 4 * tail.Q - 1) / (4 * tail.Q);
    }
    if (tail_blocks == 0) main_blocks = n_blocks;
    CU(d.partials.ensure(sizeof(double) * (size_t)std::max<int64_t>(main_blocks + tail_blocks, 1)));
    int rc = launch_sweep_transition(ctx, d, M);
    if (rc) return rc;
    rc = launch_sweep_range(ctx, d, M, G, 0, tail_blocks ? main_blocks * 4 * G.Q : d.P, 0);
    if (rc) return rc;
    if (tail_blocks) {
        rc = launch_sweep_range(ctx, d, M, tail, main_blocks * 4 * G.Q, tail_P, (size_t)main_blocks);
        if (rc) return rc;
    }
    return launch_sweep_reduce(ctx, d, (size_t)(main_blocks + tail_blocks));
}

bool host_is_pageable(const void *p);
int h2d_from_host(phb_ctx *ctx, Device &d, uint8_t *dst, const uint8_t *src, size_t bytes, cudaStream_t stream, bool pageable);
int d2h_rows_to_pageable(phb_ctx *ctx, Device &d, uint8_t *dst, size_t dst_pitch, const uint8_t *src, size_t src_pitch, size_t width,
                         int height, cudaStream_t stream);

// One full evaluation on every device.  asr=false: compute_log_likelihood; asr=true: + down pass.
int evaluate(phb_ctx *ctx, bool asr, bool want_post, double *total, bool joint = false)
{
    if (!ctx->have_aln || !ctx->have_model || !ctx->tree.valid)
        return ctx->fail(PHB_ERR_STATE, "set_alignment, set_model and set_tree must precede an evaluation");
    const TreeSched &T = ctx->tree;
    ctx->stats.clv_launches = 0;
    ctx->stats.clv_updates = 0;
    ctx->stats.algorithmic_bytes = 0;
    if (ctx->n_patterns == 0) { // empty alignment: the pattern loop never runs (likelihood.rs:185-208)
        *total = 0.0;
        return PHB_OK;
    }
    int rc = check_leaves(ctx);
    if (rc) return rc;
    if (!ctx->tree_uploaded) {
        rc = upload_tree(ctx);
        if (rc) return rc;
    }
    const ModelDev &M = asr ? ctx->model_lit : ctx->model;
    const bool use_sweep = !asr && !(ctx->flags & PHB_FLAG_LEVEL_KERNELS) && T.n_internal > 0 && M.n_clv > 0 &&
                           M.n_cat <= 4 * M.n_clv;
    // marginal ASR: the up pass is the storing sweep (each CLV written once, never read back by the up pass)
    const bool asr_sweep = asr && !joint && !(ctx->flags & PHB_FLAG_LEVEL_KERNELS) && T.n_internal > 0 && M.n_clv > 0 &&
                           M.n_clv <= SW_MAX_CONSUMERS / 8;
    // ... and its pre-order pass the on-chip pre-order sweep (every up-pass CLV read once, nothing but states written)
    const bool asr_blocked = asr_sweep && M.n_clv <= DN_MAX_K && M.n_clv == M.n_cat && !getenv("PHB_ASR_LEVEL_DOWN");
    if (use_sweep || asr_sweep) {
        rc = ensure_sweep(ctx, M, asr_sweep);
        if (rc) return rc;
    }
    if (asr_blocked) {
        rc = ensure_down(ctx, M);
        if (rc) return rc;
    }
    Device &d0 = ctx->devs[0];
    d0.ev_used = 0;
    d0.ev_log.clear();
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
    const bool timing = ctx->flags & PHB_FLAG_TIMING;
    for (Device &d : ctx->devs) {
        const int64_t wave = asr_sweep ? (int64_t)std::max(d.sm_count, 1) * 4 * (asr_blocked ? sweep_q_asr(M.n_clv) : sweep_q_max(M.n_clv)) : 0;
        rc = plan_device(ctx, d, M, asr, want_post, use_sweep, wave);
        if (rc) return rc;
    }
    if (timing) {
        CU(cudaSetDevice(d0.ordinal));
        CU(cudaEventCreate(&ev_begin));
        CU(cudaEventCreate(&ev_end));
        CU(cudaEventRecord(ev_begin, d0.stream));
    }
    // enqueue everything on every device's stream, then join
    for (Device &d : ctx->devs) {
        CU(cudaSetDevice(d.ordinal));
        if (d.P == 0) {
            d.h_total[0] = d.h_total[1] = 0.0;
            continue;
        }
        if (!use_sweep) {
            ScopedTimer tm(ctx, &d, PH_TRANSITION);
            int nthreads = T.n * M.n_cat;
            PHB_LAUNCH(transition_kernel, dim3((nthreads + 127) / 128), dim3(128), d.stream, d.blen.as<double>(), T.n, M,
                       d.Pmat.as<double>());
            CHECK_LAUNCH();
            if (joint) {
                const size_t n_ent = (size_t)16 * T.n * M.n_cat;
                PHB_LAUNCH(log_transition_kernel, dim3((unsigned)((n_ent + 255) / 256)), dim3(256), d.stream,
                           d.Pmat.as<double>(), n_ent, d.lnP.as<double>());
                CHECK_LAUNCH();
            }
        }
        if (!d.tip_constants_valid && !joint) {
            ScopedTimer tm(ctx, &d, PH_ROOT);
            int64_t groups = d.P_pad / 4;
            PHB_LAUNCH(tip_constants_kernel, dim3((unsigned)((groups * 4 + TIPC_SITES - 1) / TIPC_SITES)), dim3(TIPC_THREADS), d.stream,
                       d.tips.as<uint8_t>(), d.P_pad, d.P_pad, d.leaf_rows.as<int32_t>(), (int)T.leaf_rows.size(), M,
                       d.leaf_ls.as<double>(), d.inv_mask.as<uint8_t>());
            CHECK_LAUNCH();
            d.tip_constants_valid = true;
            d.tip_constants_rows = T.leaf_rows;
        }
        if (use_sweep) {
            rc = run_sweep(ctx, d, M);
            if (rc) return rc;
            CU(cudaMemcpyAsync(d.h_total, d.total.p, 2 * sizeof(double), cudaMemcpyDeviceToHost, d.stream));
            continue;
        }
        size_t part_off = 0;
        SweepGeom tile_geom;
        if (asr_sweep) { // geometry for one tile's worth of patterns; partial sums are per sweep block
            size_t smem_max = d.smem_optin;
            if (!tile_geom.plan(ctx->sweep, smem_max, std::min<int64_t>(d.tile_pad, d.P), std::max(d.sm_count, 1),
                                asr_blocked ? sweep_q_asr(M.n_clv) : 0))
                return ctx->fail(PHB_ERR_INVALID_MODEL_PARAMETER, "sweep: %d CLV categories do not fit one CTA", M.n_clv);
            const int64_t bs = 4 * tile_geom.Q;
            CU(d.partials.ensure(sizeof(double) * (size_t)(d.n_tiles * ((d.tile_pad + bs - 1) / bs + 1))));
            rc = launch_sweep_transition(ctx, d, M);
            if (rc) return rc;
            if (asr_blocked) {
                rc = launch_down_transition(ctx, d, M);
                if (rc) return rc;
            }
        }
        const int64_t Ptot_out = ctx->n_patterns;
        const bool out_pageable = asr && ctx->asr_states_out && (size_t)T.n * (size_t)d.P >= bounce_min_bytes() &&
                                  host_is_pageable(ctx->asr_states_out);
        int64_t prev_off = 0;
        int prev_len = 0;
        auto ship_pageable = [&](int64_t off, int len) -> int {
            int r = d2h_rows_to_pageable(ctx, d, ctx->asr_states_out + d.p_begin + off, (size_t)Ptot_out, d.states.as<uint8_t>() + off,
                                         (size_t)d.P_pad, (size_t)len, T.n, d.copy_stream);
            if (r || !(ctx->asr_post_out && want_post)) return r;
            return d2h_rows_to_pageable(ctx, d, reinterpret_cast<uint8_t *>(ctx->asr_post_out + (d.p_begin + off) * 4),
                                        (size_t)Ptot_out * 4 * sizeof(double),
                                        reinterpret_cast<const uint8_t *>(d.posteriors.as<double>() + off * 4),
                                        (size_t)d.P_pad * 4 * sizeof(double), (size_t)len * 4 * sizeof(double), T.n, d.copy_stream);
        };
        for (int t = 0; t < d.n_tiles; ++t) {
            int64_t off = (int64_t)t * d.tile_pad;
            int len = (int)std::min<int64_t>(d.tile_pad, d.P - off);
            if (asr_sweep) {
                rc = launch_sweep_range(ctx, d, M, tile_geom, off, len, part_off, d.clv.as<double>(), d.cat_ll.as<double>(),
                                        (size_t)d.tile_pad, asr_blocked);
            } else
                rc = joint ? run_tile_joint(ctx, d, M, off, len, part_off) : run_tile_up(ctx, d, M, off, len, part_off, asr);
            if (rc) return rc;
            if (asr) {
                rc = joint ? PHB_OK : asr_blocked ? launch_down_sweep(ctx, d, M, 4 * tile_geom.Q, off, len, want_post)
                                                  : run_tile_down(ctx, d, M, off, len, want_post);
                if (rc) return rc;
                if (ctx->asr_states_out && out_pageable) {
                    // a pageable destination (a caller's Vec): the tile before this one goes home through the bounce buffers
                    // while the GPU works on this one; this tile's turn comes in the next iteration
                    if (prev_len > 0) {
                        rc = ship_pageable(prev_off, prev_len);
                        if (rc) return rc;
                    }
                    const int buf = t & 1;
                    CU(cudaEventRecord(d.ev_free[buf], d.stream));
                    CU(cudaStreamWaitEvent(d.copy_stream, d.ev_free[buf], 0));
                    prev_off = off;
                    prev_len = len;
                } else if (ctx->asr_states_out) { // ship this tile's states (and posteriors) on the copy stream
                    const int buf = t & 1;
                    CU(cudaEventRecord(d.ev_free[buf], d.stream));
                    CU(cudaStreamWaitEvent(d.copy_stream, d.ev_free[buf], 0));
                    const int64_t Ptot = ctx->n_patterns;
                    CU(cudaMemcpy2DAsync(ctx->asr_states_out + d.p_begin + off, (size_t)Ptot, d.states.as<uint8_t>() + off,
                                         (size_t)d.P_pad, (size_t)len, (size_t)T.n, cudaMemcpyDeviceToHost, d.copy_stream));
                    if (ctx->asr_post_out && want_post)
                        CU(cudaMemcpy2DAsync(ctx->asr_post_out + (d.p_begin + off) * 4, (size_t)Ptot * 4 * sizeof(double),
                                             d.posteriors.as<double>() + off * 4, (size_t)d.P_pad * 4 * sizeof(double),
                                             (size_t)len * 4 * sizeof(double), (size_t)T.n, cudaMemcpyDeviceToHost,
                                             d.copy_stream));
                }
            }
            if (asr_sweep) part_off += (size_t)((len + 4 * tile_geom.Q - 1) / (4 * tile_geom.Q));
            else part_off += (size_t)((len + ROOT_THREADS - 1) / ROOT_THREADS);
        }
        if (prev_len > 0) { // the last tile's outputs
            rc = ship_pageable(prev_off, prev_len);
            if (rc) return rc;
        }
        {
            ScopedTimer tm(ctx, &d, PH_ROOT);
            PHB_LAUNCH(final_reduce_kernel, dim3(1), dim3(1024), d.stream, d.partials.as<double>(), (int)part_off,
                       d.total.as<double>(), asr_sweep ? d.sw_err.as<unsigned int>() : (unsigned int *)nullptr);
            CHECK_LAUNCH();
        }
        CU(cudaMemcpyAsync(d.h_total, d.total.p, 2 * sizeof(double), cudaMemcpyDeviceToHost, d.stream));
    }
    if (timing) {
        CU(cudaSetDevice(d0.ordinal));
        CU(cudaEventRecord(ev_end, d0.stream));
    }
    double sum = 0.0;
    for (Device &d : ctx->devs) { // rank-ordered host sum: deterministic
        CU(cudaSetDevice(d.ordinal));
        CU(cudaStreamSynchronize(d.stream));
        if (asr && ctx->asr_states_out) CU(cudaStreamSynchronize(d.copy_stream));
        sum += d.h_total[0];
        if (d.h_total[1] != 0.0)
            return ctx->fail(PHB_ERR_NUMERICAL_INSTABILITY, "fused sweep: unknown op code in the schedule (device %d)", d.ordinal);
    }
    *total = sum;
    // stats
    phb_stats &S = ctx->stats;
    S.n_patterns = ctx->n_patterns;
    S.tile_patterns = d0.tile_pad;
    S.n_tiles = d0.n_tiles;
    S.n_levels = T.n_levels;
    S.n_internal = T.n_internal;
    S.n_clv_categories = M.n_clv;
    S.kernel_launches = ctx->launches;
    S.clv_updates = (int64_t)T.n_internal * ctx->n_patterns;
    S.algorithmic_bytes = (32.0 * M.n_clv * T.sum_one_plus_i + T.sum_t) * (double)ctx->n_patterns;
    S.sweep_ops = (use_sweep || asr_sweep) ? ctx->sweep.n_ops : 0;
    S.sweep_stack_levels = (use_sweep || asr_sweep) ? ctx->sweep.n_levels : 0;
    S.sweep_smem_levels = (use_sweep || asr_sweep) ? ctx->sweep_geom.smem_levels : 0;
    S.sweep_stages = (use_sweep || asr_sweep) ? ctx->sweep_geom.n_stages : 0;
    S.sweep_block_sites = (use_sweep || asr_sweep) ? 4 * ctx->sweep_geom.Q : 0;
    S.sweep_stream_bytes = (use_sweep || asr_sweep) ? (int64_t)ctx->sweep.stream_bytes : 0;
    S.sweep_guards = (use_sweep || asr_sweep) ? ctx->sweep.n_guards : 0;
    S.sweep_normouts = (use_sweep || asr_sweep) ? ctx->sweep.n_normout : 0;
    S.sweep_fp64_instr = (use_sweep || asr_sweep) ? ctx->sweep.fp64_instr : 0;
    S.sweep_flops = (use_sweep || asr_sweep) ? ctx->sweep.flops : 0;
    if (timing) {
        S.ms_transition = S.ms_clv = S.ms_root = S.ms_down = 0;
        for (auto &pr : d0.ev_log) {
            float ms = 0;
            cudaEventElapsedTime(&ms, d0.ev_pool[pr.second].a, d0.ev_pool[pr.second].b);
            (pr.first == PH_TRANSITION ? S.ms_transition : pr.first == PH_CLV ? S.ms_clv : pr.first == PH_ROOT ? S.ms_root : S.ms_down) += ms;
        }
        float ms = 0;
        cudaEventElapsedTime(&ms, ev_begin, ev_end);
        S.ms_total = ms;
        cudaEventDestroy(ev_begin);
        cudaEventDestroy(ev_end);
    }
    return PHB_OK;
}


// ---- path-only NNI rescoring (nni.cuh) -----------------------------------------------------------
// Host side: turns each valid move into a walk v -> root over the base tree, caches the base tree's
// CLVs tile by tile with the level kernels, scores all walks of a tile with one launch per category
// count, and sums the per-tile results per neighbour.  Only strictly binary trees take this path.
bool nni_path_eligible(const phb_ctx *ctx)
{
    const TreeSched &T = ctx->tree;
    const ModelDev &M = ctx->model;
    return !(ctx->flags & PHB_FLAG_NNI_FULL) && T.gops.empty() && T.n_internal > 0 && M.n_clv >= 1 && M.n_clv <= 4;
}

int score_nni_paths(phb_ctx *ctx, int n_moves, const int32_t *node_id, const uint8_t *left_ch, double *totals)
{
    const TreeSched &T = ctx->tree;
    const ModelDev &M = ctx->model;
    int rc = check_leaves(ctx);
    if (rc) return rc;
    if (!ctx->tree_uploaded) {
        rc = upload_tree(ctx);
        if (rc) return rc;
    }
    // pre-order rank of every node: neighbours are scored in DFS order of v
    std::vector<int> rank(T.n, 0);
    {
        std::vector<int> st(1, T.root);
        int r = 0;
        while (!st.empty()) {
            int v = st.back();
            st.pop_back();
            rank[v] = r++;
            for (int k = T.child_off[v + 1] - 1; k >= T.child_off[v]; --k) st.push_back(T.children[k]);
        }
    }
    std::vector<int> order;
    order.reserve(n_moves);
    for (int m = 0; m < n_moves; ++m) {
        const int v = node_id[m];
        totals[m] = NAN; // leaf / root / unknown id: NNI::nni errors (src/tree.rs:1213-1242)
        if (v < 0 || v >= T.n || v == T.root || T.child_off[v + 1] - T.child_off[v] != 2) continue;
        totals[m] = 0.0;
        order.push_back(m);
    }
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return rank[node_id[a]] < rank[node_id[b]]; });
    std::vector<NniMove> moves;
    std::vector<NniStep> steps;
    moves.reserve(order.size());
    double read_bytes_per_pattern = 0;
    auto src_of = [&](int c) { return T.slot[c] < 0 ? T.leaf_row[c] : T.slot[c]; };
    auto cost_of = [&](int c) { return T.slot[c] < 0 ? 1.0 : 36.0 * M.n_clv; };
    auto other_child = [&](int par, int not_this) {
        for (int k = T.child_off[par]; k < T.child_off[par + 1]; ++k)
            if (T.children[k] != not_this) return T.children[k];
        return -1;
    };
    for (int m : order) {
        const int v = node_id[m];
        const int p = T.parent[v];
        const int k0 = T.children[T.child_off[v]], k1 = T.children[T.child_off[v] + 1];
        const int c = left_ch[m] ? k1 : k0; // `left_ch as usize` indexes the child list: true => children[1]
        const int a = left_ch[m] ? k0 : k1;
        const int s = other_child(p, v);    // first child of p that is not v
        NniMove mv;
        memset(&mv, 0, sizeof mv);
        mv.out = m;
        mv.a_src = src_of(a);
        mv.a_edge = a;
        mv.a_is_tip = T.slot[a] < 0;
        mv.step_begin = (int)steps.size();
        read_bytes_per_pattern += cost_of(a);
        auto push = [&](int sib, int up_node) {
            NniStep st;
            st.sib_src = src_of(sib);
            st.sib_edge = sib;
            st.sib_is_tip = T.slot[sib] < 0;
            st.up_edge = up_node == T.root ? -1 : up_node;
            steps.push_back(st);
            read_bytes_per_pattern += cost_of(sib);
        };
        push(s, v);
        push(c, p);
        for (int g = p; g != T.root; g = T.parent[g]) push(other_child(T.parent[g], g), T.parent[g]);
        mv.step_end = (int)steps.size();
        moves.push_back(mv);
    }
    // edge-local jobs (upper partials): the two neighbours of one edge (p -> v) share every load
#ifdef PHB_WITH_WALK
    const bool use_upper = !(ctx->flags & PHB_FLAG_NNI_WALK);
#else
    const bool use_upper = true; // the walk kernel is not part of the product build: PHB_FLAG_NNI_WALK is ignored
#endif
    std::vector<NniEdgeJob> jobs;
    std::vector<int32_t> job_rows; // totals index of partial row 2*job + {0,1}
    if (use_upper) {
        read_bytes_per_pattern = 0;
        std::vector<int> job_of(T.n, -1);
        for (int m : order) {
            const int v = node_id[m];
            if (job_of[v] < 0) {
                const int p = T.parent[v];
                const int k0 = T.children[T.child_off[v]], k1 = T.children[T.child_off[v] + 1];
                const int sn = other_child(p, v);
                NniEdgeJob jb;
                memset(&jb, 0, sizeof jb);
                jb.k0_edge = k0; jb.k0_is_tip = T.slot[k0] < 0; // the *_src rows are filled in once the rows are known
                jb.k1_edge = k1; jb.k1_is_tip = T.slot[k1] < 0;
                jb.s_edge = sn; jb.s_is_tip = T.slot[sn] < 0;
                jb.v_edge = v;
                jb.p_upper_slot = p == T.root ? -1 : p; // dense id for now
                jb.out0 = jb.out1 = -1;
                job_of[v] = (int)jobs.size();
                jobs.push_back(jb);
                job_rows.push_back(-1);
                job_rows.push_back(-1);
                read_bytes_per_pattern += cost_of(k0) + cost_of(k1) + cost_of(sn) + (p == T.root ? 0.0 : 36.0 * M.n_clv);
            }
            // a repeated (v, left_ch) keeps its own output row: later duplicates are filled in on the host
            const int j = job_of[v], w = left_ch[m] ? 1 : 0;
            if (job_rows[2 * j + w] < 0) {
                job_rows[2 * j + w] = m;
                (w ? jobs[j].out1 : jobs[j].out0) = m;
            }
        }
    }
    // upper partials are read at the upper end p of a scored edge and, recursively, by p's own partials:
    // only the ancestors of the scored edges get them
    std::vector<uint8_t> need_upper(std::max(T.n, 1), 0);
    if (use_upper)
        for (int m : order)
            for (int g = T.parent[node_id[m]]; g != T.root && g >= 0 && !need_upper[g]; g = T.parent[g]) need_upper[g] = 1;
    // Rows.  lrow[v]: row of v's lower partial in the cache (only the partials some scored edge or some upper partial
    // reads get one when the cache is written by the storing sweep; every internal node when the level kernels write
    // it); urow[v]: row of v's upper partial (an internal node with an internal child, on a path from a scored edge to
    // the root).
    std::vector<int> lrow(T.n, -1), urow(T.n, -1);
    int n_store = 0, n_upper = 0;
    const bool level_cache_env = getenv("PHB_NNI_LEVEL_CACHE") != nullptr; // cross-check: every node, level kernels
    const bool sweep_cache = use_upper && !level_cache_env && sweep_q_max(M.n_clv) >= sweep_q_step(M.n_clv) && T.n_internal < 0x7fff0000;
    std::vector<DownOp> ndops;
    std::vector<DownChild> ndkids;
    std::vector<int> ndlevel_off;
    if (!sweep_cache) {
        for (int v = 0; v < T.n; ++v) lrow[v] = T.slot[v];
        n_store = T.n_internal;
    }
    if (use_upper) {
        auto want_lower = [&](int c) {
            if (T.slot[c] >= 0 && lrow[c] < 0) lrow[c] = n_store++;
        };
        for (const NniEdgeJob &jb : jobs) {
            want_lower(jb.k0_edge);
            want_lower(jb.k1_edge);
            want_lower(jb.s_edge);
        }
        for (const DownChild &dc : T.dkids)
            if (dc.pad && need_upper[dc.node]) urow[dc.node] = n_upper++;
        for (const DownOp &o : T.dops) { // the sibling of a child whose upper partial is wanted is read through P
            if (o.child_end - o.child_begin != 2) continue;
            const DownChild &c0 = T.dkids[o.child_begin], &c1 = T.dkids[o.child_begin + 1];
            if (urow[c0.node] >= 0) want_lower(c1.node);
            if (urow[c1.node] >= 0) want_lower(c0.node);
        }
        ndkids = T.dkids;
        for (DownChild &dc : ndkids) {
            dc.src = dc.is_tip ? dc.src : lrow[dc.node];
            dc.pad = urow[dc.node] + 1;
        }
        // per depth, only the parents with a child whose upper partial is wanted: the launch holds no idle blocks
        ndlevel_off.assign(1, 0);
        for (int dl = 0; dl <= T.max_depth; ++dl) {
            for (int k = T.dlevel_off[dl]; k < T.dlevel_off[dl + 1]; ++k) {
                DownOp o = T.dops[k];
                bool wanted = false;
                for (int c = o.child_begin; c < o.child_end; ++c) wanted = wanted || ndkids[c].pad != 0;
                if (!wanted) continue;
                o.parent_slot = o.pad == T.root ? -1 : urow[o.pad];
                ndops.push_back(o);
            }
            ndlevel_off.push_back((int)ndops.size());
        }
        for (NniEdgeJob &jb : jobs) {
            jb.k0_src = jb.k0_is_tip ? T.leaf_row[jb.k0_edge] : lrow[jb.k0_edge];
            jb.k1_src = jb.k1_is_tip ? T.leaf_row[jb.k1_edge] : lrow[jb.k1_edge];
            jb.s_src = jb.s_is_tip ? T.leaf_row[jb.s_edge] : lrow[jb.s_edge];
            if (jb.p_upper_slot >= 0) jb.p_upper_slot = urow[jb.p_upper_slot];
        }
    }
    // The storing sweep as the cache pass: its own schedule (cherries whose CLV is wanted stay explicit), swapped into
    // ctx->sweep while the pass runs so that ensure_sweep / launch_sweep_* work on it.
    struct SweepSwap {
        phb_ctx *c;
        bool on;
        void flip()
        {
            std::swap(c->sweep, c->nni_sweep);
            std::swap(c->sweep_geom, c->nni_geom);
            std::swap(c->sweep_guards_valid, c->nni_guards_valid);
            for (Device &d : c->devs) d.sweep_uploaded = false; // the device tables belong to the other schedule now
        }
        ~SweepSwap()
        {
            if (on) flip();
        }
    } swap_guard{ctx, false};
    std::vector<int> pred_row; // per cache row: the row holding the running exponent count before its subtree began
    int n_rows = n_store;
    if (sweep_cache) {
        std::vector<char> keep(T.n, 0);
        for (int v = 0; v < T.n; ++v) keep[v] = lrow[v] >= 0;
        swap_guard.flip();
        swap_guard.on = true;
        rc = ensure_sweep(ctx, M, false, &keep);
        if (rc) return rc;
        SweepSched &S = ctx->sweep;
        std::vector<int> row_of(T.n, -1);
        for (int v = 0; v < T.n; ++v) row_of[v] = lrow[v];
        pred_row.assign(n_store, -1);
        for (int v = 0; v < T.n; ++v) {
            if (lrow[v] < 0) continue;
            const int fo = S.first_op[v];
            if (fo < 0 || S.done_op[v] < 0) return ctx->fail(PHB_ERR_NUMERICAL_INSTABILITY, "NNI cache: node %d has no op of its own", v);
            if (fo == 0) continue;
            const int y = S.op_node[fo - 1];
            if (y < 0) return ctx->fail(PHB_ERR_NUMERICAL_INSTABILITY, "NNI cache: op %d completes no node", fo - 1);
            if (row_of[y] < 0) row_of[y] = n_rows++;
            pred_row[lrow[v]] = row_of[y];
        }
        for (Device &d : ctx->devs) { // the op table is about to change under a copy that may still be in flight
            cudaSetDevice(d.ordinal);
            cudaStreamSynchronize(d.stream);
        }
        for (size_t k = 0; k < S.op_node.size(); ++k) S.ops[k].w2 = S.op_node[k] >= 0 && row_of[S.op_node[k]] >= 0 ? (uint32_t)row_of[S.op_node[k]] : SW_NO_ROW;
        for (Device &d : ctx->devs) {
            if (d.P == 0) continue;
            CU(cudaSetDevice(d.ordinal));
            CU(cudaMemcpyAsync(d.sw_ops.p, S.ops.data(), sizeof(SweepOp) * S.ops.size(), cudaMemcpyHostToDevice, d.stream));
        }
    }
    phb_stats &S = ctx->stats;
    S.nni_path_moves = (int)moves.size();
    S.nni_path_steps = use_upper ? (int64_t)jobs.size() * 3 : (int64_t)steps.size();
    S.nni_read_bytes = read_bytes_per_pattern * (double)ctx->n_patterns;
    S.nni_tiles = 0;
    S.nni_cache_bytes = 0;
    if (moves.empty() || ctx->n_patterns == 0) return PHB_OK;
    const int nm = (int)moves.size();
    static const int upper_minb = [] { const char *e = getenv("PHB_UPPER_MINB"); return e ? atoi(e) : 3; }();
    static const bool nni_minb3 = [] { const char *e = getenv("PHB_NNI_MINB"); return e && atoi(e) == 3; }();
    std::vector<std::vector<double>> dev_totals(ctx->devs.size());
    for (size_t di = 0; di < ctx->devs.size(); ++di) {
        Device &d = ctx->devs[di];
        if (d.P == 0) continue;
        CU(cudaSetDevice(d.ordinal));
        CU(d.Pmat.ensure(sizeof(double) * 16 * (size_t)T.n * M.n_cat));
        CU(d.leaf_ls.ensure(sizeof(double) * (size_t)d.P_pad));
        CU(d.inv_mask.ensure((size_t)d.P_pad));
        CU(d.nni_moves.ensure(sizeof(NniMove) * moves.size()));
        CU(d.nni_steps.ensure(sizeof(NniStep) * steps.size()));
        CU(d.nni_totals.ensure(sizeof(double) * (size_t)n_moves));
        CU(cudaMemcpyAsync(d.nni_moves.p, moves.data(), sizeof(NniMove) * moves.size(), cudaMemcpyHostToDevice, d.stream));
        CU(cudaMemcpyAsync(d.nni_steps.p, steps.data(), sizeof(NniStep) * steps.size(), cudaMemcpyHostToDevice, d.stream));
        CU(cudaMemsetAsync(d.nni_totals.p, 0, sizeof(double) * (size_t)n_moves, d.stream));
        // tile the pattern axis so that one tile's CLVs + subtree exponents fit the budget
        const size_t row_bytes = (size_t)M.n_clv * (4 * sizeof(double) + sizeof(int));
        const size_t per_site = ((size_t)n_store + (size_t)n_upper) * row_bytes + (size_t)(sweep_cache ? n_rows : 0) * M.n_clv * sizeof(int);
        size_t budget = ctx->clv_budget;
        if (!budget) { // 30 % of the card at most: the cache is scratch, and tiles cost no time
            size_t free_b = 0, total_b = 0;
            CU(cudaMemGetInfo(&free_b, &total_b));
            const size_t avail = free_b + d.clv.cap + d.nni_exps.cap + d.nni_upper.cap + d.nni_uexps.cap + d.nni_cnt.cap;
            budget = std::min((size_t)((double)avail * 0.85), total_b / 10 * 3);
        }
        int64_t tile = d.P_pad;
        if (per_site * (size_t)d.P_pad > budget) {
            tile = (int64_t)(budget / per_site) / SITES_PER_BLOCK * SITES_PER_BLOCK;
            if (tile < SITES_PER_BLOCK)
                return ctx->fail(PHB_ERR_OOM, "CLV budget %zu B cannot hold one %d-site tile of %d cached nodes", budget,
                                 SITES_PER_BLOCK, T.n_internal);
        }
        const int n_tiles = (int)((d.P + tile - 1) / tile);
        const size_t tile_pad = (size_t)tile;
        const size_t slot_stride = (size_t)M.n_clv * 4 * tile_pad;
        CU(d.clv.ensure(sizeof(double) * slot_stride * (size_t)std::max(n_store, 1)));
        CU(d.nni_exps.ensure(sizeof(int) * (size_t)M.n_clv * tile_pad * (size_t)std::max(n_store, 1)));
        SweepGeom tile_geom;
        if (use_upper) {
            CU(d.nni_upper.ensure(sizeof(double) * slot_stride * (size_t)std::max(n_upper, 1)));
            CU(d.nni_uexps.ensure(sizeof(int) * (size_t)M.n_clv * tile_pad * (size_t)std::max(n_upper, 1)));
            CU(d.nni_jobs.ensure(sizeof(NniEdgeJob) * jobs.size()));
            CU(d.nni_rows.ensure(sizeof(int32_t) * job_rows.size()));
            CU(d.nni_dops.ensure(sizeof(DownOp) * std::max<size_t>(ndops.size(), 1)));
            CU(d.nni_dkids.ensure(sizeof(DownChild) * std::max<size_t>(ndkids.size(), 1)));
            CU(cudaMemcpyAsync(d.nni_dops.p, ndops.data(), sizeof(DownOp) * ndops.size(), cudaMemcpyHostToDevice, d.stream));
            CU(cudaMemcpyAsync(d.nni_dkids.p, ndkids.data(), sizeof(DownChild) * ndkids.size(), cudaMemcpyHostToDevice, d.stream));
            CU(cudaMemcpyAsync(d.nni_jobs.p, jobs.data(), sizeof(NniEdgeJob) * jobs.size(), cudaMemcpyHostToDevice, d.stream));
            CU(cudaMemcpyAsync(d.nni_rows.p, job_rows.data(), sizeof(int32_t) * job_rows.size(), cudaMemcpyHostToDevice, d.stream));
        }
        if (sweep_cache) {
            CU(d.nni_cnt.ensure(sizeof(int) * (size_t)M.n_clv * tile_pad * (size_t)n_rows));
            CU(d.nni_pred.ensure(sizeof(int) * (size_t)std::max(n_store, 1)));
            CU(cudaMemcpyAsync(d.nni_pred.p, pred_row.data(), sizeof(int) * pred_row.size(), cudaMemcpyHostToDevice, d.stream));
            if (!tile_geom.plan(ctx->sweep, d.smem_optin, std::min<int64_t>(tile, d.P), std::max(d.sm_count, 1)))
                return ctx->fail(PHB_ERR_INVALID_MODEL_PARAMETER, "sweep: %d CLV categories do not fit one CTA", M.n_clv);
            const int64_t bs = 4 * tile_geom.Q;
            CU(d.partials.ensure(sizeof(double) * (size_t)((tile + bs - 1) / bs + 1)));
            CU(d.per_pattern.ensure(sizeof(double) * (size_t)d.P_pad));
            CU(d.cat_ll.ensure(sizeof(double) * (size_t)M.n_cat * tile_pad));
            CU(d.total.ensure(2 * sizeof(double)));
        }
        if (&d == &ctx->devs[0]) {
            d.ev_used = 0;
            d.ev_log.clear();
        }
        const int nni_bs = nni_block_sites(M.n_clv);
        const int max_sb = (int)((std::min<int64_t>(tile, d.P) + nni_bs - 1) / nni_bs);
        CU(d.nni_partials.ensure(sizeof(double) * (size_t)std::max<size_t>(nm, 2 * jobs.size()) * (size_t)max_sb));
        if (&d == &ctx->devs[0]) {
            S.nni_tiles = n_tiles;
            S.nni_cache_bytes = (double)per_site * (double)std::min<int64_t>(tile, d.P); // resident: one tile at a time
        }
        {
            ScopedTimer tm(ctx, &d, PH_TRANSITION);
            int nthreads = T.n * M.n_cat;
            PHB_LAUNCH(transition_kernel, dim3((nthreads + 127) / 128), dim3(128), d.stream, d.blen.as<double>(), T.n, M,
                       d.Pmat.as<double>());
            CHECK_LAUNCH();
        }
        if (sweep_cache) {
            rc = launch_sweep_transition(ctx, d, M);
            if (rc) return rc;
        }
        if (!d.tip_constants_valid) {
            int64_t groups = d.P_pad / 4;
            PHB_LAUNCH(tip_constants_kernel, dim3((unsigned)((groups * 4 + TIPC_SITES - 1) / TIPC_SITES)), dim3(TIPC_THREADS), d.stream,
                       d.tips.as<uint8_t>(), d.P_pad, d.P_pad, d.leaf_rows.as<int32_t>(), (int)T.leaf_rows.size(), M,
                       d.leaf_ls.as<double>(), d.inv_mask.as<uint8_t>());
            CHECK_LAUNCH();
            d.tip_constants_valid = true;
            d.tip_constants_rows = T.leaf_rows;
        }
        for (int t = 0; t < n_tiles; ++t) {
            const int64_t off = (int64_t)t * tile;
            const int len = (int)std::min<int64_t>(tile, d.P - off);
            const int bpr = (len + SITES_PER_BLOCK - 1) / SITES_PER_BLOCK;
            if (sweep_cache) { // the base tree's partials of this tile: one storing sweep + the exponent differences
                rc = launch_sweep_range(ctx, d, M, tile_geom, off, len, 0, d.clv.as<double>(), d.cat_ll.as<double>(), tile_pad, false,
                                        d.nni_cnt.as<int>(), (uint32_t)n_store);
                if (rc) return rc;
                if (n_store > 0) {
                    ScopedTimer tm(ctx, &d, PH_CLV);
                    const size_t row_elems = (size_t)M.n_clv * tile_pad;
                    PHB_LAUNCH(nni_exp_rows_kernel, dim3((unsigned)((row_elems + 255) / 256)), dim3(256), d.stream, d.nni_cnt.as<int>(),
                               d.nni_pred.as<int>(), n_store, row_elems, d.nni_exps.as<int>());
                    CHECK_LAUNCH();
                }
            }
            for (int l = 1; l <= T.n_levels && !sweep_cache; ++l) { // ... or level by level, every node
                const int nb = T.level_off[l] - T.level_off[l - 1];
                if (nb <= 0) continue;
                ScopedTimer tm(ctx, &d, PH_CLV);
                PHB_LAUNCH(clv_cache_level_kernel, dim3((unsigned)((int64_t)nb * M.n_clv * bpr)), dim3(CLV_THREADS), d.stream,
                           d.ops.as<NodeOp>() + T.level_off[l - 1], nb, M, d.Pmat.as<double>(), d.clv.as<double>(),
                           slot_stride, tile_pad, len, bpr, d.tips.as<uint8_t>(), d.P_pad, off, d.nni_exps.as<int>());
                CHECK_LAUNCH();
            }
            const int n_sb = (len + nni_bs - 1) / nni_bs;
            if (use_upper) {
                for (int dl = 0; dl <= T.max_depth; ++dl) { // upper partials, root downwards
                    const int nb = ndlevel_off[dl + 1] - ndlevel_off[dl];
                    if (nb <= 0) continue;
                    ScopedTimer tm(ctx, &d, PH_UPPER);
#define PHB_UPPER_LAUNCH(MINB)                                                                                         \
    PHB_LAUNCH(upper_level_kernel<MINB>, dim3((unsigned)((int64_t)nb * M.n_clv * bpr)), dim3(CLV_THREADS), d.stream,   \
               d.nni_dops.as<DownOp>() + ndlevel_off[dl], nb, d.nni_dkids.as<DownChild>(), T.root, M, d.Pmat.as<double>(), \
               d.clv.as<double>(), d.nni_upper.as<double>(), slot_stride, tile_pad, len, bpr, d.nni_exps.as<int>(),    \
               d.nni_uexps.as<int>(), d.tips.as<uint8_t>(), d.P_pad, off)
                    if (upper_minb == 2) PHB_UPPER_LAUNCH(2);
                    else if (upper_minb == 4) PHB_UPPER_LAUNCH(4);
                    else PHB_UPPER_LAUNCH(3);
#undef PHB_UPPER_LAUNCH
                    CHECK_LAUNCH();
                }
                const int nj = (int)jobs.size();
                const int jchunk = std::max(1, (int)std::min<int64_t>(nj, (int64_t)0x7fffffff / std::max(n_sb, 1)));
                for (int j0 = 0; j0 < nj; j0 += jchunk) {
                    const int cj = std::min(jchunk, nj - j0);
                    ScopedTimer tm(ctx, &d, PH_DOWN);
#define PHB_NNI_EDGE(NC)                                                                                               \
    PHB_LAUNCH(nni_edge_kernel<NC>, dim3((unsigned)((int64_t)n_sb * cj)), dim3(nni_threads(NC)), d.stream,             \
               d.nni_jobs.as<NniEdgeJob>() + j0, cj, M, d.Pmat.as<double>(), d.clv.as<double>(), d.nni_upper.as<double>(), \
               slot_stride, tile_pad, len, d.nni_exps.as<int>(), d.nni_uexps.as<int>(), d.tips.as<uint8_t>(), d.P_pad, off, \
               d.leaf_ls.as<double>() + off, d.inv_mask.as<uint8_t>() + off, d.mult.as<double>() + off,                \
               d.nni_partials.as<double>(), n_sb)
                    switch (M.n_clv) {
                    case 1: PHB_NNI_EDGE(1); break;
                    case 2: PHB_NNI_EDGE(2); break;
                    case 3: PHB_NNI_EDGE(3); break;
                    default: PHB_NNI_EDGE(4); break;
                    }
#undef PHB_NNI_EDGE
                    CHECK_LAUNCH();
                    PHB_LAUNCH(nni_reduce_rows_kernel, dim3(2 * cj), dim3(ROOT_THREADS), d.stream, d.nni_rows.as<int32_t>() + 2 * j0,
                               d.nni_partials.as<double>(), n_sb, d.nni_totals.as<double>());
                    CHECK_LAUNCH();
                }
                continue;
            }
#ifdef PHB_WITH_WALK
            const int chunk = std::max(1, (int)std::min<int64_t>(nm, (int64_t)0x7fffffff / std::max(n_sb, 1)));
            for (int m0 = 0; m0 < nm; m0 += chunk) { // (the grid holds n_sb * moves-in-chunk CTAs)
                const int cm = std::min(chunk, nm - m0);
                ScopedTimer tm(ctx, &d, PH_DOWN);
#define PHB_NNI_LAUNCH(NC, MINB)                                                                                       \
    PHB_LAUNCH((nni_score_kernel<NC, MINB>), dim3((unsigned)((int64_t)n_sb * cm)), dim3(nni_threads(NC)), d.stream,    \
               d.nni_moves.as<NniMove>() + m0, cm, d.nni_steps.as<NniStep>(), M, d.Pmat.as<double>(), d.clv.as<double>(), \
               slot_stride, tile_pad, len, d.nni_exps.as<int>(), d.tips.as<uint8_t>(), d.P_pad, off,                   \
               d.leaf_ls.as<double>() + off, d.inv_mask.as<uint8_t>() + off, d.mult.as<double>() + off,                \
               d.nni_partials.as<double>(), n_sb)
                switch (M.n_clv * 2 + (nni_minb3 ? 1 : 0)) {
                case 2: PHB_NNI_LAUNCH(1, 2); break;
                case 3: PHB_NNI_LAUNCH(1, 3); break;
                case 4: PHB_NNI_LAUNCH(2, 2); break;
                case 5: PHB_NNI_LAUNCH(2, 3); break;
                case 6: PHB_NNI_LAUNCH(3, 2); break;
                case 7: PHB_NNI_LAUNCH(3, 3); break;
                case 8: PHB_NNI_LAUNCH(4, 2); break;
                default: PHB_NNI_LAUNCH(4, 3); break;
                }
#undef PHB_NNI_LAUNCH
                CHECK_LAUNCH();
                PHB_LAUNCH(nni_reduce_kernel, dim3(cm), dim3(ROOT_THREADS), d.stream, d.nni_moves.as<NniMove>() + m0,
                           d.nni_partials.as<double>(), n_sb, d.nni_totals.as<double>());
                CHECK_LAUNCH();
            }
#endif
        }
        dev_totals[di].resize(n_moves);
        CU(cudaMemcpyAsync(dev_totals[di].data(), d.nni_totals.p, sizeof(double) * (size_t)n_moves, cudaMemcpyDeviceToHost,
                           d.stream));
    }
    for (size_t di = 0; di < ctx->devs.size(); ++di) { // rank-ordered host sum: deterministic
        Device &d = ctx->devs[di];
        if (d.P == 0) continue;
        CU(cudaSetDevice(d.ordinal));
        CU(cudaStreamSynchronize(d.stream));
        for (const NniMove &mv : moves) totals[mv.out] += dev_totals[di][mv.out];
    }
    if (use_upper) { // a (v, left_ch) asked for more than once was scored once
        std::vector<int> first(2 * (size_t)T.n, -1);
        for (int m : order) {
            int &f = first[2 * (size_t)node_id[m] + (left_ch[m] ? 1 : 0)];
            if (f < 0) f = m;
            else totals[m] = totals[f];
        }
    }
    S.kernel_launches = ctx->launches;
    if (ctx->flags & PHB_FLAG_TIMING) { // per-phase device time of the last call (device 0)
        Device &d0 = ctx->devs[0];
        S.ms_nni_cache = S.ms_nni_upper = S.ms_nni_edge = 0;
        for (auto &pr : d0.ev_log) {
            float ms = 0;
            cudaEventElapsedTime(&ms, d0.ev_pool[pr.second].a, d0.ev_pool[pr.second].b);
            if (pr.first == PH_CLV) S.ms_nni_cache += ms;
            else if (pr.first == PH_UPPER) S.ms_nni_upper += ms;
            else if (pr.first == PH_DOWN) S.ms_nni_edge += ms;
        }
    }
    return PHB_OK;
}

} // namespace

#include "aa_engine.cuh"

// =================================================================================================
// C ABI
// =================================================================================================
extern "C" {

const char *phb_version(void) { return "phylo_b200 0.1.0 (sm_100a)"; }

const char *phb_last_error(phb_ctx *ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int32_t phb_ctx_create(const int32_t *devices, int32_t n_dev, uint64_t flags, phb_ctx **out)
{
    if (!out) return PHB_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) return PHB_ERR_CUDA;
    std::vector<int> ords;
    if (!devices || n_dev <= 0) ords.push_back(0);
    else ords.assign(devices, devices + n_dev);
    phb_ctx *ctx = new phb_ctx();
    ctx->flags = flags;
    ctx->is_aa = (flags & PHB_FLAG_AMINO_ACID) != 0;
    if (ctx->is_aa && ords.size() != 1) { // the 20-state engines drive one device
        delete ctx;
        return PHB_ERR_INVALID_ARGUMENT;
    }
    memset(&ctx->stats, 0, sizeof ctx->stats);
    memset(&ctx->model, 0, sizeof ctx->model);
    for (int o : ords) {
        if (o < 0 || o >= count) {
            delete ctx;
            return PHB_ERR_INVALID_ARGUMENT;
        }
        int major = 0;
        cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, o);
        if (major != 10) { // the kernels exist only as sm_100a SASS; there is no other path
            delete ctx;
            return PHB_ERR_CUDA;
        }
    }
    ctx->devs.resize(ords.size());
    for (size_t i = 0; i < ords.size(); ++i) {
        Device &d = ctx->devs[i];
        d.ordinal = ords[i];
        cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, d.ordinal);
        {
            int v = 0;
            cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, d.ordinal);
            d.smem_optin = (size_t)v;
        }
        if (cudaSetDevice(d.ordinal) != cudaSuccess || cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking) != cudaSuccess ||
            cudaStreamCreateWithFlags(&d.copy_stream, cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreate(&d.ev_copied[0]) != cudaSuccess || cudaEventCreate(&d.ev_copied[1]) != cudaSuccess ||
            cudaEventCreate(&d.ev_free[0]) != cudaSuccess || cudaEventCreate(&d.ev_free[1]) != cudaSuccess ||
            cudaHostAlloc((void **)&d.h_total, 2 * sizeof(double), cudaHostAllocDefault) != cudaSuccess) {
            phb_ctx_destroy(ctx);
            return PHB_ERR_CUDA;
        }
    }
    *out = ctx;
    return PHB_OK;
}

void phb_ctx_destroy(phb_ctx *ctx)
{
    if (!ctx) return;
    for (Device &d : ctx->devs) {
        cudaSetDevice(d.ordinal);
        if (d.copy_stream) cudaStreamSynchronize(d.copy_stream);
        if (d.stream) cudaStreamSynchronize(d.stream);
        DevBuf *bufs[] = {&d.tips, &d.mult, &d.leaf_ls, &d.inv_mask, &d.per_pattern, &d.row_invalid, &d.Pmat, &d.blen,
                          &d.clv, &d.scale, &d.cat_ll, &d.partials, &d.total, &d.ops, &d.gops, &d.gkids, &d.dops,
                          &d.dkids, &d.leaf_rows, &d.states, &d.posteriors, &d.staging, &d.sim_pre, &d.sim_parent,
                          &d.sim_depth, &d.sim_leafrow, &d.sim_stack, &d.scratch, &d.sw_descs, &d.sw_blocks,
                          &d.sw_stream, &d.sw_gstack, &d.sw_err, &d.sw_opoff, &d.dn_ops, &d.dn_descs, &d.dn_blocks, &d.dn_stream, &d.dn_opoff, &d.dn_gstack, &d.sw_ops, &d.uops, &d.ukids, &d.lnP, &d.jbest,
                          &d.nni_upper, &d.nni_uexps, &d.nni_jobs, &d.nni_rows, &d.nni_exps, &d.nni_partials, &d.nni_totals, &d.nni_moves, &d.nni_steps, &d.nni_need, &d.nni_cnt, &d.nni_pred, &d.nni_dops, &d.nni_dkids};
        for (DevBuf *b : bufs) b->release();
        if (ctx->is_aa) aa_release(ctx);
        for (auto &ep : d.ev_pool) {
            cudaEventDestroy(ep.a);
            cudaEventDestroy(ep.b);
        }
        if (d.h_total) cudaFreeHost(d.h_total);
        for (int i = 0; i < 2; ++i) {
            if (d.bounce[i]) cudaFreeHost(d.bounce[i]);
            if (d.ev_bounce[i]) cudaEventDestroy(d.ev_bounce[i]);
        }
        if (d.tm_a) cudaEventDestroy(d.tm_a);
        if (d.tm_b) cudaEventDestroy(d.tm_b);
        d.staging2.release();
        for (int i = 0; i < 2; ++i) {
            if (d.ev_copied[i]) cudaEventDestroy(d.ev_copied[i]);
            if (d.ev_free[i]) cudaEventDestroy(d.ev_free[i]);
        }
        if (d.copy_stream) cudaStreamDestroy(d.copy_stream);
        if (d.stream) cudaStreamDestroy(d.stream);
    }
    delete ctx;
}

int32_t phb_set_clv_budget(phb_ctx *ctx, uint64_t bytes)
{
    if (!ctx) return PHB_ERR_INVALID_ARGUMENT;
    ctx->clv_budget = bytes;
    return PHB_OK;
}

// Shards [0, P) contiguously over the devices and sizes the per-device alignment buffers.
static int shard_and_alloc(phb_ctx *ctx, int n_rows, int64_t P)
{
    const int G = (int)ctx->devs.size();
    int64_t per = (P + G - 1) / G;
    per = round_up(per, 2);
    for (int g = 0; g < G; ++g) {
        Device &d = ctx->devs[g];
        d.p_begin = std::min<int64_t>((int64_t)g * per, P);
        d.P = std::min<int64_t>(per, P - d.p_begin);
        d.P_pad = std::max<int64_t>(round_up(d.P, SITES_PER_BLOCK), SITES_PER_BLOCK);
        d.tip_constants_valid = false;
        CU(cudaSetDevice(d.ordinal));
        CU(d.tips.ensure((size_t)std::max(n_rows, 1) * (size_t)d.P_pad + TIPS_SLACK));
        CU(cudaMemsetAsync(d.tips.as<uint8_t>() + (size_t)std::max(n_rows, 1) * (size_t)d.P_pad, CODE_GAP, TIPS_SLACK, d.stream));
        CU(d.mult.ensure(sizeof(double) * (size_t)d.P_pad));
        CU(d.row_invalid.ensure(sizeof(int) * (size_t)std::max(n_rows, 1)));
        CU(cudaMemsetAsync(d.row_invalid.p, 0, sizeof(int) * (size_t)std::max(n_rows, 1), d.stream));
    }
    ctx->n_rows = n_rows;
    ctx->n_patterns = P;
    return PHB_OK;
}

} // extern "C" (templates need C++ linkage)

namespace {

// Streams the pattern-major host bytes of device d's shard through two staging buffers: chunk c is copied
// on copy_stream while chunk c-1 is packed (and whatever `after_chunk` enqueues runs) on d.stream.
// bounds: ascending pattern offsets inside the shard, bounds.front() == 0, bounds.back() == d.P, every
// interior bound a multiple of 64.
// `begin` puts the copies of the first two chunks on the wire and returns how many it issued, so that a
// caller can do its host-side preparation (schedules, uploads) under them; the main loop then skips those.
// Is this host pointer plain (unregistered, pageable) memory?  A Rust caller's Vec<u8> is.
bool host_is_pageable(const void *p)
{
    if (const char *e = getenv("PHB_FORCE_BOUNCE")) return atoi(e) != 0; // tests: take the bounce path (1) / never (0)
#ifdef PHB_HOST_EMU
    (void)p;
    return false;
#else
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeUnregistered;
#endif
}

// dst (device) <- src (host), `bytes`, in order on `stream`.  Pinned source: one asynchronous copy.  Pageable source:
// slices are copied by the pool's threads into two pinned bounce buffers and leave from there, the host copy of slice
// k + 1 overlapping the DMA of slice k; the call returns when the last slice has been handed to the stream.
int h2d_from_host(phb_ctx *ctx, Device &d, uint8_t *dst, const uint8_t *src, size_t bytes, cudaStream_t stream, bool pageable)
{
    // small inputs are not worth the pool and the 64 MB of pinned memory: the driver's own staging is fine for them
    if (!pageable || (bytes && bytes < bounce_min_bytes() && !d.bounce[0])) {
        if (bytes) CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, stream));
        return PHB_OK;
    }
    if (!ctx->copy_pool) {
        int n = 7;
        if (const char *e = getenv("PHB_COPY_THREADS")) n = std::max(0, atoi(e) - 1);
        n = std::min<int>(n, std::max(0, (int)std::thread::hardware_concurrency() - 1));
        ctx->copy_pool.reset(new CopyPool(n));
    }
    for (int i = 0; i < 2; ++i)
        if (!d.bounce[i]) {
            CU(cudaHostAlloc((void **)&d.bounce[i], bounce_bytes(), cudaHostAllocDefault));
            CU(cudaEventCreate(&d.ev_bounce[i]));
        }
    for (size_t off = 0; off < bytes; off += bounce_bytes()) {
        const size_t n = std::min(bounce_bytes(), bytes - off);
        const int b = (int)(d.bounce_uses & 1);
        if (d.bounce_uses >= 2) CU(cudaEventSynchronize(d.ev_bounce[b])); // the DMA that last read this bounce buffer
        ctx->copy_pool->copy(d.bounce[b], src + off, n);
        CU(cudaMemcpyAsync(dst + off, d.bounce[b], n, cudaMemcpyHostToDevice, stream));
        CU(cudaEventRecord(d.ev_bounce[b], stream));
        ++d.bounce_uses;
    }
    return PHB_OK;
}

// The same for `nr` separate host rows of `W` bytes each (row r at rows(r)) going to dst + r * pitch: as many rows as fit
// are gathered into a bounce buffer, one memcpy task per row, and leave as one (2-D) copy.
template <class RowPtr>
int h2d_rows_from_host(phb_ctx *ctx, Device &d, uint8_t *dst, size_t pitch, RowPtr rows, size_t W, int nr, cudaStream_t stream,
                       bool pageable)
{
    if (!pageable || W == 0 || W > bounce_bytes() || (W * (size_t)nr < bounce_min_bytes() && !d.bounce[0])) {
        for (int r = 0; r < nr; ++r) {
            int rc = h2d_from_host(ctx, d, dst + (size_t)r * pitch, rows(r), W, stream, pageable);
            if (rc) return rc;
        }
        return PHB_OK;
    }
    int rc = h2d_from_host(ctx, d, dst, rows(0), 0, stream, true); // pool and bounce buffers exist from here on
    if (rc) return rc;
    const int per_slice = (int)std::max<size_t>(1, bounce_bytes() / W);
    std::vector<CopyPool::Task> list;
    for (int r0 = 0; r0 < nr; r0 += per_slice) {
        const int m = std::min(per_slice, nr - r0);
        const int b = (int)(d.bounce_uses & 1);
        if (d.bounce_uses >= 2) CU(cudaEventSynchronize(d.ev_bounce[b]));
        list.clear();
        for (int r = 0; r < m; ++r) list.push_back({rows(r0 + r), d.bounce[b] + (size_t)r * W, W});
        ctx->copy_pool->run(list);
        CU(cudaMemcpy2DAsync(dst + (size_t)r0 * pitch, pitch, d.bounce[b], W, W, (size_t)m, cudaMemcpyHostToDevice, stream));
        CU(cudaEventRecord(d.ev_bounce[b], stream));
        ++d.bounce_uses;
    }
    return PHB_OK;
}

// The way back: `height` device rows of `width` bytes (row r at src + r * src_pitch) into a caller's PAGEABLE buffer (row r
// at dst + r * dst_pitch).  Groups of rows arrive in the pinned bounce buffers by DMA and are scattered to the caller's
// memory by the pool, the DMA of group g + 1 overlapping the host copy of group g.  Blocks until everything is home.
int d2h_rows_to_pageable(phb_ctx *ctx, Device &d, uint8_t *dst, size_t dst_pitch, const uint8_t *src, size_t src_pitch, size_t width,
                         int height, cudaStream_t stream)
{
    if (width == 0 || height <= 0) return PHB_OK;
    if (width > bounce_bytes()) { // a row does not fit a bounce buffer: the driver's staged copy
        CU(cudaMemcpy2DAsync(dst, dst_pitch, src, src_pitch, width, (size_t)height, cudaMemcpyDeviceToHost, stream));
        CU(cudaStreamSynchronize(stream));
        return PHB_OK;
    }
    int rc = h2d_from_host(ctx, d, nullptr, nullptr, 0, stream, true); // pool and bounce buffers exist from here on
    if (rc) return rc;
    for (int b = 0; b < 2 && d.bounce_uses >= 2; ++b) CU(cudaEventSynchronize(d.ev_bounce[b])); // earlier traffic through them
    const int per = (int)std::max<size_t>(1, bounce_bytes() / width);
    const int n_groups = (height + per - 1) / per;
    auto issue = [&](int g) -> cudaError_t {
        const int r0 = g * per, m = std::min(per, height - r0), b = g & 1;
        cudaError_t e = cudaMemcpy2DAsync(d.bounce[b], width, src + (size_t)r0 * src_pitch, src_pitch, width, (size_t)m,
                                          cudaMemcpyDeviceToHost, stream);
        if (e != cudaSuccess) return e;
        return cudaEventRecord(d.ev_bounce[b], stream);
    };
    CU(issue(0));
    std::vector<CopyPool::Task> list;
    for (int g = 0; g < n_groups; ++g) {
        if (g + 1 < n_groups) CU(issue(g + 1)); // its bounce buffer was emptied by the host in iteration g - 1
        const int r0 = g * per, m = std::min(per, height - r0), b = g & 1;
        CU(cudaEventSynchronize(d.ev_bounce[b]));
        list.clear();
        for (int r = 0; r < m; ++r) list.push_back({d.bounce[b] + (size_t)r * width, dst + (size_t)(r0 + r) * dst_pitch, width});
        ctx->copy_pool->run(list);
    }
    d.bounce_uses += 2; // both events have been recorded (and waited for)
    return PHB_OK;
}

int stream_pattern_major_begin(phb_ctx *ctx, Device &d, const uint8_t *bytes, int n_rows, const std::vector<int64_t> &bounds,
                               int *issued)
{
    const bool pageable = host_is_pageable(bytes);
    size_t max_bytes = 0;
    for (size_t c = 0; c + 1 < bounds.size(); ++c) max_bytes = std::max(max_bytes, (size_t)(bounds[c + 1] - bounds[c]) * n_rows);
    CU(d.staging.ensure(max_bytes));
    if (bounds.size() > 2) CU(d.staging2.ensure(max_bytes));
    // the staging buffers may still be read by packing kernels of an earlier call
    CU(cudaEventRecord(d.ev_free[0], d.stream));
    CU(cudaStreamWaitEvent(d.copy_stream, d.ev_free[0], 0));
    *issued = 0;
    for (size_t c = 0; c + 1 < bounds.size() && c < 2; ++c) {
        const int buf = (int)(c & 1);
        const int64_t c0 = bounds[c], np = bounds[c + 1] - c0;
        if (np > 0) {
            uint8_t *st = (buf ? d.staging2 : d.staging).as<uint8_t>();
            int rc = h2d_from_host(ctx, d, st, bytes + (size_t)(d.p_begin + c0) * n_rows, (size_t)np * n_rows, d.copy_stream, pageable);
            if (rc) return rc;
            CU(cudaEventRecord(d.ev_copied[buf], d.copy_stream));
        }
        *issued = (int)c + 1;
    }
    return PHB_OK;
}

template <class F>
int stream_pattern_major(phb_ctx *ctx, Device &d, const uint8_t *bytes, int n_rows, const std::vector<int64_t> &bounds,
                         F &&after_chunk, int issued = -1)
{
    if (issued < 0) {
        int rc0 = stream_pattern_major_begin(ctx, d, bytes, n_rows, bounds, &issued);
        if (rc0) return rc0;
    }
    const bool pageable = host_is_pageable(bytes);
    for (size_t c = 0; c + 1 < bounds.size(); ++c) {
        const int buf = (int)(c & 1);
        const int64_t c0 = bounds[c], np = bounds[c + 1] - c0;
        if (np <= 0) continue;
        uint8_t *st = (buf ? d.staging2 : d.staging).as<uint8_t>();
        if ((int)c >= issued) {
            if (c >= 2) CU(cudaStreamWaitEvent(d.copy_stream, d.ev_free[buf], 0));
            int rc1 = h2d_from_host(ctx, d, st, bytes + (size_t)(d.p_begin + c0) * n_rows, (size_t)np * n_rows, d.copy_stream, pageable);
            if (rc1) return rc1;
            CU(cudaEventRecord(d.ev_copied[buf], d.copy_stream));
        }
        CU(cudaStreamWaitEvent(d.stream, d.ev_copied[buf], 0));
        dim3 grid((unsigned)((np + PK_P - 1) / PK_P), (unsigned)((n_rows + PK_R - 1) / PK_R));
        PHB_LAUNCH(pack_pattern_major_tile_kernel, grid, dim3(256), d.stream, st, n_rows, np, d.tips.as<uint8_t>(), d.P_pad,
                   c0, d.row_invalid.as<int>());
        CHECK_LAUNCH();
        CU(cudaEventRecord(d.ev_free[buf], d.stream));
        int rc = after_chunk(c0, np);
        if (rc) return rc;
    }
    return PHB_OK;
}

// padding codes + multiplicities of one device's shard (everything of set_alignment except the bytes)
int upload_pad_and_mult(phb_ctx *ctx, Device &d, int n_rows, const uint64_t *multiplicity)
{
    for (int r0 = 0; r0 < n_rows && d.P_pad > d.P; r0 += 32768) {
        dim3 grid((unsigned)((d.P_pad - d.P + 255) / 256), (unsigned)std::min(32768, n_rows - r0));
        PHB_LAUNCH(fill_pad_kernel, grid, dim3(256), d.stream, d.tips.as<uint8_t>(), r0, d.P, d.P_pad);
        CHECK_LAUNCH();
    }
    const uint64_t *dm = nullptr;
    if (multiplicity && d.P > 0) {
        CU(d.scratch.ensure(sizeof(uint64_t) * (size_t)d.P));
        CU(cudaMemcpyAsync(d.scratch.p, multiplicity + d.p_begin, sizeof(uint64_t) * (size_t)d.P, cudaMemcpyHostToDevice,
                           d.stream));
        dm = d.scratch.as<uint64_t>();
    }
    PHB_LAUNCH(mult_to_double_kernel, dim3((unsigned)std::min<int64_t>((d.P_pad + 255) / 256, 4096)), dim3(256), d.stream, dm,
               d.P, d.P_pad, d.mult.as<double>());
    CHECK_LAUNCH();
    return PHB_OK;
}

// join + gather the per-row validity flags written by the packing kernels
int gather_row_flags(phb_ctx *ctx, int n_rows)
{
    std::vector<int> tmp(std::max(n_rows, 1));
    for (Device &d : ctx->devs) {
        CU(cudaSetDevice(d.ordinal));
        CU(cudaMemcpyAsync(tmp.data(), d.row_invalid.p, sizeof(int) * (size_t)std::max(n_rows, 1), cudaMemcpyDeviceToHost,
                           d.stream));
        CU(cudaStreamSynchronize(d.stream));
        for (int r = 0; r < n_rows; ++r) ctx->row_invalid_host[r] |= tmp[r];
    }
    // the codes of global pattern 0 (first column of the first non-empty shard)
    ctx->row_first_invalid.assign(std::max(n_rows, 1), 0);
    for (Device &d : ctx->devs) {
        if (d.P == 0 || n_rows == 0) continue;
        CU(cudaSetDevice(d.ordinal));
        CU(cudaMemcpy2DAsync(ctx->row_first_invalid.data(), 1, d.tips.p, (size_t)d.P_pad, 1, (size_t)n_rows,
                             cudaMemcpyDeviceToHost, d.stream));
        CU(cudaStreamSynchronize(d.stream));
        for (int r = 0; r < n_rows; ++r) ctx->row_first_invalid[r] = ctx->row_first_invalid[r] == CODE_INVALID;
        break;
    }
    return PHB_OK;
}

} // namespace

extern "C" {

int32_t phb_set_alignment(phb_ctx *ctx, int32_t n_rows, int64_t n_patterns, const uint8_t *bytes, int32_t layout,
                          const uint64_t *multiplicity)
{
    if (!ctx) return PHB_ERR_INVALID_ARGUMENT;
    if (n_rows < 0 || n_patterns < 0 || (layout != PHB_LAYOUT_PATTERN_MAJOR && layout != PHB_LAYOUT_ROW_MAJOR) ||
        (!bytes && n_rows > 0 && n_patterns > 0))
        return ctx->fail(PHB_ERR_INVALID_ARGUMENT, "set_alignment: bad arguments");
    if (ctx->is_aa) return aa_set_alignment(ctx, n_rows, n_patterns, bytes, layout, multiplicity);
    ctx->have_aln = false;
    int rc = shard_and_alloc(ctx, n_rows, n_patterns);
    if (rc) return rc;
    ctx->row_invalid_host.assign(std::max(n_rows, 1), 0);
    ctx->row_first_invalid.clear();
    const size_t CHUNK = (size_t)256 << 20; // staging bytes per H2D copy
    for (Device &d : ctx->devs) {
        CU(cudaSetDevice(d.ordinal));
        if (d.P > 0 && n_rows > 0) {
            if (layout == PHB_LAYOUT_PATTERN_MAJOR) {
                int64_t pp = std::max<int64_t>(64, (int64_t)(CHUNK / (size_t)n_rows) / 64 * 64);
                std::vector<int64_t> bounds(1, 0);
                while (d.P - bounds.back() > pp) bounds.push_back(bounds.back() + pp);
                bounds.push_back(d.P);
                rc = stream_pattern_major(ctx, d, bytes, n_rows, bounds, [](int64_t, int64_t) { return 0; });
                if (rc) return rc;
            } else {
                int64_t rr = std::max<int64_t>(1, (int64_t)(CHUNK / (size_t)d.P));
                rr = std::min<int64_t>(std::min<int64_t>(rr, n_rows), 32768);
                CU(d.staging.ensure((size_t)rr * d.P));
                for (int64_t r0 = 0; r0 < n_rows; r0 += rr) {
                    int64_t nr = std::min<int64_t>(rr, n_rows - r0);
                    if (!host_is_pageable(bytes))
                        CU(cudaMemcpy2DAsync(d.staging.p, (size_t)d.P, bytes + (size_t)r0 * n_patterns + d.p_begin,
                                             (size_t)n_patterns, (size_t)d.P, (size_t)nr, cudaMemcpyHostToDevice, d.stream));
                    else {
                        rc = h2d_rows_from_host(ctx, d, d.staging.as<uint8_t>(), (size_t)d.P,
                                                [&](int r) { return bytes + (size_t)(r0 + r) * n_patterns + d.p_begin; }, (size_t)d.P,
                                                (int)nr, d.stream, true);
                        if (rc) return rc;
                    }
                    dim3 grid((unsigned)std::min<int64_t>((d.P + 255) / 256, 4096), (unsigned)nr);
                    PHB_LAUNCH(pack_row_major_kernel, grid, dim3(256), d.stream, d.staging.as<uint8_t>(), (int)nr, d.P,
                               d.P, (int64_t)0, d.tips.as<uint8_t>() + (size_t)r0 * d.P_pad, d.P_pad, (int64_t)0,
                               d.row_invalid.as<int>() + r0);
                    CHECK_LAUNCH();
                }
            }
        }
        rc = upload_pad_and_mult(ctx, d, n_rows, multiplicity);
        if (rc) return rc;
    }
    rc = gather_row_flags(ctx, n_rows);
    if (rc) return rc;
    ctx->have_aln = true;
    return PHB_OK;
}

// bytes: contiguous [n_rows][width], or null with row_ptrs[r] -> the `width` bytes of row r
static int32_t compress_impl(phb_ctx *ctx, int32_t n_rows, int64_t width, const uint8_t *bytes, const uint8_t *const *row_ptrs,
                             int64_t *n_patterns, uint64_t *site_to_pattern, uint64_t *multiplicity)
{
    if (!ctx || !n_patterns) return PHB_ERR_INVALID_ARGUMENT;
    if (n_rows < 0 || width < 0 || width > 0x7fffffff || (!bytes && !row_ptrs && n_rows > 0 && width > 0))
        return ctx->fail(PHB_ERR_INVALID_ARGUMENT, "compress_columns: bad arguments");
    if (!bytes && row_ptrs)
        for (int r = 0; r < n_rows && width > 0; ++r)
            if (!row_ptrs[r]) return ctx->fail(PHB_ERR_INVALID_ARGUMENT, "compress_rows: row %d is null", r);
    ctx->have_aln = false;
    *n_patterns = 0;
    if (width == 0 && ctx->is_aa) return aa_adopt_compressed(ctx, n_rows, 0, nullptr, 0, nullptr, nullptr);
    if (width == 0) { // no sites: no patterns (alignment.rs:102 never iterates)
        int rc0 = shard_and_alloc(ctx, n_rows, 0);
        if (rc0) return rc0;
        ctx->row_invalid_host.assign(std::max(n_rows, 1), 0);
        ctx->row_first_invalid.clear();
        for (Device &d : ctx->devs) {
            rc0 = upload_pad_and_mult(ctx, d, n_rows, nullptr);
            if (rc0) return rc0;
        }
        ctx->have_aln = true;
        return PHB_OK;
    }
    Device &d0 = ctx->devs[0];
    CU(cudaSetDevice(d0.ordinal));
    static const bool trace = getenv("PHB_COMPRESS_TRACE") != nullptr; // stage wall times on stderr
    auto t_prev = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        if (!trace) return;
        cudaStreamSynchronize(d0.stream);
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[phb compress] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t_prev).count());
        t_prev = now;
    };
    const int64_t pitch = round_up(width, 16) + 16; // 8-site loads of the last thread stay inside the row
    const int rows_alloc = std::max(n_rows, 1);
    DevBuf raw, hash, keys, sites, sites2, head, gid, gstart, rep, isrep, size_at, pid, replist, s2p, mult, temp, flag, stage;
    DevBuf *all[] = {&raw, &hash, &keys, &sites, &sites2, &head, &gid, &gstart, &rep, &isrep, &size_at, &pid, &replist,
                     &s2p, &mult, &temp, &flag, &stage};
    struct Cleanup {
        DevBuf **b;
        size_t n;
        ~Cleanup() { for (size_t i = 0; i < n; ++i) b[i]->release(); }
    } cleanup{all, sizeof all / sizeof all[0]};
    const size_t W = (size_t)width;
    // a caller's Vec<u8> rows are pageable memory: they go through the pinned bounce buffers (h2d_rows_from_host)
    const bool pageable = width > 0 && n_rows > 0 && host_is_pageable(bytes ? (const void *)bytes : (const void *)row_ptrs[0]);
    CU(raw.ensure((size_t)rows_alloc * (size_t)pitch));
    CU(hash.ensure(sizeof(uint64_t) * W));
    CU(keys.ensure(sizeof(uint64_t) * W));
    CU(sites.ensure(sizeof(uint32_t) * W));
    CU(sites2.ensure(sizeof(uint32_t) * W));
    CU(head.ensure(sizeof(uint32_t) * W));
    CU(gid.ensure(sizeof(uint32_t) * W));
    CU(gstart.ensure(sizeof(uint32_t) * W));
    CU(rep.ensure(sizeof(uint32_t) * W));
    CU(isrep.ensure(sizeof(uint32_t) * W));
    CU(size_at.ensure(sizeof(uint32_t) * W));
    CU(pid.ensure(sizeof(uint32_t) * W));
    CU(replist.ensure(sizeof(uint32_t) * W));
    CU(s2p.ensure(sizeof(uint64_t) * W));
    CU(mult.ensure(sizeof(uint64_t) * W));
    CU(flag.ensure(sizeof(unsigned int)));
    const size_t temp_bytes = compress_temp_bytes(width);
    CU(temp.ensure(temp_bytes));
    lap("allocations");
    // 1. bytes -> device, row chunk by row chunk on the copy stream; the first hash pass folds each chunk as it lands
    CU(cudaEventRecord(d0.ev_free[0], d0.stream));
    CU(cudaStreamWaitEvent(d0.copy_stream, d0.ev_free[0], 0));
    const int rows_per_chunk = (int)std::max<int64_t>(1, std::min<int64_t>(std::max(n_rows, 1), ((int64_t)256 << 20) / width));
    const unsigned hash_grid = (unsigned)((width + 256 * HASH_SITES_PER_THREAD - 1) / (256 * HASH_SITES_PER_THREAD));
    const unsigned w_grid = (unsigned)((width + 255) / 256);
    uint64_t seed = 0x243F6A8885A308D3ull;
    std::vector<cudaEvent_t> chunk_events;
    struct EvCleanup {
        std::vector<cudaEvent_t> &v;
        ~EvCleanup() { for (cudaEvent_t e : v) cudaEventDestroy(e); }
    } evc{chunk_events};
    int64_t P = 0;
    for (int attempt = 0; attempt < 4; ++attempt) {
        const uint64_t key = splitmix64(seed + attempt), mul = splitmix64(key) | 1ull;
        for (int r0 = 0, ci = 0; r0 < std::max(n_rows, 1); r0 += rows_per_chunk, ++ci) {
            const int nr = std::min(rows_per_chunk, n_rows - r0);
            if (attempt == 0 && nr > 0) {
                if (bytes && !pageable)
                    CU(cudaMemcpy2DAsync(raw.as<uint8_t>() + (size_t)r0 * pitch, (size_t)pitch, bytes + (size_t)r0 * W, W, W,
                                         (size_t)nr, cudaMemcpyHostToDevice, d0.copy_stream));
                else {
                    int rcc = h2d_rows_from_host(ctx, d0, raw.as<uint8_t>() + (size_t)r0 * pitch, (size_t)pitch,
                                                 [&](int r) { return bytes ? bytes + (size_t)(r0 + r) * W : row_ptrs[r0 + r]; }, W, nr,
                                                 d0.copy_stream, pageable);
                    if (rcc) return rcc;
                }
                cudaEvent_t e;
                CU(cudaEventCreate(&e));
                chunk_events.push_back(e);
                CU(cudaEventRecord(e, d0.copy_stream));
                CU(cudaStreamWaitEvent(d0.stream, e, 0));
            }
            const bool first = r0 == 0, last = r0 + rows_per_chunk >= n_rows;
            PHB_LAUNCH(column_hash_kernel, dim3(hash_grid), dim3(256), d0.stream, raw.as<uint8_t>(), pitch, r0, std::max(nr, 0),
                       width, key, mul, first ? 1 : 0, last ? 1 : 0, hash.as<uint64_t>());
            CHECK_LAUNCH();
        }
        lap("copy + column hash");
        // 2. stable sort of (hash, site); groups of equal hash
        PHB_LAUNCH(iota_kernel, dim3(w_grid), dim3(256), d0.stream, sites.as<uint32_t>(), width);
        CHECK_LAUNCH();
        CU(dev_sort_pairs(temp.p, temp_bytes, hash.as<uint64_t>(), keys.as<uint64_t>(), sites.as<uint32_t>(),
                          sites2.as<uint32_t>(), width, d0.stream));
        PHB_LAUNCH(group_head_kernel, dim3(w_grid), dim3(256), d0.stream, keys.as<uint64_t>(), width, head.as<uint32_t>());
        CHECK_LAUNCH();
        CU(dev_scan(temp.p, temp_bytes, head.as<uint32_t>(), gid.as<uint32_t>(), width, true, d0.stream));
        PHB_LAUNCH(group_start_kernel, dim3(w_grid), dim3(256), d0.stream, head.as<uint32_t>(), gid.as<uint32_t>(), width,
                   gstart.as<uint32_t>());
        CHECK_LAUNCH();
        uint32_t n_groups = 0;
        CU(cudaMemcpyAsync(&n_groups, gid.as<uint32_t>() + (W - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, d0.stream));
        CU(cudaStreamSynchronize(d0.stream));
        lap("sort + groups");
        PHB_LAUNCH(group_rep_kernel, dim3(w_grid), dim3(256), d0.stream, sites2.as<uint32_t>(), gid.as<uint32_t>(),
                   gstart.as<uint32_t>(), n_groups, width, rep.as<uint32_t>(), isrep.as<uint32_t>(), size_at.as<uint32_t>());
        CHECK_LAUNCH();
        // 3. exact verification of every merged site against the first site of its group
        CU(cudaMemsetAsync(flag.p, 0, sizeof(unsigned int), d0.stream));
        if ((int64_t)n_groups < width && n_rows > 0) {
            const int rpb = std::max(1, (n_rows + 63) / 64);
            PHB_LAUNCH(column_verify_kernel, dim3(w_grid, (unsigned)((n_rows + rpb - 1) / rpb)), dim3(256), d0.stream,
                       raw.as<uint8_t>(), pitch, n_rows, rpb, width, rep.as<uint32_t>(), flag.as<unsigned int>());
            CHECK_LAUNCH();
        }
        unsigned int mismatch = 0;
        CU(cudaMemcpyAsync(&mismatch, flag.p, sizeof(unsigned int), cudaMemcpyDeviceToHost, d0.stream));
        CU(cudaStreamSynchronize(d0.stream));
        lap("verification");
        if (mismatch) { // two different columns share a 64-bit hash: hash again under another key
            if (attempt == 3) return ctx->fail(PHB_ERR_NUMERICAL_INSTABILITY, "compress_columns: hash collisions under 4 keys");
            continue;
        }
        P = (int64_t)n_groups;
        break;
    }
    // 4. pattern ids in order of first appearance
    CU(dev_scan(temp.p, temp_bytes, isrep.as<uint32_t>(), pid.as<uint32_t>(), width, false, d0.stream));
    PHB_LAUNCH(pattern_index_kernel, dim3(w_grid), dim3(256), d0.stream, rep.as<uint32_t>(), isrep.as<uint32_t>(),
               pid.as<uint32_t>(), size_at.as<uint32_t>(), width, replist.as<uint32_t>(), s2p.as<uint64_t>(), mult.as<uint64_t>());
    CHECK_LAUNCH();
    if (site_to_pattern)
        CU(cudaMemcpyAsync(site_to_pattern, s2p.p, sizeof(uint64_t) * W, cudaMemcpyDeviceToHost, d0.stream));
    if (multiplicity)
        CU(cudaMemcpyAsync(multiplicity, mult.p, sizeof(uint64_t) * (size_t)P, cudaMemcpyDeviceToHost, d0.stream));
    CU(cudaStreamSynchronize(d0.stream));
    lap("pattern ids + results to host");
    // 5. the unique columns become the context's alignment: tip codes [row][pattern] per device shard
    if (ctx->is_aa) { // 20-state codes, one device (aa_engine.cuh)
        int rca = aa_adopt_compressed(ctx, n_rows, P, raw.as<uint8_t>(), pitch, replist.as<uint32_t>(), mult.as<uint64_t>());
        if (rca) return rca;
        lap("gather to 20-state tip codes");
        *n_patterns = P;
        return PHB_OK;
    }
    int rc = shard_and_alloc(ctx, n_rows, P);
    if (rc) return rc;
    ctx->row_invalid_host.assign(std::max(n_rows, 1), 0);
    ctx->row_first_invalid.clear();
    lap("shard_and_alloc");
    for (Device &d : ctx->devs) {
        const bool same = &d == &d0;
        CU(cudaSetDevice(d0.ordinal));
        uint8_t *dst_tips = d.tips.as<uint8_t>();
        double *dst_mult = d.mult.as<double>();
        int *dst_flags = d.row_invalid.as<int>();
        if (!same) { // gather on device 0, then one device-to-device copy per buffer
            const size_t need = (size_t)rows_alloc * (size_t)d.P_pad + sizeof(double) * (size_t)d.P_pad + sizeof(int) * (size_t)rows_alloc + 64;
            CU(stage.ensure(need));
            dst_mult = stage.as<double>();
            dst_flags = reinterpret_cast<int *>(stage.as<uint8_t>() + sizeof(double) * (size_t)d.P_pad);
            dst_tips = stage.as<uint8_t>() + sizeof(double) * (size_t)d.P_pad + sizeof(int) * (size_t)rows_alloc;
            CU(cudaMemsetAsync(dst_flags, 0, sizeof(int) * (size_t)rows_alloc, d0.stream));
            CU(cudaStreamSynchronize(d.stream)); // shard_and_alloc's memsets on the destination
        }
        const int rows_per_launch = 32768 * GATHER_ROWS;
        for (int r0 = 0; r0 < n_rows && d.P > 0; r0 += rows_per_launch) {
            const int nr = std::min(rows_per_launch, n_rows - r0);
            dim3 grid((unsigned)((d.P + 1023) / 1024), (unsigned)((nr + GATHER_ROWS - 1) / GATHER_ROWS));
            PHB_LAUNCH(gather_patterns_kernel, grid, dim3(256), d0.stream, raw.as<uint8_t>() + (size_t)r0 * pitch, pitch,
                       replist.as<uint32_t>(), d.p_begin, d.P, nr, dst_tips + (size_t)r0 * d.P_pad, d.P_pad, (int64_t)0,
                       dst_flags + r0);
            CHECK_LAUNCH();
        }
        for (int r0 = 0; r0 < n_rows && d.P_pad > d.P; r0 += 32768) {
            dim3 grid((unsigned)((d.P_pad - d.P + 255) / 256), (unsigned)std::min(32768, n_rows - r0));
            PHB_LAUNCH(fill_pad_kernel, grid, dim3(256), d0.stream, dst_tips, r0, d.P, d.P_pad);
            CHECK_LAUNCH();
        }
        PHB_LAUNCH(mult_u64_to_double_kernel, dim3((unsigned)std::min<int64_t>((d.P_pad + 255) / 256, 4096)), dim3(256), d0.stream,
                   mult.as<uint64_t>(), d.p_begin, d.P, d.P_pad, dst_mult);
        CHECK_LAUNCH();
        if (!same) {
            CU(cudaMemcpyAsync(d.tips.p, dst_tips, (size_t)rows_alloc * (size_t)d.P_pad, cudaMemcpyDeviceToDevice, d0.stream));
            CU(cudaMemcpyAsync(d.mult.p, dst_mult, sizeof(double) * (size_t)d.P_pad, cudaMemcpyDeviceToDevice, d0.stream));
            CU(cudaMemcpyAsync(d.row_invalid.p, dst_flags, sizeof(int) * (size_t)rows_alloc, cudaMemcpyDeviceToDevice, d0.stream));
        }
        CU(cudaStreamSynchronize(d0.stream));
    }
    lap("gather to tip codes");
    rc = gather_row_flags(ctx, n_rows);
    if (rc) return rc;
    *n_patterns = P;
    ctx->have_aln = true;
    return PHB_OK;
}

int32_t phb_compress_columns(phb_ctx *ctx, int32_t n_rows, int64_t width, const uint8_t *bytes, int64_t *n_patterns,
                             uint64_t *site_to_pattern, uint64_t *multiplicity)
{
    return compress_impl(ctx, n_rows, width, bytes, nullptr, n_patterns, site_to_pattern, multiplicity);
}

int32_t phb_compress_rows(phb_ctx *ctx, int32_t n_rows, int64_t width, const uint8_t *const *rows, int64_t *n_patterns,
                          uint64_t *site_to_pattern, uint64_t *multiplicity)
{
    return compress_impl(ctx, n_rows, width, nullptr, rows, n_patterns, site_to_pattern, multiplicity);
}

int32_t phb_set_model(phb_ctx *ctx, const double *lambda, const double *V, const double *pi, const double *sqrt_pi,
                      const double *inv_sqrt_pi, int32_t n_cat, const double *rates, const double *weights)
{
    if (!ctx) return PHB_ERR_INVALID_ARGUMENT;
    if (!lambda || !V || !pi || !sqrt_pi || !inv_sqrt_pi || !rates || !weights)
        return ctx->fail(PHB_ERR_INVALID_ARGUMENT, "set_model: null pointer");
    if (n_cat < 1 || n_cat > MAX_CAT)
        return ctx->fail(PHB_ERR_INVALID_MODEL_PARAMETER, "set_model: n_cat must be in [1, %d]", MAX_CAT);
    if (ctx->is_aa) return aa_set_model(ctx, lambda, V, pi, sqrt_pi, inv_sqrt_pi, n_cat, rates, weights);
    ModelDev M;
    memset(&M, 0, sizeof M);
    memcpy(M.lambda, lambda, sizeof M.lambda);
    memcpy(M.V, V, sizeof M.V);
    memcpy(M.pi, pi, sizeof M.pi);
    memcpy(M.sqrt_pi, sqrt_pi, sizeof M.sqrt_pi);
    memcpy(M.inv_sqrt_pi, inv_sqrt_pi, sizeof M.inv_sqrt_pi);
    for (int c = 0; c < n_cat; ++c) {
        M.rates[c] = rates[c];
        M.weights[c] = weights[c];
        M.ln_weights[c] = std::log(weights[c]);
    }
    // leaf Profile::scale() constants: ln(max) with max = 1, 0.5, 1/3, 0.25 (alphabet.rs:44-64)
    M.ln_tip[0] = 0.0;
    M.ln_tip[1] = 0.0;
    M.ln_tip[2] = std::log(0.5);
    M.ln_tip[3] = std::log(1.0 / 3.0);
    M.ln_tip[4] = std::log(0.25);
    M.n_cat = n_cat;
    build_model_dev(M, (ctx->flags & PHB_FLAG_LITERAL_INVARIANT) != 0, ctx->model);
    build_model_dev(M, true, ctx->model_lit);
    ctx->have_model = true;
    ctx->sweep_guards_valid = false; // the guard plan depends on the rates
    ctx->nni_guards_valid = false;
    return PHB_OK;
}

int32_t phb_set_tree(phb_ctx *ctx, int32_t n_nodes, const int32_t *postorder, const int32_t *child_offsets,
                     const int32_t *children, const double *branch_len, const int32_t *leaf_row, int32_t root)
{
    if (!ctx) return PHB_ERR_INVALID_ARGUMENT;
    if (n_nodes < 1 || !postorder || !child_offsets || !branch_len || !leaf_row || root < 0 || root >= n_nodes)
        return ctx->fail(PHB_ERR_INVALID_ARGUMENT, "set_tree: bad arguments");
    if (child_offsets[0] != 0 || child_offsets[n_nodes] != n_nodes - 1 || (n_nodes > 1 && !children))
        return ctx->fail(PHB_ERR_INVALID_ARGUMENT, "set_tree: child CSR must hold exactly n_nodes - 1 edges");
    for (int v = 0; v < n_nodes; ++v)
        if (child_offsets[v + 1] < child_offsets[v])
            return ctx->fail(PHB_ERR_INVALID_ARGUMENT, "set_tree: child offsets must be non-decreasing");
    for (Device &d : ctx->devs) { // previous uploads read the old host vectors
        cudaSetDevice(d.ordinal);
        cudaStreamSynchronize(d.stream);
    }
    TreeSched &T = ctx->tree;
    T.n = n_nodes;
    T.root = root;
    T.child_off.assign(child_offsets, child_offsets + n_nodes + 1);
    T.children.assign(children, children + (n_nodes - 1));
    T.blen.assign(branch_len, branch_len + n_nodes);
    T.leaf_row.assign(leaf_row, leaf_row + n_nodes);
    ctx->tree_uploaded = false;
    int rc = T.build(ctx->err);
    if (rc) return rc;
    // the caller's post-order must be a valid schedule (children before parents, root last)
    {
        std::vector<char> seen(n_nodes, 0);
        for (int k = 0; k < n_nodes; ++k) {
            int v = postorder[k];
            if (v < 0 || v >= n_nodes || seen[v]) {
                T.valid = false;
                return ctx->fail(PHB_ERR_NUMERICAL_INSTABILITY, "set_tree: postorder is not a permutation");
            }
            for (int c = child_offsets[v]; c < child_offsets[v + 1]; ++c)
                if (!seen[children[c]]) {
                    T.valid = false;
                    return ctx->fail(PHB_ERR_NUMERICAL_INSTABILITY, "set_tree: node %d precedes its child in postorder", v);
                }
            seen[v] = 1;
        }
    }
    if (ctx->is_aa) return aa_upload_tree(ctx);
    return upload_tree(ctx);
}

int32_t phb_log_likelihood(phb_ctx *ctx, double *total, double *per_pattern)
{
    if (!ctx || !total) return PHB_ERR_INVALID_ARGUMENT;
    int rc = ctx->is_aa ? aa_evaluate(ctx, false, nullptr, nullptr, total) : evaluate(ctx, false, false, total);
    if (rc) return rc;
    if (per_pattern) return phb_get_per_pattern(ctx, per_pattern);
    return PHB_OK;
}

int32_t phb_log_likelihood_host(phb_ctx *ctx, int32_t n_rows, int64_t n_patterns, const uint8_t *bytes, int32_t layout,
                                const uint64_t *multiplicity, double *total, double *per_pattern)
{
    if (!ctx || !total) return PHB_ERR_INVALID_ARGUMENT;
    if (n_rows < 0 || n_patterns < 0 || (layout != PHB_LAYOUT_PATTERN_MAJOR && layout != PHB_LAYOUT_ROW_MAJOR) ||
        (!bytes && n_rows > 0 && n_patterns > 0))
        return ctx->fail(PHB_ERR_INVALID_ARGUMENT, "log_likelihood_host: bad arguments");
    if (!ctx->have_model || !ctx->tree.valid)
        return ctx->fail(PHB_ERR_STATE, "set_model and set_tree must precede log_likelihood_host");
    const TreeSched &T = ctx->tree;
    const ModelDev &M = ctx->model;
    // (a 20-state context takes the two-call path: its packing kernel already runs chunk by chunk behind the copies)
    const bool fused = !ctx->is_aa && layout == PHB_LAYOUT_PATTERN_MAJOR && !(ctx->flags & PHB_FLAG_LEVEL_KERNELS) && T.n_internal > 0 &&
                       M.n_clv > 0 && M.n_cat <= 4 * M.n_clv && n_patterns > 0 && n_rows > 0;
    if (!fused) { // the general two-call path
        int rc = phb_set_alignment(ctx, n_rows, n_patterns, bytes, layout, multiplicity);
        if (rc) return rc;
        return phb_log_likelihood(ctx, total, per_pattern);
    }
    // every leaf needs a row before anything is copied (likelihood.rs:99-103)
    for (size_t i = 0; i < T.leaf_nodes.size(); ++i)
        if (T.leaf_rows[i] < 0 || T.leaf_rows[i] >= n_rows)
            return ctx->fail(PHB_ERR_INVALID_ALIGNMENT, "Leaf in tree not found in alignment leaf order (node %d)",
                             T.leaf_nodes[i]);
    ctx->have_aln = false;
    int rc = shard_and_alloc(ctx, n_rows, n_patterns);
    if (rc) return rc;
    ctx->row_invalid_host.assign(std::max(n_rows, 1), 0);
    ctx->row_first_invalid.clear();
    // Chunks of whole passes of the widest CTA (1 by default: measured 195.2 / 199.9 / 207.9 ms end to end for
    // 1 / 2 / 3 passes at 10k x 1M — the copy engine is the bottleneck, so smaller chunks only shorten the
    // un-overlapped tail); a short remainder joins the last chunk, which gets its own geometry so that it does
    // not end in a nearly empty pass.  The first two copies go on the wire BEFORE the host-side preparation
    // (tree upload, sweep schedule, plans), which then costs nothing on top of them.
    std::vector<std::vector<int64_t>> all_bounds(ctx->devs.size());
    std::vector<int> issued(ctx->devs.size(), 0);
    std::vector<int64_t> chunk_of(ctx->devs.size(), 0);
    auto drain = [&](int code) { // the caller's buffer must not be in flight when an error goes back
        for (Device &d : ctx->devs) {
            cudaSetDevice(d.ordinal);
            cudaStreamSynchronize(d.copy_stream);
        }
        return code;
    };
    for (size_t di = 0; di < ctx->devs.size(); ++di) {
        Device &d = ctx->devs[di];
        if (d.P == 0) continue;
        CU(cudaSetDevice(d.ordinal));
        const int n_sm = std::max(d.sm_count, 1);
        const int q_max = sweep_q_max(M.n_clv);
        const int64_t pass = (int64_t)n_sm * 4 * q_max;
        static const int chunk_passes = [] { const char *e = getenv("PHB_HOST_CHUNK_PASSES"); int v = e ? atoi(e) : 1; return v >= 1 && v <= 16 ? v : 1; }();
        const int64_t chunk = round_up(chunk_passes * pass, 64);
        std::vector<int64_t> &bounds = all_bounds[di];
        bounds.assign(1, 0);
        while (d.P - bounds.back() > chunk + chunk / 2) bounds.push_back(bounds.back() + chunk);
        bounds.push_back(d.P);
        chunk_of[di] = chunk;
        rc = stream_pattern_major_begin(ctx, d, bytes, n_rows, bounds, &issued[di]);
        if (rc) return drain(rc);
    }
    if (!ctx->tree_uploaded) {
        rc = upload_tree(ctx);
        if (rc) return drain(rc);
    }
    rc = ensure_sweep(ctx, M, false);
    if (rc) return drain(rc);
    ctx->stats.clv_launches = 0;
    const SweepSched &S = ctx->sweep;
    Device &d0 = ctx->devs[0];
    d0.ev_used = 0;
    d0.ev_log.clear();
    for (size_t di = 0; di < ctx->devs.size(); ++di) {
        Device &d = ctx->devs[di];
        CU(cudaSetDevice(d.ordinal));
        if (d.P == 0) {
            d.h_total[0] = d.h_total[1] = 0.0;
            continue;
        }
        rc = plan_device(ctx, d, M, false, false, true);
        if (rc) return drain(rc);
        rc = upload_pad_and_mult(ctx, d, n_rows, multiplicity);
        if (rc) return drain(rc);
        rc = launch_sweep_transition(ctx, d, M);
        if (rc) return drain(rc);
        const int n_sm = std::max(d.sm_count, 1);
        const int64_t chunk = chunk_of[di];
        const std::vector<int64_t> &bounds = all_bounds[di];
        SweepGeom g_main, g_tail;
        if (!g_main.plan(S, d.smem_optin, chunk, n_sm) || !g_tail.plan(S, d.smem_optin, d.P - bounds[bounds.size() - 2], n_sm))
            return drain(ctx->fail(PHB_ERR_INVALID_MODEL_PARAMETER, "sweep: %d CLV categories do not fit one CTA", M.n_clv));
        size_t n_part_max = 0;
        for (size_t c = 0; c + 1 < bounds.size(); ++c) {
            const SweepGeom &G = c + 2 == bounds.size() ? g_tail : g_main;
            n_part_max += (size_t)((bounds[c + 1] - bounds[c] + 4 * G.Q - 1) / (4 * G.Q));
        }
        CU(d.partials.ensure(sizeof(double) * std::max<size_t>(n_part_max, 1)));
        size_t part_off = 0;
        const int64_t last0 = bounds[bounds.size() - 2];
        rc = stream_pattern_major(ctx, d, bytes, n_rows, bounds, [&](int64_t c0, int64_t np) -> int {
            const SweepGeom &G = c0 == last0 ? g_tail : g_main;
            const int64_t np4 = round_up(np, 4);
            PHB_LAUNCH(tip_constants_kernel, dim3((unsigned)((np4 + TIPC_SITES - 1) / TIPC_SITES)), dim3(TIPC_THREADS), d.stream,
                       d.tips.as<uint8_t>() + c0, d.P_pad, np4, d.leaf_rows.as<int32_t>(), (int)T.leaf_rows.size(), M,
                       d.leaf_ls.as<double>() + c0, d.inv_mask.as<uint8_t>() + c0);
            CHECK_LAUNCH();
            int r2 = launch_sweep_range(ctx, d, M, G, c0, np, part_off);
            part_off += (size_t)((np + 4 * G.Q - 1) / (4 * G.Q));
            return r2;
        }, issued[di]);
        if (rc) return drain(rc);
        d.tip_constants_valid = true;
        d.tip_constants_rows = T.leaf_rows;
        rc = launch_sweep_reduce(ctx, d, part_off);
        if (rc) return drain(rc);
        CU(cudaMemcpyAsync(d.h_total, d.total.p, 2 * sizeof(double), cudaMemcpyDeviceToHost, d.stream));
    }
    rc = gather_row_flags(ctx, n_rows); // joins every device stream
    if (rc) return rc;
    ctx->have_aln = true;
    rc = check_leaves(ctx); // invalid byte in a leaf row (likelihood.rs:105-107): the result is discarded
    if (rc) return rc;
    double sum = 0.0;
    for (Device &d : ctx->devs) {
        sum += d.h_total[0]; // rank order: deterministic
        if (d.h_total[1] != 0.0)
            return ctx->fail(PHB_ERR_NUMERICAL_INSTABILITY, "fused sweep: unknown op code in the schedule (device %d)", d.ordinal);
    }
    *total = sum;
    phb_stats &st = ctx->stats;
    st.n_patterns = n_patterns;
    st.tile_patterns = d0.P_pad;
    st.n_tiles = 1;
    st.n_levels = T.n_levels;
    st.n_internal = T.n_internal;
    st.n_clv_categories = M.n_clv;
    st.kernel_launches = ctx->launches;
    st.clv_updates = (int64_t)T.n_internal * n_patterns;
    st.algorithmic_bytes = (32.0 * M.n_clv * T.sum_one_plus_i + T.sum_t) * (double)n_patterns;
    st.sweep_ops = S.n_ops;
    st.sweep_stack_levels = S.n_levels;
    st.sweep_smem_levels = ctx->sweep_geom.smem_levels;
    st.sweep_stages = ctx->sweep_geom.n_stages;
    st.sweep_block_sites = 4 * ctx->sweep_geom.Q;
    st.sweep_stream_bytes = (int64_t)S.stream_bytes;
    st.sweep_guards = S.n_guards;
    st.sweep_normouts = S.n_normout;
    st.sweep_fp64_instr = S.fp64_instr;
    st.sweep_flops = S.flops;
    if (per_pattern) return phb_get_per_pattern(ctx, per_pattern);
    return PHB_OK;
}

int32_t phb_get_per_pattern(phb_ctx *ctx, double *per_pattern)
{
    if (!ctx || !per_pattern) return PHB_ERR_INVALID_ARGUMENT;
    if (ctx->is_aa) return aa_get_per_pattern(ctx, per_pattern);
    for (Device &d : ctx->devs) {
        if (d.P == 0) continue;
        CU(cudaSetDevice(d.ordinal));
        if (d.per_pattern.cap < sizeof(double) * (size_t)d.P)
            return ctx->fail(PHB_ERR_STATE, "no evaluation has run yet");
        CU(cudaMemcpyAsync(per_pattern + d.p_begin, d.per_pattern.p, sizeof(double) * (size_t)d.P, cudaMemcpyDeviceToHost,
                           d.stream));
        CU(cudaStreamSynchronize(d.stream));
    }
    return PHB_OK;
}

int32_t phb_marginal_asr(phb_ctx *ctx, uint8_t *states, double *posteriors, double *total)
{
    if (!ctx || !total || !states) return PHB_ERR_INVALID_ARGUMENT;
    if (ctx->is_aa) return aa_evaluate(ctx, true, states, posteriors, total);
    // the outputs leave tile by tile on the copy stream while the next tile is computed (pinned host
    // buffers get the full PCIe rate; pageable ones work, staged by the driver)
    ctx->asr_states_out = states;
    ctx->asr_post_out = posteriors;
    int rc = evaluate(ctx, true, posteriors != nullptr, total);
    ctx->asr_states_out = nullptr;
    ctx->asr_post_out = nullptr;
    return rc;
}

int32_t phb_joint_asr(phb_ctx *ctx, uint8_t *states, double *total)
{
    if (!ctx || !total || !states) return PHB_ERR_INVALID_ARGUMENT;
    if (ctx->is_aa) return aa_joint(ctx, states, total);
    ctx->asr_states_out = states;
    ctx->asr_post_out = nullptr;
    int rc = evaluate(ctx, true, false, total, true);
    ctx->asr_states_out = nullptr;
    return rc;
}

int32_t phb_score_nni(phb_ctx *ctx, int32_t n_moves, const int32_t *node_id, const uint8_t *left_ch, double *totals)
{
    if (!ctx || n_moves < 0 || (n_moves > 0 && (!node_id || !left_ch || !totals))) return PHB_ERR_INVALID_ARGUMENT;
    if (!ctx->tree.valid) return ctx->fail(PHB_ERR_STATE, "score_nni: set_tree first");
    if (!ctx->have_aln || !ctx->have_model)
        return ctx->fail(PHB_ERR_STATE, "score_nni: set_alignment and set_model first");
    ctx->stats.nni_path_moves = 0;
    ctx->stats.nni_path_steps = 0;
    ctx->stats.nni_tiles = 0;
    ctx->stats.nni_cache_bytes = ctx->stats.nni_read_bytes = 0;
    // Binary trees: walk v -> root against the cached CLVs of the base tree (nni.cuh).  20-state contexts re-evaluate
    // the path v -> root of every neighbour against the base tree's CLVs, any arity (aa_engine.cuh).
    if (ctx->is_aa && !(ctx->flags & PHB_FLAG_NNI_FULL)) return aa_score_nni(ctx, n_moves, node_id, left_ch, totals);
    if (!ctx->is_aa && nni_path_eligible(ctx)) return score_nni_paths(ctx, n_moves, node_id, left_ch, totals);
    // Otherwise (polytomies / unifurcations, PHB_FLAG_NNI_FULL): each neighbour is re-scheduled and scored
    // from scratch, which is what the reference's caller loop does.
    TreeSched base = ctx->tree;
    int rc = PHB_OK;
    for (int m = 0; m < n_moves && rc == PHB_OK; ++m) {
        int v = node_id[m];
        totals[m] = NAN;
        // NaN: unknown node, leaf or root (the reference's Err), and the two cases its indexing panics on -- the second
        // child of a unifurcation, a node without a sibling.  A unifurcation's ONLY child (left_ch == false) does move.
        if (v < 0 || v >= base.n || v == base.root || base.child_off[v + 1] - base.child_off[v] <= (left_ch[m] ? 1 : 0)) continue;
        // NNI::nni, src/tree.rs:1213-1242 on dense child lists
        std::vector<std::vector<int>> ch(base.n);
        for (int u = 0; u < base.n; ++u)
            ch[u].assign(base.children.begin() + base.child_off[u], base.children.begin() + base.child_off[u + 1]);
        int p = base.parent[v];
        int c1 = ch[v][left_ch[m] ? 1 : 0];
        int sib = -1;
        for (int x : ch[p])
            if (x != v) {
                sib = x;
                break;
            }
        if (sib < 0) continue;
        ch[v].erase(std::find(ch[v].begin(), ch[v].end(), c1));
        ch[p].erase(std::find(ch[p].begin(), ch[p].end(), sib));
        ch[p].push_back(c1);
        ch[v].push_back(sib);
        TreeSched &T = ctx->tree;
        T.children.clear();
        for (int u = 0; u < base.n; ++u) {
            T.child_off[u] = (int)T.children.size();
            T.children.insert(T.children.end(), ch[u].begin(), ch[u].end());
        }
        T.child_off[base.n] = (int)T.children.size();
        for (Device &d : ctx->devs) {
            cudaSetDevice(d.ordinal);
            cudaStreamSynchronize(d.stream);
        }
        rc = T.build(ctx->err);
        if (rc) break;
        ctx->tree_uploaded = false;
        rc = ctx->is_aa ? aa_evaluate(ctx, false, nullptr, nullptr, &totals[m]) : evaluate(ctx, false, false, &totals[m]);
    }
    for (Device &d : ctx->devs) {
        cudaSetDevice(d.ordinal);
        cudaStreamSynchronize(d.stream);
    }
    ctx->tree = base;
    ctx->tree_uploaded = false;
    return rc;
}

// The caller loop of any other rearrangement (SPR::spr src/tree/ops.rs:34-41, rerooting, ...):
//     let mut t = base.clone(); t.spr(e1, e2)?; t.log_likelihood(&model, &aln)?
// as ONE call: every candidate arrives as a flattened tree, is scheduled and scored in full against the resident
// alignment and model (no re-compression, no re-upload of the alignment), and the context's own tree is left as it
// was.  A candidate the library cannot schedule (malformed child lists, a leaf without a row) scores NaN.
int32_t phb_score_trees(phb_ctx *ctx, int32_t n_trees, const phb_tree_desc *trees, double *totals)
{
    if (!ctx || n_trees < 0 || (n_trees > 0 && (!trees || !totals))) return PHB_ERR_INVALID_ARGUMENT;
    if (!ctx->have_aln || !ctx->have_model) return ctx->fail(PHB_ERR_STATE, "score_trees: set_alignment and set_model first");
    const bool had_tree = ctx->tree.valid;
    for (Device &d : ctx->devs) {
        cudaSetDevice(d.ordinal);
        cudaStreamSynchronize(d.stream);
    }
    const TreeSched base = ctx->tree;
    int rc = PHB_OK;
    for (int i = 0; i < n_trees && rc == PHB_OK; ++i) {
        const phb_tree_desc &t = trees[i];
        totals[i] = NAN;
        int r2 = phb_set_tree(ctx, t.n_nodes, t.postorder, t.child_offsets, t.children, t.branch_len, t.leaf_row, t.root);
        if (r2 == PHB_OK) r2 = ctx->is_aa ? aa_evaluate(ctx, false, nullptr, nullptr, &totals[i]) : evaluate(ctx, false, false, &totals[i]);
        if (r2 == PHB_ERR_CUDA || r2 == PHB_ERR_OOM) rc = r2; // device trouble ends the batch; a bad candidate does not
        else if (r2 != PHB_OK) totals[i] = NAN;
    }
    for (Device &d : ctx->devs) {
        cudaSetDevice(d.ordinal);
        cudaStreamSynchronize(d.stream);
    }
    ctx->tree = base;
    ctx->tree.valid = had_tree;
    ctx->tree_uploaded = false;
    return rc;
}

int32_t phb_simulate_alignment(phb_ctx *ctx, int64_t n_patterns, int64_t global_offset, uint64_t seed)
{
    if (!ctx || n_patterns < 0) return PHB_ERR_INVALID_ARGUMENT;
    if (ctx->is_aa) return ctx->fail(PHB_ERR_STATE, "simulate_alignment: not available in a 20-state (PHB_FLAG_AMINO_ACID) context");
    if (!ctx->have_model || !ctx->tree.valid) return ctx->fail(PHB_ERR_STATE, "simulate: set_model and set_tree first");
    TreeSched &T = ctx->tree;
    const int n_leaves = (int)T.leaf_nodes.size();
    // rows = leaves in ascending dense id
    for (int v = 0; v < T.n; ++v) T.leaf_row[v] = -1;
    for (int i = 0; i < n_leaves; ++i) T.leaf_row[T.leaf_nodes[i]] = i;
    for (Device &d : ctx->devs) {
        cudaSetDevice(d.ordinal);
        cudaStreamSynchronize(d.stream);
    }
    int rc = T.build(ctx->err);
    if (rc) return rc;
    ctx->have_aln = false;
    rc = shard_and_alloc(ctx, n_leaves, n_patterns);
    if (rc) return rc;
    rc = upload_tree(ctx);
    if (rc) return rc;
    ctx->row_invalid_host.assign(std::max(n_leaves, 1), 0);
    ctx->row_first_invalid.clear();
    // pre-order (left-to-right) for the walk
    std::vector<int> pre;
    pre.reserve(T.n);
    {
        std::vector<int> st(1, T.root);
        while (!st.empty()) {
            int v = st.back();
            st.pop_back();
            pre.push_back(v);
            for (int k = T.child_off[v + 1] - 1; k >= T.child_off[v]; --k) st.push_back(T.children[k]);
        }
    }
    const ModelDev &M = ctx->model_lit;
    for (Device &d : ctx->devs) {
        CU(cudaSetDevice(d.ordinal));
        CU(d.Pmat.ensure(sizeof(double) * 16 * (size_t)T.n * M.n_cat));
        CU(d.sim_pre.ensure(sizeof(int) * T.n));
        CU(d.sim_parent.ensure(sizeof(int) * T.n));
        CU(d.sim_depth.ensure(sizeof(int) * T.n));
        CU(d.sim_leafrow.ensure(sizeof(int) * T.n));
        CU(d.sim_stack.ensure((size_t)(T.max_depth + 1) * (size_t)d.P_pad));
        CU(cudaMemcpyAsync(d.sim_pre.p, pre.data(), sizeof(int) * T.n, cudaMemcpyHostToDevice, d.stream));
        CU(cudaMemcpyAsync(d.sim_parent.p, T.parent.data(), sizeof(int) * T.n, cudaMemcpyHostToDevice, d.stream));
        CU(cudaMemcpyAsync(d.sim_depth.p, T.depth.data(), sizeof(int) * T.n, cudaMemcpyHostToDevice, d.stream));
        CU(cudaMemcpyAsync(d.sim_leafrow.p, T.leaf_row.data(), sizeof(int) * T.n, cudaMemcpyHostToDevice, d.stream));
        {
            int nthreads = T.n * M.n_cat;
            PHB_LAUNCH(transition_kernel, dim3((nthreads + 127) / 128), dim3(128), d.stream, d.blen.as<double>(), T.n, M,
                       d.Pmat.as<double>());
            CHECK_LAUNCH();
        }
        if (d.P > 0) {
            PHB_LAUNCH(simulate_kernel, dim3((unsigned)((d.P + 127) / 128)), dim3(128), d.stream, d.sim_pre.as<int32_t>(),
                       d.sim_parent.as<int32_t>(), d.sim_depth.as<int32_t>(), d.sim_leafrow.as<int32_t>(), T.n, T.root, M,
                       d.Pmat.as<double>(), d.sim_stack.as<uint8_t>(), d.tips.as<uint8_t>(), d.P_pad, d.P,
                       global_offset + d.p_begin, seed);
            CHECK_LAUNCH();
        }
        for (int r0 = 0; r0 < n_leaves && d.P_pad > d.P; r0 += 32768) {
            dim3 grid((unsigned)((d.P_pad - d.P + 255) / 256), (unsigned)std::min(32768, n_leaves - r0));
            PHB_LAUNCH(fill_pad_kernel, grid, dim3(256), d.stream, d.tips.as<uint8_t>(), r0, d.P, d.P_pad);
            CHECK_LAUNCH();
        }
        PHB_LAUNCH(mult_to_double_kernel, dim3((unsigned)std::min<int64_t>((d.P_pad + 255) / 256, 4096)), dim3(256),
                   d.stream, (const uint64_t *)nullptr, d.P, d.P_pad, d.mult.as<double>());
        CHECK_LAUNCH();
        CU(cudaStreamSynchronize(d.stream)); // `pre` is a local
        d.sim_stack.release();
    }
    ctx->have_aln = true;
    return PHB_OK;
}

int32_t phb_get_alignment(phb_ctx *ctx, uint8_t *bytes, int32_t layout, int64_t first_pattern, int64_t n_patterns)
{
    if (!ctx || !bytes || first_pattern < 0 || n_patterns < 0) return PHB_ERR_INVALID_ARGUMENT;
    if (!ctx->have_aln) return ctx->fail(PHB_ERR_STATE, "get_alignment: no alignment set");
    if (first_pattern + n_patterns > ctx->n_patterns || (layout != 0 && layout != 1))
        return ctx->fail(PHB_ERR_INVALID_ARGUMENT, "get_alignment: range or layout out of bounds");
    if (ctx->is_aa) return aa_get_alignment(ctx, bytes, layout, first_pattern, n_patterns);
    const int n_rows = ctx->n_rows;
    for (Device &d : ctx->devs) {
        int64_t lo = std::max(first_pattern, d.p_begin), hi = std::min(first_pattern + n_patterns, d.p_begin + d.P);
        if (hi <= lo || n_rows == 0) continue;
        int64_t cnt = hi - lo;
        CU(cudaSetDevice(d.ordinal));
        CU(d.staging.ensure((size_t)cnt * n_rows));
        for (int r0 = 0; r0 < n_rows; r0 += 32768) {
            dim3 grid((unsigned)std::min<int64_t>((cnt + 255) / 256, 4096), (unsigned)std::min(32768, n_rows - r0));
            PHB_LAUNCH(unpack_kernel, grid, dim3(256), d.stream, d.tips.as<uint8_t>(), n_rows, r0, d.P_pad,
                       lo - d.p_begin, cnt, (int)layout, d.staging.as<uint8_t>());
            CHECK_LAUNCH();
        }
        if (layout == PHB_LAYOUT_PATTERN_MAJOR)
            CU(cudaMemcpyAsync(bytes + (size_t)(lo - first_pattern) * n_rows, d.staging.p, (size_t)cnt * n_rows,
                               cudaMemcpyDeviceToHost, d.stream));
        else
            CU(cudaMemcpy2DAsync(bytes + (lo - first_pattern), (size_t)n_patterns, d.staging.p, (size_t)cnt, (size_t)cnt,
                                 (size_t)n_rows, cudaMemcpyDeviceToHost, d.stream));
        CU(cudaStreamSynchronize(d.stream));
    }
    return PHB_OK;
}

int32_t phb_get_stats(phb_ctx *ctx, phb_stats *out)
{
    if (!ctx || !out) return PHB_ERR_INVALID_ARGUMENT;
    ctx->stats.kernel_launches = ctx->launches;
    *out = ctx->stats;
    return PHB_OK;
}

int32_t phb_timer_start(phb_ctx *ctx)
{
    if (!ctx) return PHB_ERR_INVALID_ARGUMENT;
    for (Device &d : ctx->devs) {
        CU(cudaSetDevice(d.ordinal));
        if (!d.tm_a) {
            CU(cudaEventCreate(&d.tm_a));
            CU(cudaEventCreate(&d.tm_b));
        }
        CU(cudaEventRecord(d.tm_a, d.stream));
    }
    return PHB_OK;
}

int32_t phb_timer_stop(phb_ctx *ctx, double *ms)
{
    if (!ctx || !ms) return PHB_ERR_INVALID_ARGUMENT;
    double worst = 0.0;
    for (Device &d : ctx->devs) {
        if (!d.tm_a) return ctx->fail(PHB_ERR_STATE, "timer_stop without timer_start");
        CU(cudaSetDevice(d.ordinal));
        CU(cudaEventRecord(d.tm_b, d.stream));
    }
    for (Device &d : ctx->devs) {
        CU(cudaSetDevice(d.ordinal));
        CU(cudaStreamSynchronize(d.stream));
        float t = 0;
        cudaEventElapsedTime(&t, d.tm_a, d.tm_b);
        worst = std::max(worst, (double)t);
    }
    *ms = worst;
    return PHB_OK;
}

int32_t phb_transition_matrices(phb_ctx *ctx, int32_t n_t, const double *t, double *out)
{
    if (!ctx || n_t < 0 || (n_t > 0 && (!t || !out))) return ctx ? ctx->fail(PHB_ERR_INVALID_ARGUMENT, "transition_matrices: null/negative argument") : PHB_ERR_INVALID_ARGUMENT;
    if (!ctx->have_model) return ctx->fail(PHB_ERR_STATE, "set_model must precede transition_matrices");
    if (n_t == 0) return PHB_OK;
    if (ctx->is_aa) return aa_transition_matrices(ctx, n_t, t, out);
    Device &d = ctx->devs[0];
    const ModelDev &M = ctx->model_lit;
    CU(cudaSetDevice(d.ordinal));
    const size_t n_out = (size_t)n_t * M.n_cat * 16;
    CU(d.scratch.ensure(sizeof(double) * ((size_t)n_t + n_out)));
    double *d_t = d.scratch.as<double>(), *d_P = d_t + n_t;
    CU(cudaMemcpyAsync(d_t, t, sizeof(double) * n_t, cudaMemcpyHostToDevice, d.stream));
    int nthreads = n_t * M.n_cat;
    PHB_LAUNCH(transition_kernel, dim3((nthreads + 127) / 128), dim3(128), d.stream, d_t, n_t, M, d_P);
    CHECK_LAUNCH();
    CU(cudaMemcpyAsync(out, d_P, sizeof(double) * n_out, cudaMemcpyDeviceToHost, d.stream));
    CU(cudaStreamSynchronize(d.stream));
    return PHB_OK;
}

// ---- allocation-free variant of phb_transition_matrices ---------------------------------------------------------
// The staging memory of transition_kernel (branch lengths in, matrices out) lives in a file-scope device array instead
// of the context's grow-on-demand scratch buffer, so the call performs no cudaMalloc / cudaFree.  Every GPU that loads
// the library holds the array until exit.  In host code the array's NAME is a host-side stand-in, not a device
// address: the kernel gets the pointer cudaGetSymbolAddress returns, resolved once per device with that device current.
static __device__ __align__(256) unsigned char g_noalloc_scratch[PHB_NOALLOC_SCRATCH_BYTES];
constexpr int NOALLOC_MAX_DEVICES = 64;
static void *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};
static std::mutex g_noalloc_mu;

// `ordinal` must be the current device
static cudaError_t noalloc_scratch_ptr(int ordinal, void **out)
{
    if (ordinal < 0 || ordinal >= NOALLOC_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[ordinal]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_noalloc_ptr[ordinal] = p;
    }
    *out = g_noalloc_ptr[ordinal];
    return cudaSuccess;
}

int32_t phb_transition_matrices_noalloc(phb_ctx *ctx, int32_t n_t, const double *t, double *out)
{
    if (!ctx || n_t < 0 || (n_t > 0 && (!t || !out)))
        return ctx ? ctx->fail(PHB_ERR_INVALID_ARGUMENT, "transition_matrices_noalloc: null/negative argument") : PHB_ERR_INVALID_ARGUMENT;
    if (ctx->is_aa) return ctx->fail(PHB_ERR_STATE, "transition_matrices_noalloc: not available in a 20-state (PHB_FLAG_AMINO_ACID) context");
    if (!ctx->have_model) return ctx->fail(PHB_ERR_STATE, "set_model must precede transition_matrices_noalloc");
    if (n_t == 0) return PHB_OK;
    Device &d = ctx->devs[0];
    const ModelDev &M = ctx->model_lit;
    const size_t n_out = (size_t)n_t * M.n_cat * 16;
    const size_t need = sizeof(double) * ((size_t)n_t + n_out);
    if (need > (size_t)PHB_NOALLOC_SCRATCH_BYTES)
        return ctx->fail(PHB_ERR_INVALID_ARGUMENT,
                         "transition_matrices_noalloc: %d branch lengths x %d categories need %zu bytes of staging, the fixed device "
                         "array holds %zu (at most %zu branch lengths for this model); use phb_transition_matrices",
                         (int)n_t, M.n_cat, need, (size_t)PHB_NOALLOC_SCRATCH_BYTES,
                         (size_t)PHB_NOALLOC_SCRATCH_BYTES / (sizeof(double) * (1 + 16 * (size_t)M.n_cat)));
    CU(cudaSetDevice(d.ordinal));
    void *base = nullptr;
    CU(noalloc_scratch_ptr(d.ordinal, &base));
    // nothing has to start at zero: the kernel reads the n_t branch lengths copied in below and writes every one of the
    // n_out matrix entries that is copied out; whatever an earlier call left beyond them is never read
    double *d_t = static_cast<double *>(base), *d_P = d_t + n_t;
    CU(cudaMemcpyAsync(d_t, t, sizeof(double) * n_t, cudaMemcpyHostToDevice, d.stream));
    int nthreads = n_t * M.n_cat;
    PHB_LAUNCH(transition_kernel, dim3((nthreads + 127) / 128), dim3(128), d.stream, d_t, n_t, M, d_P);
    CHECK_LAUNCH();
    CU(cudaMemcpyAsync(out, d_P, sizeof(double) * n_out, cudaMemcpyDeviceToHost, d.stream));
    CU(cudaStreamSynchronize(d.stream));
    return PHB_OK;
}

int32_t phb_measure_fp64_tflops(phb_ctx *ctx, double *tflops)
{
    if (!ctx || !tflops) return PHB_ERR_INVALID_ARGUMENT;
    Device &d = ctx->devs[0];
    CU(cudaSetDevice(d.ordinal));
    const int blocks = 148 * 8, threads = 256, iters = 4096;
    CU(d.scratch.ensure(sizeof(double) * blocks * threads));
    cudaEvent_t a, b;
    CU(cudaEventCreate(&a));
    CU(cudaEventCreate(&b));
    for (int rep = 0; rep < 2; ++rep) {
        CU(cudaEventRecord(a, d.stream));
        PHB_LAUNCH(fp64_fma_kernel, dim3(blocks), dim3(threads), d.stream, d.scratch.as<double>(), iters);
        CHECK_LAUNCH();
        CU(cudaEventRecord(b, d.stream));
        CU(cudaStreamSynchronize(d.stream));
    }
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    *tflops = 2.0 * 8.0 * (double)iters * blocks * threads / ((double)ms * 1e-3) / 1e12;
    return PHB_OK;
}

} // extern "C"
