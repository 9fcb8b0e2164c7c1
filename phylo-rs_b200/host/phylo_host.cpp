// phylo_host.cpp — implementation of the C++ host mirror (phylo_host.hpp) and its C facade (`phh_*`, bound
// by phylo_rs_b200/api.py).  All `file:line` citations are relative to /root/reference.
#include "phylo_host.hpp"

#include "../../include/phylo_b200.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <limits>
#include <sstream>

namespace phylo {

// =====================================================================================================
// errors
// =====================================================================================================
const char *error_kind_name(ErrorKind k)
{
    switch (k) {
    case ErrorKind::MissingBranchLength: return "MissingBranchLength";
    case ErrorKind::AlphabetMismatch: return "AlphabetMismatch";
    case ErrorKind::InvalidAlignment: return "InvalidAlignment";
    case ErrorKind::EigendecompFailure: return "EigendecompFailure";
    case ErrorKind::NumericalInstability: return "NumericalInstability";
    case ErrorKind::InvalidModelParameter: return "InvalidModelParameter";
    case ErrorKind::UnknownNode: return "UnknownNode";
    case ErrorKind::NoParent: return "NoParent";
    case ErrorKind::Newick: return "Newick";
    case ErrorKind::Backend: return "Backend";
    }
    return "Unknown";
}

// =====================================================================================================
// alphabet (src/alphabet.rs:23-65)
// =====================================================================================================
std::optional<std::size_t> Nucleotide::index_of(std::uint8_t c)
{
    switch (c) {
    case 'A': return 0;
    case 'C': return 1;
    case 'G': return 2;
    case 'T':
    case 'U': return 3;
    default: return std::nullopt;
    }
}

std::uint8_t Nucleotide::char_of(std::size_t i) { return static_cast<std::uint8_t>(CANONICAL[i]); }

std::optional<std::array<double, 4>> Nucleotide::profile(std::uint8_t c)
{
    // IUPAC code -> set of compatible states (bit i = state i); the profile spreads 1/|set| over the set,
    // except the gap which is all ones (alphabet.rs:44-64).
    unsigned mask = 0;
    switch (c) {
    case 'A': mask = 1; break;
    case 'C': mask = 2; break;
    case 'G': mask = 4; break;
    case 'T':
    case 'U': mask = 8; break;
    case 'R': mask = 1 | 4; break;
    case 'Y': mask = 2 | 8; break;
    case 'S': mask = 2 | 4; break;
    case 'W': mask = 1 | 8; break;
    case 'K': mask = 4 | 8; break;
    case 'M': mask = 1 | 2; break;
    case 'B': mask = 2 | 4 | 8; break;
    case 'D': mask = 1 | 4 | 8; break;
    case 'H': mask = 1 | 2 | 8; break;
    case 'V': mask = 1 | 2 | 4; break;
    case 'N': mask = 15; break;
    case GAP: return std::array<double, 4>{1.0, 1.0, 1.0, 1.0};
    default: return std::nullopt;
    }
    const int bits = __builtin_popcount(mask);
    const double v = bits == 1 ? 1.0 : bits == 2 ? 0.5 : bits == 3 ? 1.0 / 3.0 : 0.25;
    std::array<double, 4> out{};
    for (int i = 0; i < 4; ++i) out[i] = (mask >> i & 1u) ? v : 0.0;
    return out;
}

// =====================================================================================================
// tree
// =====================================================================================================
PhyloTree::PhyloTree(NodeID root_id) : root_(root_id)
{
    // SimpleRootedTree::new (tree.rs:196-209): an arena holding just the root
    nodes_.resize(root_id + 1);
    Node r;
    r.id = root_id;
    nodes_[root_id] = r;
}

const Node &PhyloTree::get_node(NodeID id) const
{
    if (!contains(id)) throw AsrError(ErrorKind::UnknownNode, "unknown node id " + std::to_string(id));
    return *nodes_[id];
}

Node &PhyloTree::node_mut(NodeID id)
{
    if (!contains(id)) throw AsrError(ErrorKind::UnknownNode, "unknown node id " + std::to_string(id));
    return *nodes_[id];
}

std::vector<NodeID> PhyloTree::get_node_ids() const
{
    std::vector<NodeID> out;
    for (NodeID i = 0; i < nodes_.size(); ++i)
        if (nodes_[i]) out.push_back(i);
    return out;
}

std::size_t PhyloTree::num_nodes() const
{
    std::size_t n = 0;
    for (const auto &x : nodes_) n += x.has_value();
    return n;
}

NodeID PhyloTree::next_id() const
{
    // tree.rs:234-239: the lowest vacant arena slot, else the arena length
    for (NodeID i = 0; i < nodes_.size(); ++i)
        if (!nodes_[i]) return i;
    return nodes_.size();
}

void PhyloTree::set_node(Node node)
{
    if (node.id >= nodes_.size()) nodes_.resize(node.id + 1);
    const NodeID id = node.id;
    nodes_[id] = std::move(node);
}

void PhyloTree::set_child(NodeID parent, NodeID child)
{
    node_mut(parent).children.push_back(child); // appended last (simple_rtree.rs:193-200)
    node_mut(child).parent = parent;
}

void PhyloTree::add_child(NodeID parent, Node node)
{
    const NodeID id = node.id;
    set_node(std::move(node));
    set_child(parent, id);
}

void PhyloTree::delete_edge(NodeID parent, NodeID child)
{
    auto &ch = node_mut(parent).children;
    ch.erase(std::remove(ch.begin(), ch.end(), child), ch.end());
    node_mut(child).parent.reset();
}

void PhyloTree::split_edge(std::pair<NodeID, NodeID> edge, Node node)
{
    // simple_rtree.rs:143-153: the new node takes the child's place at the END of the parent's list
    const NodeID id = node.id;
    set_node(std::move(node));
    auto &ch = node_mut(edge.first).children;
    ch.erase(std::remove(ch.begin(), ch.end(), edge.second), ch.end());
    set_child(edge.first, id);
    set_child(id, edge.second);
}

void PhyloTree::set_node_taxa(NodeID id, std::optional<std::string> taxa) { node_mut(id).taxa = std::move(taxa); }

void PhyloTree::set_edge_weight(NodeID child, std::optional<float> w) { node_mut(child).weight = w; }

std::optional<NodeID> PhyloTree::get_taxa_node_id(const std::string &taxa) const
{
    for (NodeID i = 0; i < nodes_.size(); ++i)
        if (nodes_[i] && nodes_[i]->taxa && *nodes_[i]->taxa == taxa) return i;
    return std::nullopt;
}

std::optional<float> PhyloTree::get_edge_weight(NodeID, NodeID child) const
{
    return get_node(child).weight; // the child's own weight (simple_rtree.rs:460-468)
}

std::vector<NodeID> PhyloTree::postord_ids(std::optional<NodeID> start) const
{
    // Left-to-right post-order.  The reference's DFSPostOrderIterator (iter.rs:112-138) re-pushes a node
    // in front of its untaken children; the (node, next-child) stack below visits in the same order.
    std::vector<NodeID> out;
    std::vector<std::pair<NodeID, std::size_t>> st;
    st.emplace_back(start.value_or(root_), 0);
    (void)get_node(st.back().first);
    while (!st.empty()) {
        auto &top = st.back();
        const auto &kids = get_node(top.first).children;
        if (top.second < kids.size()) {
            const NodeID c = kids[top.second++];
            st.emplace_back(c, 0);
        } else {
            out.push_back(top.first);
            st.pop_back();
        }
    }
    return out;
}

std::vector<NodeID> PhyloTree::preord_ids(std::optional<NodeID> start) const
{
    std::vector<NodeID> out, st{start.value_or(root_)};
    while (!st.empty()) {
        const NodeID v = st.back();
        st.pop_back();
        out.push_back(v);
        const auto &kids = get_node(v).children;
        for (auto it = kids.rbegin(); it != kids.rend(); ++it) st.push_back(*it);
    }
    return out;
}

void PhyloTree::nni(NodeID node_id, bool left_ch)
{
    // NNI::nni (tree.rs:1213-1242): left_ch == true moves children(v)[1]
    if (!contains(node_id)) throw AsrError(ErrorKind::UnknownNode, "unknown node id " + std::to_string(node_id));
    if (is_leaf(node_id) || node_id == root_)
        throw AsrError(ErrorKind::NoParent, "node " + std::to_string(node_id) + " is a leaf or the root");
    const NodeID parent = *get_node(node_id).parent;
    const auto kids = get_node(node_id).children;
    const std::size_t pick = left_ch ? 1 : 0;
    if (pick >= kids.size()) throw AsrError(ErrorKind::NoParent, "node has no child at the requested index");
    const NodeID moved = kids[pick];
    NodeID sibling = node_id;
    for (NodeID c : get_node(parent).children)
        if (c != node_id) {
            sibling = c;
            break;
        }
    if (sibling == node_id) throw AsrError(ErrorKind::NoParent, "node has no sibling");
    delete_edge(node_id, moved);
    delete_edge(parent, sibling);
    set_child(parent, moved);
    set_child(node_id, sibling);
}

std::string PhyloTree::to_newick() const
{
    // iterative writer: "(" children ")" label ":" weight
    std::string out;
    struct Frame {
        NodeID v;
        std::size_t k;
    };
    std::vector<Frame> st{{root_, 0}};
    while (!st.empty()) {
        Frame &f = st.back();
        const Node &n = get_node(f.v);
        if (f.k == 0 && !n.children.empty()) out += '(';
        if (f.k < n.children.size()) {
            if (f.k > 0) out += ',';
            const NodeID c = n.children[f.k++];
            st.push_back({c, 0});
            continue;
        }
        if (!n.children.empty()) out += ')';
        if (n.taxa) out += *n.taxa;
        if (n.weight) {
            std::ostringstream os;
            os.precision(9);
            os << *n.weight;
            out += ':' + os.str();
        }
        st.pop_back();
    }
    return out + ';';
}

// ---- Newick (tree/newick.rs:168-284) -------------------------------------------------------------------
namespace {
bool nwk_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\v' || c == '\f'; }
bool nwk_delim(char c)
{
    return nwk_space(c) || c == '(' || c == ')' || c == '[' || c == ']' || c == ',' || c == ':' || c == ';' || c == '\'';
}
} // namespace

PhyloTree PhyloTree::from_newick(std::string_view src)
{
    PhyloTree tree(0);
    NodeID current = tree.root_;
    std::vector<NodeID> stack;
    bool saw_content = false;
    std::size_t i = 0;
    const std::size_t n = src.size();
    auto skip_trivia = [&]() {
        for (;;) {
            while (i < n && nwk_space(src[i])) ++i;
            if (i < n && src[i] == '[') { // comments are dropped here (the annotation handler's business)
                const std::size_t e = src.find(']', i);
                if (e == std::string_view::npos) throw AsrError(ErrorKind::Newick, "unterminated comment");
                i = e + 1;
            } else
                return;
        }
    };
    auto read_token = [&]() {
        const std::size_t s = i;
        while (i < n && !nwk_delim(src[i])) ++i;
        return std::string(src.substr(s, i - s));
    };
    auto open_node = [&](NodeID parent) {
        Node nd;
        nd.id = tree.next_id();
        const NodeID id = nd.id;
        tree.set_node(std::move(nd));
        tree.set_child(parent, id);
        return id;
    };
    for (;;) {
        skip_trivia();
        if (i >= n || src[i] == ';') break;
        const char c = src[i];
        if (c == '(') {
            ++i;
            stack.push_back(current);
            current = open_node(current);
            saw_content = true;
        } else if (c == ',') {
            ++i;
            if (stack.empty()) throw AsrError(ErrorKind::Newick, "unbalanced parentheses");
            current = open_node(stack.back());
        } else if (c == ')') {
            ++i;
            if (stack.empty()) throw AsrError(ErrorKind::Newick, "unbalanced parentheses");
            current = stack.back();
            stack.pop_back();
        } else if (c == ':') {
            ++i;
            skip_trivia();
            const std::string tok = read_token();
            char *end = nullptr;
            const float w = std::strtof(tok.c_str(), &end); // branch lengths are f32 (tree.rs:45)
            if (tok.empty() || end != tok.c_str() + tok.size())
                throw AsrError(ErrorKind::Newick, "invalid branch length '" + tok + "'");
            tree.node_mut(current).weight = w;
        } else if (c == '\'') {
            ++i;
            std::string label;
            for (;;) {
                if (i >= n) throw AsrError(ErrorKind::Newick, "unterminated quote");
                if (src[i] == '\'') {
                    if (i + 1 < n && src[i + 1] == '\'') {
                        label += '\'';
                        i += 2;
                    } else {
                        ++i;
                        break;
                    }
                } else
                    label += src[i++];
            }
            tree.node_mut(current).taxa = label;
            saw_content = true;
        } else {
            const std::string label = read_token();
            if (label.empty()) throw AsrError(ErrorKind::Newick, "invalid character");
            tree.node_mut(current).taxa = label;
            saw_content = true;
        }
    }
    if (!stack.empty()) throw AsrError(ErrorKind::Newick, "unbalanced parentheses");
    if (!saw_content) throw AsrError(ErrorKind::Newick, "empty tree");
    return tree;
}

// ---- Yule (tree.rs:505-553) ----------------------------------------------------------------------------
namespace {
struct Xoshiro256ss {
    std::uint64_t s[4];
    explicit Xoshiro256ss(std::uint64_t seed)
    {
        for (auto &w : s) { // splitmix64
            seed += 0x9E3779B97F4A7C15ull;
            std::uint64_t z = seed;
            z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
            z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
            w = z ^ (z >> 31);
        }
    }
    static std::uint64_t rotl(std::uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
    std::uint64_t next()
    {
        const std::uint64_t r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
        s[2] ^= s[0];
        s[3] ^= s[1];
        s[1] ^= s[2];
        s[0] ^= s[3];
        s[2] ^= t;
        s[3] = rotl(s[3], 45);
        return r;
    }
    std::uint64_t below(std::uint64_t n) { return (std::uint64_t)(((unsigned __int128)next() * n) >> 64); }
    double unit() { return (double)(next() >> 11) * 0x1.0p-53; }
};
} // namespace

PhyloTree PhyloTree::yule(std::size_t num_taxa, std::uint64_t seed)
{
    PhyloTree tree(0);
    if (num_taxa < 3) return tree; // degenerate, as the reference
    for (NodeID id = 1; id <= 2; ++id) {
        Node nd;
        nd.id = id;
        nd.taxa = std::to_string(id - 1);
        tree.add_child(0, std::move(nd));
    }
    if (num_taxa < 4) return tree; // the reference returns the two-leaf tree for num_taxa == 3 (tree.rs:526-528)
    Xoshiro256ss rng(seed);
    std::vector<NodeID> leaves{1, 2};
    for (std::size_t i = 2; i < num_taxa; ++i) {
        const NodeID leaf = leaves[rng.below(leaves.size())];
        const NodeID parent = *tree.get_node(leaf).parent;
        Node split;
        split.id = tree.nodes_.size(); // the arena is dense here, so next_id() == its length
        const NodeID split_id = split.id;
        tree.split_edge({parent, leaf}, std::move(split));
        Node nl;
        nl.id = tree.nodes_.size();
        nl.taxa = std::to_string(i);
        const NodeID new_leaf = nl.id;
        tree.add_child(split_id, std::move(nl));
        leaves.push_back(new_leaf);
    }
    return tree;
}

void PhyloTree::assign_exponential_lengths(std::uint64_t seed, double mean, double lo, double hi)
{
    Xoshiro256ss rng(seed);
    for (NodeID id = 0; id < nodes_.size(); ++id) {
        if (!nodes_[id] || id == root_) continue;
        double x = -mean * std::log(1.0 - rng.unit());
        x = std::min(std::max(x, lo), hi);
        nodes_[id]->weight = static_cast<float>(x);
    }
}

} // namespace phylo
