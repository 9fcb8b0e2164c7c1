"""CPU ORACLE — host-side restatement (test infrastructure, NOT product code).

Pure-Python restatement of the reference's tree arena, traversals, Newick id assignment, FASTA
parsing, column compression, Yule simulator and NNI edit — the pieces of sriram98v/phylo-rs that
*produce the schedule and the inputs* for the likelihood path.  Small cases only (Python loops).
All `file:line` citations are relative to /root/reference.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module.
"""
from __future__ import annotations

import numpy as np


class Node:
    """src/node.rs:54-78 — id, parent, ordered children, optional taxa label, optional f32 weight."""

    __slots__ = ("id", "parent", "children", "taxa", "weight")

    def __init__(self, nid):
        self.id = nid
        self.parent = None
        self.children = []
        self.taxa = None
        self.weight = None  # np.float32 or None


class RefTree:
    """SimpleRootedTree<String, f32, f32> (src/tree.rs:45, :99-134): arena Vec<Option<Node>>."""

    def __init__(self, root_id=0):
        # SimpleRootedTree::new, src/tree.rs:196-209
        self.root = root_id
        self.nodes = [None] * (root_id + 1)
        self.nodes[root_id] = Node(root_id)

    # -- arena ---------------------------------------------------------------------------
    def next_id(self):
        # src/tree.rs:234-239: lowest vacant slot, else arena length
        for i, n in enumerate(self.nodes):
            if n is None:
                return i
        return len(self.nodes)

    def set_node(self, node):
        # src/tree.rs:358-376
        if node.id >= len(self.nodes):
            self.nodes.extend([None] * (node.id - len(self.nodes) + 1))
        self.nodes[node.id] = node

    def get_node_ids(self):
        # src/tree.rs:354-356
        return [i for i, n in enumerate(self.nodes) if n is not None]

    def set_child(self, parent_id, child_id):
        # src/tree/simple_rtree.rs:193-200 (child appended last)
        self.nodes[parent_id].children.append(child_id)
        self.nodes[child_id].parent = parent_id

    def add_child(self, parent_id, node):
        # src/tree/simple_rtree.rs:57-66
        self.set_node(node)
        self.set_child(parent_id, node.id)

    def delete_edge(self, parent_id, child_id):
        # src/tree/simple_rtree.rs:126-133
        self.nodes[parent_id].children.remove(child_id)
        self.nodes[child_id].parent = None

    def split_edge(self, edge, node):
        # src/tree/simple_rtree.rs:143-153
        p_id, c_id = edge
        self.set_node(node)
        self.nodes[p_id].children.remove(c_id)
        self.set_child(p_id, node.id)
        self.set_child(node.id, c_id)

    def is_leaf(self, nid):
        # src/node/simple_rnode.rs:74-76: leaf == no children
        return len(self.nodes[nid].children) == 0

    def get_taxa_node_id(self, name):
        # src/tree/simple_rtree.rs:388-390 -> src/tree.rs:452-459 (scan by label equality)
        for n in self.nodes:
            if n is not None and n.taxa == name:
                return n.id
        return None

    def get_edge_weight(self, parent_id, child_id):
        # src/tree/simple_rtree.rs:460-468: the child's own weight
        return self.nodes[child_id].weight

    def set_edge_weight(self, child_id, w):
        self.nodes[child_id].weight = np.float32(w)

    def clone(self):
        t = RefTree.__new__(RefTree)
        t.root = self.root
        t.nodes = []
        for n in self.nodes:
            if n is None:
                t.nodes.append(None)
            else:
                m = Node(n.id)
                m.parent, m.children, m.taxa, m.weight = n.parent, list(n.children), n.taxa, n.weight
                t.nodes.append(m)
        return t

    # -- traversals ----------------------------------------------------------------------
    def postord_ids(self, start=None):
        """DFSPostOrderIterator, src/iter.rs:112-138 via src/tree.rs:764-772."""
        start = self.root if start is None else start
        taken = [False] * len(self.nodes)
        taken[start] = True
        stack = [start]  # front of the deque == end of this list
        out = []
        while stack:
            node = stack.pop()
            kids = []
            for ch in self.nodes[node].children:
                if not taken[ch]:
                    taken[ch] = True
                    kids.append(ch)
            if kids:
                stack.append(node)
                for ch in reversed(kids):
                    stack.append(ch)
            else:
                out.append(node)
        return out

    def preord_ids(self, start=None):
        """PreOrder::preord_ids, src/iter/node_iter.rs:137-161."""
        start = self.root if start is None else start
        stack = [start]
        out = []
        while stack:
            x = stack.pop()
            out.append(x)
            for ch in reversed(self.nodes[x].children):
                stack.append(ch)
        return out

    # -- edits ---------------------------------------------------------------------------
    def nni(self, node_id, left_ch):
        """NNI::nni, src/tree.rs:1213-1242.  Returns an error string or None."""
        if node_id >= len(self.nodes) or self.nodes[node_id] is None:
            return "UnknownNode"
        if self.is_leaf(node_id) or node_id == self.root:
            return "NoParent"
        p = self.nodes[node_id].parent
        ch = list(self.nodes[node_id].children)
        c1 = ch[1 if left_ch else 0]
        sib = [x for x in self.nodes[p].children if x != node_id][0]
        self.delete_edge(node_id, c1)
        self.delete_edge(p, sib)
        self.set_child(p, c1)
        self.set_child(node_id, sib)
        return None

    # -- output --------------------------------------------------------------------------
    def to_newick(self):
        def rec(v):
            n = self.nodes[v]
            s = ""
            if n.children:
                s += "(" + ",".join(rec(c) for c in n.children) + ")"
            if n.taxa is not None:
                s += n.taxa
            if n.weight is not None:
                s += ":" + repr(float(n.weight))
            return s

        import sys

        sys.setrecursionlimit(max(10000, 4 * len(self.nodes)))
        return rec(self.root) + ";"


def parse_newick(src):
    """parse_newick, src/tree/newick.rs:168-284: ids in encounter order via next_id, children in
    source order, ':len' parsed as f32 (:234-248).  Comments/quotes are not needed by fixtures."""
    if isinstance(src, bytes):
        src = src.decode()
    tree = RefTree(0)
    current = tree.root
    stack = []
    i, n = 0, len(src)
    delims = set("(),:;[]'")
    while i < n:
        c = src[i]
        if c.isspace():
            i += 1
            continue
        if c == ";":
            break
        if c == "(":
            i += 1
            stack.append(current)
            child = tree.next_id()
            tree.set_node(Node(child))
            tree.set_child(current, child)
            current = child
        elif c == ",":
            i += 1
            parent = stack[-1]
            sib = tree.next_id()
            tree.set_node(Node(sib))
            tree.set_child(parent, sib)
            current = sib
        elif c == ")":
            i += 1
            current = stack.pop()
        elif c == ":":
            i += 1
            while i < n and src[i].isspace():
                i += 1
            j = i
            while j < n and src[j] not in delims and not src[j].isspace():
                j += 1
            tree.nodes[current].weight = np.float32(src[i:j])
            i = j
        else:
            j = i
            while j < n and src[j] not in delims and not src[j].isspace():
                j += 1
            tree.nodes[current].taxa = src[i:j]
            i = j
    if stack:
        raise ValueError("UnbalancedParens")
    return tree


class RefAlignment:
    """Alignment, src/alignment.rs:5-10.  `seqs` keeps insertion order (the reference's HashMap
    order is random per process and only permutes rows inside a pattern)."""

    def __init__(self, seqs, width):
        self.seqs = seqs
        self.width = width

    @staticmethod
    def from_fasta_bytes(data):
        """src/alignment.rs:14-83."""
        seqs = {}
        cur_id, cur_seq = "", bytearray()
        for line in data.split(b"\n"):
            line = line.strip(b" \t\r\n\x0b\x0c")
            if not line:
                continue
            if line[:1] == b">":
                if cur_id:
                    if cur_id in seqs:
                        raise ValueError(f"InvalidAlignment: Duplicate taxon ID: {cur_id}")
                    seqs[cur_id] = bytes(cur_seq)
                header = line[1:]
                end = len(header)
                for k, b in enumerate(header):
                    if chr(b).isspace():
                        end = k
                        break
                raw = header[:end].decode("utf-8", "replace")
                if not raw:
                    raise ValueError("InvalidAlignment: Empty taxon ID in FASTA header")
                cur_id, cur_seq = raw, bytearray()
            else:
                cur_seq.extend(line.upper())
        if cur_id:
            if cur_id in seqs:
                raise ValueError(f"InvalidAlignment: Duplicate taxon ID: {cur_id}")
            seqs[cur_id] = bytes(cur_seq)
        if not seqs:
            raise ValueError("InvalidAlignment: Empty FASTA file")
        width = len(next(iter(seqs.values())))
        if not all(len(s) == width for s in seqs.values()):
            raise ValueError("InvalidAlignment: Ragged alignment")
        return RefAlignment(seqs, width)

    def compress_columns(self):
        """src/alignment.rs:93-125 -> (patterns [P][rows] uint8, site_to_pattern, multiplicity,
        leaf_order).  Patterns in first-appearance order."""
        leaf_order = list(self.seqs.keys())
        rows = len(leaf_order)
        if rows == 0 or self.width == 0:
            return np.zeros((0, rows), np.uint8), np.zeros(0, np.int64), np.zeros(0, np.uint64), leaf_order
        mat = np.frombuffer(b"".join(self.seqs[n] for n in leaf_order), dtype=np.uint8).reshape(rows, self.width)
        cols = np.ascontiguousarray(mat.T)
        idx = {}
        patterns, mult, s2p = [], [], []
        for i in range(self.width):
            key = cols[i].tobytes()
            j = idx.get(key)
            if j is None:
                j = len(patterns)
                idx[key] = j
                patterns.append(cols[i])
                mult.append(0)
            mult[j] += 1
            s2p.append(j)
        return (np.stack(patterns).astype(np.uint8), np.asarray(s2p, np.int64),
                np.asarray(mult, np.uint64), leaf_order)


def yule(num_taxa, rng):
    """Yule::yule, src/tree.rs:505-553, with an explicit rng (reference: unseeded thread_rng).
    `rng.integers(k)` picks the index into current_leaf_ids.  No branch lengths (as the reference)."""
    tree = RefTree(0)
    if num_taxa < 3:
        return tree
    for nid, label in ((1, "0"), (2, "1")):
        nd = Node(nid)
        tree.add_child(0, nd)
        tree.nodes[nid].taxa = label
    if num_taxa < 4:
        return tree
    leaves = [1, 2]
    for i in range(2, num_taxa):
        leaf = leaves[int(rng.integers(len(leaves)))]
        par = tree.nodes[leaf].parent
        split_id = len(tree.nodes)  # arena is dense here, next_id == len
        tree.split_edge((par, leaf), Node(split_id))
        new_id = len(tree.nodes)
        tree.add_child(split_id, Node(new_id))
        tree.nodes[new_id].taxa = str(i)
        leaves.append(new_id)
    return tree


def flatten(tree, leaf_order):
    """Dense re-indexing of the arena + the schedule arrays the C oracle consumes.

    Raises the reference's errors: a row naming no node -> AlphabetMismatch (likelihood.rs:174-176);
    a leaf without a row -> InvalidAlignment (likelihood.rs:99-103, raised lazily by the C code as
    leaf_row = -1)."""
    ids = tree.get_node_ids()
    dense = {nid: k for k, nid in enumerate(ids)}
    n = len(ids)
    leaf_id_map = []
    for name in leaf_order:
        nid = tree.get_taxa_node_id(name)
        if nid is None:
            raise ValueError(f"AlphabetMismatch: Taxon {name} in alignment not found in tree")
        leaf_id_map.append(nid)
    leaf_row = np.full(n, -1, np.int32)
    for r, nid in enumerate(leaf_id_map):
        if tree.is_leaf(nid) and leaf_row[dense[nid]] < 0:  # position() returns the first match
            leaf_row[dense[nid]] = r
    child_off = np.zeros(n + 1, np.int32)
    children = []
    parent = np.full(n, -1, np.int32)
    blen = np.zeros(n, np.float64)
    for nid in ids:
        nd = tree.nodes[nid]
        k = dense[nid]
        for c in nd.children:
            children.append(dense[c])
        child_off[k + 1] = len(children)
        if nd.parent is not None and nid != tree.root:
            parent[k] = dense[nd.parent]
        # f32 -> f64 widening, None -> 0.0 (likelihood.rs:115-118)
        blen[k] = float(nd.weight) if nd.weight is not None else 0.0
    post = np.asarray([dense[x] for x in tree.postord_ids()], np.int32)
    pre = np.asarray([dense[x] for x in tree.preord_ids()], np.int32)
    return dict(
        n_nodes=n, ids=np.asarray(ids, np.int64), postorder=post, preorder=pre, child_off=child_off,
        children=np.asarray(children, np.int32), parent=parent, blen=blen, leaf_row=leaf_row,
        root=dense[tree.root],
    )
