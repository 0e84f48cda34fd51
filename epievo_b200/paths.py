"""Host-side containers for evolutionary histories, trees and the local-paths text
format.

``PathSet`` is the flat *node-major CSR* exchange format the C-ABI takes
(include/epievo_b200.h): one row per (node, site), row = node*n_sites + site, node 0
(the root) empty.  It mirrors the reference's ``vector<vector<Path>>`` of
``struct Path {init_state, tot_time, jumps}`` (reference src/libepievo/Path.hpp:33-57)
with ``tot_time`` held per node, because every row on a branch carries the same value
(asserted at reference Segment.cpp:76 and Path.cpp:168).

``read_paths`` / ``write_paths`` speak the reference's local-paths text format
(reader Path.cpp:123-148, writer Path.cpp:62-71 and
prog/epievo_est_params_histories.cpp:56-75): ``NODE:<name>`` headers, then one line
per site ``<site>\\t<init_state>\\t<tot_time>\\t<jump>\\t...`` with a trailing tab and
``max_digits10`` doubles.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Sequence

import numpy as np


@dataclass
class PathSet:
    n_sites: int
    n_nodes: int
    init_state: np.ndarray            # uint8  [n_nodes*n_sites]
    row_off: np.ndarray               # uint64 [n_nodes*n_sites + 1]
    jumps: np.ndarray                 # float64 [row_off[-1]]
    tot_time: np.ndarray              # float64 [n_nodes]
    node_names: List[str] = field(default_factory=list)

    def __post_init__(self):
        self.init_state = np.ascontiguousarray(self.init_state, dtype=np.uint8)
        self.row_off = np.ascontiguousarray(self.row_off, dtype=np.uint64)
        self.jumps = np.ascontiguousarray(self.jumps, dtype=np.float64)
        self.tot_time = np.ascontiguousarray(self.tot_time, dtype=np.float64)
        n_rows = self.n_sites * self.n_nodes
        if self.init_state.shape != (n_rows,):
            raise ValueError("init_state must have n_nodes*n_sites entries")
        if self.row_off.shape != (n_rows + 1,):
            raise ValueError("row_off must have n_nodes*n_sites+1 entries")
        if int(self.row_off[-1]) != self.jumps.shape[0]:
            raise ValueError("row_off[-1] must equal len(jumps)")
        if self.tot_time.shape != (self.n_nodes,):
            raise ValueError("tot_time must have n_nodes entries")

    # -- row access ---------------------------------------------------------------
    def row(self, node: int, site: int) -> np.ndarray:
        r = node * self.n_sites + site
        return self.jumps[int(self.row_off[r]):int(self.row_off[r + 1])]

    def init(self, node: int, site: int) -> int:
        return int(self.init_state[node * self.n_sites + site])

    def row_len(self) -> np.ndarray:
        """Jump count of every row, shape [n_nodes, n_sites]."""
        return np.diff(self.row_off).astype(np.int64).reshape(self.n_nodes, self.n_sites)

    def row_code(self) -> np.ndarray:
        """Compact per-row code of epv_upload_paths_compact: jump count | init_state << 15."""
        n = np.diff(self.row_off)
        if n.size and int(n.max()) > 0x7fff:
            raise ValueError("a row holds more than 32767 jumps: use the CSR format")
        return (n.astype(np.uint16) | (self.init_state.astype(np.uint16) << np.uint16(15)))

    @classmethod
    def from_row_code(cls, n_sites, n_nodes, code, jumps, tot_time, node_names=()):
        code = np.asarray(code, dtype=np.uint16)
        off = np.zeros(code.size + 1, np.uint64)
        np.cumsum(code & np.uint16(0x7fff), dtype=np.uint64, out=off[1:])
        return cls(n_sites, n_nodes, (code >> np.uint16(15)).astype(np.uint8), off, jumps, tot_time,
                   list(node_names))

    def end_state(self) -> np.ndarray:
        """State at the end of every row (Path::end_state, Path.hpp:53-55)."""
        st = self.init_state.reshape(self.n_nodes, self.n_sites).astype(np.int64)
        return ((st + self.row_len()) & 1).astype(np.uint8)

    def copy(self) -> "PathSet":
        return PathSet(self.n_sites, self.n_nodes, self.init_state.copy(),
                       self.row_off.copy(), self.jumps.copy(), self.tot_time.copy(),
                       list(self.node_names))

    def site_slice(self, lo: int, hi: int) -> "PathSet":
        """Sites [lo, hi) of every node, as a new PathSet (sharding helper)."""
        n = hi - lo
        init = np.empty(self.n_nodes * n, np.uint8)
        lens = np.empty(self.n_nodes * n, np.int64)
        chunks = []
        for b in range(self.n_nodes):
            r0 = b * self.n_sites + lo
            init[b * n:(b + 1) * n] = self.init_state[r0:r0 + n]
            lens[b * n:(b + 1) * n] = np.diff(self.row_off[r0:r0 + n + 1].astype(np.int64))
            chunks.append(self.jumps[int(self.row_off[r0]):int(self.row_off[r0 + n])])
        off = np.zeros(self.n_nodes * n + 1, np.uint64)
        np.cumsum(lens, out=off[1:].view(np.int64))
        return PathSet(n, self.n_nodes, init, off, np.concatenate(chunks) if chunks
                       else np.zeros(0), self.tot_time.copy(), list(self.node_names))

    @staticmethod
    def concat_sites(parts: Sequence["PathSet"]) -> "PathSet":
        """Concatenate PathSets along the site axis (same nodes / branch lengths)."""
        n_nodes = parts[0].n_nodes
        n = sum(p.n_sites for p in parts)
        init = np.concatenate([p.init_state.reshape(n_nodes, -1) for p in parts], axis=1)
        lens = np.concatenate([p.row_len() for p in parts], axis=1)
        chunks = []
        for b in range(n_nodes):
            for p in parts:
                chunks.append(p.jumps[int(p.row_off[b * p.n_sites]):int(p.row_off[(b + 1) * p.n_sites])])
        off = np.zeros(n_nodes * n + 1, np.uint64)
        np.cumsum(lens.reshape(-1), out=off[1:].view(np.int64))
        return PathSet(n, n_nodes, init.reshape(-1), off,
                       np.concatenate(chunks) if chunks else np.zeros(0),
                       parts[0].tot_time.copy(), list(parts[0].node_names))

    @staticmethod
    def from_rows(n_sites: int, n_nodes: int, init_state, rows, tot_time,
                  node_names: Sequence[str] = ()) -> "PathSet":
        """rows[node][site] is a sequence of jump times (node 0 may be empty/None)."""
        init = np.zeros(n_nodes * n_sites, np.uint8)
        off = np.zeros(n_nodes * n_sites + 1, np.uint64)
        flat: List[float] = []
        k = 0
        for b in range(n_nodes):
            for s in range(n_sites):
                r = b * n_sites + s
                off[r] = k
                if b > 0:
                    init[r] = init_state[b][s]
                    jr = rows[b][s]
                    flat.extend(jr)
                    k += len(jr)
        off[-1] = k
        return PathSet(n_sites, n_nodes, init, off, np.asarray(flat, np.float64),
                       np.asarray(tot_time, np.float64), list(node_names))

    def same_as(self, other: "PathSet", rtol: float = 0.0) -> bool:
        if (self.n_sites, self.n_nodes) != (other.n_sites, other.n_nodes):
            return False
        n = self.n_sites
        if not np.array_equal(self.init_state[n:], other.init_state[n:]):
            return False
        if not np.array_equal(self.row_off, other.row_off):
            return False
        if rtol == 0.0:
            return bool(np.array_equal(self.jumps, other.jumps)
                        and np.array_equal(self.tot_time[1:], other.tot_time[1:]))
        return bool(np.allclose(self.jumps, other.jumps, rtol=rtol, atol=0.0)
                    and np.allclose(self.tot_time[1:], other.tot_time[1:], rtol=rtol))


def read_paths(path_file: str) -> PathSet:
    """Parse a local-paths text file (reference Path.cpp:123-148)."""
    names: List[str] = []
    init: List[List[int]] = []
    rows: List[List[List[float]]] = []
    tot: List[float] = []
    with open(path_file) as fh:
        for line in fh:
            line = line.rstrip("\n")
            if len(line) > 4 and line.startswith("NODE"):
                names.append(line[line.index(":") + 1:])
                init.append([])
                rows.append([])
                tot.append(0.0)
            elif line.strip():
                tok = line.split()
                init[-1].append(int(tok[1]))
                tot[-1] = float(tok[2])
                rows[-1].append([float(x) for x in tok[3:]])
    n_nodes = len(names)
    if n_nodes < 2:
        raise RuntimeError("cannot read: " + path_file)
    n_sites = len(rows[1])
    for b in range(1, n_nodes):
        if len(rows[b]) != n_sites:
            raise RuntimeError("ragged node block in: " + path_file)
    return PathSet.from_rows(n_sites, n_nodes, init, rows, tot, names)


def read_paths_native(path_file: str, n_threads: int = 0) -> PathSet:
    """read_paths through the library's multi-threaded parser (epv_pathfile_read)."""
    import ctypes as C

    from . import _lib as L
    lib = L.load()
    h = C.c_void_p()
    ns, nn, nj = C.c_uint64(), C.c_uint64(), C.c_uint64()
    if lib.epv_pathfile_read(path_file.encode(), n_threads, C.byref(h), C.byref(ns), C.byref(nn), C.byref(nj)):
        raise RuntimeError(lib.epv_io_last_error().decode())
    try:
        n_rows = ns.value * nn.value
        init = np.zeros(n_rows, np.uint8)
        off = np.zeros(n_rows + 1, np.uint64)
        jumps = np.zeros(max(1, nj.value))
        tot = np.zeros(nn.value)
        if lib.epv_pathfile_get(h, init.ctypes.data_as(L.u8p), off.ctypes.data_as(L.u64p),
                                jumps.ctypes.data_as(L.dp), tot.ctypes.data_as(L.dp)):
            raise RuntimeError(lib.epv_io_last_error().decode())
        names = [lib.epv_pathfile_node_name(h, b).decode() for b in range(nn.value)]
    finally:
        lib.epv_pathfile_free(h)
    return PathSet(ns.value, nn.value, init, off, jumps[:nj.value].copy(), tot, names)


def write_paths_native(path_file: str, ps: PathSet, node_names: Sequence[str] = (),
                       n_threads: int = 0) -> None:
    """write_paths through the library's multi-threaded formatter (epv_pathfile_write)."""
    import ctypes as C

    from . import _lib as L
    lib = L.load()
    names = list(node_names) or list(ps.node_names) or ["node_%d" % i for i in range(ps.n_nodes)]
    arr = (C.c_char_p * ps.n_nodes)(*[n.encode() for n in names])
    j = ps.jumps if len(ps.jumps) else np.zeros(1)
    if lib.epv_pathfile_write(path_file.encode(), ps.n_sites, ps.n_nodes, arr,
                              ps.init_state.ctypes.data_as(L.u8p), ps.row_off.ctypes.data_as(L.u64p),
                              j.ctypes.data_as(L.dp), ps.tot_time.ctypes.data_as(L.dp), n_threads):
        raise RuntimeError(lib.epv_io_last_error().decode())


def write_paths_binary(path_file: str, ps: PathSet, node_names: Sequence[str] = (),
                       n_threads: int = 0) -> None:
    """The binary sidecar form of a local-paths file (epv_pathfile_write_binary; read_paths_native reads both
    forms, the read_paths drop-in prefers `<file>.epvb` next to a text file that is not newer)."""
    import ctypes as C

    from . import _lib as L
    lib = L.load()
    names = list(node_names) or list(ps.node_names) or ["node_%d" % i for i in range(ps.n_nodes)]
    arr = (C.c_char_p * ps.n_nodes)(*[n.encode() for n in names])
    j = ps.jumps if len(ps.jumps) else np.zeros(1)
    if lib.epv_pathfile_write_binary(path_file.encode(), ps.n_sites, ps.n_nodes, arr,
                                     ps.init_state.ctypes.data_as(L.u8p), ps.row_off.ctypes.data_as(L.u64p),
                                     j.ctypes.data_as(L.dp), ps.tot_time.ctypes.data_as(L.dp), n_threads):
        raise RuntimeError(lib.epv_io_last_error().decode())


def _fmt(x: float) -> str:
    # operator<<(ostream, Path) sets precision(max_digits10) = 17 significant digits in
    # the default (%g-like) float format (Path.cpp:62-71).
    return "%.17g" % x


def write_paths(path_file: str, ps: PathSet, node_names: Sequence[str] = ()) -> None:
    """Write the local-paths text format byte-compatibly with the reference writers
    (prog/epievo_est_params_histories.cpp:56-75, 281-284)."""
    names = list(node_names) or list(ps.node_names) or [
        "node_%d" % i for i in range(ps.n_nodes)]
    with open(path_file, "w") as out:
        out.write("NODE:%s\n" % names[0])
        for b in range(1, ps.n_nodes):
            out.write("NODE:%s\n" % names[b])
            tt = _fmt(float(ps.tot_time[b]))
            for s in range(ps.n_sites):
                r = b * ps.n_sites + s
                j = ps.jumps[int(ps.row_off[r]):int(ps.row_off[r + 1])]
                out.write("%d\t%d\t%s\t" % (s, ps.init_state[r], tt))
                out.write("".join(_fmt(float(x)) + "\t" for x in j))
                out.write("\n")


# -------------------------------------------------------------------------------------
@dataclass
class Tree:
    """Preorder arrays of a phylogenetic tree, as the reference's TreeHelper holds them
    (TreeHelper.hpp:38-57, TreeHelper.cpp:43-60; get_parent_id
    PhyloTreePreorder.cpp:98-107).  Branch b is the edge above node b."""
    subtree_sizes: np.ndarray         # uint64 [n_nodes]
    parent_ids: np.ndarray            # uint64 [n_nodes]
    branches: np.ndarray              # float64 [n_nodes]
    node_names: List[str]

    @property
    def n_nodes(self) -> int:
        return int(self.subtree_sizes.shape[0])

    def is_leaf(self, node: int) -> bool:
        return int(self.subtree_sizes[node]) == 1

    def children(self, node: int) -> List[int]:
        out, ch = [], 1
        while ch < int(self.subtree_sizes[node]):
            out.append(node + ch)
            ch += int(self.subtree_sizes[node + ch])
        return out

    def leaves(self) -> List[int]:
        return [i for i in range(self.n_nodes) if self.is_leaf(i)]

    @staticmethod
    def from_preorder(subtree_sizes, branches, node_names=None) -> "Tree":
        ss = np.asarray(subtree_sizes, np.uint64)
        n = ss.shape[0]
        parent = np.zeros(n, np.uint64)
        for i in range(n):
            count = 1
            while count < int(ss[i]):
                child = i + count
                parent[child] = i
                count += int(ss[child])
        names = list(node_names) if node_names else ["node_%d" % i for i in range(n)]
        return Tree(ss, parent, np.asarray(branches, np.float64), names)

    @staticmethod
    def pair(evo_time: float) -> "Tree":
        """Two-node pseudo-tree of the pairwise programs (TreeHelper.cpp:53-60)."""
        return Tree(np.array([2, 1], np.uint64), np.array([0, 0], np.uint64),
                    np.array([0.0, evo_time]), ["root", "leaf"])

    @staticmethod
    def balanced(n_leaves: int, branch_len: float) -> "Tree":
        """Balanced binary tree with every branch the same length (C5 and
        the headline 8-leaf tree)."""
        sizes: List[int] = []
        names: List[str] = []

        def rec(leaves: int) -> int:
            idx = len(sizes)
            sizes.append(1)
            names.append("")
            if leaves > 1:
                sizes[idx] += rec(leaves // 2) + rec(leaves - leaves // 2)
            return sizes[idx]

        rec(n_leaves)
        br = [0.0] + [branch_len] * (len(sizes) - 1)
        leaf = 0
        for i, s in enumerate(sizes):
            if s == 1:
                names[i] = "L%d" % leaf
                leaf += 1
            else:
                names[i] = "N%d" % i
        return Tree.from_preorder(sizes, br, names)

    def to_newick(self) -> str:
        def rec(i: int) -> str:
            ch = self.children(i)
            s = ("(" + ",".join(rec(c) for c in ch) + ")") if ch else ""
            return "%s%s:%s" % (s, self.node_names[i], repr(float(self.branches[i])))
        return rec(0) + ";"


def parse_newick(text: str) -> Tree:
    """Minimal Newick reader giving the same preorder arrays as the reference's
    PhyloTree/PhyloTreePreorder (PhyloTree.cpp:189-203, 251-295;
    PhyloTreePreorder.cpp:32-76).  Unnamed nodes are named ``node_<k>``."""
    s = text.strip()
    if s.endswith(";"):
        s = s[:-1]
    pos = 0

    def node():
        nonlocal pos
        kids = []
        if pos < len(s) and s[pos] == "(":
            pos += 1
            while True:
                kids.append(node())
                if s[pos] == ",":
                    pos += 1
                    continue
                if s[pos] == ")":
                    pos += 1
                    break
                raise RuntimeError("bad tree near: " + s[pos:pos + 10])
        start = pos
        while pos < len(s) and s[pos] not in ",():":
            pos += 1
        name = s[start:pos].strip()
        bl = 0.0
        if pos < len(s) and s[pos] == ":":
            pos += 1
            start = pos
            while pos < len(s) and s[pos] not in ",()":
                pos += 1
            bl = float(s[start:pos])
        return (name, bl, kids)

    root = node()
    sizes: List[int] = []
    br: List[float] = []
    names: List[str] = []

    def walk(n) -> int:
        idx = len(sizes)
        sizes.append(1)
        br.append(n[1])
        names.append(n[0])
        for k in n[2]:
            sizes[idx] += walk(k)
        return sizes[idx]

    walk(root)
    count = 0
    for i, nm in enumerate(names):
        if not nm:
            names[i] = "node_%d" % count
            count += 1
    return Tree.from_preorder(sizes, br, names)


def read_tree(tree_file: str) -> Tree:
    with open(tree_file) as fh:
        return parse_newick(fh.read())
