"""Synthetic inputs for the tree-likelihood path: trees, eigen-systems, rate
categories and alignments (an AliSim-equivalent sampler).

Nothing here is on the hot path: these are the *inputs* the reference's
kernels consume (SURVEY.md section 8(d) "Synthetic inputs").  The conventions
follow the reference so that arrays can be handed to either side unchanged:

* eigen-system layout: ``evec[m][x][i]`` row = state, col = eigen index;
  ``inv_evec[m][i][x]``; reversible models via a symmetric solve with
  ``U = Pi^-1/2 V`` and ``U^-1 = V^T Pi^1/2`` (model/modelmarkov.cpp:1555-1559).
* category index is mixture-major ``c = m*ncat + k`` (tree/phylokernelnew.h:1325).
* state codes: DNA 0-3, 4-17 ambiguity bit masks (code-3), 18 unknown;
  protein 0-19, 20 B, 21 Z, 22 J, 23 unknown.
* random trees: Yule-Harding topology, branch lengths Exp(mean 0.1) clamped to
  [0.001, 0.999] (utils/tools.cpp:1071-1073, 119-146).
"""
from __future__ import annotations

import math
import re
from dataclasses import dataclass, field

import numpy as np

DNA_UNKNOWN = 18
AA_UNKNOWN = 23


# --------------------------------------------------------------------------
# trees
# --------------------------------------------------------------------------
@dataclass
class Tree:
    """Unrooted tree as an adjacency structure.

    Leaves have ids ``0..ntaxa-1`` (= taxon index = alignment row), internal
    nodes ``ntaxa..``.  ``adj[u]`` is the ordered neighbour list of ``u``
    (order matters: the reference takes the first non-dad neighbour as "left",
    tree/phylokernelnew.h:1343-1350).  ``length[(u, v)]`` with ``u < v``.
    """

    ntaxa: int
    adj: list = field(default_factory=list)
    length: dict = field(default_factory=dict)
    names: list = field(default_factory=list)

    @property
    def nnodes(self) -> int:
        return len(self.adj)

    def is_leaf(self, u: int) -> bool:
        return u < self.ntaxa

    def get_length(self, u: int, v: int) -> float:
        return self.length[(u, v) if u < v else (v, u)]

    def set_length(self, u: int, v: int, t: float) -> None:
        self.length[(u, v) if u < v else (v, u)] = float(t)

    def branches(self):
        return sorted(self.length.keys())

    # -- newick ------------------------------------------------------------
    def to_newick(self, root_leaf: int = 0, digits: int = 10) -> str:
        """Newick string rooted (for display) at the neighbour of ``root_leaf``."""
        import sys
        sys.setrecursionlimit(max(10000, 4 * self.nnodes))

        def rec(u, dad):
            if self.is_leaf(u):
                s = self.names[u]
            else:
                s = "(" + ",".join(rec(v, u) for v in self.adj[u] if v != dad) + ")"
            return s + ":" + repr(round(self.get_length(u, dad), digits))

        c = self.adj[root_leaf][0]
        inner = [rec(v, c) for v in self.adj[c] if v != root_leaf]
        return ("(" + self.names[root_leaf] + ":" +
                repr(round(self.get_length(root_leaf, c), digits)) + "," + ",".join(inner) + ");")

    def to_newick_rooted(self, root_leaf: int, digits: int = 10) -> str:
        """Newick of a ROOTED tree whose root is marked by the leaf ``root_leaf``: the top node is that leaf's
        neighbour and the root leaf itself is left out (the reference adds it when it reads a tree whose top node
        has two children, tree/mtree.cpp:739-751)."""
        import sys
        sys.setrecursionlimit(max(10000, 4 * self.nnodes))

        def rec(u, dad):
            if self.is_leaf(u):
                s = self.names[u]
            else:
                s = "(" + ",".join(rec(v, u) for v in self.adj[u] if v != dad) + ")"
            return s + ":" + repr(round(self.get_length(u, dad), digits))

        c = self.adj[root_leaf][0]
        return "(" + ",".join(rec(v, c) for v in self.adj[c] if v != root_leaf) + ");"

    def parents_from(self, root_leaf: int):
        """parent[u] on the path toward ``root_leaf`` (-1 for the root leaf itself)."""
        parent = [-1] * self.nnodes
        order = [root_leaf]
        seen = {root_leaf}
        for u in order:
            for v in self.adj[u]:
                if v not in seen:
                    seen.add(v)
                    parent[v] = u
                    order.append(v)
        return parent

    @staticmethod
    def from_newick(text: str, taxon_names: list | None = None) -> "Tree":
        """Parse a Newick string.  A degree-2 root is suppressed (unrooted tree).

        ``taxon_names`` fixes leaf ids (alignment row order); if omitted leaves
        are numbered in order of appearance.
        """
        tokens = re.findall(r"[(),;:]|[^(),;:\s]+", text)
        pos = 0
        children, lens, labels = {}, {}, {}
        counter = [0]

        def new_node():
            counter[0] += 1
            children[counter[0]] = []
            return counter[0]

        def parse():
            nonlocal pos
            u = new_node()
            if tokens[pos] == "(":
                pos += 1
                while True:
                    children[u].append(parse())
                    if tokens[pos] == ",":
                        pos += 1
                        continue
                    if tokens[pos] == ")":
                        pos += 1
                        break
                    raise ValueError("bad newick near token %d" % pos)
            if pos < len(tokens) and tokens[pos] not in "(),;:":
                labels[u] = tokens[pos]
                pos += 1
            if pos < len(tokens) and tokens[pos] == ":":
                lens[u] = float(tokens[pos + 1])
                pos += 2
            return u

        import sys
        sys.setrecursionlimit(max(10000, len(tokens)))
        root = parse()
        leaves = [u for u in children if not children[u]]
        leaf_names = [labels[u] for u in leaves]
        if taxon_names is None:
            taxon_names = leaf_names
        if sorted(taxon_names) != sorted(leaf_names):
            raise ValueError("tree taxa do not match the alignment")
        ntaxa = len(leaves)
        ids = {}
        for u in leaves:
            ids[u] = taxon_names.index(labels[u])
        nxt = ntaxa
        suppress_root = len(children[root]) == 2
        for u in children:
            if children[u] and not (u == root and suppress_root):
                ids[u] = nxt
                nxt += 1
        t = Tree(ntaxa=ntaxa, adj=[[] for _ in range(nxt)], names=list(taxon_names))

        def link(a, b, ln):
            t.adj[a].append(b)
            t.adj[b].append(a)
            t.set_length(a, b, ln)

        def build(u):
            for v in children[u]:
                build(v)
            if u == root and suppress_root:
                a, b = children[u]
                link(ids[a], ids[b], lens.get(a, 0.0) + lens.get(b, 0.0))
            else:
                for v in children[u]:
                    link(ids[u], ids[v], lens.get(v, 0.0))

        build(root)
        return t


def random_tree(ntaxa: int, seed: int, mean_len: float = 0.1,
                min_len: float = 0.001, max_len: float = 0.999) -> Tree:
    """Yule-Harding unrooted binary tree: repeatedly split a uniformly chosen
    pendant edge; lengths Exp(mean) clamped to [min, max]."""
    rng = np.random.default_rng(seed)
    assert ntaxa >= 3
    t = Tree(ntaxa=ntaxa, adj=[[] for _ in range(2 * ntaxa - 2)],
             names=["T%d" % i for i in range(ntaxa)])
    order = rng.permutation(ntaxa)

    def rlen():
        return float(min(max(rng.exponential(mean_len), min_len), max_len))

    def link(a, b, ln):
        t.adj[a].append(b)
        t.adj[b].append(a)
        t.set_length(a, b, ln)

    c = ntaxa
    for i in range(3):
        link(c, int(order[i]), rlen())
    nxt = ntaxa + 1
    placed = [int(order[i]) for i in range(3)]
    for i in range(3, ntaxa):
        leaf = placed[int(rng.integers(len(placed)))]
        par = t.adj[leaf][0]
        ln = t.length.pop((leaf, par) if leaf < par else (par, leaf))
        new = nxt
        nxt += 1
        # replace leaf by new in par's list (keeps neighbour order stable)
        t.adj[par][t.adj[par].index(leaf)] = new
        t.adj[leaf] = []
        t.adj[new].append(par)
        t.set_length(new, par, ln)
        link(new, leaf, rlen())
        link(new, int(order[i]), rlen())
        placed.append(int(order[i]))
    return t


# --------------------------------------------------------------------------
# traversal lists (test-side helper; the product's own list builder is the
# C++ host mirror in csrc/host, which follows tree/phylotree.cpp:5907-6015)
# --------------------------------------------------------------------------
def subtree_sizes(tree: Tree):
    """size[(v, u)] = number of nodes in the subtree at v seen from u."""
    import sys
    sys.setrecursionlimit(max(10000, 4 * tree.nnodes))
    size = {}

    def rec(v, u):
        s = 1
        for w in tree.adj[v]:
            if w != u:
                s += rec(w, v)
        size[(v, u)] = s
        return s

    for u in range(tree.nnodes):
        for v in tree.adj[u]:
            if (v, u) not in size:
                rec(v, u)
    return size


def traversal(tree: Tree, node: int, dad: int, slot_of=None):
    """Post-order list of internal nodes of the subtree at ``node`` seen from
    ``dad`` as ``(node, dad, out_slot, [(kind, id, length), ...])`` with
    kind 0 = leaf (id = taxon), 1 = internal (id = slot).  Children are visited
    largest subtree first (tree/phylotree.cpp:5921-5944) but listed in adjacency
    order (FOR_NEIGHBOR order used for echildren, tree/phylokernelnew.h:1012).
    ``slot_of`` maps the directed edge (node, dad) to a slot; by default slots
    are handed out in list order starting at 0.
    """
    out = []
    size = subtree_sizes(tree)
    auto = {}

    def slot(v, u):
        if slot_of is not None:
            return slot_of[(v, u)]
        return auto.setdefault((v, u), len(auto))

    stack_guard = [0]

    def rec(v, u):
        if tree.is_leaf(v):
            return
        kids = [w for w in tree.adj[v] if w != u]
        for w in sorted(kids, key=lambda w: -size[(w, v)]):
            rec(w, v)
        ch = []
        for w in kids:
            if tree.is_leaf(w):
                ch.append((0, w, tree.get_length(v, w)))
            else:
                ch.append((1, slot(w, v), tree.get_length(v, w)))
        out.append((v, u, slot(v, u), ch))

    import sys
    sys.setrecursionlimit(max(10000, 4 * tree.nnodes))
    rec(node, dad)
    return out


def traversal_rooted(tree: Tree, node: int, dad: int, root_leaf: int, slot_of=None):
    """``traversal`` for the non-reversible kernel family on a rooted tree whose root is the leaf ``root_leaf``
    (the tree's last leaf): the root leaf becomes a child of kind 2 (B200LK_CHILD_ROOT) and a child that lies on the
    path to the root carries the flag 0x100 (child->direction == TOWARD_ROOT, tree/phylotree.cpp:5710-5725)."""
    parent = tree.parents_from(root_leaf)
    out = []
    for v, u, slot, ch in traversal(tree, node, dad, slot_of):
        kids = [w for w in tree.adj[v] if w != u]
        new = []
        for w, (kind, cid, ln) in zip(kids, ch):
            if w == root_leaf:
                kind, cid = 2, 0
            if parent[v] == w:
                kind |= 0x100
            new.append((kind, cid, ln))
        out.append((v, u, slot, new))
    return out


# --------------------------------------------------------------------------
# substitution models -> eigen-systems
# --------------------------------------------------------------------------
@dataclass
class EigenModel:
    """Eigen-decomposed reversible model (one or several mixture classes)."""
    nstates: int
    eval: np.ndarray       # [nmix][n]
    evec: np.ndarray       # [nmix][n][n]   row = state, col = eigen index
    inv_evec: np.ndarray   # [nmix][n][n]   row = eigen index
    freqs: np.ndarray      # [nmix][n]
    mixw: np.ndarray       # [nmix]
    Q: np.ndarray          # [nmix][n][n] normalised rate matrices

    @property
    def nmix(self) -> int:
        return int(self.eval.shape[0])


def _decompose(Q: np.ndarray, pi: np.ndarray):
    s = np.sqrt(pi)
    S = (Q * s[:, None]) / s[None, :]
    S = 0.5 * (S + S.T)
    w, V = np.linalg.eigh(S)
    evec = V / s[:, None]
    inv_evec = V.T * s[None, :]
    return w, evec, inv_evec


def reversible_model(exch: np.ndarray, freqs: np.ndarray, total_norm: float = 1.0):
    """Q_xy = r_xy * pi_y, rows sum to 0, mean rate normalised to 1."""
    n = len(freqs)
    pi = np.asarray(freqs, dtype=np.float64)
    pi = pi / pi.sum()
    R = np.asarray(exch, dtype=np.float64)
    Q = R * pi[None, :]
    np.fill_diagonal(Q, 0.0)
    np.fill_diagonal(Q, -Q.sum(axis=1))
    Q = Q / (-(pi * np.diag(Q)).sum()) * total_norm
    return Q, pi


def make_model(classes, mixw=None) -> EigenModel:
    """classes: list of (exchangeabilities[n][n] symmetric, freqs[n])."""
    ev, U, Ui, F, Qs = [], [], [], [], []
    for exch, fr in classes:
        Q, pi = reversible_model(exch, fr)
        w, evec, inv = _decompose(Q, pi)
        ev.append(w); U.append(evec); Ui.append(inv); F.append(pi); Qs.append(Q)
    nmix = len(classes)
    if mixw is None:
        mixw = np.full(nmix, 1.0 / nmix)
    mixw = np.asarray(mixw, dtype=np.float64)
    return EigenModel(nstates=len(F[0]), eval=np.ascontiguousarray(ev),
                      evec=np.ascontiguousarray(U), inv_evec=np.ascontiguousarray(Ui),
                      freqs=np.ascontiguousarray(F), mixw=mixw / mixw.sum(), Q=np.ascontiguousarray(Qs))


@dataclass
class NonrevModel:
    """What the non-reversible kernel family takes from ModelMarkov: Q, its complex eigen-system
    (P(t) = Re(cevec exp(ceval t) cinv_evec), model/modelmarkov.cpp:481-488) and the state frequencies."""
    nstates: int
    Q: np.ndarray          # [nmix][n][n]
    ceval: np.ndarray      # [nmix][n]      complex
    cevec: np.ndarray      # [nmix][n][n]   complex
    cinv_evec: np.ndarray  # [nmix][n][n]   complex
    freqs: np.ndarray      # [nmix][n]
    mixw: np.ndarray       # [nmix]

    @property
    def nmix(self) -> int:
        return int(self.Q.shape[0])


def nonrev_from_Q(Qs, freqs, mixw=None) -> NonrevModel:
    Qs = np.ascontiguousarray(Qs, dtype=np.float64)
    ev, V, Vi = [], [], []
    for Q in Qs:
        w, v = np.linalg.eig(Q)
        ev.append(w); V.append(v); Vi.append(np.linalg.inv(v))
    nmix = len(Qs)
    mixw = np.full(nmix, 1.0 / nmix) if mixw is None else np.asarray(mixw, dtype=np.float64)
    return NonrevModel(nstates=Qs.shape[1], Q=Qs, ceval=np.ascontiguousarray(ev, dtype=np.complex128),
                       cevec=np.ascontiguousarray(V, dtype=np.complex128),
                       cinv_evec=np.ascontiguousarray(Vi, dtype=np.complex128),
                       freqs=np.ascontiguousarray(freqs, dtype=np.float64), mixw=mixw / mixw.sum())


def random_nonrev_model(nstates: int, seed: int, nmix: int = 1) -> NonrevModel:
    """Seeded unrestricted (UNREST-like) rate matrices: free off-diagonal rates, the state frequencies are the
    stationary distribution of Q, mean rate normalised to 1."""
    rng = np.random.default_rng(seed)
    Qs, F = [], []
    for _ in range(nmix):
        Q = rng.gamma(1.0, 1.0, size=(nstates, nstates)) + 0.02
        np.fill_diagonal(Q, 0.0)
        np.fill_diagonal(Q, -Q.sum(axis=1))
        w, v = np.linalg.eig(Q.T)
        pi = np.real(v[:, np.argmin(np.abs(w))])
        pi = pi / pi.sum()
        Q = Q / (-(pi * np.diag(Q)).sum())
        Qs.append(Q); F.append(pi)
    w = rng.dirichlet(np.full(nmix, 5.0)) if nmix > 1 else None
    return nonrev_from_Q(Qs, F, w)


def nonrev_from_reversible(model: EigenModel) -> NonrevModel:
    """A reversible model in the form the non-reversible kernels take (params->kernel_nonrev)."""
    return NonrevModel(nstates=model.nstates, Q=np.ascontiguousarray(model.Q),
                       ceval=model.eval.astype(np.complex128), cevec=model.evec.astype(np.complex128),
                       cinv_evec=model.inv_evec.astype(np.complex128), freqs=model.freqs, mixw=model.mixw)


def gtr_model(rates=(1.0, 2.0, 1.5, 1.2, 2.5, 1.0), freqs=(0.3, 0.2, 0.2, 0.3)) -> EigenModel:
    """GTR with rates in the order AC AG AT CG CT GT (model/modeldna.cpp)."""
    R = np.zeros((4, 4))
    k = 0
    for i in range(4):
        for j in range(i + 1, 4):
            R[i, j] = R[j, i] = rates[k]
            k += 1
    return make_model([(R, np.asarray(freqs))])


_PROTEIN_DATA = None


def protein_data():
    """LG and C60 as data (iqtree2_b200/data/protein_models.json, written by tools/extract_model_data.py from the
    published tables the reference carries: model/modelprotein.cpp:287-310, model/modelmixture.cpp:917-977)."""
    global _PROTEIN_DATA
    if _PROTEIN_DATA is None:
        import json
        import os
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "protein_models.json")) as f:
            _PROTEIN_DATA = json.load(f)
    return _PROTEIN_DATA


def lg_exchangeabilities() -> np.ndarray:
    low = protein_data()["LG"]["lower_triangle"]
    R = np.zeros((20, 20))
    k = 0
    for i in range(20):
        for j in range(i):
            R[i, j] = R[j, i] = low[k]
            k += 1
    return R


def lg_model() -> EigenModel:
    """`-m LG`: LG exchangeabilities and LG's own frequencies."""
    return make_model([(lg_exchangeabilities(), np.asarray(protein_data()["LG"]["freqs"]))])


def lg_c60_model() -> EigenModel:
    """`-m LG+C60`: 60 classes over the LG exchangeabilities, class k with profile C60pi_k, rate 1, fixed weight w_k
    (`model C60 = POISSON+G+FMIX{C60pi1:1:w1,...}` with POISSON replaced by LG; every class normalised to one expected
    substitution, so the global rate is 1: ModelMixture::initMixture, model/modelmixture.cpp:1411-1437)."""
    d = protein_data()["C60"]
    R = lg_exchangeabilities()
    return make_model([(R, np.asarray(p)) for p in d["profiles"]], mixw=np.asarray(d["weights"]))


def random_reversible_model(nstates: int, seed: int, nmix: int = 1) -> EigenModel:
    """A seeded random reversible model (stand-in where a named empirical matrix
    is not needed: the kernels only see the eigen-system)."""
    rng = np.random.default_rng(seed)
    classes = []
    R = rng.gamma(1.0, 1.0, size=(nstates, nstates)) + 0.01
    R = 0.5 * (R + R.T)
    for _ in range(nmix):
        fr = rng.dirichlet(np.full(nstates, 2.0 if nmix == 1 else 0.3)) + 1e-4
        classes.append((R, fr / fr.sum()))
    w = rng.dirichlet(np.full(nmix, 5.0)) if nmix > 1 else None
    return make_model(classes, w)


def discrete_gamma(alpha: float, ncat: int = 4):
    """Mean-of-category discrete Gamma rates rescaled to mean 1
    (model/rategamma.cpp:98-170), equal proportions."""
    from scipy.special import gammainc, gammaincinv
    if ncat == 1:
        return np.ones(1), np.ones(1)
    cuts = gammaincinv(alpha, np.arange(1, ncat) / ncat) / alpha  # Gamma(alpha, beta=alpha)
    upper = np.concatenate([gammainc(alpha + 1.0, cuts * alpha), [1.0]])
    lower = np.concatenate([[0.0], upper[:-1]])
    rates = (upper - lower) * ncat
    rates = rates / rates.mean()
    return rates, np.full(ncat, 1.0 / ncat)


# --------------------------------------------------------------------------
# alignments
# --------------------------------------------------------------------------
_DNA = {c: i for i, c in enumerate("ACGT")}
_DNA_AMBIG = {"R": 1 + 4, "Y": 2 + 8, "W": 1 + 8, "S": 2 + 4, "M": 1 + 2, "K": 4 + 8,
              "B": 2 + 4 + 8, "H": 1 + 2 + 8, "D": 1 + 4 + 8, "V": 1 + 2 + 4}
_AA = {c: i for i, c in enumerate("ARNDCQEGHILKMFPSTWYV")}


def encode_dna(ch: str) -> int:
    ch = ch.upper()
    if ch == "U":
        ch = "T"
    if ch in _DNA:
        return _DNA[ch]
    if ch in _DNA_AMBIG:
        return _DNA_AMBIG[ch] + 3      # code = bitmask + 3 (model/modeldna.cpp:462-475)
    return DNA_UNKNOWN


def encode_aa(ch: str) -> int:
    ch = ch.upper()
    if ch in _AA:
        return _AA[ch]
    return {"B": 20, "Z": 21, "J": 22}.get(ch, AA_UNKNOWN)


def state_indicators(nstates: int) -> np.ndarray:
    """tipvec[s][x] = 1 if real state x is compatible with state code s, for
    s = 0..STATE_UNKNOWN (model/modelsubst.cpp:182-207, modeldna.cpp:462-475,
    modelprotein.cpp:1147-1164)."""
    if nstates == 4:
        su = DNA_UNKNOWN
        m = np.zeros((su + 1, 4))
        for s in range(4):
            m[s, s] = 1.0
        for s in range(4, su):
            bits = s - 3
            for x in range(4):
                if bits & (1 << x):
                    m[s, x] = 1.0
        m[su, :] = 1.0
        return m
    if nstates == 20:
        su = AA_UNKNOWN
        m = np.zeros((su + 1, 20))
        for s in range(20):
            m[s, s] = 1.0
        m[20, [2, 3]] = 1.0     # B = N or D
        m[21, [5, 6]] = 1.0     # Z = Q or E
        m[22, [9, 10]] = 1.0    # J = I or L
        m[su, :] = 1.0
        return m
    m = np.zeros((nstates + 1, nstates))
    m[:nstates] = np.eye(nstates)
    m[nstates] = 1.0
    return m


def tip_partial_lh(model: EigenModel) -> np.ndarray:
    """tip_partial_lh[s][m][i] = sum_x inv_evec[m][i][x] * tipvec[s][x]
    (tree/phylotreesse.cpp:368-375)."""
    ind = state_indicators(model.nstates)
    return np.ascontiguousarray(np.einsum("mix,sx->smi", model.inv_evec, ind))


def read_phylip(path: str):
    with open(path) as f:
        ntax, nsite = map(int, f.readline().split()[:2])
        names, seqs = [], []
        for _ in range(ntax):
            parts = f.readline().split()
            names.append(parts[0])
            seqs.append("".join(parts[1:]))
        # interleaved continuation blocks
        rest = [ln.strip() for ln in f if ln.strip()]
        for i, ln in enumerate(rest):
            seqs[i % ntax] += "".join(ln.split())
    assert all(len(s) == nsite for s in seqs), "sequence length mismatch"
    return names, seqs


@dataclass
class Alignment:
    """Compressed alignment: ``states[taxon][pattern]`` uint8 codes,
    ``ptn_freq[pattern]``; ``is_const``/``const_state`` per pattern as used by
    computePtnInvar (tree/phylotreesse.cpp:549-612)."""
    nstates: int
    names: list
    states: np.ndarray
    ptn_freq: np.ndarray

    @property
    def ntaxa(self) -> int:
        return int(self.states.shape[0])

    @property
    def nptn(self) -> int:
        return int(self.states.shape[1])

    @property
    def nsite(self) -> int:
        return int(round(float(self.ptn_freq.sum())))

    @property
    def state_unknown(self) -> int:
        return {4: DNA_UNKNOWN, 20: AA_UNKNOWN}.get(self.nstates, self.nstates)


def compress_patterns(cols: np.ndarray, nstates: int, names, compress: bool = True) -> Alignment:
    """cols: [ntaxa][nsite] uint8 -> unique site patterns in first-seen order."""
    cols = np.ascontiguousarray(cols, dtype=np.uint8)
    if not compress:
        return Alignment(nstates, list(names), cols, np.ones(cols.shape[1]))
    _, first, counts = np.unique(cols.T, axis=0, return_index=True, return_counts=True)
    order = np.argsort(first)
    return Alignment(nstates, list(names), np.ascontiguousarray(cols[:, first[order]]),
                     counts[order].astype(np.float64))


def alignment_from_phylip(path: str, seq_type: str = "DNA") -> Alignment:
    names, seqs = read_phylip(path)
    enc = encode_dna if seq_type == "DNA" else encode_aa
    table = np.array([enc(chr(b)) for b in range(256)], dtype=np.uint8)
    cols = np.stack([table[np.frombuffer(s.encode(), dtype=np.uint8)] for s in seqs])
    return compress_patterns(cols, 4 if seq_type == "DNA" else 20, names)


def ptn_invar(aln: Alignment, model: EigenModel, p_invar: float) -> np.ndarray:
    """ptn_invar[p] = p_inv * pi(const state) for constant patterns
    (tree/phylotreesse.cpp:549-612); 0 when p_inv == 0."""
    out = np.zeros(aln.nptn)
    if p_invar <= 0.0:
        return out
    ind = state_indicators(aln.nstates)
    compat = np.ones((aln.nptn, aln.nstates))
    for t in range(aln.ntaxa):
        compat *= ind[aln.states[t]]
    fr = (model.freqs * model.mixw[:, None]).sum(axis=0)
    return p_invar * (compat * fr[None, :]).sum(axis=1)


def simulate_alignment(tree: Tree, model: EigenModel, rates, props, nsite: int, seed: int,
                       compress: bool = False, chunk: int = 1 << 18) -> Alignment:
    """Evolve ``nsite`` columns down ``tree`` (AliSim-equivalent): root state from
    the stationary distribution of the site's mixture class, rate category
    drawn by ``props``, child state from P(r*t) = U exp(lambda r t) U^-1."""
    rng = np.random.default_rng(seed)
    n, nmix, ncat = model.nstates, model.nmix, len(rates)
    cols = np.empty((tree.ntaxa, nsite), dtype=np.uint8)
    root = tree.ntaxa  # first internal node
    # DFS order of directed edges from the root
    order, stack = [], [(root, -1)]
    while stack:
        u, d = stack.pop()
        for v in tree.adj[u]:
            if v != d:
                order.append((u, v))
                stack.append((v, u))
    for lo in range(0, nsite, chunk):
        m = min(chunk, nsite - lo)
        cls = rng.choice(nmix, size=m, p=model.mixw) if nmix > 1 else np.zeros(m, dtype=np.int64)
        cat = rng.choice(ncat, size=m, p=np.asarray(props) / np.sum(props))
        mc = cls * ncat + cat
        cumf = np.cumsum(model.freqs, axis=1)
        st = {root: (rng.random(m)[:, None] > cumf[cls][:, :-1]).sum(axis=1).astype(np.int64)}
        for u, v in order:
            t = tree.get_length(u, v)
            P = np.empty((nmix * ncat, n, n))
            for a in range(nmix):
                for k in range(ncat):
                    e = np.exp(model.eval[a] * (rates[k] * t))
                    P[a * ncat + k] = (model.evec[a] * e[None, :]) @ model.inv_evec[a]
            P = np.clip(P, 0.0, None)
            cum = np.cumsum(P, axis=2)
            cum /= cum[:, :, -1:]
            rows = cum[mc, st[u]]                      # [m][n]
            child = (rng.random(m)[:, None] > rows[:, :-1]).sum(axis=1).astype(np.int64)
            if tree.is_leaf(v):
                cols[v, lo:lo + m] = child
            else:
                st[v] = child
            # parent states no longer needed once all its children are done: keep simple
    return compress_patterns(cols, n, tree.names, compress=compress)
