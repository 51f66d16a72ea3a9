"""Prototype of a PARALLELISABLE formulation of the lake growth (erosion.rs:2066-2137), checked against the
oracle's serial heap loop.  CPU only; design work for replacing the one-warp kernel k_lake_connect.

The reference grows a forest from the boundary lakes Prim-style with a heap keyed
(pass height, cell on the side being connected, cell on the connected side).  Claim implemented here:

  1. Treat all boundary lakes as one vertex S.  Process the passes by ascending height ("weight classes"),
     merging components with a union-find: this builds the Kruskal reconstruction tree (KRT).  Which
     components merge at which height does not depend on how ties are broken.
  2. The only thing that depends on the direction of growth is, inside one weight class, WHICH of several
     equal-height passes is used and in what order components are attached.  Given the component through
     which growth enters a KRT node (top-down from S), that is a priority-first search over the node's
     class edges with key (cell on the unconnected side, cell on the connected side) - two or three edges
     in practice.
  3. Growth visits a KRT node as: everything of the entry child, then the other children in attachment
     order, each completely: the connection order is a traversal of the KRT.

Steps 1 and 3 are sorts, union-find/Boruvka rounds and tree traversals (parallel); step 2 is tiny per node.

usage: python tools/lake_growth_prototype.py [x_lg] [steps] [kind]
"""
import os
import sys
import heapq
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from veloren_b200 import synth          # noqa: E402
from oracle import pyoracle as po       # noqa: E402


def read_dump(path):
    raw = np.fromfile(path, dtype=np.int64)
    ne = int(raw[0])
    edges = raw[1:1 + 5 * ne].reshape(ne, 5)          # from, to, lake_from, lake_to, to_is_boundary
    nc = int(raw[1 + 5 * ne])
    conn = raw[2 + 5 * ne:2 + 5 * ne + 3 * nc].reshape(nc, 3)   # lake, chunk_idx, neighbor_idx (in order)
    return edges, conn


def grow(edges, h):
    """Returns [(lake, chunk_idx, neighbor_idx)] in connection order, computed through the KRT."""
    S = -1
    lf = edges[:, 2].copy()
    lt = np.where(edges[:, 4] == 1, S, edges[:, 3])
    w = np.maximum(h[edges[:, 0]], h[edges[:, 1]])
    # directed entries: key by (unconnected lake = lake_from, other side)
    # undirected weight classes over vertex pairs
    lakes = np.unique(lf)
    parent = {int(v): int(v) for v in lakes}
    parent[S] = S

    def find(v):
        while parent[v] != v:
            parent[v] = parent[parent[v]]
            v = parent[v]
        return v

    # KRT nodes: id -> dict(children=[node ids], edges=[edge indices]); leaves are vertices
    node_of_comp = {v: ("leaf", v) for v in parent}
    krt = {}
    order = np.argsort(w, kind="stable")
    i = 0
    next_id = 0
    while i < order.size:
        j = i
        while j < order.size and w[order[j]] == w[order[i]]:
            j += 1
        cls = order[i:j]
        # class quotient graph on current components
        adj = {}
        for e in cls:
            a, b = find(int(lf[e])), find(int(lt[e]))
            if a == b:
                continue
            adj.setdefault(a, []).append((b, int(e)))
            adj.setdefault(b, []).append((a, int(e)))
        seen = set()
        for c0 in list(adj):
            if c0 in seen:
                continue
            comp, stack, es = [], [c0], set()
            seen.add(c0)
            while stack:
                c = stack.pop()
                comp.append(c)
                for (d, e) in adj[c]:
                    es.add(e)
                    if d not in seen:
                        seen.add(d)
                        stack.append(d)
            nid = ("node", next_id)
            next_id += 1
            krt[nid] = dict(children=[node_of_comp[c] for c in comp], comps=comp, edges=sorted(es))
            root = comp[0]
            for c in comp[1:]:
                parent[c] = root
            if S in comp:                               # keep S as representative (cosmetic)
                parent[root] = S
                parent[S] = S
                root = S
            node_of_comp[find(root)] = nid
        i = j
    top = node_of_comp[find(S)]
    cache = {}

    def vertices(nid):
        if nid[0] == "leaf":
            return [nid[1]]
        if nid not in cache:
            out_v = []
            for c in krt[nid]["children"]:
                out_v.extend(vertices(c))
            cache[nid] = out_v
        return cache[nid]

    out = []
    _expand([("child", top, S, None)], krt, lf, lt, edges, vertices, out)
    return out


def _expand(items, krt, lf, lt, edges, vertices, out):
    """Depth-first: for each (child, entry vertex, connecting record) emit the record, then the child's subtree."""
    import heapq as hq
    stack = [iter(items)]
    while stack:
        try:
            _, child, entry, rec = next(stack[-1])
        except StopIteration:
            stack.pop()
            continue
        if rec is not None:
            out.append(rec)
        if child[0] == "leaf":
            continue
        nd = krt[child]
        child_of = {}
        for ci, c in enumerate(nd["children"]):
            for v in vertices(c):
                child_of[v] = ci
        connected = {child_of[entry]}
        seq = [("child", nd["children"][child_of[entry]], entry, None)]
        heap = []

        def push_from(ci):
            for e in nd["edges"]:
                a, b = int(lf[e]), int(lt[e])
                if child_of[b] == ci and child_of[a] not in connected:
                    hq.heappush(heap, (int(edges[e, 0]), int(edges[e, 1]), e))

        push_from(child_of[entry])
        while heap:
            x, y, e = hq.heappop(heap)
            a = int(lf[e])
            ca = child_of[a]
            if ca in connected:
                continue
            connected.add(ca)
            seq.append(("child", nd["children"][ca], a, (a, x, y)))
            push_from(ca)
        assert len(connected) == len(nd["children"]), "KRT node not fully attached"
        stack.append(iter(seq))


def main():
    x_lg = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    kind = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    sys.setrecursionlimit(1000000)
    f = synth.generate(x_lg, x_lg, kind=kind)
    o = po.Oracle(f, n_threads=8)
    path = os.path.join(tempfile.gettempdir(), "orc_lakes_dump.bin")
    for it in range(steps):
        h_before = o.get("h").copy()
        os.environ["ORC_DUMP_LAKES"] = path
        o.step()
        del os.environ["ORC_DUMP_LAKES"]
        edges, conn = read_dump(path)
        got = grow(edges, h_before)
        ref = [tuple(int(v) for v in r) for r in conn]
        same_set = sorted(got) == sorted(ref)
        same_order = got == ref
        ties = len(edges) - len(np.unique(np.maximum(h_before[edges[:, 0]], h_before[edges[:, 1]])))
        print("step %d: lakes %d, directed passes %d (%d duplicates of a height): choices %s, order %s" %
              (it, len(ref), len(edges), ties, "IDENTICAL" if same_set else "DIFFER", "IDENTICAL" if same_order else "DIFFERS"))
        if not same_order:
            for k, (g_, r_) in enumerate(zip(got, ref)):
                if g_ != r_:
                    print("  first difference at connection %d: prototype %s, reference %s" % (k, g_, r_))
                    break
            return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())
