// Junction-tree hierarchy of one graph: restates generate_junction_tree_hierarchies.py:26-116 (generate_jth)
// on top of networkx.algorithms.tree.decomposition.junction_tree, chordal.complete_to_chordal_graph (MCS-M),
// chordal.chordal_graph_cliques, tree.mst.kruskal_mst_edges and bipartite.projected_graph (networkx 3.x),
// keeping every iteration order those functions have.
#pragma once
#include <map>
#include <numeric>

#include "graph.hpp"

namespace hg_htree {

// ---- chordal.is_chordal: maximum cardinality search; the answer does not depend on tie-breaks
inline bool is_chordal(const Graph& h) {
    const size_t n = h.n();
    if (n <= 3) return true;
    std::vector<char> numbered(n, 0);
    std::vector<int> weight(n, 0);
    for (size_t step = 0; step < n; ++step) {
        int best = -1;
        for (size_t s = 0; s < n; ++s) if (!numbered[s] && (best < 0 || weight[s] > weight[best])) best = (int)s;
        std::vector<int> prev;
        for (int v : h.adj[best]) if (numbered[h.slot(v)]) prev.push_back(v);
        for (size_t a = 0; a < prev.size(); ++a)
            for (size_t b = a + 1; b < prev.size(); ++b)
                if (!h.has_edge(prev[a], prev[b])) return false;
        numbered[best] = 1;
        for (int v : h.adj[best]) if (!numbered[h.slot(v)]) ++weight[h.slot(v)];
    }
    return true;
}

// ---- chordal.complete_to_chordal_graph: H = G.copy(); chords found by MCS-M, added in the iteration order of
// the Python set of (z, y) tuples they were collected in
inline Graph complete_to_chordal(const GL& g) {
    Graph h = copy_of(g);
    if (is_chordal(h)) return h;
    const size_t n = h.n();
    std::vector<int> weight(n, 0);  // by slot of h
    std::vector<int> unnumbered = h.order;
    PySet chords;                   // key = z << 32 | y
    std::vector<char> allowed(n), reach(n);
    std::vector<int> stack;
    for (size_t i = n; i > 0; --i) {
        size_t zi = 0;
        for (size_t k = 1; k < unnumbered.size(); ++k)
            if (weight[h.slot(unnumbered[k])] > weight[h.slot(unnumbered[zi])]) zi = k;  // max(): first maximum
        const int z = unnumbered[zi];
        unnumbered.erase(unnumbered.begin() + zi);
        std::vector<int> update;
        for (int y : unnumbered) {
            if (g.has_edge(y, z)) { update.push_back(y); continue; }
            const int yw = weight[h.slot(y)];
            std::fill(allowed.begin(), allowed.end(), 0);
            for (int v : unnumbered) if (weight[h.slot(v)] < yw) allowed[h.slot(v)] = 1;
            allowed[h.slot(z)] = allowed[h.slot(y)] = 1;
            std::fill(reach.begin(), reach.end(), 0);  // nx.has_path(H.subgraph(lower + [z, y]), y, z)
            stack.assign(1, y);
            reach[h.slot(y)] = 1;
            bool found = false;
            while (!stack.empty() && !found) {
                int a = stack.back(); stack.pop_back();
                for (int b : h.nbrs(a)) {
                    int sb = h.slot(b);
                    if (!allowed[sb] || reach[sb]) continue;
                    if (b == z) { found = true; break; }
                    reach[sb] = 1; stack.push_back(b);
                }
            }
            if (found) {
                update.push_back(y);
                int zy[2] = {z, y};
                chords.add(((int64_t)z << 32) | (uint32_t)y, tuple_hash(zy, 2));
            }
        }
        for (int v : update) ++weight[h.slot(v)];
    }
    chords.each([&](int64_t k) { h.add_edge((int)(k >> 32), (int)(k & 0xffffffff)); });
    return h;
}

// ---- chordal.chordal_graph_cliques: cliques (sorted tuples) in the order the generator yields them
inline std::vector<std::vector<int>> chordal_cliques(const Graph& h) {
    std::vector<std::vector<int>> out;
    auto emit = [&](const PySet& s) { std::vector<int> c = s.items(); std::sort(c.begin(), c.end()); out.push_back(c); };
    GL hv = whole(h);
    for (const PySet& comp : connected_components(hv)) {
        Graph c = copy_of(subview(hv, comp.items()));
        if (c.n() == 1) { out.push_back({c.order[0]}); continue; }
        PySet unnumbered, numbered, wanna;
        for (int k : c.order) unnumbered.add(k);
        int v = c.order[0];  // arbitrary_element(C)
        unnumbered.discard(v);
        numbered.add(v);
        wanna.add(v);
        while (!unnumbered.empty()) {
            int best = -1, best_n = -1;  // _max_cardinality_node: first strict maximum in set order
            unnumbered.each([&](int64_t x) {
                int cnt = 0;
                for (int y : c.nbrs((int)x)) if (numbered.has(y)) ++cnt;
                if (cnt > best_n) { best_n = cnt; best = (int)x; }
            });
            v = best;
            unnumbered.discard(v);
            numbered.add(v);
            PySet nb;
            for (int y : c.nbrs(v)) nb.add(y);
            PySet next = PySet::intersection(nb, numbered);
            next.add(v);
            bool superset = true;
            wanna.each([&](int64_t k) { if (!next.has((int)k)) superset = false; });
            if (!superset) emit(wanna);
            wanna = next;
        }
        emit(wanna);
    }
    return out;
}

struct Jth {
    Graph g;
    std::vector<int> roots;
    bool has_roots = true;  // generate_jth returns (JTG, None) for a single original node
};

inline std::vector<int> sorted_intersection(const std::vector<int>& a, const std::vector<int>& b) {
    std::vector<int> r;
    std::set_intersection(a.begin(), a.end(), b.begin(), b.end(), std::back_inserter(r));
    return r;
}

// junction_tree(G) followed by bipartite.projected_graph onto the clique nodes; returns the cliques and the
// projected graph on clique indices 0..K-1
inline Graph clique_tree(const GL& g, std::vector<std::vector<int>>& cliques) {
    Graph chordal = complete_to_chordal(g);
    cliques = chordal_cliques(chordal);
    const int K = (int)cliques.size();
    Graph cg;  // clique graph; edge weight = |sepset|
    for (int i = 0; i < K; ++i) cg.add_node(i);
    std::map<std::pair<int, int>, std::vector<int>> sepset;
    for (int i = 0; i < K; ++i)
        for (int j = i + 1; j < K; ++j) {
            std::vector<int> s = sorted_intersection(cliques[i], cliques[j]);
            if (s.empty()) continue;
            cg.add_edge(i, j);
            sepset[{i, j}] = s;
        }
    auto sep_of = [&](int a, int b) -> const std::vector<int>& { return sepset[{std::min(a, b), std::max(a, b)}]; };
    // kruskal_mst_edges(minimum=False): stable sort by weight, descending
    std::vector<std::pair<int, int>> edges = cg.edges();
    std::stable_sort(edges.begin(), edges.end(), [&](const std::pair<int, int>& x, const std::pair<int, int>& y) {
        return sep_of(x.first, x.second).size() > sep_of(y.first, y.second).size();
    });
    std::vector<int> parent(K);
    std::iota(parent.begin(), parent.end(), 0);
    auto find = [&](int x) { while (parent[x] != x) x = parent[x] = parent[parent[x]]; return x; };
    Graph t;  // maximum_spanning_tree: all nodes, tree edges in the order Kruskal yields them
    for (int i = 0; i < K; ++i) t.add_node(i);
    for (auto& e : edges) {
        int a = find(e.first), b = find(e.second);
        if (a == b) continue;
        parent[a] = b;
        t.add_edge(e.first, e.second);
    }
    // sepset nodes: key K + index of the distinct sepset tuple
    std::vector<std::vector<int>> seps;
    for (auto& e : t.edges()) {
        const std::vector<int>& s = sep_of(e.first, e.second);
        int id = (int)(std::find(seps.begin(), seps.end(), s) - seps.begin());
        if (id == (int)seps.size()) seps.push_back(s);
        t.add_node(K + id);
        t.add_edge(e.first, K + id);
        t.add_edge(e.second, K + id);
        t.remove_edge(e.first, e.second);
    }
    // projected_graph(B, clique nodes): nbrs2 is a Python set of clique TUPLES
    std::vector<int64_t> chash(K);
    for (int i = 0; i < K; ++i) chash[i] = tuple_hash(cliques[i].data(), cliques[i].size());
    Graph pg;
    for (int i = 0; i < K; ++i) pg.add_node(i);
    for (int u = 0; u < K; ++u) {
        PySet nbrs2;
        for (int s : t.nbrs(u)) for (int v : t.nbrs(s)) if (v != u) nbrs2.add(v, chash[v]);
        nbrs2.each([&](int64_t v) { pg.add_edge(u, (int)v); });
    }
    return pg;
}

// generate_jth(G, zero_feature, remove_edges_every_layer=True); `original` is G.graph["original"]
inline Jth generate_jth(const GL& g, bool original) {
    Jth out;
    std::vector<int> gnodes = g.node_list();
    if (gnodes.size() == 1 && original) {
        out.g.add_node(0, g.root->at(gnodes[0]));
        out.has_roots = false;
        return out;
    }
    std::vector<std::vector<int>> cliques;
    Graph pg = clique_tree(g, cliques);
    const int K = (int)cliques.size();
    // clique_has = list(nx.subgraph(G, clique_tuple).nodes()); relabel cliques to 0..K-1
    Graph jt = relabel_copy(pg, [](int k) { return k; });
    for (int i = 0; i < K; ++i) {
        Attr& a = jt.at(i);
        a.cls = kClique; a.kind = kNoKind; a.is_list = true;
        a.list = subview(g, cliques[i]).node_list();
    }
    if (K == 1 && !original) {  // G is a clique and not the original graph: G without links, every node a root
        Graph ge;
        for (int k : gnodes) ge.add_node(k, g.root->at(k));
        out.g = std::move(ge);
        out.roots = gnodes;
        return out;
    }
    out.roots.resize(K);
    std::iota(out.roots.begin(), out.roots.end(), 0);
    int next_index = K;
    for (int anode = 0; anode < K; ++anode) {
        GL sg = subview(g, std::vector<int>(jt.at(anode).list));
        std::vector<int> members = sg.node_list();
        if (members.size() == 1) continue;
        Graph u;
        if (members.size() == 2) {
            for (int m : members) u.add_node(next_index++, g.root->at(m));
            std::vector<int> fresh = u.order;
            for (int k : fresh) u.add_edge(k, anode);
        } else {
            Jth sub = generate_jth(sg, false);
            std::vector<char> is_root(sub.g.key_span(), 0);
            for (int r : sub.roots) is_root[r] = 1;
            for (int r : sub.roots) {  // remove the tree structure among the roots of this layer
                std::vector<int> nb = sub.g.nbrs(r).to_vector();
                for (int v : nb) if (is_root[v]) sub.g.remove_edge(r, v);
            }
            std::vector<int> new_index(sub.g.key_span(), -1), roots_new;
            for (int k : sub.g.order) {
                new_index[k] = next_index;
                if (is_root[k]) roots_new.push_back(next_index);
                ++next_index;
            }
            u = relabel_copy(sub.g, [&](int k) { return new_index[k]; });
            for (int r : roots_new) u.add_edge(r, anode);
        }
        absorb(jt, u, false);  // JTG.update(U)
    }
    out.g = std::move(jt);
    return out;
}

}  // namespace hg_htree
