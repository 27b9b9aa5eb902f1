// networkx.Graph as the reference uses it: insertion-ordered node dict, insertion-ordered adjacency
// dicts, sub-graph VIEWS whose iteration order follows networkx 3.x coreviews.FilterAtlas (set order when
// the node set is less than half of the underlying graph, else the underlying order), and the handful of
// copy / relabel / union operations that rebuild adjacency from an edge iteration.
#pragma once
#include <algorithm>
#include <cstdint>
#include <stdexcept>
#include <vector>

#include "pyorder.hpp"

namespace hg_htree {

enum NodeClass : uint8_t { kLeaf = 0, kClique = 1, kVirtual = 2 };           // attribute "type"
enum NodeKind : uint8_t { kObject = 0, kRoom = 1, kObjectRoom = 2, kRoomRoom = 3,
                          kObjectVirtual = 4, kRoomVirtual = 5, kNoKind = 255 };  // attribute "node_type"

struct Attr {
    uint8_t cls = kLeaf;
    uint8_t kind = kNoKind;
    bool is_list = false;      // clique_has: int (leaf / virtual) or list (clique)
    int single = -1;
    std::vector<int> list;
};

struct Graph {
    std::vector<int> order;              // node keys in dict order (nodes are never removed)
    std::vector<int> slot_of;            // key -> slot, -1 when absent (keys are small non-negative ints)
    std::vector<std::vector<int>> adj;   // per slot: neighbour keys in dict order
    std::vector<Attr> attr;              // per slot

    size_t n() const { return order.size(); }
    bool has_node(int k) const { return k >= 0 && (size_t)k < slot_of.size() && slot_of[k] >= 0; }
    int slot(int k) const { return slot_of[k]; }
    int add_node(int k) {
        if ((size_t)k >= slot_of.size()) slot_of.resize(std::max<size_t>(k + 1, slot_of.size() * 2), -1);
        if (slot_of[k] < 0) {
            slot_of[k] = (int)order.size();
            order.push_back(k);
            adj.emplace_back();
            attr.emplace_back();
        }
        return slot_of[k];
    }
    int add_node(int k, const Attr& a) { int s = add_node(k); attr[s] = a; return s; }
    const Attr& at(int k) const { return attr[slot_of[k]]; }
    Attr& at(int k) { return attr[slot_of[k]]; }
    const std::vector<int>& nbrs(int k) const { return adj[slot_of[k]]; }
    bool has_edge(int u, int v) const {
        if (!has_node(u)) return false;
        const auto& a = adj[slot_of[u]];
        return std::find(a.begin(), a.end(), v) != a.end();
    }
    void add_edge(int u, int v) {  // Graph.add_edge: u then v enter the node dict; existing keys keep their place
        add_node(u); add_node(v);
        auto& a = adj[slot_of[u]];
        if (std::find(a.begin(), a.end(), v) == a.end()) a.push_back(v);
        auto& b = adj[slot_of[v]];
        if (std::find(b.begin(), b.end(), u) == b.end()) b.push_back(u);
    }
    void remove_edge(int u, int v) {
        auto& a = adj[slot_of[u]];
        a.erase(std::remove(a.begin(), a.end(), v), a.end());
        if (u != v) {
            auto& b = adj[slot_of[v]];
            b.erase(std::remove(b.begin(), b.end(), u), b.end());
        }
    }
    // EdgeView.__iter__: node order x adjacency order, an edge reported from its first endpoint only
    template <class F> void each_edge(F&& f) const {
        std::vector<char> seen(slot_of.size(), 0);
        for (size_t s = 0; s < order.size(); ++s) {
            for (int v : adj[s]) if (!seen[v]) f(order[s], v);
            seen[order[s]] = 1;
        }
    }
    std::vector<std::pair<int, int>> edges() const {
        std::vector<std::pair<int, int>> e;
        each_edge([&](int u, int v) { e.emplace_back(u, v); });
        return e;
    }
};

// A graph or a node-induced view of one (nx.subgraph / Graph.subgraph).  `root` is always a real graph:
// networkx does not chain views ("if already a subgraph, don't make a chain").
struct GL {
    const Graph* root = nullptr;
    bool all = true;
    PySet nodes;  // show_nodes.nodes when !all

    bool has(int k) const { return root->has_node(k) && (all || nodes.has(k)); }
    bool set_order() const { return !all && 2 * nodes.size() < root->n(); }
    template <class F> void each_node(F&& f) const {  // FilterAtlas.__iter__
        if (set_order()) nodes.each([&](int64_t k) { if (root->has_node((int)k)) f((int)k); });
        else for (int k : root->order) if (all || nodes.has(k)) f(k);
    }
    template <class F> void each_nbr(int u, F&& f) const {  // FilterAdjacency[u]: underlying order, filtered
        for (int v : root->nbrs(u)) if (all || nodes.has(v)) f(v);
    }
    std::vector<int> node_list() const { std::vector<int> v; each_node([&](int k) { v.push_back(k); }); return v; }
    size_t count() const { size_t c = 0; each_node([&](int) { ++c; }); return c; }
    bool has_edge(int u, int v) const { return has(u) && has(v) && root->has_edge(u, v); }
};

inline GL whole(const Graph& g) { GL v; v.root = &g; v.all = true; return v; }

// G.subgraph(seq): show_nodes(set(nbunch_iter(seq))) -- members of G, inserted in the order seq yields them
inline GL subview(const GL& g, const std::vector<int>& seq) {
    GL v; v.root = g.root; v.all = false;
    for (int k : seq) if (g.has(k)) v.nodes.add(k);
    return v;
}

// Graph.copy() of a graph or view: nodes in view order, then add_edge(u, v) for u in view order, v in adj order
inline Graph copy_of(const GL& g) {
    Graph h;
    std::vector<int> nodes = g.node_list();
    for (int k : nodes) h.add_node(k, g.root->at(k));
    for (int u : nodes) g.each_nbr(u, [&](int v) { h.add_edge(u, v); });
    return h;
}

// relabel_nodes(G, mapping, copy=True) for a total mapping given as a function
template <class M> inline Graph relabel_copy(const Graph& g, M&& map) {
    Graph h;
    for (int k : g.order) h.add_node(map(k));
    for (int k : g.order) h.at(map(k)) = g.at(k);
    g.each_edge([&](int u, int v) { h.add_edge(map(u), map(v)); });
    return h;
}

// R.add_nodes_from(G.nodes(data=True)); R.add_edges_from(G.edges(data=True))  (Graph.update / compose / union)
inline void absorb(Graph& r, const Graph& g, bool overwrite_attrs_of_existing) {
    for (int k : g.order) {
        bool had = r.has_node(k);
        int s = r.add_node(k);
        if (!had || overwrite_attrs_of_existing) r.attr[s] = g.at(k);
    }
    g.each_edge([&](int u, int v) { r.add_edge(u, v); });
}

// nx.connected_components(G): one PySet per component, filled in BFS order (components._plain_bfs)
inline std::vector<PySet> connected_components(const GL& g) {
    std::vector<PySet> out;
    std::vector<char> seen(g.root->slot_of.size(), 0);
    size_t n = g.count(), n_seen = 0;
    std::vector<int> nodes = g.node_list();
    for (int v : nodes) {
        if (seen[v]) continue;
        size_t target = n - n_seen;
        PySet c;
        c.add(v);
        std::vector<int> next{v}, level;
        bool done = false;
        while (!next.empty() && !done) {
            level.swap(next);
            next.clear();
            for (int a : level) {
                g.each_nbr(a, [&](int w) { if (!c.has(w)) { c.add(w); next.push_back(w); } });
                if (c.size() == target) { done = true; break; }
            }
        }
        c.each([&](int64_t k) { seen[k] = 1; });
        n_seen += c.size();
        out.push_back(std::move(c));
    }
    return out;
}

}  // namespace hg_htree
