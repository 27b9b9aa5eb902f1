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

// Adjacency list with inline storage for 8 neighbours: H-tree and scene-graph degrees are small and the builder makes
// ~200 graph copies per scene graph, so a heap allocation per adjacency list dominated its time.
struct NbrList {
    static constexpr uint32_t kInline = 8;
    int inl[kInline];
    int* heap = nullptr;
    uint32_t sz = 0, cap = kInline;

    NbrList() = default;
    NbrList(const NbrList& o) { assign(o); }
    NbrList(NbrList&& o) noexcept { steal(o); }
    NbrList& operator=(const NbrList& o) { if (this != &o) { release(); assign(o); } return *this; }
    NbrList& operator=(NbrList&& o) noexcept { if (this != &o) { release(); steal(o); } return *this; }
    ~NbrList() { release(); }

    int* data() { return heap ? heap : inl; }
    const int* data() const { return heap ? heap : inl; }
    int* begin() { return data(); }
    int* end() { return data() + sz; }
    const int* begin() const { return data(); }
    const int* end() const { return data() + sz; }
    size_t size() const { return sz; }
    bool empty() const { return sz == 0; }
    int operator[](size_t i) const { return data()[i]; }
    void push_back(int v) {
        if (sz == cap) grow(cap * 2);
        data()[sz++] = v;
    }
    void erase(int* first, int* last) {  // only the erase(remove(...), end()) idiom is used
        (void)last;
        sz = static_cast<uint32_t>(first - data());
    }
    std::vector<int> to_vector() const { return std::vector<int>(begin(), end()); }

 private:
    void grow(uint32_t ncap) {
        int* p = new int[ncap];
        std::copy(begin(), end(), p);
        delete[] heap;
        heap = p;
        cap = ncap;
    }
    void assign(const NbrList& o) {
        heap = nullptr; cap = kInline; sz = 0;
        if (o.sz > kInline) { heap = new int[o.sz]; cap = o.sz; }
        std::copy(o.begin(), o.end(), data());
        sz = o.sz;
    }
    void steal(NbrList& o) {
        heap = o.heap; sz = o.sz; cap = o.cap;
        if (!heap) std::copy(o.inl, o.inl + o.sz, inl);
        o.heap = nullptr; o.sz = 0; o.cap = kInline;
    }
    void release() { delete[] heap; heap = nullptr; }
};

struct Attr {
    uint8_t cls = kLeaf;
    uint8_t kind = kNoKind;
    bool is_list = false;      // clique_has: int (leaf / virtual) or list (clique)
    int single = -1;
    std::vector<int> list;
};

// key -> slot tables are as long as the largest key (H-tree node ids reach the hundreds) while most of the ~200 graphs
// built per scene graph have a handful of nodes: the tables are recycled through a thread-local pool (all entries -1
// when pooled; a graph resets only the keys it used), which removes a multi-KB allocation + fill per graph.
inline std::vector<std::vector<int>>& slot_pool() {
    thread_local std::vector<std::vector<int>> pool;
    return pool;
}

struct Graph {
    std::vector<int> order;              // node keys in dict order (nodes are never removed)
    std::vector<int> slot_of;            // key -> slot, -1 when absent (keys are small non-negative ints)
    int key_hi = 0;                      // 1 + largest key present

    Graph() { take_slots(); }
    Graph(const Graph& o) : order(o.order), key_hi(o.key_hi), adj(o.adj), attr(o.attr) {
        take_slots();
        if (slot_of.size() < (size_t)key_hi) slot_of.resize(key_hi, -1);
        for (size_t s = 0; s < order.size(); ++s) slot_of[order[s]] = (int)s;
    }
    Graph(Graph&& o) noexcept
        : order(std::move(o.order)), slot_of(std::move(o.slot_of)), key_hi(o.key_hi), adj(std::move(o.adj)),
          attr(std::move(o.attr)) {
        o.order.clear(); o.slot_of.clear(); o.key_hi = 0;
    }
    Graph& operator=(const Graph& o) {
        if (this != &o) { Graph t(o); *this = std::move(t); }
        return *this;
    }
    Graph& operator=(Graph&& o) noexcept {
        if (this != &o) {
            give_slots();
            order = std::move(o.order); slot_of = std::move(o.slot_of); key_hi = o.key_hi;
            adj = std::move(o.adj); attr = std::move(o.attr);
            o.order.clear(); o.slot_of.clear(); o.key_hi = 0;
        }
        return *this;
    }
    ~Graph() { give_slots(); }
    size_t key_span() const { return (size_t)key_hi; }  // size of a scratch array indexed by node key

 private:
    void take_slots() {
        auto& pool = slot_pool();
        if (!pool.empty()) { slot_of = std::move(pool.back()); pool.pop_back(); }
    }
    void give_slots() {
        if (slot_of.empty()) return;
        for (int k : order) slot_of[k] = -1;
        auto& pool = slot_pool();
        if (pool.size() < 64) pool.push_back(std::move(slot_of));
        slot_of.clear();
    }

 public:
    std::vector<NbrList> adj;            // per slot: neighbour keys in dict order
    std::vector<Attr> attr;              // per slot

    size_t n() const { return order.size(); }
    void reserve(size_t n) { order.reserve(n); adj.reserve(n); attr.reserve(n); }
    bool has_node(int k) const { return k >= 0 && k < key_hi && slot_of[k] >= 0; }
    int slot(int k) const { return slot_of[k]; }
    int add_node(int k) {
        if ((size_t)k >= slot_of.size()) slot_of.resize(std::max<size_t>(k + 1, slot_of.size() * 2), -1);
        if (slot_of[k] < 0) {
            if (k >= key_hi) key_hi = k + 1;
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
    const NbrList& nbrs(int k) const { return adj[slot_of[k]]; }
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
        std::vector<char> seen(key_span(), 0);
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
    h.reserve(nodes.size());
    for (int k : nodes) h.add_node(k, g.root->at(k));
    for (int u : nodes) g.each_nbr(u, [&](int v) { h.add_edge(u, v); });
    return h;
}

// relabel_nodes(G, mapping, copy=True) for a total mapping given as a function
template <class M> inline Graph relabel_copy(const Graph& g, M&& map) {
    Graph h;
    h.reserve(g.n());
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
    std::vector<char> seen(g.root->key_span(), 0);
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
