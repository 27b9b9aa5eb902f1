// H-tree builder: construct.py:241-371 (generate_htree + add_virtual_nodes_to_htree) restated natively, in the
// integer form nx_htree_to_torch (construct.py:374-483) consumes.  Host-only C ABI, see include/hydra_htree.h.
#include "../../../include/hydra_htree.h"

#include <atomic>
#include <cstring>
#include <string>
#include <thread>

#include "jth.hpp"

using namespace hg_htree;

namespace {

struct Unsupported : std::runtime_error { using std::runtime_error::runtime_error; };

// one augmented H-tree in homogeneous form; nodes are indexed by KEY (nx_htree_to_torch reads nodes[i])
struct Tree {
    std::vector<int32_t> order;       // node keys in dict order (for the signature)
    std::vector<uint8_t> kind;        // by key
    std::vector<int32_t> clique_has;  // by key, -1 for cliques
    std::vector<int32_t> list_ptr, list_items;  // by key: clique_has lists of cliques (empty for the others)
    std::vector<int32_t> src, dst;    // list(htree.edges) of the directed augmented tree
};

const int kRelation[6][6] = {
    // dst:  object  room  object-room  room-room  object_virtual  room_virtual     (HTREE_EDGE_TYPES + POOL + INIT)
    /* object         */ {-1, -1, 0, -1, 10, -1},
    /* room           */ {-1, -1, 2, 4, -1, 11},
    /* object-room    */ {1, 3, 8, 6, -1, -1},
    /* room-room      */ {-1, 5, 7, 9, -1, -1},
    /* object_virtual */ {-1, -1, 12, -1, -1, -1},
    /* room_virtual   */ {-1, -1, 13, 14, -1, -1},
};

// HTree helper class of construct.py:186-238
struct HTreeAcc {
    Graph jth;
    std::vector<int> roots;

    void add_object_jth(Graph& ojth, const std::vector<int>& oroots, int room) {
        std::vector<char> is_root(ojth.key_span(), 0);
        for (int r : oroots) is_root[r] = 1;
        for (int r : oroots) {
            std::vector<int> nb = ojth.nbrs(r).to_vector();
            for (int v : nb) if (is_root[v]) ojth.remove_edge(r, v);
        }
        for (int root : roots) {
            Attr& a = jth.at(root);
            if (!a.is_list) { a.is_list = true; a.list.assign(1, a.single); }
            if (std::find(a.list.begin(), a.list.end(), room) == a.list.end()) continue;
            // g = ojth.copy(); relabel to base.. (new_labels = dict(zip(range(n), range(base, base + n))));
            // self.jth = nx.compose(self.jth, g).  The reference rebuilds the whole accumulated tree here; every such
            // rebuild (copy / relabel / compose) re-adds edges in EdgeView order, which keeps, for every node, the
            // relative order of its LATER neighbours (later in node order) and nothing else of the adjacency order --
            // and that is all the final node / edge lists depend on (they are produced by one more EdgeView pass in
            // disjoint_union).  Appending the relabelled nodes and edges in place preserves exactly that, without
            // the O(|jth|) rebuild per object component.
            const int base = (int)jth.n();
            const int n = (int)ojth.n();
            auto map = [&](int k) { return k < n ? base + k : k; };
            for (int k : ojth.order) {
                const int mk = map(k);
                const bool had = jth.has_node(mk);
                const int s = jth.add_node(mk);
                if (!had) jth.attr[s] = ojth.at(k);
            }
            ojth.each_edge([&](int u, int v) { jth.add_edge(map(u), map(v)); });
            for (int r : oroots) jth.add_edge(root, base + r);
        }
    }
};

Tree build_tree(int32_t n_nodes, const uint8_t* is_room, const int64_t* edge_index, int64_t n_edges) {
    // torch_dsg_to_nx: to_networkx (DiGraph, nodes 0..N-1, edges in column order) then to_undirected()
    std::vector<std::vector<int>> succ(n_nodes);
    for (int64_t e = 0; e < n_edges; ++e) {
        int64_t u = edge_index[e], v = edge_index[n_edges + e];
        if (u < 0 || v < 0 || u >= n_nodes || v >= n_nodes) throw std::out_of_range("edge endpoint out of range");
        auto& s = succ[u];
        if (std::find(s.begin(), s.end(), (int)v) == s.end()) s.push_back((int)v);
    }
    Graph dsg;
    for (int k = 0; k < n_nodes; ++k) {
        Attr a; a.cls = kLeaf; a.kind = is_room[k] ? kRoom : kObject; a.single = k;  // generate_node_labels
        dsg.add_node(k, a);
    }
    for (int u = 0; u < n_nodes; ++u) for (int v : succ[u]) dsg.add_edge(u, v);

    Graph htree;
    GL dsg_all = whole(dsg);
    for (const PySet& comp_set : connected_components(dsg_all)) {
        Graph comp = copy_of(subview(dsg_all, comp_set.items()));
        if (comp.n() > 1000) throw Unsupported("component above the treewidth-sampling bound (1000 nodes)");
        GL comp_all = whole(comp);
        std::vector<int> room_idx;
        for (int k : comp.order) if (comp.at(k).kind == kRoom) room_idx.push_back(k);
        if (room_idx.empty()) throw Unsupported("connected component without a room node");
        // H-tree of the room graph
        Jth rooms = generate_jth(subview(comp_all, room_idx), true);
        HTreeAcc acc;
        acc.jth = std::move(rooms.g);
        acc.roots = rooms.has_roots ? rooms.roots : std::vector<int>{0};
        if (room_idx.size() == 1) acc.roots = {0};
        for (Attr& a : acc.jth.attr) if (a.cls == kClique) a.kind = kRoomRoom;
        // H-trees of the object graphs, one per room and connected component
        for (int r : room_idx) {
            std::vector<int> object_idx;
            for (int v : comp.nbrs(r)) if (comp.at(v).kind == kObject) object_idx.push_back(v);
            GL oview = subview(comp_all, object_idx);
            for (const PySet& oc : connected_components(oview)) {
                Graph ogc = copy_of(subview(dsg_all, oc.items()));
                Jth oj = generate_jth(whole(ogc), true);
                Graph& jg = oj.g;
                std::vector<int> oroots = oj.roots;
                if (ogc.n() == 1) {
                    Attr c; c.cls = kClique; c.is_list = true; c.list.assign(1, jg.at(0).single);
                    jg.add_node(1, c);
                    jg.add_edge(0, 1);
                    oroots = {1};
                }
                for (Attr& a : jg.attr) if (a.cls == kClique) { a.kind = kObjectRoom; a.list.push_back(r); }
                PySet lowest;  // set(sum([[n for n in jth.neighbors(leaf)] for leaf in leaf_nodes], []))
                for (int k : jg.order) if (jg.at(k).cls == kLeaf) for (int v : jg.nbrs(k)) lowest.add(v);
                int idx = (int)jg.n();
                for (int c : lowest.items()) {
                    Attr a; a.cls = kLeaf; a.kind = kRoom; a.single = r;
                    jg.add_node(idx, a);
                    jg.add_edge(c, idx);
                    ++idx;
                }
                acc.add_object_jth(jg, oroots, r);
            }
        }
        if (room_idx.size() == 1 && acc.jth.n() > 1) {
            Attr& a = acc.jth.at(0);
            if (a.cls != kLeaf || !a.is_list || a.list.size() != 1) throw std::logic_error("single-room root");
            a.is_list = false; a.single = a.list[0]; a.list.clear();
        }
        // nx.disjoint_union(htree, acc.jth): both relabelled to consecutive integers, then union_all
        const int first = (int)htree.n();
        std::vector<int> pos_h(htree.key_span(), -1), pos_j(acc.jth.key_span(), -1);
        for (size_t s = 0; s < htree.n(); ++s) pos_h[htree.order[s]] = (int)s;
        for (size_t s = 0; s < acc.jth.n(); ++s) pos_j[acc.jth.order[s]] = first + (int)s;
        Graph a = relabel_copy(htree, [&](int k) { return pos_h[k]; });
        Graph b = relabel_copy(acc.jth, [&](int k) { return pos_j[k]; });
        Graph merged;
        absorb(merged, a, false);
        absorb(merged, b, false);
        htree = std::move(merged);
    }

    // add_virtual_nodes_to_htree: directed copy (succ[u] = adj[u]), pooling edges leaf -> virtual, init edges
    // virtual -> clique
    const int offset = (int)htree.n();
    for (int k : htree.order) if (k >= offset) throw std::logic_error("H-tree nodes are not consecutive");
    std::vector<int> order = htree.order;
    std::vector<Attr> attr(offset);
    std::vector<std::vector<int>> out_nbrs(offset);
    for (size_t s = 0; s < htree.n(); ++s) { attr[htree.order[s]] = htree.attr[s]; out_nbrs[htree.order[s]] = htree.adj[s].to_vector(); }
    auto ensure = [&](int key) {
        if ((size_t)key >= attr.size()) { attr.resize(key + 1); out_nbrs.resize(key + 1); }
    };
    std::vector<char> present;
    auto add_edge = [&](int u, int v) {
        auto& s = out_nbrs[u];
        if (std::find(s.begin(), s.end(), v) == s.end()) s.push_back(v);
    };
    std::vector<int> leaves, cliques;
    for (int k : order) (attr[k].cls == kLeaf ? leaves : cliques).push_back(k);
    present.assign(offset, 1);
    for (int leaf : leaves) {
        const int vkey = attr[leaf].single + offset;
        ensure(vkey);
        if (present.size() <= (size_t)vkey) present.resize(vkey + 1, 0);
        if (!present[vkey]) {
            present[vkey] = 1;
            Attr a; a.cls = kVirtual; a.kind = attr[leaf].kind + 4; a.single = attr[leaf].single;
            attr[vkey] = a;
            order.push_back(vkey);
        }
        add_edge(leaf, vkey);
    }
    for (int c : cliques)
        for (int member : attr[c].list) {
            const int vkey = member + offset;
            if ((size_t)vkey >= present.size() || !present[vkey])
                throw Unsupported("clique member that is no leaf of the H-tree (single-node clique of a sub-graph)");
            add_edge(vkey, c);
        }
    const int total = (int)order.size();
    for (int k : order) if (k >= total) throw Unsupported("augmented H-tree nodes are not consecutive integers");

    Tree t;
    t.order.assign(order.begin(), order.end());
    t.kind.resize(total);
    t.clique_has.resize(total);
    t.list_ptr.assign(total + 1, 0);
    for (int k = 0; k < total; ++k) {
        const Attr& a = attr[k];
        if (a.kind > kRoomVirtual) throw std::logic_error("node without node_type");
        t.kind[k] = a.kind;
        t.clique_has[k] = a.cls == kClique ? -1 : a.single;
        if (a.cls == kClique) t.list_items.insert(t.list_items.end(), a.list.begin(), a.list.end());
        else if (a.is_list) {  // single-room root left as a list (graph of one room and nothing else)
            t.list_items.insert(t.list_items.end(), a.list.begin(), a.list.end());
            t.clique_has[k] = a.list.size() == 1 ? a.list[0] : -1;
        }
        t.list_ptr[k + 1] = (int32_t)t.list_items.size();
    }
    for (int u : order)
        for (int v : out_nbrs[u]) {
            if (kRelation[t.kind[u]][t.kind[v]] < 0) throw std::logic_error("edge between unrelated node types");
            t.src.push_back(u);
            t.dst.push_back(v);
        }
    return t;
}

}  // namespace

struct hg_htree_batch {
    std::vector<Tree> trees;
};

thread_local std::string g_err;

extern "C" {

HG_HTREE_API const char* hg_htree_last_error(void) { return g_err.c_str(); }

HG_HTREE_API int hg_htree_build(int32_t n_graphs, const int64_t* node_ptr, const uint8_t* is_room,
                                const int64_t* edge_ptr, const int64_t* const* edge_index, int32_t n_threads,
                                hg_htree_batch** out) {
    if (!out || n_graphs < 0 || (n_graphs > 0 && (!node_ptr || !is_room || !edge_ptr || !edge_index))) return HG_HTREE_EINVAL;
    hg_htree_batch* b = new hg_htree_batch;
    b->trees.resize(n_graphs);
    std::atomic<int> next{0}, status{0};
    std::string err;
    std::atomic_flag err_lock = ATOMIC_FLAG_INIT;
    auto work = [&]() {
        for (;;) {
            int g = next.fetch_add(1);
            if (g >= n_graphs || status.load() != 0) return;
            int code = 0;
            std::string msg;
            try {
                const int64_t n = node_ptr[g + 1] - node_ptr[g], e = edge_ptr[g + 1] - edge_ptr[g];
                if (n < 0 || e < 0 || n > INT32_MAX / 4) throw std::out_of_range("bad graph extents");
                b->trees[g] = build_tree((int32_t)n, is_room + node_ptr[g], edge_index[g], e);
            } catch (const Unsupported& ex) { code = HG_HTREE_EUNSUPPORTED; msg = ex.what();
            } catch (const std::out_of_range& ex) { code = HG_HTREE_EINVAL; msg = ex.what();
            } catch (const std::exception& ex) { code = HG_HTREE_EINTERNAL; msg = ex.what(); }
            if (code) {
                while (err_lock.test_and_set()) {}
                if (status.load() == 0) { status.store(code); err = "graph " + std::to_string(g) + ": " + msg; }
                err_lock.clear();
            }
        }
    };
    int nt = n_threads > 0 ? n_threads : (int)std::thread::hardware_concurrency();
    nt = std::max(1, std::min(nt, n_graphs));
    if (nt == 1) work();
    else {
        std::vector<std::thread> pool;
        for (int i = 0; i < nt; ++i) pool.emplace_back(work);
        for (auto& th : pool) th.join();
    }
    if (status.load() != 0) { g_err = err; delete b; *out = nullptr; return status.load(); }
    *out = b;
    return 0;
}

HG_HTREE_API int hg_htree_sizes(const hg_htree_batch* b, int64_t* n_nodes, int64_t* n_edges, int64_t* n_list_items) {
    if (!b) return HG_HTREE_EINVAL;
    int64_t n = 0, e = 0, l = 0;
    for (const Tree& t : b->trees) { n += t.kind.size(); e += t.src.size(); l += t.list_items.size(); }
    if (n_nodes) *n_nodes = n;
    if (n_edges) *n_edges = e;
    if (n_list_items) *n_list_items = l;
    return 0;
}

HG_HTREE_API int hg_htree_export(const hg_htree_batch* b, int64_t* node_ptr, uint8_t* node_type, int32_t* clique_has,
                                 int32_t* node_order, int64_t* edge_ptr, int32_t* edge_src, int32_t* edge_dst,
                                 uint8_t* edge_type, int64_t* list_ptr, int32_t* list_items) {
    if (!b || !node_ptr || !edge_ptr) return HG_HTREE_EINVAL;
    int64_t n = 0, e = 0, l = 0;
    node_ptr[0] = 0;
    edge_ptr[0] = 0;
    if (list_ptr) list_ptr[0] = 0;
    for (size_t g = 0; g < b->trees.size(); ++g) {
        const Tree& t = b->trees[g];
        const size_t tn = t.kind.size(), te = t.src.size();
        if (node_type) std::memcpy(node_type + n, t.kind.data(), tn);
        if (clique_has) std::memcpy(clique_has + n, t.clique_has.data(), tn * 4);
        if (node_order) std::memcpy(node_order + n, t.order.data(), tn * 4);
        if (edge_src) std::memcpy(edge_src + e, t.src.data(), te * 4);
        if (edge_dst) std::memcpy(edge_dst + e, t.dst.data(), te * 4);
        if (edge_type) for (size_t k = 0; k < te; ++k) edge_type[e + k] = (uint8_t)kRelation[t.kind[t.src[k]]][t.kind[t.dst[k]]];
        if (list_ptr) for (size_t k = 0; k < tn; ++k) list_ptr[n + k + 1] = l + t.list_ptr[k + 1];
        if (list_items) std::memcpy(list_items + l, t.list_items.data(), t.list_items.size() * 4);
        n += tn; e += te; l += t.list_items.size();
        node_ptr[g + 1] = n;
        edge_ptr[g + 1] = e;
    }
    return 0;
}

HG_HTREE_API void hg_htree_free(hg_htree_batch* b) { delete b; }

// Test hook: replays a sequence of set operations and returns the iteration order (tests/test_htree.py
// checks it against the interpreter's own set).  op: 0 add key, 1 discard key; returns the item count.
HG_HTREE_API int64_t hg_pyset_trace(const int32_t* ops, const int64_t* keys, const int64_t* hashes, int64_t n_ops,
                                    int64_t* out_keys, int64_t out_cap) {
    PySet s;
    for (int64_t i = 0; i < n_ops; ++i) {
        if (ops[i] == 0) s.add(keys[i], hashes[i]);
        else s.discard(keys[i], hashes[i]);
    }
    int64_t n = 0;
    s.each([&](int64_t k) { if (n < out_cap) out_keys[n] = k; ++n; });
    return n;
}

HG_HTREE_API int64_t hg_pytuple_hash(const int32_t* items, int64_t n) { return tuple_hash(items, (size_t)n); }

}  // extern "C"
