/* libhydra_htree.so -- native H-tree builder (host-only C ABI; no CUDA, no torch types).
 *
 * Replaces, for a batch of scene graphs, the reference's networkx pipeline
 *     generate_htree(dsg_torch)              neural_tree/construct.py:241-310
 *       -> sample_and_generate_jth / generate_jth   neural_tree/generate_junction_tree_hierarchies.py:26-116,248-333
 *     add_virtual_nodes_to_htree(htree_nx)   neural_tree/construct.py:313-371
 * and hands back the integer content nx_htree_to_torch (construct.py:374-483) turns into tensors: node types,
 * `clique_has`, the directed edge list in `list(htree_nx.edges)` order and the edge types.  Results are
 * bit-identical to the reference run on networkx 3.x / CPython 3.12 (node numbering included; see
 * hydra_gnn_b200/csrc/htree/pyorder.hpp for why that needs a model of Python's set iteration order).
 *
 * The reference-side binding a maintainer would add is shown in INTEGRATION.md (ctypes, from
 * Hydra_mp3d_htree_data.compute_torch_data, mp3d_dataset.py:437-458).
 */
#ifndef HYDRA_HTREE_H_
#define HYDRA_HTREE_H_

#include <stdint.h>

#if defined(__GNUC__)
#define HG_HTREE_API __attribute__((visibility("default")))
#else
#define HG_HTREE_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define HG_HTREE_EINVAL (-1)       /* bad pointer / extent / edge endpoint */
#define HG_HTREE_EUNSUPPORTED (-2) /* input the reference handles on another path (graph component of more than
                                      1000 nodes -> bounded-treewidth subsampling; component without a room) or
                                      fails on itself (clique member that never becomes a leaf) */
#define HG_HTREE_EINTERNAL (-3)

typedef struct hg_htree_batch hg_htree_batch;

/* Text of the last error raised on the calling thread ("" if none). */
HG_HTREE_API const char* hg_htree_last_error(void);

/* Builds the augmented H-trees of `n_graphs` independent scene graphs.
 *   node_ptr   [n_graphs+1]  graph g owns nodes node_ptr[g]..node_ptr[g+1]-1 of `is_room`
 *   is_room    [total nodes] 0 = object, 1 = room, in the node order of HeteroData.to_homogeneous()
 *                            (construct.py:46 -- objects first, then rooms, for the reference's datasets)
 *   edge_ptr   [n_graphs+1]  edge_ptr[g+1]-edge_ptr[g] = E_g
 *   edge_index [n_graphs]    per graph a pointer to its [2, E_g] int64 edge index (row 0 sources, row 1 targets),
 *                            node ids local to the graph, in to_homogeneous() column order; direction is ignored
 *                            the way to_networkx(...).to_undirected() ignores it (construct.py:52)
 *   n_threads                worker threads over graphs; <= 0 = hardware concurrency
 * Returns 0 and a handle in *out, or a negative code (*out = NULL). */
HG_HTREE_API int hg_htree_build(int32_t n_graphs, const int64_t* node_ptr, const uint8_t* is_room,
                                const int64_t* edge_ptr, const int64_t* const* edge_index, int32_t n_threads,
                                hg_htree_batch** out);

/* Totals over the batch: H-tree nodes, directed edges, and entries of the cliques' `clique_has` lists. */
HG_HTREE_API int hg_htree_sizes(const hg_htree_batch* b, int64_t* n_nodes, int64_t* n_edges, int64_t* n_list_items);

/* Copies the batch out.  Any pointer except node_ptr / edge_ptr may be NULL.
 *   node_ptr   [n_graphs+1]   H-tree g owns nodes node_ptr[g]..; node k of the tree is the networkx node with key k
 *   node_type  [nodes]        index into HTREE_NODE_TYPES + HTREE_VIRTUAL_NODE_TYPES (construct.py:14,29):
 *                             0 object, 1 room, 2 object-room, 3 room-room, 4 object_virtual, 5 room_virtual
 *   clique_has [nodes]        source node id (local to the input graph) of a leaf / virtual node, -1 for cliques
 *   node_order [nodes]        node keys in networkx dict order (what `htree_nx.nodes` iterates)
 *   edge_ptr   [n_graphs+1]
 *   edge_src, edge_dst [edges] node keys local to the tree, in `list(htree_nx.edges)` order
 *   edge_type  [edges]        index into HTREE_EDGE_TYPES + HTREE_POOL_EDGE_TYPES + HTREE_INIT_EDGE_TYPES
 *                             (construct.py:15-38, the default edge_type_names of nx_htree_to_torch)
 *   list_ptr   [nodes+1], list_items: the `clique_has` lists of the clique nodes (empty ranges elsewhere) */
HG_HTREE_API int hg_htree_export(const hg_htree_batch* b, int64_t* node_ptr, uint8_t* node_type, int32_t* clique_has,
                                 int32_t* node_order, int64_t* edge_ptr, int32_t* edge_src, int32_t* edge_dst,
                                 uint8_t* edge_type, int64_t* list_ptr, int32_t* list_items);

HG_HTREE_API void hg_htree_free(hg_htree_batch* b);

/* Test hooks for the container-order model (pyorder.hpp): replay add (op 0) / discard (op 1) on a Python-set
 * model and return its iteration order; CPython's hash of a tuple of non-negative ints. */
HG_HTREE_API int64_t hg_pyset_trace(const int32_t* ops, const int64_t* keys, const int64_t* hashes, int64_t n_ops,
                                    int64_t* out_keys, int64_t out_cap);
HG_HTREE_API int64_t hg_pytuple_hash(const int32_t* items, int64_t n);

#ifdef __cplusplus
}
#endif
#endif /* HYDRA_HTREE_H_ */
