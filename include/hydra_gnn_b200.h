/*
 * hydra_gnn_b200 -- C ABI of the sm_100a message-passing kernels.
 *
 * Drop-in boundary (inner).  The reference (MIT-SPARK/Hydra-GNN) has no FFI of its own:
 * every arithmetic step of its forward is a call into torch_geometric (PyG) operators,
 * which in turn run gather / scatter kernels from ATen / torch_scatter / pyg_lib.  Each
 * entry point below replaces one such operator family; the comment above it cites the
 * reference call site(s) (paths relative to /root/reference/src/hydra_gnn) and the PyG
 * operator it stands in for.  INTEGRATION.md shows the ctypes binding.
 *
 * Conventions
 *  - plain C: pointers and sizes only, no torch types.  All pointers are DEVICE pointers
 *    unless the name ends in `_host`.  The caller owns every buffer (inputs, outputs,
 *    workspace); the library never allocates, never synchronises, keeps no global state.
 *  - every launch goes to `stream` (a cudaStream_t passed as void*), so calls are
 *    CUDA-graph capturable and thread-safe per stream.
 *  - return value: 0 (HG_OK) or a negative HG_E* code; hg_strerror() gives the text.
 *  - features are fp32, row-major, with an explicit leading dimension (`ld*`, in floats);
 *    indices inside the library are int32 (edge_index arrives as int64 and is narrowed by
 *    hg_plan_build, which reports HG_ERANGE if a count does not fit).
 *  - CSR: rowptr[n_dst+1], col[E'] (source ids), perm[E'] (original edge id, -1 for an
 *    added self loop).  Rows are ordered by original edge id, added loops last: identical
 *    to a stable sort by destination of PyG's (remove_self_loops -> add_self_loops) list.
 */
#ifndef HYDRA_GNN_B200_H_
#define HYDRA_GNN_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HG_ABI_VERSION 2

#if defined(__GNUC__)
#define HG_API __attribute__((visibility("default")))
#else
#define HG_API
#endif

enum {
    HG_OK = 0,
    HG_EINVAL = -1,   /* bad argument (null pointer, negative size, unsupported shape) */
    HG_ERANGE = -2,   /* a size does not fit int32 / an index is outside [0, n) */
    HG_EWORKSPACE = -3, /* workspace too small */
    HG_ECUDA = -4,    /* a CUDA runtime call failed; see hg_last_cuda_error() */
    HG_EALIGN = -5,   /* pointer / leading dimension violates the alignment contract */
    HG_EUNSUPPORTED = -6,
    HG_ECAPACITY = -7 /* problem larger than the fixed capacity of a *_static entry point */
};

enum { HG_REDUCE_SUM = 0, HG_REDUCE_MEAN = 1, HG_REDUCE_MAX = 2 };
enum { HG_ACT_NONE = 0, HG_ACT_RELU = 1, HG_ACT_ELU = 2 };
enum { HG_GROUP_SUM = 0, HG_GROUP_MEAN = 1 };
enum { HG_PREC_FP32 = 0, HG_PREC_TF32X3 = 1 };

#define HG_MAX_RELATIONS 8   /* relations fused per destination type (H-tree needs 4) */
#define HG_MAX_SEGMENTS 8    /* K-segments of one fused projection */
#define HG_MAX_PLAN_RELATIONS 16 /* relations per plan-build call (H-tree has 15) */
#define HG_TILE_ROWS 32      /* destination rows per tile of hg_plan_rel.tile_win / hg_gat_fused */

HG_API const char* hg_strerror(int code);
HG_API int hg_abi_version(void);
/* last cudaError_t seen by this library on the calling thread (0 = none) */
HG_API int hg_last_cuda_error(void);
/* 0 if no usable device / kernel image for it; otherwise SM count of the current device */
HG_API int hg_device_sm_count(void);
/* sizeof of ABI struct #which (0 plan_rel, 1 seg_args, 2 gat_rel, 3 gat_args, 4 lin_seg, 5 linear_args,
 * 6 gat_train_rel, 7 gat_train_args, 8 htree_split_args) */
HG_API int hg_sizeof(int which);

/* ------------------------------------------------------------------------------------
 * K1  Graph plan: CSR by destination for up to HG_MAX_PLAN_RELATIONS relations at once.
 *
 * Replaces: the implicit per-layer `index_select` + `scatter` bookkeeping of PyG
 * MessagePassing.propagate, plus `remove_self_loops`/`add_self_loops` inside GATConv
 * (built by models/utils.py:49-86 with add_self_loops=(source == target), :124) and
 * `add_remaining_self_loops` inside gcn_norm (GCNConv, models/utils.py:16).  Built once
 * per batch, reused by every layer.
 *
 * self_loops = 0: edges kept as they are.
 * self_loops = 1: edges with src == dst dropped, one loop (k,k) appended for every
 *                 k < min(n_src, n_dst) (perm = -1).  This is both GATConv's and (for unit
 *                 edge weights) gcn_norm's edge list.
 * Out-of-range endpoints are skipped and counted in status[0] (never dereferenced).
 * E' = rowptr[n_dst] stays on the device; col/perm must hold E + (self_loops ? min(n_src,
 * n_dst) : 0) entries.  Integer work: bit-exact with the stable-sort oracle.
 * ---------------------------------------------------------------------------------- */
typedef struct hg_plan_rel {
    const int64_t* src;   /* [E] source ids      (edge_index[0]) */
    const int64_t* dst;   /* [E] destination ids (edge_index[1]) */
    int64_t n_edges;      /* E */
    int32_t n_src;
    int32_t n_dst;
    int32_t self_loops;   /* 0 | 1 */
    int32_t reserved;
    int32_t* rowptr;      /* out [n_dst + 1] */
    int32_t* col;         /* out [E + loops] */
    int32_t* perm;        /* out [E + loops], may be NULL */
    float* inv_sqrt_deg;  /* out [n_dst] deg^-1/2 (0 when deg == 0), may be NULL (gcn_norm) */
    int32_t* tile_win;    /* out [ceil(n_dst / HG_TILE_ROWS)][2], may be NULL: for every tile of HG_TILE_ROWS consecutive
                             destination rows, {lowest source id, number of source rows spanned} over the tile's CSR
                             entries ({0, 0} for a tile without edges).  A batch of scene graphs is a disjoint union in
                             graph order, so the span of a tile is a few dozen rows: hg_gat_fused stages it in shared
                             memory with bulk copies (hg_gat_rel.tile_win) */
} hg_plan_rel;

HG_API size_t hg_plan_workspace_bytes(const hg_plan_rel* rels_host, int n_rel);
HG_API int hg_plan_build(const hg_plan_rel* rels_host, int n_rel, int32_t* status /* [1], zeroed by the call */,
                  void* workspace, size_t workspace_bytes, void* stream);

/* hg_plan_build without a caller workspace and without any allocation (no cudaMalloc / cudaFree, no allocating
 * library routine) -- for code under CUDA-graph capture, tight serving loops and harnesses that forbid allocation.
 * NO counterpart in the reference.
 * The scratch is a file-scope `static __device__` array of HG_PLAN_STATIC_SCRATCH_BYTES inside the library, held by
 * every GPU that loads it until process exit.  Same outputs, bit for bit, as hg_plan_build.  Returns HG_ECAPACITY
 * (before any CUDA call) when hg_plan_workspace_bytes(rels, n_rel) - 256 exceeds HG_PLAN_STATIC_SCRATCH_BYTES
 * (16 MiB holds one 4096-graph MP3D-shaped DSG batch; an H-tree batch of that size does not fit).
 * All calls on a device share the one array, which keeps the previous call's contents (the plan zeroes what it
 * needs zeroed): calls that overlap on the GPU -- different streams, different threads -- RACE; serialise them. */
#define HG_PLAN_STATIC_SCRATCH_BYTES (16u << 20)
HG_API int hg_plan_build_static(const hg_plan_rel* rels_host, int n_rel, int32_t* status /* [1], zeroed by the call */,
                  void* stream);

/* ------------------------------------------------------------------------------------
 * K2 / K5  Segmented reduction over CSR rows (sum / mean / max), warp-per-destination,
 * vectorised feature loads, neighbour lists staged through shared memory.
 *
 *   out[r, :] = act( dst_scale[i] * reduce_{e in row i} src_scale[col[e]] * x[col[e], :]
 *                    + self_coef * x_self[i, :] + bias[:] ),   i = row_map ? row_map[r] : r
 *
 * Replaces: MessagePassing.aggregate -> scatter(reduce) of SAGEConv (mean, models/utils.py:14),
 * GCNConv (sum of D^-1/2 (A+I) D^-1/2 messages, :16; src_scale = dst_scale = deg^-1/2),
 * GINConv (sum + (1+eps) x_i, :18-26; self_coef_ptr = eps, self_coef_add = 1),
 * and LeafPool (mean; heterogeneous_neural_tree_network.py:18-31,181-185;
 * homogeneous_neural_tree_network.py:100 with row_map = indices of room_mask).
 * Empty rows reduce to 0 (PyG semantics) before the self / bias terms.  `x` may be NULL only when the
 * source set is empty (n_src == 0: every row is empty and x is never read); the same holds for
 * `xs` / `a_src` of hg_gat_fused.
 * ---------------------------------------------------------------------------------- */
typedef struct hg_seg_args {
    const float* x;  int64_t ldx;          /* [n_src, F] gathered operand */
    const int32_t* rowptr; const int32_t* col;
    const float* src_scale;                 /* NULL | [n_src] */
    const float* dst_scale;                 /* NULL | [n_dst] */
    const int32_t* row_map;                 /* NULL | [n_out] CSR row of output row r */
    const float* x_self; int64_t ldself;    /* NULL | [n_dst, F] indexed by i */
    const float* self_coef_ptr;             /* NULL | device scalar c; coefficient = self_coef_add + c */
    float self_coef_add;
    const float* bias;                      /* NULL | [F] */
    float* out; int64_t ldout;              /* [n_out, F] */
    int32_t n_out; int32_t F;
    int32_t op;  /* HG_REDUCE_* */
    int32_t act; /* HG_ACT_* */
} hg_seg_args;

HG_API int hg_seg_reduce(const hg_seg_args* args_host, void* stream);

/* ------------------------------------------------------------------------------------
 * K3 + K4  Fused GAT for ONE destination node type: for every relation r targeting it,
 * edge score -> segment softmax -> weighted aggregation -> head concat/mean -> +bias_r;
 * then the HeteroConv cross-relation group (sum | mean) and the model's activation, all in
 * one pass; `out` is written exactly once.
 *
 *   s_e,h  = leaky_relu(a_src[col[e], h] + a_dst[i, h] (+ a_edge[e, h]), neg_slope)
 *   w_e,h  = exp(s_e,h - max_row) / (sum_row exp(.) + 1e-16)
 *   o_r[i] = concat_h | mean_h ( sum_e w_e,h * xs[col[e], h, :] ) + bias_r
 *   out[i] = act( group_r o_r[i] + addend[i] )
 *
 * Replaces: GATConv.edge_update + torch_geometric.utils.softmax (4 scatter passes) +
 * propagate + concat/mean + bias (models/utils.py:31-87,103-140), HeteroConv's
 * `torch.stack(outs).sum(0)` / `.mean(0)` (models/utils.py:132-138;
 * heterogeneous_neural_tree_network.py:103), and the following F.elu / F.relu
 * (heterogeneous_network.py:108-119; homogeneous_network.py:135).
 * xs / a_src / a_dst are columns of the fused projection output (hg_linear), hence the
 * leading dimensions.  a_edge is in CSR order (one row per CSR entry).
 * ---------------------------------------------------------------------------------- */
typedef struct hg_gat_rel {
    const float* xs;    int64_t ldxs;   /* [n_src, H*C] projected sources */
    const float* a_src; int64_t lda_src;/* [n_src, H] */
    const float* a_dst; int64_t lda_dst;/* [n_dst, H] */
    const float* a_edge;                /* NULL | [E', H] contiguous, CSR order */
    const int32_t* rowptr; const int32_t* col;
    const float* bias;                  /* NULL | [H*C] (concat) or [C] */
    const int32_t* tile_win;            /* NULL | [ceil(n_dst / HG_TILE_ROWS)][2] from hg_plan_build: source-row window of
                                           every destination tile.  A hint: when EVERY relation of the call carries one
                                           (head-mean, C = 64, H <= 4, n_rel <= 4, no alpha_saved / dropout) the kernel
                                           bulk-copies the windows of xs into shared memory and gathers from there; sources
                                           outside the staged rows are read from global memory, so any window that stays
                                           inside [0, n_src) gives the same result.  Pass it for relations whose sources
                                           are local (a few dozen rows per tile, <= ~9 edges per row) */
    int32_t n_src;                      /* source rows of xs: required (> 0) with tile_win -- the window kernel describes xs
                                           to the TMA unit as a [n_src, H*C] tensor so that a box reaching past the last row
                                           is zero-filled instead of read; ignored otherwise */
    int32_t reserved;
} hg_gat_rel;

typedef struct hg_gat_args {
    hg_gat_rel rel[HG_MAX_RELATIONS];
    int32_t n_rel;
    int32_t n_dst;
    int32_t H; int32_t C;
    int32_t concat;     /* 1: out width H*C; 0: head mean, out width C */
    int32_t group;      /* HG_GROUP_* over relations */
    int32_t act;        /* HG_ACT_* */
    float neg_slope;
    float* out; int64_t ldout;
    float* alpha_saved[HG_MAX_RELATIONS]; /* NULL | [E', H] softmax weights (for hg_gat_train_bwd) */
    const float* addend; int64_t ldadd;   /* NULL | [n_dst, out width]: added after the group, before act
                                             (output of aggregate-first relations of the same dst type) */
    int32_t src_shared; /* 1: aggregate-first form -- xs is [n_src, C], ONE row shared by all heads;
                           out[i, h, :] = sum_e w_eh xs[col[e], :], width H*C (requires concat = 1) */
    float dropout_p;    /* training forward only (0 for inference): attention dropout, GATConv(dropout=p) --
                           w_e,h is zeroed with probability p (counter-based hash of seed, relation slot, CSR
                           position, head) and scaled by 1/(1-p) otherwise; alpha_saved then holds the softmax weight
                           BEFORE dropout with the decision in its sign bit (negative = dropped), the encoding
                           hg_gat_train_bwd reads */
    uint64_t dropout_seed;
    const uint64_t* dropout_seed_dev; /* NULL | device scalar added to dropout_seed when the kernel runs: a captured
                                         CUDA graph of a training step draws a fresh mask on every replay (the
                                         caller advances the scalar inside the graph) */
} hg_gat_args;

HG_API int hg_gat_fused(const hg_gat_args* args_host, void* stream);

/* ------------------------------------------------------------------------------------
 * K6  Fused projection:  y[M, N] = act( (sum_s x_s[M, K_s] * W_s[N, K_s]^T + bias) * scale + shift )
 *
 * The K dimension is a virtual concatenation of up to HG_MAX_SEGMENTS (x_s, W_s) pairs, so
 * SAGEConv's lin_l(mean) + lin_r(x) (+ the HeteroConv sum over relations) is ONE call.  The
 * attention logits a = (xW^T).att are produced by the same call: the caller appends the
 * folded rows (W^T att)^T to W (see hydra_gnn_b200/nn.py), so they are just extra columns.
 *
 * Replaces: torch.nn.Linear inside GATConv.lin_src/lin_dst/lin_edge, SAGEConv.lin_l/lin_r,
 * GCNConv.lin, GINConv.nn and post_mp_room/object (models/utils.py:9-28,49-86;
 * homogeneous_network.py:119-120) -- cuBLAS SGEMM in the reference.
 * precision: HG_PREC_FP32 = CUDA-core fp32 FMA; HG_PREC_TF32X3 = tcgen05 kind::tf32 with
 * 3-term error compensation (fp32-class accuracy on the tensor pipe).
 * ---------------------------------------------------------------------------------- */
typedef struct hg_lin_seg {
    const float* x; int64_t ldx;   /* [M, K] */
    const float* w; int64_t ldw;   /* [N, K] */
    int32_t K; int32_t reserved;
} hg_lin_seg;

typedef struct hg_linear_args {
    hg_lin_seg seg[HG_MAX_SEGMENTS];
    int32_t n_seg;
    int32_t M; int32_t N;
    const float* bias;   /* NULL | [N] */
    const float* scale;  /* NULL | [N]  (eval-mode BatchNorm fold) */
    const float* shift;  /* NULL | [N] */
    int32_t act;         /* HG_ACT_* */
    int32_t precision;   /* HG_PREC_* */
    float* y; int64_t ldy;
    int32_t w_prepacked; /* 1: `workspace` already holds the image written by hg_linear_pack_weights for these
                            weights (same N, K segmentation); the call then launches ONE kernel */
    int32_t reserved;
} hg_linear_args;

/* Bytes of the split + swizzled weight image the tensor-core path reads (0 for HG_PREC_FP32): the size of
 * `workspace` for hg_linear and of `dst` for hg_linear_pack_weights.  Depends on N and the K segments only. */
HG_API size_t hg_linear_workspace_bytes(const hg_linear_args* args_host);
/* Weights are constant across the batches of an inference run: pack them once (x / y / M are ignored). */
HG_API int hg_linear_pack_weights(const hg_linear_args* args_host, void* dst, size_t dst_bytes, void* stream);
HG_API int hg_linear(const hg_linear_args* args_host, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------
 * Edge-attribute helpers for GAT_edge (homogeneous_network.py:72-82;
 * heterogeneous_network.py:68-79; mp3d_dataset.py:298-319).
 *   hg_edge_logits: a_edge[p, h] = perm[p] >= 0 ? dot(edge_attr[perm[p], :D], v[h, :D]) : fill_h
 *   with v = folded (lin_edge^T att_edge) [H, D]; fill_h = dot(fill[:D], v[h]) (loops).
 *   hg_relative_pos: edge_attr[e, :] = pos_dst[dst[e], :3] - pos_src[src[e], :3]   (compute_relative_pos,
 *   mp3d_dataset.py:316-319; exact fp32 subtraction, bit-identical to the torch expression)
 * ---------------------------------------------------------------------------------- */
HG_API int hg_edge_logits(const float* edge_attr, int64_t ld_attr, int32_t D, const int32_t* perm, const int32_t* rowptr_last,
                   int64_t cap, const float* v /* [H, D] */, const float* fill /* NULL | [D] */, int32_t H,
                   float* a_edge /* [cap, H] */, void* stream);
HG_API int hg_relative_pos(const float* pos_src /* [n_src, 3] */, int64_t n_src, const float* pos_dst /* [n_dst, 3] */,
                    int64_t n_dst, const int64_t* src, const int64_t* dst, int64_t E, float* edge_attr /* [E, 3] */,
                    int32_t* status /* NULL | [1]: += number of edges with an endpoint outside [0, n) -- their rows are
                                       NaN, nothing is dereferenced (the caller zeroes it) */, void* stream);

/* out[r, :F] = act(x[idx[r], :F]) -- the `x[room_mask, :]` / `x[object_mask, :]` row selection of
 * homogeneous_network.py:139-145 and homogeneous_neural_tree_network.py:103-109 (idx = nonzero(mask),
 * computed once per batch with the plan). */
HG_API int hg_gather_rows(const float* x, int64_t ldx, const int32_t* idx, int64_t n_out, int32_t F, int32_t act,
                          float* out, int64_t ldout, void* stream);

/* Collate (integer, bit-exact): the index half of PyG `Batch.from_data_list` (torch_geometric is what
 * base_training_job.py:164-178's DataLoader calls; restated in hydra_gnn_b200/data.py).  `index` is the
 * plain concatenation [2, E] of the per-graph edge lists (graph-local node ids); entries of graph g
 * (columns edge_ptr[g] .. edge_ptr[g+1]) get src_ptr[g] added to row 0 and dst_ptr[g] to row 1, in place.
 * All pointers are device int64; `*_ptr` have n_graphs + 1 entries (exclusive prefix sums of the counts). */
HG_API int hg_collate_offsets(int64_t* index, int64_t n_edges, const int64_t* edge_ptr, const int64_t* src_ptr,
                              const int64_t* dst_ptr, int32_t n_graphs, void* stream);

/* batch[n] = g for ptr[g] <= n < ptr[g+1]: the `batch` vector of a collated node type (device int64). */
HG_API int hg_batch_vector(int64_t* batch, int64_t n_nodes, const int64_t* ptr, int32_t n_graphs, void* stream);

/* ------------------------------------------------------------------------------------
 * hg_htree_split -- the integer half of `heterogeneous_htree_to_homogeneous` (mp3d_dataset.py:72-119) applied to every
 * graph of a batched heterogeneous H-tree and collated (PyG `Batch.from_data_list`, base_training_job.py:164-178), as
 * one pass on the device: per graph the node types in store order (PyG `to_homogeneous`), every node's homogeneous
 * index (`place`), `node_type`, `batch`, the homogeneous `ptr`, `room_mask` / `object_mask`, and the edges of all relations
 * -- per graph in relation store order, endpoints mapped to homogeneous ids -- split by endpoint type into
 * init (source is a virtual node), pool (target is a virtual node, source is not) and tree (neither) lists.
 * Integer work, bit-exact with the reference's per-graph conversion + collate (tests/golden/convert_golden.npz).
 * Every relation's edge list must be grouped by graph in graph order (what collate produces); edges that are not, or
 * whose endpoints lie outside their graph's node range, are counted in status[0] and their slots hold -1.
 * Output sizes are host-known: N = sum of the types' node counts; the three edge lists hold the summed edge counts of
 * the relations of each class.
 * ---------------------------------------------------------------------------------- */
#define HG_MAX_HTREE_TYPES 8
enum { HG_EDGE_TREE = 0, HG_EDGE_INIT = 1, HG_EDGE_POOL = 2 };

typedef struct hg_htree_type {
    const int64_t* ptr;      /* [n_graphs + 1] first node of every graph within this type (PyG Batch `ptr`) */
    int64_t n_nodes;         /* = ptr[n_graphs] */
    int64_t* place;          /* out, may be NULL: [n_nodes] homogeneous index of every node of this type */
    int32_t is_room_virtual; /* 1: nodes of this type set room_mask */
    int32_t is_object_virtual;
} hg_htree_type;

typedef struct hg_htree_rel {
    const int64_t* src;      /* [n_edges] ids within src_type (batched) */
    const int64_t* dst;      /* [n_edges] ids within dst_type */
    int64_t n_edges;
    int32_t src_type, dst_type; /* indices into type[] */
} hg_htree_rel;

typedef struct hg_htree_split_args {
    hg_htree_type type[HG_MAX_HTREE_TYPES];
    hg_htree_rel rel[HG_MAX_PLAN_RELATIONS];
    int32_t n_types, n_rel, n_graphs, reserved;
    int64_t* ptr;            /* out [n_graphs + 1] homogeneous node offsets */
    int64_t* node_type;      /* out [N] index into type[] */
    int64_t* batch;          /* out [N] graph of every node */
    uint8_t* room_mask;      /* out [N] 0 / 1 (torch.bool storage) */
    uint8_t* object_mask;    /* out [N] */
    int64_t* edges[3];       /* out, by HG_EDGE_*: [2, n_class] row-major (row 0 = sources), n_class = summed n_edges */
    int32_t* status;         /* [1], zeroed by the call */
} hg_htree_split_args;

HG_API size_t hg_htree_split_workspace_bytes(const hg_htree_split_args* args_host);
HG_API int hg_htree_split(const hg_htree_split_args* args_host, void* workspace, size_t workspace_bytes, void* stream);

/* ====================================================================================
 * Training step (BASELINE.json configs[4]; reference: base_training_job.py:203-219 --
 * `loss.backward()` through the PyG layers, then `optimizer.step()` of torch.optim.Adam).
 * The entry points below are the backward mirrors of K3/K4 and K6 plus the optimizer update.
 * ==================================================================================== */

/* ------------------------------------------------------------------------------------
 * hg_gat_train_fwd / hg_gat_train_bwd -- GAT aggregation of every relation targeting ONE
 * destination type, training form: same arithmetic as hg_gat_fused (project-first), plus
 * attention dropout (GATConv(dropout=p): F.dropout on the softmax weights, models/utils.py:49-86)
 * and the softmax weights saved for the backward.
 *
 *   forward :  w_e,h (softmax, saved in alpha)  ->  w'_e,h = w_e,h * keep(seed, r, e, h) / (1 - p)
 *              out[i] = act( group_r ( concat_h|mean_h sum_e w'_e,h xs[col[e], h, :] + bias_r ) )
 *   backward:  g[i]   = dout[i] * act'(out[i]) * (group == MEAN ? 1 / n_rel : 1)        (pass 1, by destination)
 *              dbias += sum_i g[i]                      (same vector for every relation of the group)
 *              dw'_e,h = <g_h[i], xs[col[e], h, :]>;  dw = dw' * keep / (1 - p)
 *              ds_e,h = w_e,h (dw_e,h - sum_e' w_e',h dw_e',h) * leaky'(raw score)       -> dlogit[e, h]
 *              da_dst[i, h] = sum_e ds_e,h
 *              dxs[j, h, :] = sum_{e: col[e] = j} w'_e,h g_h[dst(e)]                      (pass 2, by source, over the
 *              da_src[j, h] = sum_{e: col[e] = j} ds_e,h                                  transposed CSR; no atomics)
 * keep() is a counter-based hash of (seed, relation slot, CSR position, head): the backward regenerates it.
 * Every output column is written exactly once (plain stores).
 * Limits of one call: H <= 4 and H * C <= 256 (HG_EUNSUPPORTED beyond).  Heads attend independently and, with one head, the
 * softmax backward is linear in dw, so a caller covers wider layers exactly by head groups / channel chunks over its own
 * rows of the parameters (hydra_gnn_b200/train.py: _run_head_groups, _run_wide_single_head).
 * ---------------------------------------------------------------------------------- */
typedef struct hg_gat_train_rel {
    const float* xs;    int64_t ldxs;    /* [n_src, H*C] */
    const float* a_src; int64_t lda_src; /* [n_src, H] */
    const float* a_dst; int64_t lda_dst; /* [n_dst, H] */
    const int32_t* rowptr; const int32_t* col; /* CSR by destination */
    const float* bias;                   /* NULL | [out width] (forward only) */
    float* alpha;                        /* [E', H]: written by fwd (before dropout), read by bwd */
    /* ---- backward only ---- */
    float* dlogit;                       /* [E', H] scratch: written by pass 1, read by pass 2 */
    const int32_t* rowptr_t; const int32_t* col_t; /* CSR of the reversed relation (rows = sources, col = destination) */
    const int32_t* pos_t;                /* [E'] position in the forward CSR of every entry of the reversed one */
    float* dxs;    int64_t lddxs;        /* [n_src, H*C] */
    float* da_src; int64_t ldda_src;     /* [n_src, H] */
    float* da_dst; int64_t ldda_dst;     /* [n_dst, H] */
    int32_t n_src; int32_t reserved;
    const float* a_edge;                 /* bwd: NULL | [E', H] in CSR order, the edge term of the raw score (GAT_edge: the
                                          * forward is hg_gat_fused with the same a_edge); its gradient IS dlogit */
} hg_gat_train_rel;

typedef struct hg_gat_train_args {
    hg_gat_train_rel rel[HG_MAX_RELATIONS];
    int32_t n_rel; int32_t n_dst;
    int32_t H; int32_t C;
    int32_t concat; int32_t group; int32_t act; float neg_slope;
    float dropout_p; int32_t reserved;
    uint64_t seed;
    float* out; int64_t ldout;           /* fwd: written; bwd: read (activation derivative) */
    const float* dout; int64_t lddout;   /* bwd: gradient w.r.t. out */
    float* g; int64_t ldg;               /* bwd scratch [n_dst, out width] */
    float* dbias;                        /* bwd: NULL | [out width], accumulated with atomics (caller zeroes) */
} hg_gat_train_args;

HG_API int hg_gat_train_fwd(const hg_gat_train_args* args_host, void* stream);
HG_API int hg_gat_train_bwd(const hg_gat_train_args* args_host, void* stream);

/* pos_t[q] = position in the forward CSR (perm / col, E' = *n_live entries) of entry q of the reversed CSR
 * (perm_t / col_t, *n_live_t entries); both built by hg_plan_build from the same edge list with the same
 * self_loops flag and perm requested.  scratch: int32 [n_edges + n_loop_nodes]. */
HG_API int hg_csr_transpose_map(const int32_t* perm, const int32_t* col, const int32_t* n_live, int64_t cap,
                                const int32_t* perm_t, const int32_t* col_t, const int32_t* n_live_t,
                                int64_t n_edges, int32_t n_loop_nodes, int32_t* scratch, int32_t* pos_t, void* stream);

/* ------------------------------------------------------------------------------------
 * hg_linear_wgrad -- weight gradient of K6:  dW[N, K] (+)= dY[M, N]^T X[M, K]  (fp32 FMA, split over M,
 * partial tiles added with atomics; `accumulate` = 0 zeroes dW first).  The input gradient dX = dY W is
 * hg_linear with the transposed weight.  Replaces the autograd of torch.nn.Linear (cuBLAS SGEMM).
 * ---------------------------------------------------------------------------------- */
HG_API int hg_linear_wgrad(const float* dy, int64_t lddy, const float* x, int64_t ldx, int32_t M, int32_t N, int32_t K,
                           float* dw, int64_t lddw, int32_t accumulate, void* stream);

/* ------------------------------------------------------------------------------------
 * hg_adam_step -- torch.optim.Adam (base_training_job.py:77-79: lr, weight_decay; betas (0.9, 0.999),
 * eps 1e-8, amsgrad off) over ONE flat parameter buffer:
 *   g = grad * grad_scale + weight_decay * p;  m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2
 *   p -= lr / (1 - b1^t) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
 * `step` (t >= 1) is a host value.  grad_scale folds the 1 / world_size of a gradient all-reduce.
 * ---------------------------------------------------------------------------------- */
HG_API int hg_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                        float beta1, float beta2, float eps, float weight_decay, float grad_scale, int32_t step,
                        const int32_t* step_dev /* NULL | device scalar t >= 1 used instead of `step` (a captured
                        CUDA graph of the training step advances it inside the graph) */, void* stream);

/* g[m, c] = dy[m, c] * act'(y[m, c]) and dbias[c] += sum_m g[m, c] (either output may be NULL; `g` may alias `dy`).
 * y is the activation's OUTPUT (ELU: y > 0 ? 1 : y + 1; ReLU: y > 0).  Backward of the fused bias + activation
 * epilogues of hg_linear / hg_seg_reduce (autograd of F.relu / F.elu + the bias broadcast in the reference). */
HG_API int hg_act_bwd(const float* dy, int64_t lddy, const float* y, int64_t ldy, int64_t M, int32_t N, int32_t act,
                      float* g, int64_t ldg, float* dbias /* NULL | [N], accumulated (caller zeroes) */, void* stream);

/* ------------------------------------------------------------------------------------
 * hg_synth_mp3d_* -- the synthetic "MP3D-shaped" scene-graph batch of hydra_gnn_b200/synth.py generated on the
 * device (BASELINE.json configs[3] streams 1 M graphs in micro-batches of 16,384 that are
 * generated where they are consumed -- ~75 GB of fp32 inputs do not fit host memory).  No counterpart in the
 * reference, whose inputs are preprocessed Matterport3D pickles (mp3d_dataset.py:135-296); the node / edge layout
 * is that of its HeteroData (EDGE_TYPES order, objects grouped by room).  Every draw is the counter hash
 * splitmix64(seed, graph id, stream, index) of synth.py, so graph g is the same whichever rank, micro-batch or side
 * (host / device) generates it: the integer structure (counts, edge lists, labels, batch vectors) and every feature
 * that is an affine map of a uniform draw are BIT-IDENTICAL to the host generator; the object positions go through
 * log / cos / sqrt in double (Box-Muller) and agree to the last float32 bit except where libm and CUDA round the
 * double differently (~1 value in 2^29).
 *
 *   hg_synth_mp3d_counts: prefix[g] = exclusive prefix sums over the graphs of {rooms, objects, room-room pairs,
 *                         object-object pairs}, prefix[n_graphs] = totals (the caller reads them to size the outputs:
 *                         one host sync per micro-batch, or ONE for all micro-batches of a job).
 *   hg_synth_mp3d_fill:   features, labels, batch vectors and the four edge lists ([2, E] int64, row 0 = source);
 *                         every undirected pair gives the two directed edges a->b, b->a next to each other.
 * ---------------------------------------------------------------------------------- */
#define HG_SYNTH_MAX_ROOMS 12
#define HG_SYNTH_MAX_OBJECTS 40
#define HG_SYNTH_W2V_DIM 300
#define HG_SYNTH_OBJECT_LABELS 28
#define HG_SYNTH_ROOM_LABELS 26
#define HG_SYNTH_CDF_ROOMS 65      /* entries of the Poisson(5) cdf table */
#define HG_SYNTH_CDF_OBJECTS 129   /* entries of the Poisson(12) cdf table */

typedef struct hg_synth_args {
    int64_t graph_id0;          /* the batch is graphs graph_id0 .. graph_id0 + n_graphs - 1 */
    int32_t n_graphs;
    int32_t reserved;
    uint64_t seed;
    const double* cdf_rooms;    /* device [HG_SYNTH_CDF_ROOMS]:   cumulative Poisson(5) table of synth._poisson_icdf */
    const double* cdf_objects;  /* device [HG_SYNTH_CDF_OBJECTS]: cumulative Poisson(12) table */
    int64_t* prefix;            /* device [n_graphs + 1][4]: written by _counts, read by _fill */
    /* ---- _fill only; sizes are the totals read from prefix[n_graphs] */
    int64_t n_rooms, n_objects, n_rr_pairs, n_oo_pairs;
    const float* word2vec;      /* NULL | device [HG_SYNTH_OBJECT_LABELS, HG_SYNTH_W2V_DIM] */
    float* x_rooms;             /* [n_rooms, 6]  position | size */
    float* x_objects; int64_t ld_objects; /* [n_objects, 6 (+ 300 with word2vec)] position | size | word2vec[label] */
    int64_t* rooms_to_rooms;    /* [2, 2 * n_rr_pairs] */
    int64_t* objects_to_objects;/* [2, 2 * n_oo_pairs] */
    int64_t* objects_to_rooms;  /* [2, n_objects] */
    int64_t* rooms_to_objects;  /* [2, n_objects] */
    int64_t* room_label;        /* NULL | [n_rooms] */
    int64_t* object_label;      /* NULL | [n_objects] */
    int64_t* room_batch;        /* NULL | [n_rooms]   graph slot (0 .. n_graphs - 1) of every room */
    int64_t* object_batch;      /* NULL | [n_objects] */
    int32_t* status;            /* device scalar, set to 1 when the sizes above disagree with prefix[n_graphs] (nothing
                                   is written outside the declared sizes) */
} hg_synth_args;

HG_API int hg_synth_mp3d_counts(const hg_synth_args* args_host, void* stream);
HG_API int hg_synth_mp3d_fill(const hg_synth_args* args_host, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* HYDRA_GNN_B200_H_ */
