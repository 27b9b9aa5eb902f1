// hg_htree_split -- batched heterogeneous H-tree -> homogeneous form (integer work, bit-exact).
//
// Reference: heterogeneous_htree_to_homogeneous (mp3d_dataset.py:72-119) per graph + PyG's collate.  Per graph the
// homogeneous node order is "node types in store order"; the edge order is "relations in store order", each relation's
// edges in their original order; an edge goes to the init list when its source type is virtual, else to the pool list
// when its target type is virtual, else to the tree list.  Everything below is counting and prefix sums:
//   k_node_ptr    per graph: nodes per type -> offset of every type inside the graph; scan over graphs -> ptr
//   k_nodes       thread per node: graph by binary search in the type's ptr -> place, node_type, batch, masks
//   k_edge_count  thread per edge: graph of the source -> cnt[relation][graph] (integer atomics); order check
//   k_edge_scan   one block per relation: first edge of every graph inside the relation (exclusive scan of cnt);
//                 one block per class: first slot of every graph inside the class list
//   k_edge_off    thread per graph: slot of every (relation, graph) run inside its class list
//   k_edge_put    thread per edge: slot = off[relation][graph] + rank inside the run; endpoints -> homogeneous ids
#include "hg_common.cuh"

namespace hg {
namespace {

constexpr int kThreads = 256;
constexpr int kScanThreads = 1024;

struct SplitParams {
    hg_htree_split_args a;
    int64_t* type_off;  // [n_types][G]  offset of type t inside graph g (homogeneous order)
    int32_t* cnt;       // [n_rel][G]    edges of relation k in graph g
    int64_t* first;     // [n_rel][G]    first edge of graph g inside relation k
    int64_t* base;      // [3][G]        first slot of graph g inside the class list
    int64_t* off;       // [n_rel][G]    first slot of the (k, g) run inside its class list
    int32_t cls[HG_MAX_PLAN_RELATIONS];
};

// graph g with ptr[g] <= i < ptr[g + 1] (ptr non-decreasing, ptr[0] = 0, i < ptr[G])
__device__ __forceinline__ int graph_of(const int64_t* __restrict__ ptr, int G, int64_t i) {
    int lo = 0, hi = G;  // invariant: ptr[lo] <= i < ptr[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(ptr + mid) <= i) lo = mid; else hi = mid;
    }
    return lo;
}

// block-wide exclusive scan of one int64 per thread (kScanThreads threads); returns the prefix, total in `total`
__device__ __forceinline__ int64_t block_scan(int64_t v, int64_t& total) {
    __shared__ int64_t warp_tot[kScanThreads / 32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int64_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int64_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warp_tot[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int64_t w = warp_tot[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int64_t t = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += t;
        }
        warp_tot[lane] = w;
    }
    __syncthreads();
    const int64_t pre = (wid ? warp_tot[wid - 1] : 0) + inc - v;
    total = warp_tot[kScanThreads / 32 - 1];
    __syncthreads();
    return pre;
}

__global__ void __launch_bounds__(kScanThreads) k_node_ptr(const __grid_constant__ SplitParams P) {
    const hg_htree_split_args& A = P.a;
    const int G = A.n_graphs;
    int64_t carry = 0;
    for (int g0 = 0; g0 < G; g0 += kScanThreads) {
        const int g = g0 + threadIdx.x;
        int64_t n = 0;
        if (g < G) {
            for (int t = 0; t < A.n_types; ++t) {
                P.type_off[static_cast<int64_t>(t) * G + g] = n;
                n += __ldg(A.type[t].ptr + g + 1) - __ldg(A.type[t].ptr + g);
            }
        }
        int64_t total;
        const int64_t pre = block_scan(n, total);
        if (g < G) A.ptr[g] = carry + pre;
        carry += total;
    }
    if (threadIdx.x == 0) A.ptr[G] = carry;
}

__global__ void __launch_bounds__(kThreads) k_nodes(const __grid_constant__ SplitParams P) {
    const hg_htree_split_args& A = P.a;
    const int t = blockIdx.y, G = A.n_graphs;
    const hg_htree_type& T = A.type[t];
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < T.n_nodes; i += stride) {
        const int g = graph_of(T.ptr, G, i);
        const int64_t at = A.ptr[g] + P.type_off[static_cast<int64_t>(t) * G + g] + (i - __ldg(T.ptr + g));
        if (T.place) T.place[i] = at;
        A.node_type[at] = t;
        A.batch[at] = g;
        A.room_mask[at] = T.is_room_virtual ? 1 : 0;
        A.object_mask[at] = T.is_object_virtual ? 1 : 0;
    }
}

__global__ void __launch_bounds__(kThreads) k_edge_count(const __grid_constant__ SplitParams P) {
    const hg_htree_split_args& A = P.a;
    const int k = blockIdx.y, G = A.n_graphs;
    const hg_htree_rel& R = A.rel[k];
    const hg_htree_type& S = A.type[R.src_type];
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t e = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; e < R.n_edges; e += stride) {
        const int64_t s = R.src[e];
        if (static_cast<uint64_t>(s) >= static_cast<uint64_t>(S.n_nodes)) {
            atomicAdd(A.status, 1);
            continue;
        }
        const int g = graph_of(S.ptr, G, s);
        atomicAdd(&P.cnt[static_cast<int64_t>(k) * G + g], 1);
        if (e > 0) {  // grouped by graph, in graph order
            const int64_t sp = R.src[e - 1];
            if (static_cast<uint64_t>(sp) < static_cast<uint64_t>(S.n_nodes) && graph_of(S.ptr, G, sp) > g) atomicAdd(A.status, 1);
        }
    }
}

// blocks 0 .. n_rel - 1: first[k][.] = exclusive scan of cnt[k][.]; blocks n_rel .. n_rel + 2: base[c][.] of class c
__global__ void __launch_bounds__(kScanThreads) k_edge_scan(const __grid_constant__ SplitParams P) {
    const hg_htree_split_args& A = P.a;
    const int G = A.n_graphs, b = blockIdx.x;
    int64_t carry = 0;
    for (int g0 = 0; g0 < G; g0 += kScanThreads) {
        const int g = g0 + threadIdx.x;
        int64_t n = 0;
        if (g < G) {
            if (b < A.n_rel) {
                n = P.cnt[static_cast<int64_t>(b) * G + g];
            } else {
                for (int k = 0; k < A.n_rel; ++k)
                    if (P.cls[k] == b - A.n_rel) n += P.cnt[static_cast<int64_t>(k) * G + g];
            }
        }
        int64_t total;
        const int64_t pre = block_scan(n, total);
        if (g < G) (b < A.n_rel ? P.first[static_cast<int64_t>(b) * G + g] : P.base[static_cast<int64_t>(b - A.n_rel) * G + g]) = carry + pre;
        carry += total;
    }
}

__global__ void __launch_bounds__(kThreads) k_edge_off(const __grid_constant__ SplitParams P) {
    const hg_htree_split_args& A = P.a;
    const int G = A.n_graphs;
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G) return;
    int64_t run[3] = {P.base[g], P.base[static_cast<int64_t>(G) + g], P.base[2 * static_cast<int64_t>(G) + g]};
    for (int k = 0; k < A.n_rel; ++k) {
        const int c = P.cls[k];
        int64_t r = c == 0 ? run[0] : (c == 1 ? run[1] : run[2]);
        P.off[static_cast<int64_t>(k) * G + g] = r;
        r += P.cnt[static_cast<int64_t>(k) * G + g];
        if (c == 0) run[0] = r; else if (c == 1) run[1] = r; else run[2] = r;
    }
}

struct ClassSizes { int64_t n[3]; };

__global__ void __launch_bounds__(kThreads) k_edge_put(const __grid_constant__ SplitParams P, const __grid_constant__ ClassSizes Z) {
    const hg_htree_split_args& A = P.a;
    const int k = blockIdx.y, G = A.n_graphs;
    const hg_htree_rel& R = A.rel[k];
    const hg_htree_type& S = A.type[R.src_type];
    const hg_htree_type& D = A.type[R.dst_type];
    const int c = P.cls[k];
    int64_t* const out = A.edges[c];
    const int64_t n_class = Z.n[c];
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t e = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; e < R.n_edges; e += stride) {
        const int64_t s = R.src[e], d = R.dst[e];
        if (static_cast<uint64_t>(s) >= static_cast<uint64_t>(S.n_nodes)) continue;  // counted by k_edge_count
        const int g = graph_of(S.ptr, G, s);
        const int64_t slot = P.off[static_cast<int64_t>(k) * G + g] + (e - P.first[static_cast<int64_t>(k) * G + g]);
        if (slot < 0 || slot >= n_class) continue;  // only when the list is not grouped by graph (counted)
        const int64_t d0 = __ldg(D.ptr + g), d1 = __ldg(D.ptr + g + 1);
        if (d < d0 || d >= d1) {  // target outside the source's graph
            atomicAdd(A.status, 1);
            out[slot] = -1;
            out[n_class + slot] = -1;
            continue;
        }
        const int64_t g0 = A.ptr[g];
        out[slot] = g0 + P.type_off[static_cast<int64_t>(R.src_type) * G + g] + (s - __ldg(S.ptr + g));
        out[n_class + slot] = g0 + P.type_off[static_cast<int64_t>(R.dst_type) * G + g] + (d - d0);
    }
}

inline size_t up(size_t v) { return (v + 255) / 256 * 256; }

struct Layout { size_t type_off, cnt, first, base, off, total; };

int layout(const hg_htree_split_args* a, Layout& L) {
    if (!a || a->n_types < 1 || a->n_types > HG_MAX_HTREE_TYPES || a->n_rel < 0 || a->n_rel > HG_MAX_PLAN_RELATIONS ||
        a->n_graphs < 1)
        return HG_EINVAL;
    const size_t G = static_cast<size_t>(a->n_graphs);
    size_t o = 0;
    L.cnt = o;      o = up(o + 4 * G * a->n_rel);  // first: the region that must start at zero
    L.type_off = o; o = up(o + 8 * G * a->n_types);
    L.first = o;    o = up(o + 8 * G * a->n_rel);
    L.base = o;     o = up(o + 8 * G * 3);
    L.off = o;      o = up(o + 8 * G * a->n_rel);
    L.total = o;
    return HG_OK;
}

}  // namespace
}  // namespace hg

extern "C" size_t hg_htree_split_workspace_bytes(const hg_htree_split_args* a) {
    hg::Layout L;
    return hg::layout(a, L) == HG_OK ? L.total + 256 : 0;
}

extern "C" int hg_htree_split(const hg_htree_split_args* a, void* workspace, size_t workspace_bytes, void* stream_) {
    using namespace hg;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    Layout L;
    const int rc = layout(a, L);
    if (rc != HG_OK) return rc;
    if (!a->ptr || !a->node_type || !a->batch || !a->room_mask || !a->object_mask || !a->status) return HG_EINVAL;
    if (!workspace || workspace_bytes < L.total) return HG_EWORKSPACE;
    if (!aligned(workspace, 256)) return HG_EALIGN;
    SplitParams P;
    P.a = *a;
    char* ws = static_cast<char*>(workspace);
    P.cnt = reinterpret_cast<int32_t*>(ws + L.cnt);
    P.type_off = reinterpret_cast<int64_t*>(ws + L.type_off);
    P.first = reinterpret_cast<int64_t*>(ws + L.first);
    P.base = reinterpret_cast<int64_t*>(ws + L.base);
    P.off = reinterpret_cast<int64_t*>(ws + L.off);
    ClassSizes Z = {{0, 0, 0}};
    int64_t max_nodes = 0, max_edges = 0;
    for (int t = 0; t < a->n_types; ++t) {
        if (!a->type[t].ptr || a->type[t].n_nodes < 0) return HG_EINVAL;
        max_nodes = a->type[t].n_nodes > max_nodes ? a->type[t].n_nodes : max_nodes;
    }
    for (int k = 0; k < HG_MAX_PLAN_RELATIONS; ++k) P.cls[k] = 0;
    for (int k = 0; k < a->n_rel; ++k) {
        const hg_htree_rel& R = a->rel[k];
        if (R.n_edges < 0 || R.src_type < 0 || R.src_type >= a->n_types || R.dst_type < 0 || R.dst_type >= a->n_types) return HG_EINVAL;
        if (R.n_edges > 0 && (!R.src || !R.dst)) return HG_EINVAL;
        const hg_htree_type& s = a->type[R.src_type];
        const hg_htree_type& d = a->type[R.dst_type];
        if ((s.is_room_virtual || s.is_object_virtual) && (d.is_room_virtual || d.is_object_virtual))
            return HG_EUNSUPPORTED;  // the reference would list such an edge under init AND pool; its schema has none
        P.cls[k] = (s.is_room_virtual || s.is_object_virtual) ? HG_EDGE_INIT
                   : ((d.is_room_virtual || d.is_object_virtual) ? HG_EDGE_POOL : HG_EDGE_TREE);
        Z.n[P.cls[k]] += R.n_edges;
        max_edges = R.n_edges > max_edges ? R.n_edges : max_edges;
    }
    for (int c = 0; c < 3; ++c)
        if (Z.n[c] > 0 && !a->edges[c]) return HG_EINVAL;

    HG_CUDA_TRY(cudaMemsetAsync(ws + L.cnt, 0, L.type_off - L.cnt, stream));
    HG_CUDA_TRY(cudaMemsetAsync(a->status, 0, sizeof(int32_t), stream));
    const int sms = sm_count() > 0 ? sm_count() : 148;
    auto grid = [&](int64_t n) {
        int64_t g = ceil_div<int64_t>(n > 0 ? n : 1, kThreads);
        const int64_t cap = static_cast<int64_t>(sms) * 8 * 4;
        return static_cast<unsigned>(g < cap ? g : cap);
    };
    k_node_ptr<<<1, kScanThreads, 0, stream>>>(P);
    if (max_nodes > 0) k_nodes<<<dim3(grid(max_nodes), a->n_types), kThreads, 0, stream>>>(P);
    if (a->n_rel > 0) {
        if (max_edges > 0) k_edge_count<<<dim3(grid(max_edges), a->n_rel), kThreads, 0, stream>>>(P);
        k_edge_scan<<<a->n_rel + 3, kScanThreads, 0, stream>>>(P);
        k_edge_off<<<static_cast<unsigned>(ceil_div(a->n_graphs, kThreads)), kThreads, 0, stream>>>(P);
        if (max_edges > 0) k_edge_put<<<dim3(grid(max_edges), a->n_rel), kThreads, 0, stream>>>(P, Z);
    }
    return check_launch();
}
