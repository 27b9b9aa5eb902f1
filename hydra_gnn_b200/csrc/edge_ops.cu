// Edge-attribute helpers of the GAT_edge path (integer gathers + exact fp32 subtraction).
#include "hg_common.cuh"

namespace hg {
namespace {

constexpr int kThreads = 256;

// a_edge[p, h] = perm[p] >= 0 ? <edge_attr[perm[p], :D], v[h, :D]> : <fill, v[h]>, p < E' = *n_entries
__global__ void __launch_bounds__(kThreads)
k_edge_logits(const float* __restrict__ attr, int64_t ld, int D, const int32_t* __restrict__ perm,
              const int32_t* __restrict__ n_entries, int64_t cap, const float* __restrict__ v,
              const float* __restrict__ fill, int H, float* __restrict__ out) {
    const int64_t total = min(static_cast<int64_t>(__ldg(n_entries)), cap) * H;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int64_t p = t / H;
        const int h = static_cast<int>(t % H);
        const int32_t e = __ldg(perm + p);
        float acc = 0.f;
        if (e >= 0) {
            for (int d = 0; d < D; ++d) acc = fmaf(__ldg(attr + e * ld + d), __ldg(v + h * D + d), acc);
        } else if (fill) {
            for (int d = 0; d < D; ++d) acc = fmaf(__ldg(fill + d), __ldg(v + h * D + d), acc);
        }
        out[t] = acc;
    }
}

__global__ void __launch_bounds__(kThreads)
k_relative_pos(const float* __restrict__ ps, const float* __restrict__ pd, const int64_t* __restrict__ src,
               const int64_t* __restrict__ dst, int64_t E, int64_t n_src, int64_t n_dst, float* __restrict__ out,
               int32_t* __restrict__ status) {
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t e = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; e < E; e += stride) {
        const int64_t s = src[e], d = dst[e];
        if (static_cast<uint64_t>(s) >= static_cast<uint64_t>(n_src) ||
            static_cast<uint64_t>(d) >= static_cast<uint64_t>(n_dst)) {  // never dereferenced: NaN row + counted
#pragma unroll
            for (int c = 0; c < 3; ++c) out[e * 3 + c] = __int_as_float(0x7fc00000);
            if (status) atomicAdd(status, 1);
            continue;
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) out[e * 3 + c] = __ldg(pd + d * 3 + c) - __ldg(ps + s * 3 + c);
    }
}

// out[r, :F] = act(x[idx[r], :F])  (x[room_mask] of the homogeneous models; idx from the plan)
__global__ void __launch_bounds__(kThreads)
k_gather_rows(const float* __restrict__ x, int64_t ldx, const int32_t* __restrict__ idx, int64_t n_out, int F,
              int act, float* __restrict__ out, int64_t ldout) {
    const int64_t total = n_out * F;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int64_t r = t / F;
        const int f = static_cast<int>(t % F);
        out[r * ldout + f] = apply_act(__ldg(x + static_cast<int64_t>(__ldg(idx + r)) * ldx + f), act);
    }
}

unsigned grid_for(int64_t n) {
    const int sms = sm_count() > 0 ? sm_count() : 148;
    int64_t g = ceil_div<int64_t>(n > 0 ? n : 1, kThreads);
    const int64_t cap = static_cast<int64_t>(sms) * 8 * 8;
    return static_cast<unsigned>(g < cap ? g : cap);
}

// graph of element e: largest g with ptr[g] <= e (ptr ascending, ptr[0] = 0, empty graphs allowed)
__device__ __forceinline__ int32_t graph_of(const int64_t* __restrict__ ptr, int32_t n_graphs, int64_t e) {
    int32_t lo = 0, hi = n_graphs;  // invariant: ptr[lo] <= e < ptr[hi]
    while (hi - lo > 1) {
        const int32_t mid = (lo + hi) >> 1;
        if (__ldg(ptr + mid) <= e) lo = mid; else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(kThreads) k_collate_offsets(int64_t* __restrict__ index, int64_t E,
                                                              const int64_t* __restrict__ edge_ptr,
                                                              const int64_t* __restrict__ src_ptr,
                                                              const int64_t* __restrict__ dst_ptr, int32_t n_graphs) {
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t e = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; e < E; e += stride) {
        const int32_t g = graph_of(edge_ptr, n_graphs, e);
        index[e] += __ldg(src_ptr + g);
        index[E + e] += __ldg(dst_ptr + g);
    }
}

__global__ void __launch_bounds__(kThreads) k_batch_vector(int64_t* __restrict__ batch, int64_t N,
                                                           const int64_t* __restrict__ ptr, int32_t n_graphs) {
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t n = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; n < N; n += stride)
        batch[n] = graph_of(ptr, n_graphs, n);
}

}  // namespace
}  // namespace hg

extern "C" int hg_edge_logits(const float* edge_attr, int64_t ld_attr, int32_t D, const int32_t* perm,
                              const int32_t* rowptr_last, int64_t cap, const float* v, const float* fill, int32_t H,
                              float* a_edge, void* stream_) {
    using namespace hg;
    if (cap < 0 || D < 0 || H < 1) return HG_EINVAL;
    if (cap == 0) return HG_OK;
    if (!perm || !rowptr_last || !v || !a_edge || (D > 0 && !edge_attr)) return HG_EINVAL;
    k_edge_logits<<<grid_for(cap * H), kThreads, 0, static_cast<cudaStream_t>(stream_)>>>(
        edge_attr, ld_attr, D, perm, rowptr_last, cap, v, fill, H, a_edge);
    return check_launch();
}

extern "C" int hg_relative_pos(const float* pos_src, int64_t n_src, const float* pos_dst, int64_t n_dst,
                               const int64_t* src, const int64_t* dst, int64_t E, float* edge_attr, int32_t* status,
                               void* stream_) {
    using namespace hg;
    if (E < 0 || n_src < 0 || n_dst < 0) return HG_EINVAL;
    if (E == 0) return HG_OK;
    if (!src || !dst || !edge_attr || (n_src > 0 && !pos_src) || (n_dst > 0 && !pos_dst)) return HG_EINVAL;
    k_relative_pos<<<grid_for(E), kThreads, 0, static_cast<cudaStream_t>(stream_)>>>(pos_src, pos_dst, src, dst, E, n_src,
                                                                                      n_dst, edge_attr, status);
    return check_launch();
}

extern "C" int hg_gather_rows(const float* x, int64_t ldx, const int32_t* idx, int64_t n_out, int32_t F, int32_t act,
                              float* out, int64_t ldout, void* stream_) {
    using namespace hg;
    if (n_out < 0 || F < 0) return HG_EINVAL;
    if (n_out == 0 || F == 0) return HG_OK;
    if (!x || !idx || !out) return HG_EINVAL;
    k_gather_rows<<<grid_for(n_out * F), kThreads, 0, static_cast<cudaStream_t>(stream_)>>>(x, ldx, idx, n_out, F, act,
                                                                                            out, ldout);
    return check_launch();
}

extern "C" int hg_collate_offsets(int64_t* index, int64_t n_edges, const int64_t* edge_ptr, const int64_t* src_ptr,
                                  const int64_t* dst_ptr, int32_t n_graphs, void* stream_) {
    using namespace hg;
    if (n_edges < 0 || n_graphs < 0) return HG_EINVAL;
    if (n_edges == 0) return HG_OK;
    if (!index || !edge_ptr || !src_ptr || !dst_ptr || n_graphs < 1) return HG_EINVAL;
    k_collate_offsets<<<grid_for(n_edges), kThreads, 0, static_cast<cudaStream_t>(stream_)>>>(index, n_edges, edge_ptr,
                                                                                              src_ptr, dst_ptr, n_graphs);
    return check_launch();
}

extern "C" int hg_batch_vector(int64_t* batch, int64_t n_nodes, const int64_t* ptr, int32_t n_graphs, void* stream_) {
    using namespace hg;
    if (n_nodes < 0 || n_graphs < 0) return HG_EINVAL;
    if (n_nodes == 0) return HG_OK;
    if (!batch || !ptr || n_graphs < 1) return HG_EINVAL;
    k_batch_vector<<<grid_for(n_nodes), kThreads, 0, static_cast<cudaStream_t>(stream_)>>>(batch, n_nodes, ptr, n_graphs);
    return check_launch();
}
