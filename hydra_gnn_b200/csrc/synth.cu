// hg_synth_mp3d_* -- hydra_gnn_b200/synth.py's scene-graph generator on the device (include/hydra_gnn_b200.h).
//
// Every draw is u01(seed, graph id, stream, index) = splitmix64 chain -> 53-bit uniform, a pure function, so the
// integer structure is reproduced bit for bit: counts by table look-up of the SAME float64 cdf tables the host
// builds, pair selection by exact float64 compares, ordered edge lists by ballot compaction in the host's
// enumeration order.  Double arithmetic that the host does as separate numpy operations is written with
// __dmul_rn / __dadd_rn so that no fused multiply-add changes a rounding.
//   k_counts  thread per graph: rooms, objects, selected room pairs, selected object pairs
//   k_prefix  one CTA: exclusive prefix sums of the four counts over the graphs (+ totals)
//   k_fill    warp per graph: room rows, object rows (6 computed columns + the label's word2vec row), labels,
//             batch vectors, rooms<->rooms, objects<->objects, objects<->rooms edge lists
// HBM-bound on the object feature rows it writes (1,224 B per object).
#include "hg_common.cuh"

namespace hg {
namespace {

constexpr int kMaxRooms = HG_SYNTH_MAX_ROOMS, kMaxObj = HG_SYNTH_MAX_OBJECTS;

__device__ __forceinline__ uint64_t mix(uint64_t z) {  // synth._mix
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// synth._u01: the graph key is hoisted (key = mix(seed ^ mix(gid)))
__device__ __forceinline__ uint64_t graph_key(uint64_t seed, int64_t gid) { return mix(seed ^ mix(static_cast<uint64_t>(gid))); }
__device__ __forceinline__ double u01(uint64_t gkey, uint32_t stream, uint64_t idx) {
    uint64_t k = mix(gkey ^ (static_cast<uint64_t>(stream) * 0xD1B54A32D192ED03ull));
    k = mix(k ^ mix(idx));
    return static_cast<double>(k >> 11) * (1.0 / 9007199254740992.0);
}

// np.searchsorted(cdf, u, side="right"): number of table entries <= u
__device__ __forceinline__ int icdf(const double* __restrict__ cdf, int n, double u) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(cdf + mid) <= u) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

__device__ __forceinline__ int n_rooms_of(const hg_synth_args& A, uint64_t gkey) {
    return clampi(icdf(A.cdf_rooms, HG_SYNTH_CDF_ROOMS, u01(gkey, 1, 0)), 1, kMaxRooms);
}
__device__ __forceinline__ int n_objects_of(const hg_synth_args& A, uint64_t gkey, int room) {
    return clampi(icdf(A.cdf_objects, HG_SYNTH_CDF_OBJECTS, u01(gkey, 2, static_cast<uint64_t>(room))), 1, kMaxObj);
}

// rooms a < b of one graph: spanning link to a random earlier room w.p. 0.8, any other pair w.p. 0.05
__device__ __forceinline__ bool room_pair(uint64_t gkey, int a, int b) {
    const int parent = static_cast<int>(floor(__dmul_rn(u01(gkey, 3, b), static_cast<double>(b))));
    const bool tree = (u01(gkey, 4, b) < 0.8) && (a == parent);
    const bool extra = u01(gkey, 5, static_cast<uint64_t>(a * kMaxRooms + b)) < 0.05;
    return tree || extra;
}

// objects a < b of room `room` holding n objects: each pair w.p. min(1, 4 / (n - 1))
__device__ __forceinline__ bool object_pair(uint64_t gkey, int room, int a, int b, double p) {
    return u01(gkey, 6, static_cast<uint64_t>((room * kMaxObj + a) * kMaxObj + b)) < p;
}
__device__ __forceinline__ double pair_prob(int n) {
    const double q = 4.0 / static_cast<double>(n - 1 > 1 ? n - 1 : 1);
    return q < 1.0 ? q : 1.0;
}

__global__ void __launch_bounds__(128) k_counts(const __grid_constant__ hg_synth_args A) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= A.n_graphs) return;
    const uint64_t gkey = graph_key(A.seed, A.graph_id0 + g);
    const int R = n_rooms_of(A, gkey);
    int64_t objects = 0, oo = 0, rr = 0;
    for (int r = 0; r < R; ++r) {
        const int n = n_objects_of(A, gkey, r);
        objects += n;
        const double p = pair_prob(n);
        for (int a = 0; a < n; ++a)
            for (int b = a + 1; b < n; ++b) oo += object_pair(gkey, r, a, b, p) ? 1 : 0;
    }
    for (int a = 0; a < R; ++a)
        for (int b = a + 1; b < R; ++b) rr += room_pair(gkey, a, b) ? 1 : 0;
    int64_t* out = A.prefix + 4 * static_cast<int64_t>(g);
    out[0] = R; out[1] = objects; out[2] = rr; out[3] = oo;
}

// In-place exclusive prefix sums of the four columns, one CTA of 1024 threads, chunks of 1024 graphs with a carry.
__global__ void __launch_bounds__(1024) k_prefix(const __grid_constant__ hg_synth_args A) {
    __shared__ int64_t warp_tot[32][4];
    __shared__ int64_t carry[4];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x < 4) carry[threadIdx.x] = 0;
    __syncthreads();
    for (int g0 = 0; g0 < A.n_graphs; g0 += 1024) {
        const int g = g0 + threadIdx.x;
        int64_t v[4], inc[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) v[c] = g < A.n_graphs ? A.prefix[4 * static_cast<int64_t>(g) + c] : 0, inc[c] = v[c];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int64_t t = __shfl_up_sync(0xffffffffu, inc[c], o);
                if (lane >= o) inc[c] += t;
            }
        }
        if (lane == 31) {
#pragma unroll
            for (int c = 0; c < 4; ++c) warp_tot[wid][c] = inc[c];
        }
        __syncthreads();
        if (wid == 0) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                int64_t w = warp_tot[lane][c];
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int64_t t = __shfl_up_sync(0xffffffffu, w, o);
                    if (lane >= o) w += t;
                }
                warp_tot[lane][c] = w;  // inclusive over warps
            }
        }
        __syncthreads();
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int64_t ex = carry[c] + (wid ? warp_tot[wid - 1][c] : 0) + inc[c] - v[c];
            if (g < A.n_graphs) A.prefix[4 * static_cast<int64_t>(g) + c] = ex;
        }
        __syncthreads();
        if (threadIdx.x < 4) carry[threadIdx.x] += warp_tot[31][threadIdx.x];
        __syncthreads();
    }
    if (threadIdx.x < 4) A.prefix[4 * static_cast<int64_t>(A.n_graphs) + threadIdx.x] = carry[threadIdx.x];
}

__device__ __forceinline__ double normal01(uint64_t gkey, uint32_t stream, uint64_t idx) {  // synth._normal (Box-Muller)
    const double u1 = u01(gkey, stream, 2 * idx), u2 = u01(gkey, stream, 2 * idx + 1);
    const double r = sqrt(__dmul_rn(-2.0, log(__dsub_rn(1.0, u1))));
    return __dmul_rn(r, cos(__dmul_rn(6.283185307179586, u2)));
}

__device__ __forceinline__ void put_pair(int64_t* list, int64_t n_edges, int64_t k, int64_t a, int64_t b) {
    list[2 * k] = a;               // row 0 (sources): a, b
    list[2 * k + 1] = b;
    list[n_edges + 2 * k] = b;     // row 1 (targets): b, a
    list[n_edges + 2 * k + 1] = a;
}

__global__ void __launch_bounds__(256) k_fill(const __grid_constant__ hg_synth_args A) {
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (g >= A.n_graphs) return;
    const int64_t* tot = A.prefix + 4 * static_cast<int64_t>(A.n_graphs);
    if (tot[0] != A.n_rooms || tot[1] != A.n_objects || tot[2] != A.n_rr_pairs || tot[3] != A.n_oo_pairs) {
        if (lane == 0 && A.status) *A.status = 1;
        return;
    }
    const int64_t* pre = A.prefix + 4 * static_cast<int64_t>(g);
    const int64_t room0 = pre[0], obj0 = pre[1], rr0 = pre[2], oo0 = pre[3];
    const uint64_t gkey = graph_key(A.seed, A.graph_id0 + g);
    const int R = n_rooms_of(A, gkey);

    // ---- rooms: lane r owns room r
    const int n_mine = lane < R ? n_objects_of(A, gkey, lane) : 0;
    int inc = n_mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    const int start_mine = inc - n_mine;  // index inside the graph of the room's first object
    double rpos[3] = {0.0, 0.0, 0.0};
    if (lane < R) {
        float* xr = A.x_rooms + (room0 + lane) * 6;
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            rpos[d] = __dadd_rn(-10.0, __dmul_rn(20.0, u01(gkey, 7, static_cast<uint64_t>(lane * 3 + d))));
            xr[d] = __double2float_rn(rpos[d]);
            xr[3 + d] = __double2float_rn(__dadd_rn(2.0, __dmul_rn(6.0, u01(gkey, 8, static_cast<uint64_t>(lane * 3 + d)))));
        }
        if (A.room_label)
            A.room_label[room0 + lane] = static_cast<int64_t>(floor(__dmul_rn(u01(gkey, 12, lane), static_cast<double>(HG_SYNTH_ROOM_LABELS))));
        if (A.room_batch) A.room_batch[room0 + lane] = g;
    }

    // ---- objects, room by room (objects of a room are consecutive rows)
    for (int r = 0; r < R; ++r) {
        const int n = __shfl_sync(0xffffffffu, n_mine, r), s = __shfl_sync(0xffffffffu, start_mine, r);
        double rp[3];
#pragma unroll
        for (int d = 0; d < 3; ++d) rp[d] = __shfl_sync(0xffffffffu, rpos[d], r);
        for (int o0 = 0; o0 < n; o0 += 32) {
            const int o = o0 + lane;
            int label = 0;
            if (o < n) {
                const int oig = s + o;  // object index inside its graph
                const int64_t row = obj0 + oig;
                float* xo = A.x_objects + row * A.ld_objects;
#pragma unroll
                for (int d = 0; d < 3; ++d) {
                    const uint64_t idx = static_cast<uint64_t>(oig * 3 + d);
                    xo[d] = __double2float_rn(__dadd_rn(rp[d], __dmul_rn(1.5, normal01(gkey, 9, idx))));
                    xo[3 + d] = __double2float_rn(__dadd_rn(0.1, __dmul_rn(1.9, u01(gkey, 10, idx))));
                }
                label = static_cast<int>(floor(__dmul_rn(u01(gkey, 11, oig), static_cast<double>(HG_SYNTH_OBJECT_LABELS))));
                if (A.object_label) A.object_label[row] = label;
                if (A.object_batch) A.object_batch[row] = g;
                A.objects_to_rooms[row] = row;
                A.objects_to_rooms[A.n_objects + row] = room0 + r;
                A.rooms_to_objects[row] = room0 + r;
                A.rooms_to_objects[A.n_objects + row] = row;
            }
            if (A.word2vec) {  // the label's table row, copied by the whole warp
                const int cnt = n - o0 < 32 ? n - o0 : 32;
                for (int t = 0; t < cnt; ++t) {
                    const int lbl = __shfl_sync(0xffffffffu, label, t);
                    const float* src = A.word2vec + lbl * HG_SYNTH_W2V_DIM;
                    float* dst = A.x_objects + (obj0 + s + o0 + t) * A.ld_objects + 6;
                    for (int c = lane; c < HG_SYNTH_W2V_DIM; c += 32) dst[c] = __ldg(src + c);
                }
            }
        }
    }

    // ---- rooms <-> rooms: pairs in (a, b) order, a < b < R <= 12 (one chunk of lanes per a)
    {
        const int64_t n_edges = 2 * A.n_rr_pairs;
        int64_t k = rr0;
        for (int a = 0; a + 1 < R; ++a) {
            const int b = a + 1 + lane;
            const bool hit = b < R && room_pair(gkey, a, b);
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (hit) put_pair(A.rooms_to_rooms, n_edges, k + __popc(m & ((1u << lane) - 1)), room0 + a, room0 + b);
            k += __popc(m);
        }
    }

    // ---- objects <-> objects inside every room, pairs in (a, b) order
    {
        const int64_t n_edges = 2 * A.n_oo_pairs;
        int64_t k = oo0;
        for (int r = 0; r < R; ++r) {
            const int n = __shfl_sync(0xffffffffu, n_mine, r), s = __shfl_sync(0xffffffffu, start_mine, r);
            const double p = pair_prob(n);
            const int64_t base = obj0 + s;
            for (int a = 0; a + 1 < n; ++a)
                for (int b0 = a + 1; b0 < n; b0 += 32) {
                    const int b = b0 + lane;
                    const bool hit = b < n && object_pair(gkey, r, a, b, p);
                    const unsigned m = __ballot_sync(0xffffffffu, hit);
                    if (hit) put_pair(A.objects_to_objects, n_edges, k + __popc(m & ((1u << lane) - 1)), base + a, base + b);
                    k += __popc(m);
                }
        }
    }
}

int check_common(const hg_synth_args* a) {
    if (!a || a->n_graphs < 0 || a->graph_id0 < 0) return HG_EINVAL;
    if (a->n_graphs > 0 && (!a->cdf_rooms || !a->cdf_objects || !a->prefix)) return HG_EINVAL;
    return HG_OK;
}

}  // namespace
}  // namespace hg

extern "C" int hg_synth_mp3d_counts(const hg_synth_args* a, void* stream_) {
    using namespace hg;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int rc = check_common(a);
    if (rc != HG_OK) return rc;
    if (!a->prefix) return HG_EINVAL;
    if (a->n_graphs > 0) k_counts<<<ceil_div(a->n_graphs, 128), 128, 0, stream>>>(*a);
    k_prefix<<<1, 1024, 0, stream>>>(*a);
    return check_launch();
}

extern "C" int hg_synth_mp3d_fill(const hg_synth_args* a, void* stream_) {
    using namespace hg;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int rc = check_common(a);
    if (rc != HG_OK) return rc;
    if (a->n_graphs == 0) return HG_OK;
    if (a->n_rooms < 0 || a->n_objects < 0 || a->n_rr_pairs < 0 || a->n_oo_pairs < 0) return HG_EINVAL;
    if (!a->x_rooms || !a->x_objects || !a->objects_to_rooms || !a->rooms_to_objects) return HG_EINVAL;
    if ((a->n_rr_pairs > 0 && !a->rooms_to_rooms) || (a->n_oo_pairs > 0 && !a->objects_to_objects)) return HG_EINVAL;
    if (a->ld_objects < (a->word2vec ? 6 + HG_SYNTH_W2V_DIM : 6)) return HG_EINVAL;
    if (a->status) HG_CUDA_TRY(cudaMemsetAsync(a->status, 0, sizeof(int32_t), stream));
    k_fill<<<ceil_div(a->n_graphs, 8), 256, 0, stream>>>(*a);
    return check_launch();
}
