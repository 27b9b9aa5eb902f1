// hg_synth_mp3d_* -- hydra_gnn_b200/synth.py's scene-graph generator on the device (include/hydra_gnn_b200.h).
//
// Every draw is u01(seed, graph id, stream, index) = splitmix64 chain -> 53-bit uniform, a pure function, so the
// integer structure is reproduced bit for bit: counts by table look-up of the SAME float64 cdf tables the host
// builds, pair selection by exact float64 compares, ordered edge lists by ballot compaction in the host's
// enumeration order.  Double arithmetic that the host does as separate numpy operations is written with
// __dmul_rn / __dadd_rn so that no fused multiply-add changes a rounding.
//   k_counts  warp per graph: rooms, objects, selected room pairs, selected object pairs
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

// Warp per graph.  Counting needs no order, so the lanes stride over the flattened a x b grid of a room (the hash is
// evaluated for a < b only): ~5 iterations per 12-object room instead of 11 partly filled ones.
__global__ void __launch_bounds__(256) k_counts(const __grid_constant__ hg_synth_args A) {
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (g >= A.n_graphs) return;
    const uint64_t gkey = graph_key(A.seed, A.graph_id0 + g);
    const int R = n_rooms_of(A, gkey);
    const int n_mine = lane < R ? n_objects_of(A, gkey, lane) : 0;
    int objects = n_mine, rr = 0, oo = 0;
    for (int k = lane; k < R * R; k += 32) {
        const int a = k / R, b = k - a * R;
        rr += (a < b && room_pair(gkey, a, b)) ? 1 : 0;
    }
    for (int r = 0; r < R; ++r) {
        const int n = __shfl_sync(0xffffffffu, n_mine, r);
        const double p = pair_prob(n);
        for (int k = lane; k < n * n; k += 32) {
            const int a = k / n, b = k - a * n;
            oo += (a < b && object_pair(gkey, r, a, b, p)) ? 1 : 0;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        objects += __shfl_xor_sync(0xffffffffu, objects, o);
        rr += __shfl_xor_sync(0xffffffffu, rr, o);
        oo += __shfl_xor_sync(0xffffffffu, oo, o);
    }
    if (lane == 0) {
        int64_t* out = A.prefix + 4 * static_cast<int64_t>(g);
        out[0] = R; out[1] = objects; out[2] = rr; out[3] = oo;
    }
}

// In-place exclusive prefix sums of the four columns: one CTA of 512 threads, 8 consecutive graphs per thread (4,096
// graphs per pass, all loads of a pass in flight together; the per-graph counts fit 32 bits), a carry between passes.
constexpr int kPrefixItems = 8, kPrefixThreads = 512;
__global__ void __launch_bounds__(kPrefixThreads) k_prefix(const __grid_constant__ hg_synth_args A) {
    constexpr int kWarps = kPrefixThreads / 32;
    __shared__ int64_t warp_tot[32][4];
    __shared__ int64_t carry[4];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x < 4) carry[threadIdx.x] = 0;
    __syncthreads();
    for (int64_t g0 = 0; g0 < A.n_graphs; g0 += kPrefixThreads * kPrefixItems) {
        const int64_t base = g0 + static_cast<int64_t>(threadIdx.x) * kPrefixItems;
        int4 v[kPrefixItems];
        int64_t tot[4] = {0, 0, 0, 0};
#pragma unroll
        for (int t = 0; t < kPrefixItems; ++t) {
            v[t] = make_int4(0, 0, 0, 0);
            if (base + t < A.n_graphs) {
                const longlong4 w = *reinterpret_cast<const longlong4*>(A.prefix + 4 * (base + t));
                v[t] = make_int4(static_cast<int>(w.x), static_cast<int>(w.y), static_cast<int>(w.z), static_cast<int>(w.w));
            }
            tot[0] += v[t].x; tot[1] += v[t].y; tot[2] += v[t].z; tot[3] += v[t].w;
        }
        int64_t inc[4] = {tot[0], tot[1], tot[2], tot[3]};
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
                int64_t w = lane < kWarps ? warp_tot[lane][c] : 0;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int64_t t = __shfl_up_sync(0xffffffffu, w, o);
                    if (lane >= o) w += t;
                }
                warp_tot[lane][c] = w;  // inclusive over warps
            }
        }
        __syncthreads();
        int64_t run[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) run[c] = carry[c] + (wid ? warp_tot[wid - 1][c] : 0) + inc[c] - tot[c];
#pragma unroll
        for (int t = 0; t < kPrefixItems; ++t) {
            if (base + t < A.n_graphs)
                *reinterpret_cast<longlong4*>(A.prefix + 4 * (base + t)) = make_longlong4(run[0], run[1], run[2], run[3]);
            run[0] += v[t].x; run[1] += v[t].y; run[2] += v[t].z; run[3] += v[t].w;
        }
        __syncthreads();
        if (threadIdx.x < 4) carry[threadIdx.x] += warp_tot[31][threadIdx.x];  // lanes >= kWarps scanned zeros: row 31 = the total
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

// Pair q of the row-major upper triangle of an n x n grid (row a holds the n - 1 - a pairs (a, a + 1 .. n - 1)).
__device__ __forceinline__ void pair_of(int q, int n, int& a, int& b) {
    const float t = static_cast<float>(2 * n - 1);
    a = static_cast<int>((t - sqrtf(t * t - 8.f * static_cast<float>(q))) * 0.5f);
    a = a < 0 ? 0 : (a > n - 2 ? n - 2 : a);
    while (a > 0 && a * (2 * n - a - 1) / 2 > q) --a;           // the float estimate is off by at most one row
    while ((a + 1) * (2 * n - a - 2) / 2 <= q) ++a;
    b = a + 1 + (q - a * (2 * n - a - 1) / 2);
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

    const bool vec2 = (A.ld_objects % 2 == 0) && (reinterpret_cast<uintptr_t>(A.x_objects) % 8 == 0) &&
                      (reinterpret_cast<uintptr_t>(A.word2vec) % 8 == 0);
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
                    if (vec2) {  // rows of 306 floats: column 6 of every row is 8-byte aligned, and so are the table rows
#pragma unroll
                        for (int c = lane; c < HG_SYNTH_W2V_DIM / 2; c += 32)
                            reinterpret_cast<float2*>(dst)[c] = __ldg(reinterpret_cast<const float2*>(src) + c);
                    } else {
                        for (int c = lane; c < HG_SYNTH_W2V_DIM; c += 32) dst[c] = __ldg(src + c);
                    }
                }
            }
        }
    }

    // ---- rooms <-> rooms: pairs in (a, b) order -- the lanes take 32 consecutive pairs of the row-major upper triangle
    {
        const int64_t n_edges = 2 * A.n_rr_pairs;
        int64_t k = rr0;
        const int total = R * (R - 1) / 2;
        for (int q0 = 0; q0 < total; q0 += 32) {
            int a = 0, b = 0;
            const bool valid = q0 + lane < total;
            if (valid) pair_of(q0 + lane, R, a, b);
            const bool hit = valid && room_pair(gkey, a, b);
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
            const int total = n * (n - 1) / 2;
            for (int q0 = 0; q0 < total; q0 += 32) {
                int a = 0, b = 0;
                const bool valid = q0 + lane < total;
                if (valid) pair_of(q0 + lane, n, a, b);
                const bool hit = valid && object_pair(gkey, r, a, b, p);
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
    if (!a->prefix || !aligned(a->prefix, 32)) return HG_EINVAL;  // rows of four int64 are read as one 32-byte vector
    if (a->n_graphs > 0) k_counts<<<ceil_div(a->n_graphs, 8), 256, 0, stream>>>(*a);
    k_prefix<<<1, kPrefixThreads, 0, stream>>>(*a);
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
