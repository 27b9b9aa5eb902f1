// K1 -- graph plan: CSR-by-destination for up to 16 relations in four launches.
//
// Integer work, bit-exact with a stable sort by destination of the (self-loop adjusted) edge
// list.  Stability without a radix sort: count -> scan -> unordered atomic fill of edge ids
// -> every edge ranks itself inside its (short) row.  All relations of a batch share each
// launch (blockIdx.y = relation) so a 4-relation DSG plan or a 15-relation H-tree plan costs
// the same number of launches.  HBM-bound: 16 B/edge in, 8 B/edge + 4 B/row out.
//   k_count  histogram of the kept edges + (for relations that ask for tile windows) the min / max source row of
//            every 32-row destination tile, warp-aggregated (match_any + reduce) before the two atomics
//   k_scan   ONE pass: chained scan with decoupled look-back (ticketed tiles of 2048 rows, a warp looks back 32
//            predecessors at a time); writes rowptr, the GCN degree^-1/2, the tile windows, clears the cursors
//   k_fill   unordered atomic fill of edge ids;   k_rank   stable rank inside the row, col / perm
#include <mutex>

#include "hg_common.cuh"

namespace hg {
namespace {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;  // rows per scan block
constexpr int kEdgeThreads = 256;

struct RelDev {
    const int64_t* src;
    const int64_t* dst;
    int64_t n_edges;
    int32_t n_src, n_dst, self_loops, n_loop;
    int32_t* rowptr;
    int32_t* col;
    int32_t* perm;
    float* inv_sqrt_deg;
    int32_t* cnt;      // [n_dst]  histogram, then fill cursor
    int32_t* ids_tmp;  // [E + n_loop] edge ids, row-bucketed, unordered inside a row
    unsigned long long* scan_state;  // [n_scan_blocks] (flag << 32 | value) + [1] ticket counter; zeroed by the plan
    int32_t n_scan_blocks;
    int32_t pad;
    int32_t* tile_win; // [ceil(n_dst / HG_TILE_ROWS)][2] or NULL
    int32_t* tw_tmp;   // [ceil(n_dst / HG_TILE_ROWS)][2] {max(INT32_MAX - src), max(src + 1)}, 0 = no edge; zeroed by the plan
};

struct PlanParams {
    RelDev r[HG_MAX_PLAN_RELATIONS];
    int32_t* status;
};

__device__ __forceinline__ bool edge_kept(const RelDev& R, int64_t e, int32_t& s, int32_t& d, int32_t* status) {
    const int64_t s64 = R.src[e], d64 = R.dst[e];
    if (s64 < 0 || s64 >= R.n_src || d64 < 0 || d64 >= R.n_dst) {
        if (status) atomicAdd(status, 1);
        return false;
    }
    s = static_cast<int32_t>(s64);
    d = static_cast<int32_t>(d64);
    return !(R.self_loops && s == d);
}

__global__ void __launch_bounds__(kEdgeThreads) k_count(const __grid_constant__ PlanParams P) {
    const RelDev& R = P.r[blockIdx.y];
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    const int lane = threadIdx.x & 31;
    // warp-uniform trip count: the tile-window part uses warp collectives
    for (int64_t b = static_cast<int64_t>(blockIdx.x) * blockDim.x + (threadIdx.x - lane); b < R.n_edges; b += stride) {
        const int64_t e = b + lane;
        int32_t s = 0, d = 0;
        const bool kept = e < R.n_edges && edge_kept(R, e, s, d, P.status);
        if (kept) atomicAdd(&R.cnt[d], 1);
        if (R.tw_tmp) {  // edges of one graph are neighbours in the list: a warp usually touches one or two tiles
            const unsigned live = __ballot_sync(0xffffffffu, kept);
            if (kept) {
                const int32_t tile = d / HG_TILE_ROWS;
                const unsigned grp = __match_any_sync(live, tile);
                const int32_t lo = __reduce_min_sync(grp, s), hi = __reduce_max_sync(grp, s);
                if (lane == __ffs(grp) - 1) {
                    atomicMax(&R.tw_tmp[2 * tile], INT32_MAX - lo);
                    atomicMax(&R.tw_tmp[2 * tile + 1], hi + 1);
                }
            }
        }
    }
}

// Block-wide exclusive scan of one int per thread; returns the exclusive prefix, total in `total`.
__device__ __forceinline__ int32_t block_exclusive_scan(int32_t v, int32_t& total) {
    __shared__ int32_t warp_tot[kScanThreads / kWarp];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warp_tot[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int32_t w = lane < kScanThreads / kWarp ? warp_tot[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int32_t t = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += t;
        }
        if (lane < kScanThreads / kWarp) warp_tot[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const int32_t warp_off = wid == 0 ? 0 : warp_tot[wid - 1];
    total = warp_tot[kScanThreads / kWarp - 1];
    __syncthreads();
    return warp_off + inc - v;
}

constexpr unsigned long long kAggregate = 1ull << 32, kInclusive = 2ull << 32;

__device__ __forceinline__ unsigned long long ld_state(const unsigned long long* p) {
    return *reinterpret_cast<const volatile unsigned long long*>(p);
}
__device__ __forceinline__ void st_state(unsigned long long* p, unsigned long long v) {
    *reinterpret_cast<volatile unsigned long long*>(p) = v;
}

// Sum of the totals of tiles [0, tile): warp 0 walks back 32 predecessors at a time until it meets an inclusive prefix.
// A predecessor holds a smaller TICKET, so it is running or done: the wait ends.
__device__ __forceinline__ int32_t look_back(const unsigned long long* state, int32_t tile, int lane) {
    int32_t prefix = 0;
    for (int32_t p = tile - 1; p >= 0; p -= 32) {
        const int32_t q = p - lane;
        unsigned long long w = kInclusive;  // lanes before tile 0 read as "inclusive prefix 0"
        if (q >= 0) {
            do { w = ld_state(state + q); } while ((w >> 32) == 0);
        }
        const unsigned inc = __ballot_sync(0xffffffffu, (w >> 32) == 2);
        const int stop = inc ? __ffs(inc) - 1 : 31;  // nearest inclusive prefix ends the walk
        int32_t v = lane <= stop ? static_cast<int32_t>(static_cast<uint32_t>(w)) : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        prefix += v;
        if (inc) break;
    }
    return prefix;
}

__global__ void __launch_bounds__(kScanThreads) k_scan(const __grid_constant__ PlanParams P) {
    const RelDev& R = P.r[blockIdx.y];
    if (R.n_scan_blocks == 0) {  // no destination rows: rowptr = [0]
        if (blockIdx.x == 0 && threadIdx.x == 0 && R.rowptr) R.rowptr[0] = 0;
        return;
    }
    if (static_cast<int>(blockIdx.x) >= R.n_scan_blocks) return;
    __shared__ int32_t s_tile, s_prefix;
    __shared__ int32_t s_out[kScanTile + 1];
    if (threadIdx.x == 0) s_tile = static_cast<int32_t>(atomicAdd(R.scan_state + R.n_scan_blocks, 1ull));  // ticket
    __syncthreads();
    const int32_t tile = s_tile;
    const int32_t base = tile * kScanTile + threadIdx.x * kScanItems;
    int32_t c[kScanItems];
    if (base < R.n_dst) {  // cnt is 256-byte aligned and padded to 256 bytes: whole groups of 8 are readable
        const int4 lo = *reinterpret_cast<const int4*>(R.cnt + base), hi = *reinterpret_cast<const int4*>(R.cnt + base + 4);
        c[0] = lo.x; c[1] = lo.y; c[2] = lo.z; c[3] = lo.w; c[4] = hi.x; c[5] = hi.y; c[6] = hi.z; c[7] = hi.w;
    }
    int32_t v = 0;
#pragma unroll
    for (int t = 0; t < kScanItems; ++t) {
        const int32_t k = base + t;
        c[t] = k < R.n_dst ? c[t] + (k < R.n_loop ? 1 : 0) : 0;
        v += c[t];
    }
    int32_t total;
    int32_t run = block_exclusive_scan(v, total);
    if (threadIdx.x < 32) {
        if (threadIdx.x == 0) st_state(R.scan_state + tile, (tile == 0 ? kInclusive : kAggregate) | static_cast<uint32_t>(total));
        const int32_t prefix = tile == 0 ? 0 : look_back(R.scan_state, tile, threadIdx.x);
        if (threadIdx.x == 0) {
            if (tile != 0) st_state(R.scan_state + tile, kInclusive | static_cast<uint32_t>(prefix + total));
            s_prefix = prefix;
            s_out[kScanTile] = prefix + total;
            if (tile == R.n_scan_blocks - 1) R.rowptr[R.n_dst] = prefix + total;  // E'
        }
    }
    __syncthreads();
    run += s_prefix;
#pragma unroll
    for (int t = 0; t < kScanItems; ++t) {
        s_out[threadIdx.x * kScanItems + t] = run;
        run += c[t];
    }
    __syncthreads();
    const int32_t k0 = tile * kScanTile;
    for (int i = threadIdx.x; i < kScanTile; i += kScanThreads) {
        const int32_t k = k0 + i;
        if (k >= R.n_dst) break;
        R.rowptr[k] = s_out[i];
        R.cnt[k] = 0;  // becomes the fill cursor
        if (R.inv_sqrt_deg) {
            const int32_t deg = s_out[i + 1] - s_out[i];
            R.inv_sqrt_deg[k] = deg > 0 ? 1.0f / sqrtf(static_cast<float>(deg)) : 0.f;
        }
        if (R.tile_win && (i % HG_TILE_ROWS) == 0) {  // kScanTile is a multiple of the tile: tiles do not straddle blocks
            const int32_t t = k / HG_TILE_ROWS;
            const int32_t enc_lo = R.tw_tmp[2 * t], enc_hi = R.tw_tmp[2 * t + 1];
            int32_t lo = enc_hi ? INT32_MAX - enc_lo : INT32_MAX, hi = enc_hi - 1;
            if (k < R.n_loop) {  // appended self loops (j, j) of this tile's rows
                const int32_t r1 = k + HG_TILE_ROWS < R.n_dst ? k + HG_TILE_ROWS : R.n_dst;
                const int32_t last = (r1 < R.n_loop ? r1 : R.n_loop) - 1;
                lo = k < lo ? k : lo;
                hi = last > hi ? last : hi;
            }
            R.tile_win[2 * t] = hi >= 0 ? lo : 0;
            R.tile_win[2 * t + 1] = hi >= 0 ? hi - lo + 1 : 0;
        }
    }
}

__global__ void __launch_bounds__(kEdgeThreads) k_fill(const __grid_constant__ PlanParams P) {
    const RelDev& R = P.r[blockIdx.y];
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t e = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; e < R.n_edges; e += stride) {
        int32_t s, d;
        if (edge_kept(R, e, s, d, nullptr)) {
            const int32_t slot = R.rowptr[d] + atomicAdd(&R.cnt[d], 1);
            R.ids_tmp[slot] = static_cast<int32_t>(e);
        }
    }
}

// Every kept edge ranks itself among the edge ids of its row => stable order; loops go last.
__global__ void __launch_bounds__(kEdgeThreads) k_rank(const __grid_constant__ PlanParams P) {
    const RelDev& R = P.r[blockIdx.y];
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    const int64_t total = R.n_edges + R.n_loop;
    for (int64_t e = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; e < total; e += stride) {
        if (e >= R.n_edges) {  // appended self loop (k, k)
            const int32_t k = static_cast<int32_t>(e - R.n_edges);
            const int32_t pos = R.rowptr[k + 1] - 1;
            R.col[pos] = k;
            if (R.perm) R.perm[pos] = -1;
            continue;
        }
        int32_t s, d;
        if (!edge_kept(R, e, s, d, nullptr)) continue;
        const int32_t base = R.rowptr[d];
        const int32_t n_real = R.rowptr[d + 1] - base - (d < R.n_loop ? 1 : 0);
        const int32_t me = static_cast<int32_t>(e);
        int32_t rank = 0;
        for (int32_t t = 0; t < n_real; ++t) rank += (R.ids_tmp[base + t] < me) ? 1 : 0;
        R.col[base + rank] = s;
        if (R.perm) R.perm[base + rank] = me;
    }
}

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct WsLayout {
    size_t cnt_off[HG_MAX_PLAN_RELATIONS], ids_off[HG_MAX_PLAN_RELATIONS], state_off[HG_MAX_PLAN_RELATIONS],
        tw_off[HG_MAX_PLAN_RELATIONS];
    size_t cnt_total;  // bytes of the leading region that must be zeroed: histograms, scan states + tickets, window scratch
    size_t total;
};

int layout(const hg_plan_rel* rels, int n_rel, WsLayout& L) {
    if (!rels || n_rel < 0 || n_rel > HG_MAX_PLAN_RELATIONS) return HG_EINVAL;
    size_t off = 0;
    for (int r = 0; r < n_rel; ++r) {
        if (rels[r].n_edges < 0 || rels[r].n_src < 0 || rels[r].n_dst < 0) return HG_EINVAL;
        const int64_t loops = rels[r].self_loops ? (rels[r].n_src < rels[r].n_dst ? rels[r].n_src : rels[r].n_dst) : 0;
        if (rels[r].n_edges + loops >= (int64_t(1) << 31)) return HG_ERANGE;
        L.cnt_off[r] = off;
        off = align_up(off + sizeof(int32_t) * static_cast<size_t>(rels[r].n_dst), 256);
        L.state_off[r] = off;
        off = align_up(off + sizeof(unsigned long long) * static_cast<size_t>(ceil_div<int64_t>(rels[r].n_dst, kScanTile) + 1), 256);
        L.tw_off[r] = off;
        if (rels[r].tile_win)
            off = align_up(off + 2 * sizeof(int32_t) * static_cast<size_t>(ceil_div<int64_t>(rels[r].n_dst, HG_TILE_ROWS)), 256);
    }
    L.cnt_total = off;
    for (int r = 0; r < n_rel; ++r) {
        const int64_t loops = rels[r].self_loops ? (rels[r].n_src < rels[r].n_dst ? rels[r].n_src : rels[r].n_dst) : 0;
        L.ids_off[r] = off;
        off = align_up(off + sizeof(int32_t) * static_cast<size_t>(rels[r].n_edges + loops), 256);
    }
    L.total = off;
    return HG_OK;
}

}  // namespace
}  // namespace hg

extern "C" size_t hg_plan_workspace_bytes(const hg_plan_rel* rels, int n_rel) {
    hg::WsLayout L;
    if (hg::layout(rels, n_rel, L) != HG_OK) return 0;
    return L.total + 256;
}

extern "C" int hg_plan_build(const hg_plan_rel* rels, int n_rel, int32_t* status, void* workspace,
                             size_t workspace_bytes, void* stream_) {
    using namespace hg;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    WsLayout L;
    int rc = layout(rels, n_rel, L);
    if (rc != HG_OK) return rc;
    if (n_rel == 0) return HG_OK;
    if (!status) return HG_EINVAL;
    if (!workspace || workspace_bytes < L.total) return HG_EWORKSPACE;
    if (!aligned(workspace, 256)) return HG_EALIGN;
    char* ws = static_cast<char*>(workspace);

    PlanParams P;
    P.status = status;
    int64_t max_edges = 0, max_total = 0;
    int32_t max_blocks = 0;
    for (int r = 0; r < n_rel; ++r) {
        const hg_plan_rel& h = rels[r];
        if (h.n_edges > 0 && (!h.src || !h.dst)) return HG_EINVAL;
        if (!h.rowptr || (!h.col && (h.n_edges > 0 || h.self_loops))) return HG_EINVAL;
        RelDev& d = P.r[r];
        d.src = h.src; d.dst = h.dst; d.n_edges = h.n_edges;
        d.n_src = h.n_src; d.n_dst = h.n_dst; d.self_loops = h.self_loops ? 1 : 0;
        d.n_loop = h.self_loops ? (h.n_src < h.n_dst ? h.n_src : h.n_dst) : 0;
        d.rowptr = h.rowptr; d.col = h.col; d.perm = h.perm; d.inv_sqrt_deg = h.inv_sqrt_deg;
        d.cnt = reinterpret_cast<int32_t*>(ws + L.cnt_off[r]);
        d.ids_tmp = reinterpret_cast<int32_t*>(ws + L.ids_off[r]);
        d.scan_state = reinterpret_cast<unsigned long long*>(ws + L.state_off[r]);
        d.n_scan_blocks = static_cast<int32_t>(ceil_div<int64_t>(h.n_dst, kScanTile));
        d.pad = 0;
        d.tile_win = h.tile_win;
        d.tw_tmp = h.tile_win ? reinterpret_cast<int32_t*>(ws + L.tw_off[r]) : nullptr;
        max_edges = h.n_edges > max_edges ? h.n_edges : max_edges;
        max_total = h.n_edges + d.n_loop > max_total ? h.n_edges + d.n_loop : max_total;
        max_blocks = d.n_scan_blocks > max_blocks ? d.n_scan_blocks : max_blocks;
    }
    for (int r = n_rel; r < HG_MAX_PLAN_RELATIONS; ++r) P.r[r] = RelDev{};

    HG_CUDA_TRY(cudaMemsetAsync(ws, 0, L.cnt_total, stream));
    HG_CUDA_TRY(cudaMemsetAsync(status, 0, sizeof(int32_t), stream));

    // Edge passes: grid-stride, capped at 8 waves of 148 SMs x 8 resident CTAs.
    const int sms = sm_count() > 0 ? sm_count() : 148;
    auto edge_grid = [&](int64_t n) {
        int64_t g = ceil_div<int64_t>(n > 0 ? n : 1, kEdgeThreads);
        const int64_t cap = static_cast<int64_t>(sms) * 8 * 8;
        return static_cast<unsigned>(g < cap ? g : cap);
    };
    if (max_edges > 0) k_count<<<dim3(edge_grid(max_edges), n_rel), kEdgeThreads, 0, stream>>>(P);
    k_scan<<<dim3(max_blocks > 0 ? max_blocks : 1, n_rel), kScanThreads, 0, stream>>>(P);
    if (max_edges > 0) k_fill<<<dim3(edge_grid(max_edges), n_rel), kEdgeThreads, 0, stream>>>(P);
    if (max_total > 0) k_rank<<<dim3(edge_grid(max_total), n_rel), kEdgeThreads, 0, stream>>>(P);
    return check_launch();
}

// ---------------------------------------------------------------------------------------------------
// hg_plan_build_static -- hg_plan_build for callers that cannot call cudaMalloc (code under CUDA-graph
// capture, tight serving loops, harnesses that forbid allocation).  The workspace is a file-scope device
// array of HG_PLAN_STATIC_SCRATCH_BYTES: every GPU that loads the library holds it until exit.  Every
// call shares the array (it keeps the last call's contents; the plan clears the part that must start at
// zero itself), so calls that overlap on the GPU -- two streams, two threads -- race.
// ---------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_plan_scratch[HG_PLAN_STATIC_SCRATCH_BYTES];

namespace hg {
namespace {

constexpr int kMaxDevices = 64;
std::mutex g_scratch_mu;
void* g_scratch_ptr[kMaxDevices] = {};  // device address of g_plan_scratch, per device ordinal

// In host code `g_plan_scratch` is a host-side stand-in, not a device address: ask the runtime for the
// address on the CURRENT device, once per device.
int static_scratch(void** out) {
    int dev = -1;
    HG_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return HG_EUNSUPPORTED;
    std::lock_guard<std::mutex> lock(g_scratch_mu);
    if (!g_scratch_ptr[dev]) {
        void* p = nullptr;
        HG_CUDA_TRY(cudaGetSymbolAddress(&p, g_plan_scratch));
        if (!p) return HG_ECUDA;
        g_scratch_ptr[dev] = p;
    }
    *out = g_scratch_ptr[dev];
    return HG_OK;
}

}  // namespace
}  // namespace hg

extern "C" int hg_plan_build_static(const hg_plan_rel* rels, int n_rel, int32_t* status, void* stream_) {
    using namespace hg;
    WsLayout L;
    const int rc = layout(rels, n_rel, L);
    if (rc != HG_OK) return rc;
    if (L.total > static_cast<size_t>(HG_PLAN_STATIC_SCRATCH_BYTES)) return HG_ECAPACITY;  // before any CUDA call
    if (n_rel == 0) return HG_OK;
    void* ws = nullptr;
    const int rs = static_scratch(&ws);
    if (rs != HG_OK) return rs;
    return hg_plan_build(rels, n_rel, status, ws, HG_PLAN_STATIC_SCRATCH_BYTES, stream_);
}
