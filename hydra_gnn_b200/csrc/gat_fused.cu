// K3 + K4 -- fused GAT for one destination node type.
//
// A CTA owns a tile of kTile CONSECUTIVE destination rows (consecutive rows belong to the same
// scene graph, so their neighbour sets overlap and the gathers of one tile hit L1/L2).  Per tile:
//   stage   every relation's rowptr slice, neighbour ids, source logits a_src[j,h](+a_edge) and the
//           tile's destination logits go to shared memory in bulk-parallel sweeps (all 256 threads,
//           coalesced on col) -- this removes the rowptr -> col -> a_src dependent-load chain from
//           the per-row critical path;
//   row     one warp per row: per relation, lanes = edges turn the staged logits into softmax
//           weights in place (leaky-relu, max, exp, sum + 1e-16), then lanes = feature chunks sweep
//           the staged edges, each lane gathering its K vector chunks of xs[j] (128-bit loads when the
//           width allows, U rows in flight) and accumulating in registers;
//   out     head concat / head mean (through a shared-memory row) + bias_r into the cross-relation
//           total; after the last relation: group mean, ELU/ReLU, ONE store.
// No atomics, no read-modify-write across relations, no separate softmax / stack+sum / activation
// passes.  Rows whose tile does not fit the staging buffer (hub nodes) take a direct-from-global path.
#include "hg_common.cuh"

namespace hg {
namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / kWarp;
constexpr int kTile = 32;  // max destination rows per CTA tile (4 per warp); fewer when n_dst is small
constexpr int kMaxHeads = 8;

struct GatParams {
    hg_gat_args a;
    int32_t cap;        // staged edges per relation per tile (multiple of 32)
    int32_t tile_rows;  // destination rows per tile: 8, 16 or 32
};

__device__ __forceinline__ float leaky(float v, float slope) { return v > 0.f ? v : v * slope; }

template <int VEC, int K>
__global__ void __launch_bounds__(kThreads, 3) k_gat_fused(const __grid_constant__ GatParams P) {
    constexpr int U = (K * VEC >= 16) ? 2 : 4;
    extern __shared__ __align__(16) uint8_t s_raw[];
    const hg_gat_args& A = P.a;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int H = A.H, C = A.C, F = H * C, R = A.n_rel, cap = P.cap;
    const int nc = F / VEC;
    const int nco = (A.concat ? F : C) / VEC;  // output vector chunks
    const bool head_mean = !(A.concat || H == 1);

    // shared-memory carve-up (all 4-byte items; sizes in words)
    float* s_stage = reinterpret_cast<float*>(s_raw);                   // [kWarps][F]   head-mean staging (16 B aligned)
    int32_t* s_rp = reinterpret_cast<int32_t*>(s_stage + kWarps * F);   // [R][kTile + 1]
    int32_t* s_col = s_rp + R * (kTile + 1);                            // [R][cap]
    float* s_p = reinterpret_cast<float*>(s_col + R * cap);             // [R][H][cap]  logits -> weights
    float* s_ad = s_p + static_cast<size_t>(R) * H * cap;               // [R][kTile][H] destination logits

    int hk[K];
    bool valid[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
        const int chunk = lane + 32 * k;
        valid[k] = chunk < nc;
        hk[k] = valid[k] ? (chunk * VEC) / C : 0;
    }
    float* stage = s_stage + warp * F;
    const int tile_rows = P.tile_rows;
    const int64_t n_tiles = ceil_div<int64_t>(A.n_dst, tile_rows);

    // sum of the relations' biases for this lane's output chunks (added once per row)
    float btot[K][VEC];
#pragma unroll
    for (int k = 0; k < K; ++k)
#pragma unroll
        for (int v = 0; v < VEC; ++v) btot[k][v] = 0.f;
    for (int r = 0; r < R; ++r) {
        if (!A.rel[r].bias) continue;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int chunk = lane + 32 * k;
            if (chunk >= nco) continue;
            float bv[VEC];
            ldg_vec<VEC>(A.rel[r].bias + chunk * VEC, bv);
#pragma unroll
            for (int v = 0; v < VEC; ++v) btot[k][v] += bv[v];
        }
    }

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t t0 = tile * tile_rows;
        const int rows = (A.n_dst - t0 < tile_rows) ? static_cast<int>(A.n_dst - t0) : tile_rows;

        // ---- stage 1: rowptr slices + destination logits
        for (int idx = threadIdx.x; idx < R * (kTile + 1); idx += kThreads) {
            const int r = idx / (kTile + 1), t = idx % (kTile + 1);
            s_rp[idx] = __ldg(A.rel[r].rowptr + t0 + min(t, rows));
        }
        for (int r = 0; r < R; ++r)
            for (int idx = threadIdx.x; idx < rows * H; idx += kThreads) {
                const int t = idx / H, h = idx - t * H;
                s_ad[r * kTile * H + idx] = __ldg(A.rel[r].a_dst + (t0 + t) * A.rel[r].lda_dst + h);
            }
        __syncthreads();

        // ---- stage 2: neighbour ids + source logits of every relation that fits
        for (int r = 0; r < R; ++r) {
            const hg_gat_rel& Rl = A.rel[r];
            const int32_t e0 = s_rp[r * (kTile + 1)], e1 = s_rp[r * (kTile + 1) + rows];
            if (e1 - e0 > cap) continue;
            for (int32_t e = e0 + threadIdx.x; e < e1; e += kThreads) {
                const int32_t j = __ldg(Rl.col + e);
                s_col[r * cap + (e - e0)] = j;
                for (int h = 0; h < H; ++h) {
                    float s = __ldg(Rl.a_src + static_cast<int64_t>(j) * Rl.lda_src + h);
                    if (Rl.a_edge) s += __ldg(Rl.a_edge + static_cast<int64_t>(e) * H + h);
                    s_p[(r * H + h) * cap + (e - e0)] = s;
                }
            }
        }
        __syncthreads();

        // ---- rows: one warp per row
        for (int t = warp; t < rows; t += kWarps) {
            const int64_t i = t0 + t;
            float acc[K][VEC];  // sum over relations, still in [H*C] layout (head mean is linear)
#pragma unroll
            for (int k = 0; k < K; ++k)
#pragma unroll
                for (int v = 0; v < VEC; ++v) acc[k][v] = 0.f;

            for (int r = 0; r < R; ++r) {
                const hg_gat_rel& Rl = A.rel[r];
                const int32_t e0 = s_rp[r * (kTile + 1)], e1 = s_rp[r * (kTile + 1) + rows];
                const int32_t beg = s_rp[r * (kTile + 1) + t], end = s_rp[r * (kTile + 1) + t + 1];
                const int n = end - beg;
                if (n <= 0) continue;
                if ((e1 - e0) <= cap) {
                    const int lo = beg - e0;
                    const int32_t* colp = s_col + r * cap + lo;
                    float* wbase = s_p + static_cast<size_t>(r) * H * cap + lo;
                    // logits -> softmax weights, in place (lanes = edges)
                    if (n == 1) {  // softmax over one edge: exp(0) / (1 + 1e-16) == 1
                        if (lane < H) {
                            wbase[lane * cap] = 1.f;
                            if (A.alpha_saved[r]) A.alpha_saved[r][static_cast<int64_t>(beg) * H + lane] = 1.f;
                        }
                    } else if (n <= 32) {  // whole row in registers: one pass per head
                        for (int h = 0; h < H; ++h) {
                            float* p = wbase + h * cap;
                            const float ad = s_ad[(r * kTile + t) * H + h];
                            const float sc = lane < n ? leaky(p[lane] + ad, A.neg_slope) : -INFINITY;
                            const float m = warp_max(sc);
                            const float ex = lane < n ? __expf(sc - m) : 0.f;
                            const float w = __fdividef(ex, warp_sum(ex) + 1e-16f);
                            if (lane < n) {
                                p[lane] = w;
                                if (A.alpha_saved[r]) A.alpha_saved[r][static_cast<int64_t>(beg + lane) * H + h] = w;
                            }
                        }
                    } else {
                        for (int h = 0; h < H; ++h) {
                            float* p = wbase + h * cap;
                            const float ad = s_ad[(r * kTile + t) * H + h];
                            float m = -INFINITY;
                            for (int e = lane; e < n; e += 32) {
                                const float sc = leaky(p[e] + ad, A.neg_slope);
                                p[e] = sc;
                                m = fmaxf(m, sc);
                            }
                            m = warp_max(m);
                            float l = 0.f;
                            for (int e = lane; e < n; e += 32) {
                                const float ex = __expf(p[e] - m);
                                p[e] = ex;
                                l += ex;
                            }
                            l = warp_sum(l) + 1e-16f;
                            for (int e = lane; e < n; e += 32) {
                                const float w = __fdividef(p[e], l);
                                p[e] = w;
                                if (A.alpha_saved[r]) A.alpha_saved[r][static_cast<int64_t>(beg + e) * H + h] = w;
                            }
                        }
                    }
                    __syncwarp();
                    // gather-accumulate (lanes = feature chunks)
                    for (int e = 0; e < n; e += U) {
                        float val[U][K][VEC];
                        float w[U][K];
#pragma unroll
                        for (int u = 0; u < U; ++u) {
                            const bool live = e + u < n;
                            const int32_t j = live ? colp[e + u] : 0;
                            const float* __restrict__ row = Rl.xs + static_cast<int64_t>(j) * Rl.ldxs;
#pragma unroll
                            for (int k = 0; k < K; ++k) {
                                if (live && valid[k]) {
                                    ldg_vec<VEC>(row + (lane + 32 * k) * VEC, val[u][k]);
                                    w[u][k] = wbase[hk[k] * cap + e + u];
                                } else {
                                    w[u][k] = 0.f;
#pragma unroll
                                    for (int v = 0; v < VEC; ++v) val[u][k][v] = 0.f;
                                }
                            }
                        }
#pragma unroll
                        for (int u = 0; u < U; ++u)
#pragma unroll
                            for (int k = 0; k < K; ++k)
#pragma unroll
                                for (int v = 0; v < VEC; ++v) acc[k][v] = fmaf(w[u][k], val[u][k][v], acc[k][v]);
                    }
                    __syncwarp();
                } else {
                    // hub path: the tile's edges do not fit the staging buffer -> stream from global
                    for (int h = 0; h < H; ++h) {
                        const float ad = s_ad[(r * kTile + t) * H + h];
                        float m = -INFINITY, l = 0.f;
                        for (int32_t e = beg + lane; e < end; e += 32) {
                            float sc = __ldg(Rl.a_src + static_cast<int64_t>(__ldg(Rl.col + e)) * Rl.lda_src + h) + ad;
                            if (Rl.a_edge) sc += __ldg(Rl.a_edge + static_cast<int64_t>(e) * H + h);
                            m = fmaxf(m, leaky(sc, A.neg_slope));
                        }
                        m = warp_max(m);
                        for (int32_t e = beg + lane; e < end; e += 32) {
                            float sc = __ldg(Rl.a_src + static_cast<int64_t>(__ldg(Rl.col + e)) * Rl.lda_src + h) + ad;
                            if (Rl.a_edge) sc += __ldg(Rl.a_edge + static_cast<int64_t>(e) * H + h);
                            l += __expf(leaky(sc, A.neg_slope) - m);
                        }
                        l = warp_sum(l) + 1e-16f;
                        for (int32_t e = beg; e < end; ++e) {  // warp-uniform sweep, one edge at a time
                            const int32_t j = __ldg(Rl.col + e);
                            float sc = __ldg(Rl.a_src + static_cast<int64_t>(j) * Rl.lda_src + h) + ad;
                            if (Rl.a_edge) sc += __ldg(Rl.a_edge + static_cast<int64_t>(e) * H + h);
                            const float w = __fdividef(__expf(leaky(sc, A.neg_slope) - m), l);
                            if (A.alpha_saved[r] && lane == 0) A.alpha_saved[r][static_cast<int64_t>(e) * H + h] = w;
                            const float* __restrict__ row = Rl.xs + static_cast<int64_t>(j) * Rl.ldxs;
#pragma unroll
                            for (int k = 0; k < K; ++k) {
                                if (valid[k] && hk[k] == h) {
                                    float val[VEC];
                                    ldg_vec<VEC>(row + (lane + 32 * k) * VEC, val);
#pragma unroll
                                    for (int v = 0; v < VEC; ++v) acc[k][v] = fmaf(w, val[v], acc[k][v]);
                                }
                            }
                        }
                    }
                }
            }

            // ---- row epilogue: head concat | head mean, + summed biases, group mean, activation, ONE store
            const float gscale = (A.group == HG_GROUP_MEAN) ? 1.f / static_cast<float>(R) : 1.f;
            if (head_mean) {
#pragma unroll
                for (int k = 0; k < K; ++k)
                    if (valid[k]) st_vec<VEC>(stage + (lane + 32 * k) * VEC, acc[k]);
                __syncwarp();
                const float inv_h = 1.f / static_cast<float>(H);
#pragma unroll
                for (int k = 0; k < K; ++k) {
                    const int chunk = lane + 32 * k;
                    if (chunk >= nco) continue;
                    float sum[VEC];
#pragma unroll
                    for (int v = 0; v < VEC; ++v) sum[v] = 0.f;
                    for (int h = 0; h < H; ++h) {
                        float hv[VEC];
#pragma unroll
                        for (int v = 0; v < VEC; ++v) hv[v] = stage[h * C + chunk * VEC + v];
#pragma unroll
                        for (int v = 0; v < VEC; ++v) sum[v] += hv[v];
                    }
#pragma unroll
                    for (int v = 0; v < VEC; ++v) acc[k][v] = sum[v] * inv_h;
                }
                __syncwarp();
            }
#pragma unroll
            for (int k = 0; k < K; ++k) {
                const int chunk = lane + 32 * k;
                if (chunk >= nco) continue;
                float o[VEC];
#pragma unroll
                for (int v = 0; v < VEC; ++v) {
                    float y = (acc[k][v] + btot[k][v]) * gscale;
                    if (A.act == HG_ACT_RELU) y = fmaxf(y, 0.f);
                    else if (A.act == HG_ACT_ELU) y = y > 0.f ? y : __expf(y) - 1.f;  // F.elu: exp(x) - 1
                    o[v] = y;
                }
                st_vec<VEC>(A.out + i * A.ldout + chunk * VEC, o);
            }
        }
        __syncthreads();  // staging buffers are reused by the next tile
    }
}

size_t smem_words(const hg_gat_args& A, int cap) {
    const size_t R = A.n_rel, H = A.H;
    return R * (kTile + 1) + R * cap + R * H * cap + R * kTile * H + static_cast<size_t>(kWarps) * H * A.C;
}

template <int VEC, int K>
int launch(GatParams& P, cudaStream_t stream) {
    const hg_gat_args& A = P.a;
    // staging capacity: as much as ~40 KB of shared memory holds, 64..1024 edges per relation per tile
    const size_t budget_words = 40 * 1024 / 4;
    const size_t fixed = smem_words(A, 0);
    int cap = fixed < budget_words ? static_cast<int>((budget_words - fixed) / (A.n_rel * (1 + A.H))) : 0;
    cap = cap / 32 * 32;
    cap = cap < 64 ? 64 : (cap > 1024 ? 1024 : cap);
    P.cap = cap;
    const size_t smem = smem_words(A, cap) * 4;
    if (smem > 200 * 1024) return HG_EUNSUPPORTED;
    if (smem > 48 * 1024) {
        HG_CUDA_TRY(cudaFuncSetAttribute(k_gat_fused<VEC, K>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem)));
    }
    const int sms = sm_count() > 0 ? sm_count() : 148;
    // small destination sets (rooms) get smaller tiles so that every SM still has several CTAs
    int tile_rows = kTile;
    while (tile_rows > 8 && ceil_div<int64_t>(A.n_dst, tile_rows) < static_cast<int64_t>(sms) * 12) tile_rows >>= 1;
    P.tile_rows = tile_rows;
    const int64_t need = ceil_div<int64_t>(A.n_dst, tile_rows);
    const int64_t cap_grid = static_cast<int64_t>(sms) * 8 * 16;
    const unsigned grid = static_cast<unsigned>(need < cap_grid ? need : cap_grid);
    k_gat_fused<VEC, K><<<grid, kThreads, smem, stream>>>(P);
    return check_launch();
}

template <int VEC>
int dispatch(GatParams& P, int nc, cudaStream_t stream) {
    const int k = ceil_div(nc, 32);
    switch (k) {
        case 1: return launch<VEC, 1>(P, stream);
        case 2: return launch<VEC, 2>(P, stream);
        case 3: return launch<VEC, 3>(P, stream);
        case 4: return launch<VEC, 4>(P, stream);
        case 5: return launch<VEC, 5>(P, stream);
        case 6: return launch<VEC, 6>(P, stream);
        case 7:
        case 8: return launch<VEC, 8>(P, stream);
        default: return HG_EUNSUPPORTED;
    }
}

}  // namespace
}  // namespace hg

extern "C" int hg_gat_fused(const hg_gat_args* a, void* stream_) {
    using namespace hg;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!a) return HG_EINVAL;
    if (a->n_rel < 1 || a->n_rel > HG_MAX_RELATIONS || a->n_dst < 0) return HG_EINVAL;
    if (a->H < 1 || a->H > kMaxHeads || a->C < 1) return HG_EINVAL;
    if (a->n_dst == 0) return HG_OK;
    if (!a->out) return HG_EINVAL;
    const int F = a->H * a->C;
    for (int r = 0; r < a->n_rel; ++r) {
        const hg_gat_rel& R = a->rel[r];
        // xs / a_src may be NULL only for an empty source set (every row of that relation is then empty)
        if (!R.a_dst || !R.rowptr || !R.col) return HG_EINVAL;
    }
    // Widest vector that divides C (a chunk must not straddle two heads) and fits all alignments;
    // prefer the widest one that still gives every lane of the first chunk work.
    auto ok = [&](int vec) {
        const size_t bytes = sizeof(float) * vec;
        if (a->C % vec || a->ldout % vec || !aligned(a->out, bytes)) return false;
        for (int r = 0; r < a->n_rel; ++r) {
            const hg_gat_rel& R = a->rel[r];
            if (R.ldxs % vec || !aligned(R.xs, bytes)) return false;
            if (R.bias && !aligned(R.bias, bytes)) return false;
        }
        return true;
    };
    int vec = 1;
    if (ok(4) && F / 4 >= 32) vec = 4;
    else if (ok(2) && F / 2 >= 32) vec = 2;
    else if (F >= 32) vec = 1;
    else vec = ok(4) ? 4 : (ok(2) ? 2 : 1);
    GatParams P;
    P.a = *a;
    P.cap = 0;
    P.tile_rows = kTile;
    const int nc = F / vec;
    return vec == 4 ? dispatch<4>(P, nc, stream) : (vec == 2 ? dispatch<2>(P, nc, stream) : dispatch<1>(P, nc, stream));
}
