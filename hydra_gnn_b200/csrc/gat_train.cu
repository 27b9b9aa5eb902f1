// Training form of the GAT aggregation (include/hydra_gnn_b200.h, "Training step"): forward with attention
// dropout and saved softmax weights, and the two-pass backward (by destination, then by source over the
// reversed CSR -- every gradient column is written once with plain stores, no atomics on the feature rows).
//
// Warp per row; lanes own the elements e = lane + 32 k of an H*C-wide row (coalesced 128-byte gathers).
// HBM-bound: the forward and pass 1 gather xs[col[e]] (4 H C bytes per edge), pass 2 gathers g[dst(e)]
// (4 * out-width bytes per edge).
#include "hg_common.cuh"

namespace hg {
namespace {

constexpr int kWarps = 8;
constexpr int kThreads = kWarps * kWarp;
constexpr int kMaxH = 4;

__device__ __forceinline__ float keep_scale(uint64_t seed, int r, int64_t idx, float p, float inv_keep) {
    // splitmix64 of (seed, relation slot, CSR position * H + head): regenerated identically by the backward
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (static_cast<uint64_t>(idx) * HG_MAX_RELATIONS + r + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    const float u = static_cast<float>(z >> 40) * (1.0f / 16777216.0f);
    return u >= p ? inv_keep : 0.f;
}

__device__ __forceinline__ float sel(const float (&v)[kMaxH], int h) {
    return h == 0 ? v[0] : (h == 1 ? v[1] : (h == 2 ? v[2] : v[3]));
}

__device__ __forceinline__ float leaky(float s, float slope) { return s > 0.f ? s : s * slope; }

template <int NE>
__global__ void __launch_bounds__(kThreads) k_fwd(const __grid_constant__ hg_gat_train_args A) {
    __shared__ float s_acc[kWarps][NE * kWarp];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int H = A.H, C = A.C, HC = H * C;
    const int width = A.concat ? HC : C;
    const float slope = A.neg_slope, p_drop = A.dropout_p;
    const bool drop = p_drop > 0.f;
    const float inv_keep = drop ? 1.f / (1.f - p_drop) : 1.f;
    int hk[NE];
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        const int e = lane + 32 * k;
        hk[k] = e < HC ? e / C : 0;
    }
    for (int i = blockIdx.x * kWarps + w; i < A.n_dst; i += gridDim.x * kWarps) {
        float acc[NE];
#pragma unroll
        for (int k = 0; k < NE; ++k) acc[k] = 0.f;
        for (int r = 0; r < A.n_rel; ++r) {
            const hg_gat_train_rel& R = A.rel[r];
            const int beg = R.rowptr[i], end = R.rowptr[i + 1];
            if (beg == end) continue;
            float ad[kMaxH], m[kMaxH], sum[kMaxH];
#pragma unroll
            for (int h = 0; h < kMaxH; ++h) {
                ad[h] = h < H ? __ldg(R.a_dst + static_cast<int64_t>(i) * R.lda_dst + h) : 0.f;
                m[h] = -INFINITY;
                sum[h] = 0.f;
            }
            for (int p = beg + lane; p < end; p += 32) {
                const float* as = R.a_src + static_cast<int64_t>(__ldg(R.col + p)) * R.lda_src;
#pragma unroll
                for (int h = 0; h < kMaxH; ++h)
                    if (h < H) m[h] = fmaxf(m[h], leaky(__ldg(as + h) + ad[h], slope));
            }
#pragma unroll
            for (int h = 0; h < kMaxH; ++h) m[h] = warp_max(m[h]);
            for (int p = beg + lane; p < end; p += 32) {
                const float* as = R.a_src + static_cast<int64_t>(__ldg(R.col + p)) * R.lda_src;
#pragma unroll
                for (int h = 0; h < kMaxH; ++h)
                    if (h < H) sum[h] += expf(leaky(__ldg(as + h) + ad[h], slope) - m[h]);
            }
#pragma unroll
            for (int h = 0; h < kMaxH; ++h) sum[h] = warp_sum(sum[h]) + 1e-16f;
            for (int p = beg + lane; p < end; p += 32) {
                const float* as = R.a_src + static_cast<int64_t>(__ldg(R.col + p)) * R.lda_src;
#pragma unroll
                for (int h = 0; h < kMaxH; ++h)
                    if (h < H)
                        R.alpha[static_cast<int64_t>(p) * H + h] = expf(leaky(__ldg(as + h) + ad[h], slope) - m[h]) / sum[h];
            }
            __syncwarp();
            for (int p = beg; p < end; ++p) {
                const int j = __ldg(R.col + p);
                float wv[kMaxH];
#pragma unroll
                for (int h = 0; h < kMaxH; ++h) {
                    wv[h] = 0.f;
                    if (h < H) {
                        const int64_t idx = static_cast<int64_t>(p) * H + h;
                        wv[h] = R.alpha[idx] * (drop ? keep_scale(A.seed, r, idx, p_drop, inv_keep) : 1.f);
                    }
                }
                const float* xr = R.xs + static_cast<int64_t>(j) * R.ldxs;
#pragma unroll
                for (int k = 0; k < NE; ++k) {
                    const int e = lane + 32 * k;
                    if (e < HC) acc[k] = fmaf(sel(wv, hk[k]), __ldg(xr + e), acc[k]);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < NE; ++k) s_acc[w][lane + 32 * k] = acc[k];
        __syncwarp();
        for (int c = lane; c < width; c += 32) {
            float v;
            if (A.concat) {
                v = s_acc[w][c];
            } else {
                v = 0.f;
                for (int h = 0; h < H; ++h) v += s_acc[w][h * C + c];
                v /= static_cast<float>(H);
            }
            for (int r = 0; r < A.n_rel; ++r)
                if (A.rel[r].bias) v += __ldg(A.rel[r].bias + c);
            if (A.group == HG_GROUP_MEAN) v /= static_cast<float>(A.n_rel);
            A.out[static_cast<int64_t>(i) * A.ldout + c] = apply_act(v, A.act);
        }
        __syncwarp();
    }
}

// Pass 1 of the backward: by destination row.
template <int NE>
__global__ void __launch_bounds__(kThreads) k_bwd_dst(const __grid_constant__ hg_gat_train_args A) {
    __shared__ float s_g[kWarps][NE * kWarp];
    __shared__ float s_db[NE * kWarp];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int H = A.H, C = A.C, HC = H * C;
    const int width = A.concat ? HC : C;
    const float slope = A.neg_slope, p_drop = A.dropout_p;
    const bool drop = p_drop > 0.f;
    const float inv_keep = drop ? 1.f / (1.f - p_drop) : 1.f;
    const float gscale = A.group == HG_GROUP_MEAN ? 1.f / static_cast<float>(A.n_rel) : 1.f;
    const float invH = A.concat ? 1.f : 1.f / static_cast<float>(H);
    for (int c = threadIdx.x; c < NE * kWarp; c += kThreads) s_db[c] = 0.f;
    __syncthreads();
    int hk[NE], ce[NE];
    float db[NE];
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        const int e = lane + 32 * k;
        hk[k] = e < HC ? e / C : 0;
        ce[k] = A.concat ? e : e - hk[k] * C;
        db[k] = 0.f;
    }
    for (int i = blockIdx.x * kWarps + w; i < A.n_dst; i += gridDim.x * kWarps) {
#pragma unroll
        for (int k = 0; k < NE; ++k) {
            const int c = lane + 32 * k;
            if (c < width) {
                const float o = A.out[static_cast<int64_t>(i) * A.ldout + c];
                float d = __ldg(A.dout + static_cast<int64_t>(i) * A.lddout + c);
                if (A.act == HG_ACT_ELU) d *= o > 0.f ? 1.f : o + 1.f;
                else if (A.act == HG_ACT_RELU) d = o > 0.f ? d : 0.f;
                d *= gscale;
                A.g[static_cast<int64_t>(i) * A.ldg + c] = d;
                s_g[w][c] = d;
                db[k] += d;
            }
        }
        __syncwarp();
        float gh[NE];
#pragma unroll
        for (int k = 0; k < NE; ++k) gh[k] = (lane + 32 * k) < HC ? s_g[w][ce[k]] * invH : 0.f;
        for (int r = 0; r < A.n_rel; ++r) {
            const hg_gat_train_rel& R = A.rel[r];
            const int beg = R.rowptr[i], end = R.rowptr[i + 1];
            float ad[kMaxH], t[kMaxH], dad[kMaxH];
#pragma unroll
            for (int h = 0; h < kMaxH; ++h) {
                ad[h] = h < H ? __ldg(R.a_dst + static_cast<int64_t>(i) * R.lda_dst + h) : 0.f;
                t[h] = 0.f;
                dad[h] = 0.f;
            }
            for (int p = beg; p < end; ++p) {
                const int j = __ldg(R.col + p);
                const float* xr = R.xs + static_cast<int64_t>(j) * R.ldxs;
                float part[kMaxH] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int k = 0; k < NE; ++k) {
                    const int e = lane + 32 * k;
                    if (e < HC) {
                        const float v = gh[k] * __ldg(xr + e);
#pragma unroll
                        for (int h = 0; h < kMaxH; ++h) part[h] += hk[k] == h ? v : 0.f;
                    }
                }
#pragma unroll
                for (int h = 0; h < kMaxH; ++h) {
                    if (h < H) {
                        const int64_t idx = static_cast<int64_t>(p) * H + h;
                        float da = warp_sum(part[h]);
                        if (drop) da *= keep_scale(A.seed, r, idx, p_drop, inv_keep);
                        t[h] = fmaf(R.alpha[idx], da, t[h]);
                        if (lane == 0) R.dlogit[idx] = da;
                    }
                }
            }
            __syncwarp();
            for (int p = beg + lane; p < end; p += 32) {
                const float* as = R.a_src + static_cast<int64_t>(__ldg(R.col + p)) * R.lda_src;
#pragma unroll
                for (int h = 0; h < kMaxH; ++h) {
                    if (h < H) {
                        const int64_t idx = static_cast<int64_t>(p) * H + h;
                        float ds = R.alpha[idx] * (R.dlogit[idx] - t[h]);
                        ds *= (__ldg(as + h) + ad[h]) > 0.f ? 1.f : slope;
                        R.dlogit[idx] = ds;
                        dad[h] += ds;
                    }
                }
            }
#pragma unroll
            for (int h = 0; h < kMaxH; ++h) {
                if (h < H) {
                    const float v = warp_sum(dad[h]);
                    if (lane == 0) R.da_dst[static_cast<int64_t>(i) * R.ldda_dst + h] = v;
                }
            }
            __syncwarp();
        }
    }
    if (A.dbias) {
#pragma unroll
        for (int k = 0; k < NE; ++k)
            if (lane + 32 * k < width) atomicAdd(&s_db[lane + 32 * k], db[k]);
        __syncthreads();
        for (int c = threadIdx.x; c < width; c += kThreads) atomicAdd(A.dbias + c, s_db[c]);
    }
}

// Pass 2 of the backward: by source row over the reversed CSR; blockIdx.y = relation slot.
template <int NE>
__global__ void __launch_bounds__(kThreads) k_bwd_src(const __grid_constant__ hg_gat_train_args A) {
    const int r = blockIdx.y;
    const hg_gat_train_rel& R = A.rel[r];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int H = A.H, C = A.C, HC = H * C;
    const float p_drop = A.dropout_p;
    const bool drop = p_drop > 0.f;
    const float inv_keep = drop ? 1.f / (1.f - p_drop) : 1.f;
    const float invH = A.concat ? 1.f : 1.f / static_cast<float>(H);
    int hk[NE], ce[NE];
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        const int e = lane + 32 * k;
        hk[k] = e < HC ? e / C : 0;
        ce[k] = A.concat ? e : e - hk[k] * C;
    }
    for (int j = blockIdx.x * kWarps + w; j < R.n_src; j += gridDim.x * kWarps) {
        float acc[NE];
#pragma unroll
        for (int k = 0; k < NE; ++k) acc[k] = 0.f;
        float das[kMaxH] = {0.f, 0.f, 0.f, 0.f};
        const int beg = R.rowptr_t[j], end = R.rowptr_t[j + 1];
        for (int q = beg; q < end; ++q) {
            const int i = __ldg(R.col_t + q);
            const int p = __ldg(R.pos_t + q);
            float wv[kMaxH];
#pragma unroll
            for (int h = 0; h < kMaxH; ++h) {
                wv[h] = 0.f;
                if (h < H) {
                    const int64_t idx = static_cast<int64_t>(p) * H + h;
                    wv[h] = R.alpha[idx] * (drop ? keep_scale(A.seed, r, idx, p_drop, inv_keep) : 1.f) * invH;
                    das[h] += R.dlogit[idx];
                }
            }
            const float* gr = A.g + static_cast<int64_t>(i) * A.ldg;
#pragma unroll
            for (int k = 0; k < NE; ++k)
                if (lane + 32 * k < HC) acc[k] = fmaf(sel(wv, hk[k]), __ldg(gr + ce[k]), acc[k]);
        }
        float* dx = R.dxs + static_cast<int64_t>(j) * R.lddxs;
#pragma unroll
        for (int k = 0; k < NE; ++k)
            if (lane + 32 * k < HC) dx[lane + 32 * k] = acc[k];
        if (lane < H) R.da_src[static_cast<int64_t>(j) * R.ldda_src + lane] = sel(das, lane);
    }
}

__global__ void k_map_a(const int32_t* __restrict__ perm, const int32_t* __restrict__ col, const int32_t* n_live,
                        int64_t n_edges, int32_t* __restrict__ scratch) {
    const int64_t n = *n_live;
    for (int64_t p = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; p < n;
         p += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int32_t e = perm[p];
        scratch[e >= 0 ? e : n_edges + col[p]] = static_cast<int32_t>(p);
    }
}

__global__ void k_map_b(const int32_t* __restrict__ perm_t, const int32_t* __restrict__ col_t, const int32_t* n_live_t,
                        int64_t n_edges, const int32_t* __restrict__ scratch, int32_t* __restrict__ pos_t) {
    const int64_t n = *n_live_t;
    for (int64_t q = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; q < n;
         q += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int32_t e = perm_t[q];
        pos_t[q] = scratch[e >= 0 ? e : n_edges + col_t[q]];
    }
}

int validate(const hg_gat_train_args* a, bool bwd) {
    if (!a || a->n_rel < 1 || a->n_rel > HG_MAX_RELATIONS || a->n_dst < 0 || a->H < 1 || a->C < 1) return HG_EINVAL;
    if (a->H > kMaxH || a->H * a->C > 256) return HG_EUNSUPPORTED;
    if (!(a->dropout_p >= 0.f && a->dropout_p < 1.f)) return HG_EINVAL;
    if (a->n_dst > 0 && !a->out) return HG_EINVAL;
    for (int r = 0; r < a->n_rel; ++r) {
        const hg_gat_train_rel& R = a->rel[r];
        if (!R.rowptr || (!R.a_dst && a->n_dst > 0)) return HG_EINVAL;
        if (bwd && (!R.rowptr_t || R.n_src < 0)) return HG_EINVAL;
    }
    if (bwd && a->n_dst > 0 && (!a->dout || !a->g)) return HG_EINVAL;
    return HG_OK;
}

unsigned row_grid(int64_t rows, int ctas_per_sm) {
    const int sms = sm_count() > 0 ? sm_count() : 148;
    int64_t g = ceil_div<int64_t>(rows > 0 ? rows : 1, kWarps);
    const int64_t cap = static_cast<int64_t>(sms) * ctas_per_sm;
    return static_cast<unsigned>(g < cap ? g : cap);
}

#define HG_NE_SWITCH(ne, CALL)          \
    switch (ne) {                       \
        case 1: { CALL(1); break; }     \
        case 2: { CALL(2); break; }     \
        case 3: { CALL(3); break; }     \
        case 4: { CALL(4); break; }     \
        case 5: case 6: { CALL(6); break; } \
        default: { CALL(8); break; }    \
    }

}  // namespace
}  // namespace hg

extern "C" int hg_gat_train_fwd(const hg_gat_train_args* a, void* stream_) {
    using namespace hg;
    int rc = validate(a, false);
    if (rc != HG_OK) return rc;
    if (a->n_dst == 0) return HG_OK;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int ne = ceil_div(a->H * a->C, 32);
    const unsigned grid = row_grid(a->n_dst, 64);
#define CALL(NE) k_fwd<NE><<<grid, kThreads, 0, stream>>>(*a)
    HG_NE_SWITCH(ne, CALL)
#undef CALL
    return check_launch();
}

extern "C" int hg_gat_train_bwd(const hg_gat_train_args* a, void* stream_) {
    using namespace hg;
    int rc = validate(a, true);
    if (rc != HG_OK) return rc;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int ne = ceil_div(a->H * a->C, 32);
    if (a->n_dst > 0) {
        const unsigned grid = row_grid(a->n_dst, 8);  // grid-stride: the bias gradient is reduced per block
#define CALL(NE) k_bwd_dst<NE><<<grid, kThreads, 0, stream>>>(*a)
        HG_NE_SWITCH(ne, CALL)
#undef CALL
        rc = check_launch();
        if (rc != HG_OK) return rc;
    }
    int32_t max_src = 0;
    for (int r = 0; r < a->n_rel; ++r) max_src = a->rel[r].n_src > max_src ? a->rel[r].n_src : max_src;
    if (max_src > 0) {
        const dim3 grid(row_grid(max_src, 64), a->n_rel);
#define CALL(NE) k_bwd_src<NE><<<grid, kThreads, 0, stream>>>(*a)
        HG_NE_SWITCH(ne, CALL)
#undef CALL
        rc = check_launch();
    }
    return rc;
}

extern "C" int hg_csr_transpose_map(const int32_t* perm, const int32_t* col, const int32_t* n_live, int64_t cap,
                                    const int32_t* perm_t, const int32_t* col_t, const int32_t* n_live_t,
                                    int64_t n_edges, int32_t n_loop_nodes, int32_t* scratch, int32_t* pos_t,
                                    void* stream_) {
    using namespace hg;
    if (cap < 0 || n_edges < 0 || n_loop_nodes < 0) return HG_EINVAL;
    if (cap == 0) return HG_OK;
    if (!perm || !col || !n_live || !perm_t || !col_t || !n_live_t || !scratch || !pos_t) return HG_EINVAL;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int sms = sm_count() > 0 ? sm_count() : 148;
    int64_t g = ceil_div<int64_t>(cap, 256);
    const unsigned grid = static_cast<unsigned>(g < sms * 16 ? g : sms * 16);
    k_map_a<<<grid, 256, 0, stream>>>(perm, col, n_live, n_edges, scratch);
    k_map_b<<<grid, 256, 0, stream>>>(perm_t, col_t, n_live_t, n_edges, scratch, pos_t);
    return check_launch();
}
