// K3 + K4 -- fused GAT for one destination node type.
//
// One warp owns one destination row and walks every relation that targets the row's type:
//   phase 1  per head: row max and sum(exp) of the leaky-relu'd edge scores (lanes = edges),
//   phase 2  32 edges at a time: lanes turn scores into softmax weights and stage
//            (neighbour id, weight per head) in shared memory; then all lanes sweep the staged
//            edges, each lane gathering its K vector chunks of xs[j] (128-bit loads when the
//            width allows, U edges in flight) and accumulating in registers,
//   epilogue head concat / head mean (through a shared-memory row) + bias_r, added to the
//            cross-relation total; after the last relation: group mean, ELU/ReLU, one store.
// `out` is written exactly once per row: no atomics, no read-modify-write across relations, no
// separate softmax / stack+sum / activation passes.  HBM/L2-bound (gathers of contiguous graphs).
#include "hg_common.cuh"

namespace hg {
namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / kWarp;
constexpr int kMaxHeads = 8;

struct GatParams {
    hg_gat_args a;
};

__device__ __forceinline__ float leaky(float v, float slope) { return v > 0.f ? v : v * slope; }

template <int VEC, int K>
__global__ void __launch_bounds__(kThreads) k_gat_fused(const __grid_constant__ GatParams P) {
    constexpr int U = (K * VEC >= 16) ? 1 : (K * VEC >= 8 ? 2 : 4);
    __shared__ int32_t s_col[kWarps][kWarp];
    __shared__ float s_w[kWarps][kMaxHeads][kWarp + 1];  // +1: lanes reading different heads hit different banks
    __shared__ float s_m[kWarps][kMaxHeads];
    __shared__ float s_l[kWarps][kMaxHeads];
    extern __shared__ float s_stage[];  // [kWarps][F], only used for the head mean

    const hg_gat_args& A = P.a;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int H = A.H, C = A.C, F = H * C;
    const int nc = F / VEC;
    const int nco = (A.concat ? F : C) / VEC;  // output vector chunks
    float* stage = s_stage + warp * F;

    int hk[K];
    bool valid[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
        const int chunk = lane + 32 * k;
        valid[k] = chunk < nc;
        hk[k] = valid[k] ? (chunk * VEC) / C : 0;
    }

    for (int64_t i = static_cast<int64_t>(blockIdx.x) * kWarps + warp; i < A.n_dst;
         i += static_cast<int64_t>(gridDim.x) * kWarps) {
        float tot[K][VEC];
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int v = 0; v < VEC; ++v) tot[k][v] = 0.f;

        for (int r = 0; r < A.n_rel; ++r) {
            const hg_gat_rel& R = A.rel[r];
            const int32_t beg = __ldg(R.rowptr + i), end = __ldg(R.rowptr + i + 1);
            float acc[K][VEC];
#pragma unroll
            for (int k = 0; k < K; ++k)
#pragma unroll
                for (int v = 0; v < VEC; ++v) acc[k][v] = 0.f;

            if (end > beg) {
                // ---- phase 1: per-head max and sum of exp over the whole row
                for (int h = 0; h < H; ++h) {
                    const float ad = __ldg(R.a_dst + i * R.lda_dst + h);
                    float m = -INFINITY;
                    for (int32_t c0 = beg; c0 < end; c0 += 32) {
                        const int32_t e = c0 + lane;
                        if (e < end) {
                            const int32_t j = __ldg(R.col + e);
                            float s = __ldg(R.a_src + static_cast<int64_t>(j) * R.lda_src + h) + ad;
                            if (R.a_edge) s += __ldg(R.a_edge + static_cast<int64_t>(e) * H + h);
                            m = fmaxf(m, leaky(s, A.neg_slope));
                        }
                    }
                    m = warp_max(m);
                    float l = 0.f;
                    for (int32_t c0 = beg; c0 < end; c0 += 32) {
                        const int32_t e = c0 + lane;
                        if (e < end) {
                            const int32_t j = __ldg(R.col + e);
                            float s = __ldg(R.a_src + static_cast<int64_t>(j) * R.lda_src + h) + ad;
                            if (R.a_edge) s += __ldg(R.a_edge + static_cast<int64_t>(e) * H + h);
                            l += expf(leaky(s, A.neg_slope) - m);
                        }
                    }
                    l = warp_sum(l);
                    if (lane == 0) {
                        s_m[warp][h] = m;
                        s_l[warp][h] = l + 1e-16f;
                    }
                }
                __syncwarp();

                // ---- phase 2: stage 32 edges (id + softmax weights), then gather-accumulate
                for (int32_t c0 = beg; c0 < end; c0 += 32) {
                    const int n = min(32, end - c0);
                    if (lane < n) {
                        const int32_t e = c0 + lane;
                        const int32_t j = __ldg(R.col + e);
                        s_col[warp][lane] = j;
                        for (int h = 0; h < H; ++h) {
                            float s = __ldg(R.a_src + static_cast<int64_t>(j) * R.lda_src + h) +
                                      __ldg(R.a_dst + i * R.lda_dst + h);
                            if (R.a_edge) s += __ldg(R.a_edge + static_cast<int64_t>(e) * H + h);
                            const float w = expf(leaky(s, A.neg_slope) - s_m[warp][h]) / s_l[warp][h];
                            s_w[warp][h][lane] = w;
                            if (A.alpha_saved[r]) A.alpha_saved[r][static_cast<int64_t>(e) * H + h] = w;
                        }
                    }
                    __syncwarp();
                    for (int t = 0; t < n; t += U) {
                        float val[U][K][VEC];
                        float w[U][K];
#pragma unroll
                        for (int u = 0; u < U; ++u) {
                            const bool live = t + u < n;
                            const int32_t j = live ? s_col[warp][t + u] : 0;
                            const float* __restrict__ row = R.xs + static_cast<int64_t>(j) * R.ldxs;
#pragma unroll
                            for (int k = 0; k < K; ++k) {
                                if (live && valid[k]) {
                                    ldg_vec<VEC>(row + (lane + 32 * k) * VEC, val[u][k]);
                                    w[u][k] = s_w[warp][hk[k]][t + u];
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
                }
            }

            // ---- per-relation epilogue: concat | head mean, + bias_r, into the relation total
            if (A.concat || H == 1) {
#pragma unroll
                for (int k = 0; k < K; ++k) {
                    if (!valid[k]) continue;
                    float bv[VEC];
#pragma unroll
                    for (int v = 0; v < VEC; ++v) bv[v] = 0.f;
                    if (R.bias) ldg_vec<VEC>(R.bias + (lane + 32 * k) * VEC, bv);
#pragma unroll
                    for (int v = 0; v < VEC; ++v) tot[k][v] += acc[k][v] + bv[v];
                }
            } else {
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
                    for (int h = 0; h < H; ++h)
#pragma unroll
                        for (int v = 0; v < VEC; ++v) sum[v] += stage[h * C + chunk * VEC + v];
                    float bv[VEC];
#pragma unroll
                    for (int v = 0; v < VEC; ++v) bv[v] = 0.f;
                    if (R.bias) ldg_vec<VEC>(R.bias + chunk * VEC, bv);
#pragma unroll
                    for (int v = 0; v < VEC; ++v) tot[k][v] += sum[v] * inv_h + bv[v];
                }
                __syncwarp();
            }
        }

        // ---- cross-relation group + activation, single store
        const float gscale = (A.group == HG_GROUP_MEAN) ? 1.f / static_cast<float>(A.n_rel) : 1.f;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int chunk = lane + 32 * k;
            if (chunk >= nco) continue;
            float o[VEC];
#pragma unroll
            for (int v = 0; v < VEC; ++v) o[v] = apply_act(tot[k][v] * gscale, A.act);
            st_vec<VEC>(A.out + i * A.ldout + chunk * VEC, o);
        }
    }
}

template <int VEC, int K>
int launch(const GatParams& P, cudaStream_t stream) {
    const hg_gat_args& A = P.a;
    const size_t smem = (A.concat || A.H == 1) ? 0 : sizeof(float) * kWarps * A.H * A.C;
    if (smem > 24 * 1024) {  // static staging (~11 KB) + dynamic must stay under the opt-in threshold
        HG_CUDA_TRY(cudaFuncSetAttribute(k_gat_fused<VEC, K>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem)));
    }
    const int64_t need = ceil_div<int64_t>(A.n_dst, kWarps);
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int64_t cap = static_cast<int64_t>(sms) * 8 * 16;
    const unsigned grid = static_cast<unsigned>(need < cap ? need : cap);
    k_gat_fused<VEC, K><<<grid, kThreads, smem, stream>>>(P);
    return check_launch();
}

template <int VEC>
int dispatch(const GatParams& P, int nc, cudaStream_t stream) {
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
        if (!R.xs || !R.a_src || !R.a_dst || !R.rowptr || !R.col) return HG_EINVAL;
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
    const int nc = F / vec;
    return vec == 4 ? dispatch<4>(P, nc, stream) : (vec == 2 ? dispatch<2>(P, nc, stream) : dispatch<1>(P, nc, stream));
}
