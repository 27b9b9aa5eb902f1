// K2 / K5 -- segmented reduction over CSR rows (sum / mean / max).
//
// One group of G lanes (G = 4..32, a whole warp for rows >= 32 vector chunks) owns one
// destination row; each lane keeps K vector chunks (VEC = 1/2/4 floats, 128-bit loads when
// the row width allows) of the row in registers.  Neighbour ids (and their scale factors) are
// staged through shared memory G at a time, then the gathers of U consecutive neighbours are
// issued back to back so each lane has U*K independent loads in flight.  HBM/L2-bound: rows of
// a scene graph are contiguous in the batch, so repeated gathers of a source row hit L2.
#include "hg_common.cuh"

namespace hg {
namespace {

constexpr int kThreads = 256;

struct SegParams {
    hg_seg_args a;
    int32_t f_begin;  // first feature handled by this launch (feature-block loop for very wide rows)
    int32_t f_count;  // features handled by this launch (multiple of VEC)
};

template <int VEC, int G, int K>
__global__ void __launch_bounds__(kThreads) k_seg_reduce(const __grid_constant__ SegParams P) {
    constexpr int kGroups = kThreads / G;
    constexpr int U = (K * VEC >= 16) ? 1 : (K * VEC >= 8 ? 2 : 4);
    __shared__ int32_t s_col[kThreads];
    __shared__ float s_w[kThreads];

    const hg_seg_args& A = P.a;
    const int tid = threadIdx.x;
    const int g = tid % G;
    const int grp = tid / G;
    const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << (((tid & 31) / G) * G));
    int32_t* my_col = s_col + grp * G;
    float* my_w = s_w + grp * G;
    const int nc = P.f_count / VEC;  // vector chunks in this feature block
    const float* __restrict__ x = A.x + P.f_begin;
    const bool is_max = A.op == HG_REDUCE_MAX;

    for (int64_t r = static_cast<int64_t>(blockIdx.x) * kGroups + grp; r < A.n_out;
         r += static_cast<int64_t>(gridDim.x) * kGroups) {
        const int32_t i = A.row_map ? __ldg(A.row_map + r) : static_cast<int32_t>(r);
        const int32_t beg = __ldg(A.rowptr + i), end = __ldg(A.rowptr + i + 1);
        float acc[K][VEC];
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int v = 0; v < VEC; ++v) acc[k][v] = is_max ? -INFINITY : 0.f;

        for (int32_t c0 = beg; c0 < end; c0 += G) {
            const int n = min(G, end - c0);
            if (g < n) {
                const int32_t j = __ldg(A.col + c0 + g);
                my_col[g] = j;
                my_w[g] = A.src_scale ? __ldg(A.src_scale + j) : 1.f;
            }
            __syncwarp(gmask);
            for (int t = 0; t < n; t += U) {
                float val[U][K][VEC];
                float w[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const bool live = t + u < n;
                    const int32_t j = live ? my_col[t + u] : 0;
                    w[u] = live ? my_w[t + u] : 0.f;
                    const float* __restrict__ row = x + static_cast<int64_t>(j) * A.ldx;
#pragma unroll
                    for (int k = 0; k < K; ++k) {
                        const int chunk = g + G * k;
                        if (live && chunk < nc) {
                            ldg_vec<VEC>(row + chunk * VEC, val[u][k]);
                        } else {
#pragma unroll
                            for (int v = 0; v < VEC; ++v) val[u][k][v] = is_max ? -INFINITY : 0.f;
                        }
                    }
                }
#pragma unroll
                for (int u = 0; u < U; ++u)
#pragma unroll
                    for (int k = 0; k < K; ++k)
#pragma unroll
                        for (int v = 0; v < VEC; ++v) {
                            if (is_max) {
                                if (t + u < n) acc[k][v] = fmaxf(acc[k][v], w[u] * val[u][k][v]);
                            } else {
                                acc[k][v] = fmaf(w[u], val[u][k][v], acc[k][v]);
                            }
                        }
            }
            __syncwarp(gmask);
        }

        const int32_t deg = end - beg;
        const float dscale = A.dst_scale ? __ldg(A.dst_scale + i) : 1.f;
        float coef = 0.f;
        if (A.x_self) coef = A.self_coef_add + (A.self_coef_ptr ? __ldg(A.self_coef_ptr) : 0.f);
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int chunk = g + G * k;
            if (chunk >= nc) continue;
            const int f = P.f_begin + chunk * VEC;
            float o[VEC];
#pragma unroll
            for (int v = 0; v < VEC; ++v) {
                float a = acc[k][v];
                if (deg == 0) a = 0.f;  // PyG: empty segments reduce to 0 (also for max)
                if (A.op == HG_REDUCE_MEAN && deg > 0) a = a / static_cast<float>(deg);
                o[v] = a * dscale;
            }
            if (A.x_self) {
                float sv[VEC];
                ldg_vec<VEC>(A.x_self + static_cast<int64_t>(i) * A.ldself + f, sv);
#pragma unroll
                for (int v = 0; v < VEC; ++v) o[v] = fmaf(coef, sv[v], o[v]);
            }
            if (A.bias) {
                float bv[VEC];
                ldg_vec<VEC>(A.bias + f, bv);
#pragma unroll
                for (int v = 0; v < VEC; ++v) o[v] += bv[v];
            }
#pragma unroll
            for (int v = 0; v < VEC; ++v) o[v] = apply_act(o[v], A.act);
            st_vec<VEC>(A.out + r * A.ldout + f, o);
        }
    }
}

template <int VEC, int G, int K>
int launch(const SegParams& P, cudaStream_t stream) {
    constexpr int kGroups = kThreads / G;
    const int64_t need = ceil_div<int64_t>(P.a.n_out, kGroups);
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int64_t cap = static_cast<int64_t>(sms) * 8 * 16;  // 16 waves of 148 SMs x 8 resident CTAs
    const unsigned grid = static_cast<unsigned>(need < cap ? need : cap);
    k_seg_reduce<VEC, G, K><<<grid, kThreads, 0, stream>>>(P);
    return check_launch();
}

template <int VEC>
int dispatch(const SegParams& P, cudaStream_t stream) {
    const int nc = P.f_count / VEC;
    if (nc <= 4) return launch<VEC, 4, 1>(P, stream);
    if (nc <= 8) return launch<VEC, 8, 1>(P, stream);
    if (nc <= 16) return launch<VEC, 16, 1>(P, stream);
    if (nc <= 32) return launch<VEC, 32, 1>(P, stream);
    if (nc <= 64) return launch<VEC, 32, 2>(P, stream);
    if (nc <= 96) return launch<VEC, 32, 3>(P, stream);
    if (nc <= 128) return launch<VEC, 32, 4>(P, stream);
    if (nc <= 192) return launch<VEC, 32, 6>(P, stream);
    return launch<VEC, 32, 8>(P, stream);
}

}  // namespace
}  // namespace hg

extern "C" int hg_seg_reduce(const hg_seg_args* a, void* stream_) {
    using namespace hg;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!a) return HG_EINVAL;
    if (a->n_out < 0 || a->F < 0) return HG_EINVAL;
    if (a->n_out == 0 || a->F == 0) return HG_OK;
    // x may be NULL only for an empty source set (n_src == 0): every row is then empty and x is never read
    if (!a->rowptr || !a->out) return HG_EINVAL;
    if (a->op < HG_REDUCE_SUM || a->op > HG_REDUCE_MAX) return HG_EINVAL;
    if (a->act < HG_ACT_NONE || a->act > HG_ACT_ELU) return HG_EINVAL;

    // Widest vector the row width, leading dimensions and base pointers all allow.
    auto ok = [&](int vec) {
        const size_t bytes = sizeof(float) * vec;
        if (a->F % vec) return false;
        if (a->ldx % vec || a->ldout % vec || !aligned(a->x, bytes) || !aligned(a->out, bytes)) return false;
        if (a->x_self && (a->ldself % vec || !aligned(a->x_self, bytes))) return false;
        if (a->bias && !aligned(a->bias, bytes)) return false;
        return true;
    };
    const int vec = ok(4) ? 4 : (ok(2) ? 2 : 1);
    const int block = 32 * 8 * vec;  // widest feature block one launch covers
    for (int f0 = 0; f0 < a->F; f0 += block) {
        SegParams P;
        P.a = *a;
        P.f_begin = f0;
        P.f_count = (a->F - f0 < block) ? a->F - f0 : block;
        int rc = vec == 4 ? dispatch<4>(P, stream) : (vec == 2 ? dispatch<2>(P, stream) : dispatch<1>(P, stream));
        if (rc != HG_OK) return rc;
    }
    return HG_OK;
}
