// K2 / K5 -- segmented reduction over CSR rows (sum / mean / max).
//
// Two kernels.  k_seg_tile (the common case: output row r = CSR row r, rows up to 160 vector chunks):
// a 128-thread CTA owns 32 consecutive destination rows; the neighbour ids (and source scales) of the
// whole tile are staged in shared memory by threads = edges (coalesced), then threads = (row, chunk)
// gather with one IMAD.WIDE.U32 per address and 4 neighbours in flight per thread, and the epilogue
// (mean, destination scale, self term, bias, activation) is applied in registers before ONE store per
// chunk.  Tiles whose edge list does not fit the staging buffer (hub rows) take a warp-per-row path.
// k_seg_reduce (row_map readouts and very wide rows): one group of G lanes (G = 4..32, a whole warp for rows >= 32 vector chunks) owns one
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

// ------------------------------------------------------------------------------ tile kernel
#ifndef HG_SEG_KC2
#define HG_SEG_KC2 0
#endif
#ifndef HG_SEG_OCC
#define HG_SEG_OCC 8
#endif
constexpr int kTileThreads = 128;
constexpr int kTileRows = 32;
constexpr int kTileCap = 1024;  // staged edges per tile (8 KB of shared memory)

template <int VEC>
__device__ __forceinline__ void seg_finish(const hg_seg_args& A, float (&acc)[VEC], int64_t i, int f, int32_t deg, float coef) {
    const float dscale = A.dst_scale ? __ldg(A.dst_scale + i) : 1.f;
    float o[VEC];
#pragma unroll
    for (int v = 0; v < VEC; ++v) {
        float a = acc[v];
        if (deg == 0) a = 0.f;  // PyG: empty segments reduce to 0 (also for max)
        if (A.op == HG_REDUCE_MEAN && deg > 0) a = a / static_cast<float>(deg);
        o[v] = a * dscale;
    }
    if (A.x_self) {
        float sv[VEC];
        ldg_vec<VEC>(A.x_self + i * A.ldself + f, sv);
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
    st_vec<VEC>(A.out + i * A.ldout + f, o);
}

// MAXOP: max reduction (else sum / mean); SCALED: per-source scale factors (GCN normalisation)
template <int VEC, int TPR, int KC, bool MAXOP, bool SCALED>
__global__ void __launch_bounds__(kTileThreads, HG_SEG_OCC) k_seg_tile(const __grid_constant__ SegParams P) {
    __shared__ int32_t s_col[kTileCap];
    __shared__ float s_w[SCALED ? kTileCap : 1];
    __shared__ int32_t s_rp[kTileRows + 1];
    const hg_seg_args& A = P.a;
    const int tid = threadIdx.x;
    constexpr int kGroups = kTileThreads / TPR;
    const int q = tid % TPR, grp = tid / TPR;
    const int nc = P.f_count / VEC;
    const int f0 = P.f_begin + q * VEC;
    const uint32_t ldb = static_cast<uint32_t>(A.ldx) * 4u;  // row pitch in bytes (host checks < 4 GB)
    const char* xb = reinterpret_cast<const char*>(A.x + f0);
    bool live[KC];
#pragma unroll
    for (int k = 0; k < KC; ++k) live[k] = q + k * TPR < nc;
    float coef = 0.f;
    if (A.x_self) coef = A.self_coef_add + (A.self_coef_ptr ? __ldg(A.self_coef_ptr) : 0.f);
    const int64_t n_tiles = ceil_div<int64_t>(A.n_out, kTileRows);

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t t0 = tile * kTileRows;
        const int rows = (A.n_out - t0 < kTileRows) ? static_cast<int>(A.n_out - t0) : kTileRows;
        __syncthreads();  // previous tile fully consumed
        if (tid <= kTileRows) s_rp[tid] = __ldg(A.rowptr + t0 + min(tid, rows));
        const int32_t e0 = __ldg(A.rowptr + t0), ne = __ldg(A.rowptr + t0 + rows) - e0;  // uniform
        if (ne > kTileCap) {  // hub tile: warp per row, straight from global memory
            for (int t = tid >> 5; t < rows; t += kTileThreads / 32) {
                const int64_t i = t0 + t;
                const int32_t beg = __ldg(A.rowptr + i), end = __ldg(A.rowptr + i + 1);
                for (int c = tid & 31; c < nc; c += 32) {
                    float acc[VEC];
#pragma unroll
                    for (int v = 0; v < VEC; ++v) acc[v] = MAXOP ? -INFINITY : 0.f;
                    for (int32_t e = beg; e < end; ++e) {
                        const int32_t j = __ldg(A.col + e);
                        const float w = SCALED ? __ldg(A.src_scale + j) : 1.f;
                        float val[VEC];
                        ldg_vec<VEC>(A.x + static_cast<int64_t>(j) * A.ldx + P.f_begin + c * VEC, val);
#pragma unroll
                        for (int v = 0; v < VEC; ++v) acc[v] = MAXOP ? fmaxf(acc[v], w * val[v]) : fmaf(w, val[v], acc[v]);
                    }
                    seg_finish<VEC>(A, acc, i, P.f_begin + c * VEC, end - beg, coef);
                }
            }
            continue;
        }
        for (int e = tid; e < ne; e += kTileThreads) {  // threads = edges: coalesced
            const int32_t j = __ldg(A.col + e0 + e);
            s_col[e] = j;
            if (SCALED) s_w[e] = __ldg(A.src_scale + j);
        }
        __syncthreads();

        if (KC == 1 && !live[0]) continue;  // no barrier below this point
        for (int t = grp; t < rows; t += kGroups) {
            const int beg = s_rp[t] - e0, end = s_rp[t + 1] - e0;
            float acc[KC][VEC];
#pragma unroll
            for (int k = 0; k < KC; ++k)
#pragma unroll
                for (int v = 0; v < VEC; ++v) acc[k][v] = MAXOP ? -INFINITY : 0.f;
#pragma unroll(KC <= 2 ? 4 : 2)
            for (int e = beg; e < end; ++e) {
                const uint32_t j = static_cast<uint32_t>(s_col[e]);
                const float w = SCALED ? s_w[e] : 1.f;
                const float* row = reinterpret_cast<const float*>(xb + static_cast<uint64_t>(j) * ldb);
                float val[KC][VEC];
#pragma unroll
                for (int k = 0; k < KC; ++k) {
                    if (KC == 1 || live[k]) {
                        ldg_vec<VEC>(row + k * TPR * VEC, val[k]);
                    } else {
#pragma unroll
                        for (int v = 0; v < VEC; ++v) val[k][v] = MAXOP ? -INFINITY : 0.f;
                    }
                }
#pragma unroll
                for (int k = 0; k < KC; ++k)
#pragma unroll
                    for (int v = 0; v < VEC; ++v) {
                        if (MAXOP) acc[k][v] = fmaxf(acc[k][v], SCALED ? w * val[k][v] : val[k][v]);
                        else acc[k][v] = SCALED ? fmaf(w, val[k][v], acc[k][v]) : acc[k][v] + val[k][v];
                    }
            }
#pragma unroll
            for (int k = 0; k < KC; ++k)
                if (KC == 1 || live[k]) seg_finish<VEC>(A, acc[k], t0 + t, f0 + k * TPR * VEC, end - beg, coef);
        }
    }
}

template <int VEC, int TPR, int KC>
int launch_tile(const SegParams& P, cudaStream_t stream) {
    const int64_t need = ceil_div<int64_t>(P.a.n_out, kTileRows);
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int64_t cap = static_cast<int64_t>(sms) * 8 * 8;  // 8 waves of 148 SMs x 8 resident CTAs
    const unsigned grid = static_cast<unsigned>(need < cap ? need : cap);
    const bool mx = P.a.op == HG_REDUCE_MAX, sc = P.a.src_scale != nullptr;
    if (mx && sc) k_seg_tile<VEC, TPR, KC, true, true><<<grid, kTileThreads, 0, stream>>>(P);
    else if (mx) k_seg_tile<VEC, TPR, KC, true, false><<<grid, kTileThreads, 0, stream>>>(P);
    else if (sc) k_seg_tile<VEC, TPR, KC, false, true><<<grid, kTileThreads, 0, stream>>>(P);
    else k_seg_tile<VEC, TPR, KC, false, false><<<grid, kTileThreads, 0, stream>>>(P);
    return check_launch();
}

template <int VEC>
int dispatch_tile(const SegParams& P, cudaStream_t stream) {
    const int nc = P.f_count / VEC;
#if HG_SEG_KC2
    if (nc <= 4) return launch_tile<VEC, 4, 1>(P, stream);
    if (nc <= 8) return launch_tile<VEC, 4, 2>(P, stream);
    if (nc <= 16) return launch_tile<VEC, 8, 2>(P, stream);
    if (nc <= 32) return launch_tile<VEC, 16, 2>(P, stream);
#else
    if (nc <= 4) return launch_tile<VEC, 4, 1>(P, stream);
    if (nc <= 8) return launch_tile<VEC, 8, 1>(P, stream);
    if (nc <= 16) return launch_tile<VEC, 16, 1>(P, stream);
    if (nc <= 32) return launch_tile<VEC, 32, 1>(P, stream);
#endif
    if (nc <= 64) return launch_tile<VEC, 32, 2>(P, stream);
    if (nc <= 96) return launch_tile<VEC, 32, 3>(P, stream);
    if (nc <= 128) return launch_tile<VEC, 32, 4>(P, stream);
    return launch_tile<VEC, 32, 5>(P, stream);  // nc <= 160 (checked by the caller)
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
    if (!a->row_map && a->F / vec <= 160 && a->ldx <= (INT32_MAX >> 2)) {  // tile kernel: out row r = CSR row r
        SegParams P;
        P.a = *a;
        P.f_begin = 0;
        P.f_count = a->F;
        return vec == 4 ? dispatch_tile<4>(P, stream) : (vec == 2 ? dispatch_tile<2>(P, stream) : dispatch_tile<1>(P, stream));
    }
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
