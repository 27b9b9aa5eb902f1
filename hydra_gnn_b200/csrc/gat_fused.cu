// K3 + K4 -- fused GAT for one destination node type (v5).
//
// A CTA (128 threads, up to 6 resident per SM) owns a tile of 16-32 consecutive destination rows
// (consecutive rows belong to the same scene graph, so their neighbour sets overlap and the gathers
// of a tile hit L1/L2).  Per tile:
//   stage    rowptr slices and destination logits; then -- threads = edges, coalesced on col -- the
//            neighbour ids, the tile-local row of every edge and the raw logits a_src[j,h](+a_edge)
//            of every relation go to shared memory;
//   softmax  threads = (relation, row, head): max and 1/(sum + 1e-16) of the row; then threads =
//            edges turn the staged logits into softmax weights in place.  No shuffles, no idle lanes
//            on short rows (mean in-degree of a scene graph is ~5), no runtime integer divisions
//            (all index maps are shifts and masks over padded layouts);
//   gather   threads = (row, output chunk): TPR threads cooperate on a row, each owning KC vector
//            chunks of the OUTPUT row.  For head-mean the thread walks the H head slices of xs[j]
//            itself, so the head reduction never leaves registers; the weights of an edge are one
//            broadcast LDS.128.  Relations are summed in the same registers;
//   out      + summed biases (+ addend), group mean, ELU/ReLU, ONE vector store per chunk.
// No atomics, no read-modify-write across relations, no separate softmax / stack+sum / activation
// passes.  Tiles whose edges do not fit the staging buffers (hub nodes) take a warp-per-row path
// that streams from global memory.
//
// src_shared = 1 is the aggregate-first form: xs is [n_src, C] (ONE row shared by all heads) and the
// output is [n_dst, H*C] with out[i, h, :] = sum_e w_eh xs[j, :]; the projection is applied afterwards
// to the aggregated rows (used when a relation has far fewer destinations than sources).
#include "hg_common.cuh"

namespace hg {
namespace {

constexpr int kThreads = 128;
constexpr int kWarps = kThreads / kWarp;
constexpr int kMaxHeads = 8;
constexpr int kMaxTile = 32;  // rows per tile: 16 (32 when a row needs only 4 threads)

struct GatParams {
    hg_gat_args a;
    int32_t cap;        // staged edges per relation per tile (multiple of 32)
    int32_t tile_rows;  // 16 or 32
    int32_t hs;         // padded head stride in shared memory: 4 or 8
    int32_t f_begin;    // first OUTPUT column of this launch (wide rows are split over launches)
    int32_t f_count;    // output columns of this launch
};

__device__ __forceinline__ float leaky(float v, float slope) { return v > 0.f ? v : v * slope; }

struct Smem {
    float* w;        // [R][cap][hs]        logits -> weights (16-byte aligned rows)
    float* ad;       // [R][kMaxTile][hs]   destination logits
    float* mx;       // [R][kMaxTile][hs]   row max
    float* inv;      // [R][kMaxTile][hs]   1 / (row sum + 1e-16)
    float* bias;     // [f_count]           summed biases of this launch's output columns
    int32_t* col;    // [R][cap]
    int32_t* rp;     // [R][kMaxTile + 1]
    uint8_t* erow;   // [R][cap]            tile-local row of the edge
};

__host__ __device__ inline size_t smem_bytes(int R, int cap, int hs, int f_count) {
    size_t words = static_cast<size_t>(R) * cap * hs + 3 * static_cast<size_t>(R) * kMaxTile * hs +
                   static_cast<size_t>((f_count + 3) / 4 * 4) + static_cast<size_t>(R) * cap +
                   static_cast<size_t>(R) * (kMaxTile + 1);
    return words * 4 + static_cast<size_t>(R) * cap + 16;
}

__device__ __forceinline__ Smem carve(uint8_t* raw, int R, int cap, int hs, int f_count) {
    Smem s;
    float* p = reinterpret_cast<float*>(raw);
    s.w = p;      p += static_cast<size_t>(R) * cap * hs;
    s.ad = p;     p += R * kMaxTile * hs;
    s.mx = p;     p += R * kMaxTile * hs;
    s.inv = p;    p += R * kMaxTile * hs;
    s.bias = p;   p += (f_count + 3) / 4 * 4;
    s.col = reinterpret_cast<int32_t*>(p);  p += static_cast<size_t>(R) * cap;
    s.rp = reinterpret_cast<int32_t*>(p);   p += R * (kMaxTile + 1);
    s.erow = reinterpret_cast<uint8_t*>(p);
    return s;
}

// Hub path: one warp per row, lanes = output chunks, everything streamed from global memory.
template <int VEC>
__device__ __noinline__ void slow_row(const GatParams& P, int64_t i, int lane) {
    const hg_gat_args& A = P.a;
    const int H = A.H, C = A.C, R = A.n_rel;
    const bool head_mean = !A.concat && !A.src_shared;
    const int nco = P.f_count / VEC;
    const float gscale = (A.group == HG_GROUP_MEAN) ? 1.f / static_cast<float>(R) : 1.f;
    for (int oc0 = 0; oc0 < nco; oc0 += 32) {
        const int oc = oc0 + lane;
        const bool live = oc < nco;
        const int f = P.f_begin + oc * VEC;  // output column
        float acc[VEC];
#pragma unroll
        for (int v = 0; v < VEC; ++v) acc[v] = 0.f;
        for (int r = 0; r < R; ++r) {
            const hg_gat_rel& Rl = A.rel[r];
            const int32_t beg = __ldg(Rl.rowptr + i), end = __ldg(Rl.rowptr + i + 1);
            if (end <= beg) continue;
            const int h_lo = head_mean ? 0 : (live ? f / C : 0);
            const int h_hi = head_mean ? H : h_lo + 1;
            for (int h = h_lo; h < h_hi; ++h) {
                const float ad = __ldg(Rl.a_dst + i * Rl.lda_dst + h);
                float m = -INFINITY, l = 0.f;
                for (int32_t e = beg; e < end; ++e) {
                    float sc = __ldg(Rl.a_src + static_cast<int64_t>(__ldg(Rl.col + e)) * Rl.lda_src + h) + ad;
                    if (Rl.a_edge) sc += __ldg(Rl.a_edge + static_cast<int64_t>(e) * H + h);
                    m = fmaxf(m, leaky(sc, A.neg_slope));
                }
                for (int32_t e = beg; e < end; ++e) {
                    float sc = __ldg(Rl.a_src + static_cast<int64_t>(__ldg(Rl.col + e)) * Rl.lda_src + h) + ad;
                    if (Rl.a_edge) sc += __ldg(Rl.a_edge + static_cast<int64_t>(e) * H + h);
                    l += __expf(leaky(sc, A.neg_slope) - m);
                }
                const float inv = 1.f / (l + 1e-16f);
                const int src_f = head_mean ? h * C + f : (A.src_shared ? f % C : f);
                for (int32_t e = beg; e < end; ++e) {
                    const int32_t j = __ldg(Rl.col + e);
                    float sc = __ldg(Rl.a_src + static_cast<int64_t>(j) * Rl.lda_src + h) + ad;
                    if (Rl.a_edge) sc += __ldg(Rl.a_edge + static_cast<int64_t>(e) * H + h);
                    const float w = __expf(leaky(sc, A.neg_slope) - m) * inv;
                    if (A.alpha_saved[r] && (head_mean ? (lane == 0 && oc0 == 0 && P.f_begin == 0) : (live && f % C == 0)))
                        A.alpha_saved[r][static_cast<int64_t>(e) * H + h] = w;
                    if (live) {
                        float val[VEC];
                        ldg_vec<VEC>(Rl.xs + static_cast<int64_t>(j) * Rl.ldxs + src_f, val);
#pragma unroll
                        for (int v = 0; v < VEC; ++v) acc[v] = fmaf(w, val[v], acc[v]);
                    }
                }
            }
        }
        if (!live) continue;
        const float hscale = head_mean ? 1.f / static_cast<float>(H) : 1.f;
        float o[VEC];
#pragma unroll
        for (int v = 0; v < VEC; ++v) o[v] = acc[v] * hscale;
        for (int r = 0; r < R; ++r)
            if (A.rel[r].bias) {
#pragma unroll
                for (int v = 0; v < VEC; ++v) o[v] += __ldg(A.rel[r].bias + f + v);
            }
#pragma unroll
        for (int v = 0; v < VEC; ++v) o[v] *= gscale;
        if (A.addend) {
#pragma unroll
            for (int v = 0; v < VEC; ++v) o[v] += __ldg(A.addend + i * A.ldadd + f + v);
        }
#pragma unroll
        for (int v = 0; v < VEC; ++v) o[v] = apply_act(o[v], A.act);
        st_vec<VEC>(A.out + i * A.ldout + f, o);
    }
}
template <int VEC, int TPR, int KC>
__global__ void __launch_bounds__(kThreads, (KC <= 2 ? 6 : 4)) k_gat_fused(const __grid_constant__ GatParams P) {
    extern __shared__ __align__(16) uint8_t s_raw[];
    const hg_gat_args& A = P.a;
    const int H = A.H, C = A.C, R = A.n_rel, cap = P.cap, hs = P.hs;
    const int hsh = hs == 4 ? 2 : 3;          // log2(hs)
    const Smem S = carve(s_raw, R, cap, hs, P.f_count);
    const int tid = threadIdx.x;
    const bool head_mean = !A.concat && !A.src_shared;
    const int nco = P.f_count / VEC;          // output chunks of this launch
    constexpr int kGroups = kThreads / TPR;   // rows gathered concurrently
    const int q = tid % TPR, grp = tid / TPR;
    const int tile_rows = P.tile_rows;
    const int64_t n_tiles = ceil_div<int64_t>(A.n_dst, tile_rows);
    const float gscale = (A.group == HG_GROUP_MEAN) ? 1.f / static_cast<float>(R) : 1.f;
    const float hscale = head_mean ? 1.f / static_cast<float>(H) : 1.f;

    for (int f = tid; f < P.f_count; f += kThreads) {
        float b = 0.f;
        for (int r = 0; r < R; ++r)
            if (A.rel[r].bias) b += __ldg(A.rel[r].bias + P.f_begin + f);
        S.bias[f] = b;
    }

    // per-thread constants of its KC output chunks
    int fo[KC], fsrc[KC], hk[KC];
#pragma unroll
    for (int k = 0; k < KC; ++k) {
        const int oc = q + k * TPR;
        const bool valid = oc < nco;
        fo[k] = valid ? P.f_begin + oc * VEC : -1;
        hk[k] = (valid && !head_mean) ? fo[k] / C : 0;
        fsrc[k] = valid ? (A.src_shared ? fo[k] % C : fo[k]) : 0;
    }

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t t0 = tile * tile_rows;
        const int rows = (A.n_dst - t0 < tile_rows) ? static_cast<int>(A.n_dst - t0) : tile_rows;
        __syncthreads();  // previous tile fully consumed (also orders the bias fill before its first use)

        // ---- stage 1: rowptr slices + destination logits (padded [R][32][hs] layout: shifts only)
        for (int idx = tid; idx < R * (kMaxTile + 1); idx += kThreads) {
            const int r = idx / (kMaxTile + 1), t = idx - r * (kMaxTile + 1);
            S.rp[idx] = __ldg(A.rel[r].rowptr + t0 + min(t, rows));
        }
        for (int idx = tid; idx < (R * kMaxTile) << hsh; idx += kThreads) {
            const int h = idx & (hs - 1), t = (idx >> hsh) & (kMaxTile - 1), r = idx >> (hsh + 5);
            if (t < rows && h < H) S.ad[idx] = __ldg(A.rel[r].a_dst + (t0 + t) * A.rel[r].lda_dst + h);
        }
        __syncthreads();
        bool fits = true;
        for (int r = 0; r < R; ++r) fits = fits && (S.rp[r * (kMaxTile + 1) + rows] - S.rp[r * (kMaxTile + 1)] <= cap);
        if (!fits) {  // hub tile (uniform branch): warp per row straight from global memory
            for (int t = tid >> 5; t < rows; t += kWarps) slow_row<VEC>(P, t0 + t, tid & 31);
            continue;
        }

        // ---- stage 2: neighbour ids, edge -> row map, raw logits (threads = edges)
        for (int r = 0; r < R; ++r) {
            const hg_gat_rel& Rl = A.rel[r];
            const int32_t* rp = S.rp + r * (kMaxTile + 1);
            const int32_t e0 = rp[0], ne = rp[rows] - e0;
            const float* __restrict__ a_src = Rl.a_src;
            const float* __restrict__ a_edge = Rl.a_edge;
            const int64_t lda = Rl.lda_src;
            for (int e = tid; e < ne; e += kThreads) {
                const int32_t j = __ldg(Rl.col + e0 + e);
                S.col[r * cap + e] = j;
                int lo = 0, hi = rows;  // row t with rp[t] <= e0 + e < rp[t + 1]
                while (hi - lo > 1) {
                    const int mid = (lo + hi) >> 1;
                    if (rp[mid] <= e0 + e) lo = mid; else hi = mid;
                }
                S.erow[r * cap + e] = static_cast<uint8_t>(lo);
                float* wrow = S.w + ((static_cast<size_t>(r) * cap + e) << hsh);
                const float* src = a_src + static_cast<int64_t>(j) * lda;
#pragma unroll
                for (int h = 0; h < kMaxHeads; ++h) {
                    if (h < hs) {
                        float s = 0.f;
                        if (h < H) {
                            s = __ldg(src + h);
                            if (a_edge) s += __ldg(a_edge + static_cast<int64_t>(e0 + e) * H + h);
                        }
                        wrow[h] = s;
                    }
                }
            }
        }
        __syncthreads();

        // ---- softmax statistics: threads = (relation, row, head) over the padded layout
        for (int idx = tid; idx < (R * kMaxTile) << hsh; idx += kThreads) {
            const int h = idx & (hs - 1), t = (idx >> hsh) & (kMaxTile - 1), r = idx >> (hsh + 5);
            if (t >= rows || h >= H) continue;
            const int32_t* rp = S.rp + r * (kMaxTile + 1);
            const int beg = rp[t] - rp[0], end = rp[t + 1] - rp[0];
            const float ad = S.ad[idx];
            const float* w = S.w + ((static_cast<size_t>(r) * cap) << hsh) + h;
            float m = -INFINITY;
            for (int e = beg; e < end; ++e) m = fmaxf(m, leaky(w[e << hsh] + ad, A.neg_slope));
            float l = 0.f;
            for (int e = beg; e < end; ++e) l += __expf(leaky(w[e << hsh] + ad, A.neg_slope) - m);
            S.mx[idx] = m;
            S.inv[idx] = 1.f / (l + 1e-16f);
        }
        __syncthreads();
        // ---- logits -> weights in place: threads = edges
        for (int r = 0; r < R; ++r) {
            const int32_t e0 = S.rp[r * (kMaxTile + 1)], ne = S.rp[r * (kMaxTile + 1) + rows] - e0;
            float* alpha = (P.f_begin == 0) ? A.alpha_saved[r] : nullptr;
            for (int e = tid; e < ne; e += kThreads) {
                const int t = S.erow[r * cap + e];
                float* wrow = S.w + ((static_cast<size_t>(r) * cap + e) << hsh);
                const int base = (r * kMaxTile + t) << hsh;
#pragma unroll
                for (int h = 0; h < kMaxHeads; ++h) {
                    if (h < H) {
                        const float w = __expf(leaky(wrow[h] + S.ad[base + h], A.neg_slope) - S.mx[base + h]) * S.inv[base + h];
                        wrow[h] = w;
                        if (alpha) alpha[static_cast<int64_t>(e0 + e) * H + h] = w;
                    }
                }
            }
        }
        __syncthreads();

        // ---- gather-accumulate: threads = (row, output chunk)
        for (int t = grp; t < rows; t += kGroups) {
            const int64_t i = t0 + t;
            float acc[KC][VEC];
#pragma unroll
            for (int k = 0; k < KC; ++k)
#pragma unroll
                for (int v = 0; v < VEC; ++v) acc[k][v] = 0.f;
            for (int r = 0; r < R; ++r) {
                const float* __restrict__ xsb = A.rel[r].xs + fsrc[0];
                const int64_t ld = A.rel[r].ldxs;
                const int32_t* rp = S.rp + r * (kMaxTile + 1);
                const int beg = rp[t] - rp[0], end = rp[t + 1] - rp[0];
                const int32_t* colp = S.col + r * cap;
                const float* wp = S.w + ((static_cast<size_t>(r) * cap) << hsh);
                if (head_mean) {
                    for (int hb = 0; hb < H; hb += 4) {  // heads in blocks of 4: their weights are one LDS.128
                        const float* __restrict__ xsh = xsb + hb * C;
                        const int hn = H - hb;  // live heads of this block (>= 1)
#pragma unroll(KC <= 2 ? 2 : 1)
                        for (int e = beg; e < end; ++e) {
                            const float* __restrict__ rowh = xsh + static_cast<int64_t>(colp[e]) * ld;
                            const float4 wa = *reinterpret_cast<const float4*>(wp + (e << hsh) + hb);
                            const float wv[4] = {wa.x, wa.y, wa.z, wa.w};
                            float val[4][KC][VEC];
#pragma unroll
                            for (int hh = 0; hh < 4; ++hh)  // all loads of the edge first: independent, in flight together
#pragma unroll
                                for (int k = 0; k < KC; ++k) {
                                    if (hh < hn && fo[k] >= 0) {
                                        ldg_vec<VEC>(rowh + hh * C + (fsrc[k] - fsrc[0]), val[hh][k]);
                                    } else {
#pragma unroll
                                        for (int v = 0; v < VEC; ++v) val[hh][k][v] = 0.f;
                                    }
                                }
#pragma unroll
                            for (int hh = 0; hh < 4; ++hh)
#pragma unroll
                                for (int k = 0; k < KC; ++k)
#pragma unroll
                                    for (int v = 0; v < VEC; ++v) acc[k][v] = fmaf(wv[hh], val[hh][k][v], acc[k][v]);
                        }
                    }
                } else {
#pragma unroll 2
                    for (int e = beg; e < end; ++e) {
                        const float* __restrict__ row = xsb + static_cast<int64_t>(colp[e]) * ld;
                        const float* we = wp + (e << hsh);
#pragma unroll
                        for (int k = 0; k < KC; ++k)
                            if (fo[k] >= 0) {
                                float val[VEC];
                                ldg_vec<VEC>(row + (fsrc[k] - fsrc[0]), val);
                                const float w = we[hk[k]];
#pragma unroll
                                for (int v = 0; v < VEC; ++v) acc[k][v] = fmaf(w, val[v], acc[k][v]);
                            }
                    }
                }
            }
            // ---- epilogue: head-mean scale, summed biases, group mean, addend, activation, ONE store
#pragma unroll
            for (int k = 0; k < KC; ++k) {
                if (fo[k] < 0) continue;
                float o[VEC];
#pragma unroll
                for (int v = 0; v < VEC; ++v) o[v] = (acc[k][v] * hscale + S.bias[fo[k] - P.f_begin + v]) * gscale;
                if (A.addend) {
                    float av[VEC];
                    ldg_vec<VEC>(A.addend + i * A.ldadd + fo[k], av);
#pragma unroll
                    for (int v = 0; v < VEC; ++v) o[v] += av[v];
                }
#pragma unroll
                for (int v = 0; v < VEC; ++v) {
                    float y = o[v];
                    if (A.act == HG_ACT_RELU) y = fmaxf(y, 0.f);
                    else if (A.act == HG_ACT_ELU) y = y > 0.f ? y : __expf(y) - 1.f;  // F.elu: exp(x) - 1
                    o[v] = y;
                }
                st_vec<VEC>(A.out + i * A.ldout + fo[k], o);
            }
        }
    }
}

template <int VEC, int TPR, int KC>
int launch(GatParams& P, cudaStream_t stream) {
    const hg_gat_args& A = P.a;
    constexpr int kGroups = kThreads / TPR;
    P.tile_rows = kGroups > 16 ? kMaxTile : 16;
    P.hs = A.H <= 4 ? 4 : 8;
    // staging capacity: what ~24 KB of shared memory holds, 128..2048 edges per relation per tile
    const size_t fixed = smem_bytes(A.n_rel, 0, P.hs, P.f_count);
    const size_t per_edge = static_cast<size_t>(A.n_rel) * (4 + 4 * P.hs + 1);
    int cap = fixed < 24 * 1024 ? static_cast<int>((24 * 1024 - fixed) / per_edge) : 0;
    cap = cap / 32 * 32;
    cap = cap < 128 ? 128 : (cap > 2048 ? 2048 : cap);
    P.cap = cap;
    const size_t smem = smem_bytes(A.n_rel, cap, P.hs, P.f_count);
    if (smem > 200 * 1024) return HG_EUNSUPPORTED;
    if (smem > 48 * 1024) {
        HG_CUDA_TRY(cudaFuncSetAttribute(k_gat_fused<VEC, TPR, KC>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem)));
    }
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int64_t need = ceil_div<int64_t>(A.n_dst, P.tile_rows);
    const int64_t cap_grid = static_cast<int64_t>(sms) * 6 * 8;  // 8 waves of 148 SMs x 6 resident CTAs
    const unsigned grid = static_cast<unsigned>(need < cap_grid ? need : cap_grid);
    k_gat_fused<VEC, TPR, KC><<<grid, kThreads, smem, stream>>>(P);
    return check_launch();
}

template <int VEC>
int dispatch(GatParams& P, cudaStream_t stream) {
    const int nco = P.f_count / VEC;
    if (nco <= 4) return launch<VEC, 4, 1>(P, stream);
    if (nco <= 8) return launch<VEC, 8, 1>(P, stream);
    if (nco <= 16) return launch<VEC, 16, 1>(P, stream);
    if (nco <= 32) return launch<VEC, 32, 1>(P, stream);
    if (nco <= 64) return launch<VEC, 32, 2>(P, stream);
    if (nco <= 96) return launch<VEC, 32, 3>(P, stream);
    if (nco <= 128) return launch<VEC, 32, 4>(P, stream);
    if (nco <= 192) return launch<VEC, 32, 6>(P, stream);
    return launch<VEC, 32, 8>(P, stream);
}

}  // namespace
}  // namespace hg

extern "C" int hg_gat_fused(const hg_gat_args* a, void* stream_) {
    using namespace hg;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!a) return HG_EINVAL;
    if (a->n_rel < 1 || a->n_rel > HG_MAX_RELATIONS || a->n_dst < 0) return HG_EINVAL;
    if (a->H < 1 || a->H > kMaxHeads || a->C < 1) return HG_EINVAL;
    if (a->act < HG_ACT_NONE || a->act > HG_ACT_ELU) return HG_EINVAL;
    if (a->src_shared && !a->concat) return HG_EINVAL;  // aggregate-first output is always [n_dst, H*C]
    if (a->n_dst == 0) return HG_OK;
    if (!a->out) return HG_EINVAL;
    for (int r = 0; r < a->n_rel; ++r) {
        const hg_gat_rel& R = a->rel[r];
        // xs / a_src may be NULL only for an empty source set (every row of that relation is then empty)
        if (!R.a_dst || !R.rowptr || !R.col) return HG_EINVAL;
    }
    const bool head_mean = !a->concat && !a->src_shared;
    const int width = head_mean ? a->C : a->H * a->C;
    // Widest vector that divides C (a chunk must not straddle two heads) and fits all alignments.
    auto ok = [&](int vec) {
        const size_t bytes = sizeof(float) * vec;
        if (a->C % vec || a->ldout % vec || !aligned(a->out, bytes)) return false;
        if (a->addend && (a->ldadd % vec || !aligned(a->addend, bytes))) return false;
        for (int r = 0; r < a->n_rel; ++r) {
            const hg_gat_rel& R = a->rel[r];
            if (R.ldxs % vec || !aligned(R.xs, bytes)) return false;
            if (R.bias && !aligned(R.bias, bytes)) return false;
        }
        return true;
    };
    const int vec = ok(4) ? 4 : (ok(2) ? 2 : 1);
    const int block = 32 * 8 * vec;  // widest output block one launch covers
    for (int f0 = 0; f0 < width; f0 += block) {
        GatParams P;
        P.a = *a;
        P.f_begin = f0;
        P.f_count = (width - f0 < block) ? width - f0 : block;
        const int rc = vec == 4 ? dispatch<4>(P, stream) : (vec == 2 ? dispatch<2>(P, stream) : dispatch<1>(P, stream));
        if (rc != HG_OK) return rc;
    }
    return HG_OK;
}
