// K3 + K4 -- fused GAT for one destination node type (v6).
//
// A CTA (128 threads, 6-8 resident per SM) owns a tile of 32 consecutive destination rows
// (consecutive rows belong to the same scene graph, so their neighbour sets overlap and the gathers
// of a tile hit L1/L2).  Per tile, three barriers:
//   stage    rowptr slices and destination logits; and -- threads = edges, coalesced on col -- the
//            neighbour ids and the raw logits a_src[j,h](+a_edge) of every relation go to shared memory;
//   softmax  threads = (relation, row, head): leaky-relu + max, exp + sum, normalise -- three passes over
//            the row's staged logits, in place.  No shuffles, no idle lanes on short rows (mean in-degree
//            of a scene graph is ~5), no runtime integer divisions (index maps are shifts and masks over
//            padded layouts), no edge -> row map;
//   gather   threads = (row, chunk): TPR threads cooperate on a row, each owning KC vector chunks.
//            Head-mean (MODE 1, H compile-time): the thread walks the H head slices of xs[j] itself, so the
//            head reduction never leaves registers.  Aggregate-first (MODE 2): the thread owns a chunk of
//            the SOURCE row and accumulates the H weighted copies, so xs[j] is read once, not H times.
//            The weights of an edge are one broadcast LDS.128; addresses are one IMAD.WIDE per load.
//            Relations are summed in the same registers;
//   out      + summed biases (+ addend), group mean, ELU/ReLU, ONE vector store per chunk.
// No atomics, no read-modify-write across relations, no separate softmax / stack+sum / activation
// passes.  Tiles whose edges do not fit the staging buffers (hub nodes) take a warp-per-row path
// that streams from global memory.
//
// src_shared = 1 is the aggregate-first form: xs is [n_src, C] (ONE row shared by all heads) and the
// output is [n_dst, H*C] with out[i, h, :] = sum_e w_eh xs[j, :]; the projection is applied afterwards
// to the aggregated rows (used when a relation has far fewer destinations than sources).
#include <cuda.h>

#include <cstdlib>

#include "tc_common.cuh"

namespace hg {
namespace {

#ifndef HG_GAT_SMEM_KB
#define HG_GAT_SMEM_KB 24
#endif
constexpr size_t kStageBudget = HG_GAT_SMEM_KB * 1024;  // shared memory per CTA for staging (the rest of the SM's 256 KB is L1)
#ifndef HG_GAT_OCC_LONG
#define HG_GAT_OCC_LONG 8
#endif
#ifndef HG_GAT_MIN_TILE
#define HG_GAT_MIN_TILE 32  // smaller tiles on short grids measured slower (8, 16)
#endif
constexpr int kThreads = 128;
constexpr int kWarps = kThreads / kWarp;
constexpr int kMaxHeads = 8;
constexpr int kMaxTile = 32;  // rows per tile

struct GatParams {
    hg_gat_args a;
    int32_t cap;        // staged edges per relation per tile (multiple of 32)
    int32_t tile_rows;  // kMaxTile
    int32_t hs;         // padded head stride in shared memory: 4 or 8
    int32_t f_begin;    // first column of this launch (wide rows are split over launches);
    int32_t f_count;    // columns of this launch.  OUTPUT columns, except MODE 2: columns of the source row
    int32_t n_bias;     // staged bias values: f_count (MODE 2: H * f_count, head-major)
    int32_t wrows;      // window kernel: source rows the shared-memory window holds (all relations of a tile)
};

__device__ __forceinline__ float leaky(float v, float slope) { return v > 0.f ? v : v * slope; }

struct Smem {
    float* w;        // [R][cap][hs]        logits -> weights (16-byte aligned rows)
    float* ad;       // [R][kMaxTile][hs]   destination logits
    float* bias;     // [n_bias]            summed biases of this launch's output columns
    int32_t* col;    // [R][cap]
    int32_t* rp;     // [R][kMaxTile + 1]
};

__host__ __device__ inline size_t smem_bytes(int R, int cap, int hs, int n_bias) {
    size_t words = static_cast<size_t>(R) * cap * hs + static_cast<size_t>(R) * kMaxTile * hs +
                   static_cast<size_t>((n_bias + 3) / 4 * 4) + static_cast<size_t>(R) * cap +
                   static_cast<size_t>(R) * (kMaxTile + 1);
    return words * 4 + 16;
}

__device__ __forceinline__ Smem carve(uint8_t* raw, int R, int cap, int hs, int n_bias) {
    Smem s;
    float* p = reinterpret_cast<float*>(raw);
    s.w = p;      p += static_cast<size_t>(R) * cap * hs;
    s.ad = p;     p += R * kMaxTile * hs;
    s.bias = p;   p += (n_bias + 3) / 4 * 4;
    s.col = reinterpret_cast<int32_t*>(p);  p += static_cast<size_t>(R) * cap;
    s.rp = reinterpret_cast<int32_t*>(p);
    return s;
}

// Hub path: one warp per row, lanes = output chunks, everything streamed from global memory.
template <int VEC>
__device__ __noinline__ void slow_row(const GatParams& P, int64_t i, int lane, int f_begin, int f_count) {
    const hg_gat_args& A = P.a;
    const int H = A.H, C = A.C, R = A.n_rel;
    const bool head_mean = !A.concat && !A.src_shared;
    const int nco = f_count / VEC;
    const float gscale = (A.group == HG_GROUP_MEAN) ? 1.f / static_cast<float>(R) : 1.f;
    for (int oc0 = 0; oc0 < nco; oc0 += 32) {
        const int oc = oc0 + lane;
        const bool live = oc < nco;
        const int f = f_begin + oc * VEC;  // output column
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
                    float w = __expf(leaky(sc, A.neg_slope) - m) * inv;
                    float a_saved = w;
                    if (A.dropout_p > 0.f) {  // training forward: decision in the sign bit of the saved weight
                        w *= dropout_keep_scale(A.dropout_seed + (A.dropout_seed_dev ? *A.dropout_seed_dev : 0ull), r,
                                                static_cast<int64_t>(e) * H + h, A.dropout_p,
                                                1.f / (1.f - A.dropout_p));
                        if (w == 0.f) a_saved = -a_saved;
                    }
                    if (A.alpha_saved[r] && (head_mean ? (lane == 0 && oc0 == 0 && f_begin == 0) : (live && f % C == 0)))
                        A.alpha_saved[r][static_cast<int64_t>(e) * H + h] = a_saved;
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


// MODE 0: generic (concat / src_shared / head-mean with any H <= 8; thread chunks are OUTPUT chunks).
// MODE 1: head-mean, H == HT (<= 4).   MODE 2: aggregate-first (src_shared), H == HT: chunks of the SOURCE row.
// OCC: resident CTAs per SM the kernel is compiled for (register cap); 8 pays on grids of many waves.
template <int VEC, int TPR, int KC, int MODE, int HT, int OCC>
__global__ void __launch_bounds__(kThreads, OCC)
k_gat_fused(const __grid_constant__ GatParams P) {
    extern __shared__ __align__(16) uint8_t s_raw[];
    const hg_gat_args& A = P.a;
    const int H = MODE ? HT : A.H;
    const int C = A.C, R = A.n_rel, cap = P.cap;
    const int hs = MODE ? 4 : P.hs;
    const int hsh = MODE ? 2 : (P.hs == 4 ? 2 : 3);  // log2(hs)
    const Smem S = carve(s_raw, R, cap, hs, P.n_bias);
    const int tid = threadIdx.x;
    const bool head_mean = MODE == 1 || (MODE == 0 && !A.concat && !A.src_shared);
    const int nco = P.f_count / VEC;          // chunks of this launch
    constexpr int kGroups = kThreads / TPR;   // rows gathered concurrently
    const int q = tid % TPR, grp = tid / TPR;
    const int tile_rows = P.tile_rows;
    const int64_t n_tiles = ceil_div<int64_t>(A.n_dst, tile_rows);
    const float gscale = (A.group == HG_GROUP_MEAN) ? 1.f / static_cast<float>(R) : 1.f;
    const float hscale = head_mean ? 1.f / static_cast<float>(H) : 1.f;

    for (int f = tid; f < P.n_bias; f += kThreads) {
        // MODE 2 stages [H][f_count] (head-major); the others [f_count]
        const int col = (MODE == 2) ? (f / P.f_count) * C + P.f_begin + f % P.f_count : P.f_begin + f;
        float b = 0.f;
        for (int r = 0; r < R; ++r)
            if (A.rel[r].bias) b += __ldg(A.rel[r].bias + col);
        S.bias[f] = b;
    }

    // per-thread constants: chunk k of this thread starts at column f0 + k * TPR * VEC of the launch
    const int f0 = P.f_begin + q * VEC;
    bool live[KC];
    int hk[KC], fsrc[KC];  // MODE 0: head of the chunk (concat / src_shared) and its column in the source row
#pragma unroll
    for (int k = 0; k < KC; ++k) {
        live[k] = q + k * TPR < nco;
        const int fo = f0 + k * TPR * VEC;
        hk[k] = (MODE == 0 && live[k] && !head_mean) ? fo / C : 0;
        fsrc[k] = (MODE == 0 && A.src_shared && live[k]) ? fo % C : fo;
    }

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t t0 = tile * tile_rows;
        const int rows = (A.n_dst - t0 < tile_rows) ? static_cast<int>(A.n_dst - t0) : tile_rows;
        __syncthreads();  // previous tile fully consumed (also orders the bias fill before its first use)

        // ---- stage: rowptr slices + destination logits (padded [R][32][hs] layout: shifts only) ...
        for (int idx = tid; idx < R * (kMaxTile + 1); idx += kThreads) {
            const int r = idx / (kMaxTile + 1), t = idx - r * (kMaxTile + 1);
            S.rp[idx] = __ldg(A.rel[r].rowptr + t0 + min(t, rows));
        }
        for (int idx = tid; idx < (R * kMaxTile) << hsh; idx += kThreads) {
            const int h = idx & (hs - 1), t = (idx >> hsh) & (kMaxTile - 1), r = idx >> (hsh + 5);
            if (t < rows && h < H) S.ad[idx] = __ldg(A.rel[r].a_dst + (t0 + t) * A.rel[r].lda_dst + h);
        }
        // ... the tile's edge range per relation straight from global memory (uniform loads), so the
        // edge staging does not have to wait for the rowptr slices behind a barrier
        bool fits = true;
        for (int r = 0; r < R; ++r)
            fits = fits && (__ldg(A.rel[r].rowptr + t0 + rows) - __ldg(A.rel[r].rowptr + t0) <= cap);
        if (!fits) {  // hub tile (uniform branch): warp per row straight from global memory
            for (int t = tid >> 5; t < rows; t += kWarps) {
                if (MODE == 2) {
                    for (int h = 0; h < H; ++h) slow_row<VEC>(P, t0 + t, tid & 31, h * C + P.f_begin, P.f_count);
                } else {
                    slow_row<VEC>(P, t0 + t, tid & 31, P.f_begin, P.f_count);
                }
            }
            continue;
        }
        // ---- ... neighbour ids and raw logits (threads = edges)
        for (int r = 0; r < R; ++r) {
            const hg_gat_rel& Rl = A.rel[r];
            const int32_t e0 = __ldg(Rl.rowptr + t0), ne = __ldg(Rl.rowptr + t0 + rows) - e0;
            const float* __restrict__ a_src = Rl.a_src;
            const float* __restrict__ a_edge = Rl.a_edge;
            const int lda = static_cast<int>(Rl.lda_src);
            for (int e = tid; e < ne; e += kThreads) {
                const int32_t j = __ldg(Rl.col + e0 + e);
                S.col[r * cap + e] = j;
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

        // ---- softmax in place: threads = (relation, row, head) over the padded layout
        for (int idx = tid; idx < (R * kMaxTile) << hsh; idx += kThreads) {
            const int h = idx & (hs - 1), t = (idx >> hsh) & (kMaxTile - 1), r = idx >> (hsh + 5);
            if (t >= rows || h >= H) continue;
            const int32_t* rp = S.rp + r * (kMaxTile + 1);
            const int e0 = rp[0], beg = rp[t] - e0, end = rp[t + 1] - e0;
            const float ad = S.ad[idx];
            float* w = S.w + ((static_cast<size_t>(r) * cap) << hsh) + h;
            float m = -INFINITY;
            for (int e = beg; e < end; ++e) {
                const float v = leaky(w[e << hsh] + ad, A.neg_slope);
                w[e << hsh] = v;
                m = fmaxf(m, v);
            }
            float l = 0.f;
            for (int e = beg; e < end; ++e) {
                const float p = __expf(w[e << hsh] - m);
                w[e << hsh] = p;
                l += p;
            }
            const float inv = 1.f / (l + 1e-16f);
            float* alpha = (P.f_begin == 0) ? A.alpha_saved[r] : nullptr;
            if (A.dropout_p > 0.f) {  // training forward (uniform branch; inference passes 0)
                const float inv_keep = 1.f / (1.f - A.dropout_p);
                const uint64_t seed = A.dropout_seed + (A.dropout_seed_dev ? *A.dropout_seed_dev : 0ull);
                for (int e = beg; e < end; ++e) {
                    const float a = w[e << hsh] * inv;
                    const int64_t idx = static_cast<int64_t>(e0 + e) * H + h;
                    const float ks = dropout_keep_scale(seed, r, idx, A.dropout_p, inv_keep);
                    w[e << hsh] = a * ks;
                    if (alpha) alpha[idx] = ks == 0.f ? -a : a;
                }
                continue;
            }
            for (int e = beg; e < end; ++e) {
                const float a = w[e << hsh] * inv;
                w[e << hsh] = a;
                if (alpha) alpha[static_cast<int64_t>(e0 + e) * H + h] = a;
            }
        }
        __syncthreads();

        // ---- gather-accumulate: threads = (row, chunk)
        if (KC == 1 && !live[0]) continue;  // no barrier below this point
        for (int t = grp; t < rows; t += kGroups) {
            const int64_t i = t0 + t;
            constexpr int NH = (MODE == 2) ? HT : 1;  // accumulators per chunk
            float acc[NH][KC][VEC];
#pragma unroll
            for (int h = 0; h < NH; ++h)
#pragma unroll
                for (int k = 0; k < KC; ++k)
#pragma unroll
                    for (int v = 0; v < VEC; ++v) acc[h][k][v] = 0.f;
            for (int r = 0; r < R; ++r) {
                const float* __restrict__ xsb = A.rel[r].xs + fsrc[0];
                const int ld = static_cast<int>(A.rel[r].ldxs);
                const uint32_t ldb = static_cast<uint32_t>(ld) * 4u;  // row pitch in bytes (< 4 GB, checked by the host)
                const int32_t* rp = S.rp + r * (kMaxTile + 1);
                const int beg = rp[t] - rp[0], end = rp[t + 1] - rp[0];
                const int32_t* colp = S.col + r * cap;
                const float* wp = S.w + ((static_cast<size_t>(r) * cap) << hsh);
                if constexpr (MODE == 1) {
                    const char* xh[HT];  // one byte base per head: an address is then ONE IMAD.WIDE.U32
#pragma unroll
                    for (int hh = 0; hh < HT; ++hh) xh[hh] = reinterpret_cast<const char*>(xsb + hh * C);
#pragma unroll(OCC >= 7 ? 1 : 2)
                    for (int e = beg; e < end; ++e) {
                        const uint32_t j = static_cast<uint32_t>(colp[e]);
                        const float4 wa = *reinterpret_cast<const float4*>(wp + (e << 2));
                        const float wv[4] = {wa.x, wa.y, wa.z, wa.w};
                        float val[HT][KC][VEC];
#pragma unroll
                        for (int hh = 0; hh < HT; ++hh)  // all loads of the edge first: independent, in flight together
#pragma unroll
                            for (int k = 0; k < KC; ++k) {
                                if (KC == 1 || live[k]) {
                                    ldg_vec<VEC>(reinterpret_cast<const float*>(xh[hh] + static_cast<uint64_t>(j) * ldb) + k * TPR * VEC, val[hh][k]);
                                } else {
#pragma unroll
                                    for (int v = 0; v < VEC; ++v) val[hh][k][v] = 0.f;
                                }
                            }
#pragma unroll
                        for (int hh = 0; hh < HT; ++hh)
#pragma unroll
                            for (int k = 0; k < KC; ++k)
#pragma unroll
                                for (int v = 0; v < VEC; ++v) acc[0][k][v] = fmaf(wv[hh], val[hh][k][v], acc[0][k][v]);
                    }
                } else if constexpr (MODE == 2) {
#pragma unroll(KC <= 2 ? 4 : 2)
                    for (int e = beg; e < end; ++e) {
                        const uint32_t j = static_cast<uint32_t>(colp[e]);
                        const float4 wa = *reinterpret_cast<const float4*>(wp + (e << 2));
                        const float wv[4] = {wa.x, wa.y, wa.z, wa.w};
                        float val[KC][VEC];
#pragma unroll
                        for (int k = 0; k < KC; ++k) {
                            if (KC == 1 || live[k]) {
                                ldg_vec<VEC>(reinterpret_cast<const float*>(reinterpret_cast<const char*>(xsb) + static_cast<uint64_t>(j) * ldb) + k * TPR * VEC, val[k]);
                            } else {
#pragma unroll
                                for (int v = 0; v < VEC; ++v) val[k][v] = 0.f;
                            }
                        }
#pragma unroll
                        for (int hh = 0; hh < HT; ++hh)
#pragma unroll
                            for (int k = 0; k < KC; ++k)
#pragma unroll
                                for (int v = 0; v < VEC; ++v) acc[hh][k][v] = fmaf(wv[hh], val[k][v], acc[hh][k][v]);
                    }
                } else if (head_mean) {
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
                            for (int hh = 0; hh < 4; ++hh)
#pragma unroll
                                for (int k = 0; k < KC; ++k) {
                                    if (hh < hn && live[k]) {
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
                                    for (int v = 0; v < VEC; ++v) acc[0][k][v] = fmaf(wv[hh], val[hh][k][v], acc[0][k][v]);
                        }
                    }
                } else {
#pragma unroll 2
                    for (int e = beg; e < end; ++e) {
                        const float* __restrict__ row = xsb + static_cast<int64_t>(colp[e]) * ld;
                        const float* we = wp + (e << hsh);
#pragma unroll
                        for (int k = 0; k < KC; ++k)
                            if (live[k]) {
                                float val[VEC];
                                ldg_vec<VEC>(row + (fsrc[k] - fsrc[0]), val);
                                const float w = we[hk[k]];
#pragma unroll
                                for (int v = 0; v < VEC; ++v) acc[0][k][v] = fmaf(w, val[v], acc[0][k][v]);
                            }
                    }
                }
            }
            // ---- epilogue: head-mean scale, summed biases, group mean, addend, activation, ONE store per chunk
            const float* __restrict__ addp = A.addend ? A.addend + i * A.ldadd : nullptr;
            float* __restrict__ outp = A.out + i * A.ldout;
            const int act = A.act;
#pragma unroll
            for (int h = 0; h < NH; ++h)
#pragma unroll
                for (int k = 0; k < KC; ++k) {
                    if (KC > 1 && !live[k]) continue;
                    const int lc = (q + k * TPR) * VEC;  // column within the launch
                    const int oc = (MODE == 2 ? h * C : 0) + P.f_begin + lc;
                    const float* bp = S.bias + (MODE == 2 ? h * P.f_count : 0) + lc;
                    float o[VEC];
#pragma unroll
                    for (int v = 0; v < VEC; ++v) o[v] = __fmul_rn(__fadd_rn(__fmul_rn(acc[h][k][v], hscale), bp[v]), gscale);
                    if (addp) {
                        float av[VEC];
                        ldg_vec<VEC>(addp + oc, av);
#pragma unroll
                        for (int v = 0; v < VEC; ++v) o[v] += av[v];
                    }
                    if (act == HG_ACT_RELU) {
#pragma unroll
                        for (int v = 0; v < VEC; ++v) o[v] = fmaxf(o[v], 0.f);
                    } else if (act == HG_ACT_ELU) {
#pragma unroll
                        for (int v = 0; v < VEC; ++v) o[v] = o[v] > 0.f ? o[v] : __expf(o[v]) - 1.f;  // F.elu: exp(x) - 1
                    }
                    st_vec<VEC>(outp + oc, o);
                }
        }
    }
}

template <int VEC, int TPR, int KC, int MODE, int HT>
int launch(GatParams& P, cudaStream_t stream) {
    const hg_gat_args& A = P.a;
    // short grids (the rooms side of a scene-graph batch) take smaller tiles: more CTAs in flight per SM
    const int sms0 = sm_count() > 0 ? sm_count() : 148;
    P.tile_rows = kMaxTile;
    // wide aggregate-first rows (5 float2 chunks per lane: the 306-d input layer) do long, latency-bound rows on a grid of
    // ~1 wave: 16-row tiles balance the tail (measured 134 -> 115 us at 20,277 rows; 8-row tiles the same; a warp-per-row
    // streaming kernel without shared memory 114-130 us: dropped).  HG_GAT_WIDE_TILE overrides, for A/B.
    static const int wide_tile = [] {
        const char* e = std::getenv("HG_GAT_WIDE_TILE");
        const int v = e ? std::atoi(e) : 16;
        return v == 8 || v == 16 || v == 32 ? v : 16;
    }();
    if (MODE == 2 && VEC == 2) {
        if (ceil_div<int64_t>(A.n_dst, kMaxTile) < static_cast<int64_t>(sms0) * 6) P.tile_rows = wide_tile;
    } else {
        while (P.tile_rows > HG_GAT_MIN_TILE && ceil_div<int64_t>(A.n_dst, P.tile_rows) < static_cast<int64_t>(sms0) * 6) P.tile_rows >>= 1;
    }
    P.hs = A.H <= 4 ? 4 : 8;
    P.n_bias = MODE == 2 ? A.H * P.f_count : P.f_count;
    // staging capacity: what the shared-memory budget holds, 128..2048 edges per relation per tile
    const size_t fixed = smem_bytes(A.n_rel, 0, P.hs, P.n_bias);
    const size_t per_edge = static_cast<size_t>(A.n_rel) * (4 + 4 * P.hs);
    int cap = fixed < kStageBudget ? static_cast<int>((kStageBudget - fixed) / per_edge) : 0;
    cap = cap / 32 * 32;
    cap = cap < 128 ? 128 : (cap > 2048 ? 2048 : cap);
    P.cap = cap;
    const size_t smem = smem_bytes(A.n_rel, cap, P.hs, P.n_bias);
    if (smem > 200 * 1024) return HG_EUNSUPPORTED;
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int64_t need = ceil_div<int64_t>(A.n_dst, P.tile_rows);
    const int64_t cap_grid = static_cast<int64_t>(sms) * 6 * 8;  // 8 waves of 148 SMs x 6 resident CTAs
    const unsigned grid = static_cast<unsigned>(need < cap_grid ? need : cap_grid);
    auto go = [&](auto kernel) -> int {
        if (smem > 48 * 1024) {
            HG_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        }
        kernel<<<grid, kThreads, smem, stream>>>(P);
        return check_launch();
    };
    // wide aggregate-first rows (VEC 2): 5 resident CTAs (96 registers, no spill) 117 -> 97 us alone at 20,277 x 306; 6 spills
    constexpr int kOcc = (MODE == 2 && VEC == 2) ? 5 : (KC * (MODE == 2 ? HT : 1) <= 4 ? 6 : 4);
    if constexpr (MODE == 1) {
        // latency-bound on long grids: more resident warps beat more loads in flight per warp
        if (need >= static_cast<int64_t>(sms) * 6 * 4) return go(k_gat_fused<VEC, TPR, KC, MODE, HT, HG_GAT_OCC_LONG>);
    }
    return go(k_gat_fused<VEC, TPR, KC, MODE, HT, kOcc>);
}

template <int VEC>
int dispatch(GatParams& P, cudaStream_t stream) {
    const int nco = P.f_count / VEC;
    if (nco <= 4) return launch<VEC, 4, 1, 0, 0>(P, stream);
    if (nco <= 8) return launch<VEC, 8, 1, 0, 0>(P, stream);
    if (nco <= 16) return launch<VEC, 16, 1, 0, 0>(P, stream);
    if (nco <= 32) return launch<VEC, 32, 1, 0, 0>(P, stream);
    if (nco <= 64) return launch<VEC, 32, 2, 0, 0>(P, stream);
    if (nco <= 96) return launch<VEC, 32, 3, 0, 0>(P, stream);
    if (nco <= 128) return launch<VEC, 32, 4, 0, 0>(P, stream);
    if (nco <= 192) return launch<VEC, 32, 6, 0, 0>(P, stream);
    return launch<VEC, 32, 8, 0, 0>(P, stream);
}

// Specialised forms (float4 chunks, H <= 4 compile-time): MODE 1 covers C <= 128, MODE 2 C <= 256.
template <int MODE, int HT>
int dispatch_special(GatParams& P, cudaStream_t stream) {
    const int nco = P.f_count / 4;
    if (nco <= 4) return launch<4, 4, 1, MODE, HT>(P, stream);
    if (nco <= 8) return launch<4, 8, 1, MODE, HT>(P, stream);
    if constexpr (MODE == 1) {
        // long grids (the 8-CTA/SM build): two chunks per thread amortise the per-edge address/weight work
        const int sms = sm_count() > 0 ? sm_count() : 148;
        const bool long_grid = ceil_div<int64_t>(P.a.n_dst, kMaxTile) >= static_cast<int64_t>(sms) * 6 * 4;
        if (long_grid && nco <= 16 && nco > 8) return launch<4, 8, 2, MODE, HT>(P, stream);
    }
    if (nco <= 16) return launch<4, 16, 1, MODE, HT>(P, stream);
    if (nco <= 32) return launch<4, 32, 1, MODE, HT>(P, stream);
    if constexpr (MODE == 2) {
        if (nco <= 64) return launch<4, 32, 2, MODE, HT>(P, stream);
    }
    return HG_EUNSUPPORTED;
}

// ---------------------------------------------------------------------------------------------------
// Window variant (v9) -- head-mean, C = 64, H <= 4, R <= 4, inference (no alpha_saved / dropout).
//
// A batch of scene graphs is a disjoint union in graph order and object-object edges stay inside a room, so the
// sources of a tile of 32 consecutive destination rows are a window of a few dozen consecutive rows of xs (the plan
// records {first row, rows} per tile: hg_plan_rel.tile_win).  The v6 kernel above gathers those rows with
// per-thread LDG.128 and is bound by instruction issue and by the latency of its L1 misses (ncu, round 1: 80 M
// warp instructions, IPC 1.9, 45 % of the warps active, DRAM traffic 1.07x the algorithmic bytes).  A first window
// kernel with CTA-wide phases (stage / softmax / gather between __syncthreads) reached only 143 us: 39 % of its warp
// samples sat at barriers and 24 % behind the staging loads (col -> a_src: dependent DRAM round trips per tile).
// Here one persistent CTA per SM is warp-specialised and nothing in it meets at a CTA barrier:
//   window warp     per tile, in tile order: waits for a free window stage (ring of 3) and issues one cp.async.bulk
//                   per window row (H*C*4 = 768 B each; rows of a column slice of the projection output: pitch ldxs,
//                   16-byte aligned) onto the stage's `win_full` mbarrier: the HBM/L2 stream of a tile costs ~50
//                   instructions instead of 300 loads and runs up to three tiles ahead of the gather;
//   edge producers  6 warps, each with its OWN edge stage and every 6th tile (so six tiles are being staged while one
//                   is gathered: the DRAM round trips overlap): lanes = edges: neighbour id -> byte offset of its row
//                   in the window (or -1 - id for a source outside the staged rows: that edge is gathered from global
//                   memory, so ANY graph is handled) and the raw logits of all heads as one float4; then lanes =
//                   rows: softmax in three passes over the row's float4 logits; arrive on the stage's `edge_full`;
//   consumers       8 warps, tile by tile: wait `edge_full` and `win_full`; 8 threads per row, each owning float4
//                   chunks q and q + 8 of the 64 output columns: per edge one LDS of the offset, one LDS.128 of the
//                   weights, 2 H LDS.128 of the source row, 8 H FFMA (same accumulation order as v6: the two kernels
//                   agree bit for bit); one lane per warp releases both stages, then the epilogue runs from registers.
// All relations of a tile share one edge pool and one window pool (packed per tile).
// Tiles whose edges exceed the pool (hub rows) take v6's warp-per-row path in the consumer warps.
// ---------------------------------------------------------------------------------------------------
#ifndef HG_WIN_PRODUCERS
#define HG_WIN_PRODUCERS 6
#endif
#ifndef HG_WIN_STAGES
#define HG_WIN_STAGES 3
#endif
#ifndef HG_WIN_BYTES
#define HG_WIN_BYTES 49152  // 64 rows at H = 3 (98.5 % of the 32-row tiles of an MP3D-shaped batch fit)
#endif
#ifndef HG_WIN_PREDICATE
#define HG_WIN_PREDICATE 1   // gather loads predicated on i < n_edges (0: clamped index, unconditional loads)
#endif
#ifndef HG_WIN_SLEEP
#define HG_WIN_SLEEP 128    // ns of back-off in the producers' stage waits (0: plain spin)
#endif
#ifndef HG_WIN_KC
#define HG_WIN_KC 2  // float4 chunks per gather thread: 1 = 16 threads per row, 16 gather warps; 2 = 8 and 8
#endif
#ifndef HG_WIN_GROUPS
#define HG_WIN_GROUPS 1  // gather warps form this many groups that take alternate tiles (each tile in 32 / rows-per-pass passes)
#endif
constexpr int kWinKC = HG_WIN_KC, kWinTPR = 16 / kWinKC, kWinGroups = HG_WIN_GROUPS;
#ifndef HG_WIN_ESTAGES
#define HG_WIN_ESTAGES 8  // edge stages: more than producer warps, so a producer never waits for the gather of its last tile
#endif
constexpr int kWinConsumerWarps = kWinTPR, kWinProducerWarps = HG_WIN_PRODUCERS, kWinStages = HG_WIN_STAGES;
constexpr int kWinEdgeStages = HG_WIN_ESTAGES;
constexpr int kWinThreads = (kWinConsumerWarps + kWinProducerWarps + 1) * 32;  // 736 (KC = 1) or 480
constexpr int kWinMaxRegs = 65536 / kWinThreads / 8 * 8 > 128 ? 128 : 65536 / kWinThreads / 8 * 8;  // one CTA per SM
constexpr int kWinC = 64;               // channels per head (8 threads x 2 float4 chunks)
constexpr int kWinBytes = HG_WIN_BYTES;   // window pool of one stage
constexpr int kWinEdges = 320;          // edge pool of one stage (all relations of a tile)
constexpr int kWinMaxRel = 4;
constexpr int kWinBox = 4;             // source rows per TMA box
constexpr int kWinBars = 2 * kWinEdgeStages + 2 * kWinStages;
constexpr int kWinGroupWarps = kWinConsumerWarps / kWinGroups;     // gather warps per group
constexpr int kWinPassRows = kWinGroupWarps * 32 / kWinTPR;        // destination rows a group gathers at a time
constexpr int kWinPasses = kMaxTile / kWinPassRows;
static_assert(kWinConsumerWarps % kWinGroups == 0 && kMaxTile % kWinPassRows == 0, "groups must tile the 32-row tile");

// one edge stage: weights float4[ecap] | destination logits float4[rmax][32] | offsets int[ecap] |
// pool-relative rowptr int[rmax][33] (padded to 16 B) | control int[4] {hub, window rows, any source outside}
__host__ __device__ inline uint32_t win_edge_stage_bytes(int ecap, int rmax) {
    return static_cast<uint32_t>(ecap * 16 + rmax * kMaxTile * 16 + ecap * 4 + (rmax * (kMaxTile + 1) * 4 + 15) / 16 * 16 + 16);
}
__host__ __device__ inline size_t win_smem_bytes(int wrows, int rowb, int ecap, int rmax) {
    return static_cast<size_t>(kWinStages) * wrows * rowb + static_cast<size_t>(kWinEdgeStages) * win_edge_stage_bytes(ecap, rmax) +
           kWinC * 4 + kWinBars * 8 + 16;
}

__device__ __forceinline__ float ex2_ftz(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// edge ranges and window slices of a tile, packed (uniform loads; every role computes the same values)
template <int ROWB, int RMAX>
struct WinTile {
    int e0[RMAX], ne[RMAX], eb[RMAX], wlo[RMAX], wn[RMAX], wb[RMAX];
    int tot, used;
    __device__ __forceinline__ void load(const hg_gat_args& A, int64_t tile, int64_t t0, int rows, int wrows) {
        tot = 0;
        used = 0;
#pragma unroll
        for (int r = 0; r < RMAX; ++r) {
            e0[r] = ne[r] = eb[r] = wlo[r] = wn[r] = wb[r] = 0;
            if (r < A.n_rel) {
                const hg_gat_rel& Rl = A.rel[r];
                e0[r] = __ldg(Rl.rowptr + t0);
                ne[r] = __ldg(Rl.rowptr + t0 + rows) - e0[r];
                eb[r] = tot;
                tot += ne[r];
                if (Rl.tile_win) {
                    const int2 w = __ldg(reinterpret_cast<const int2*>(Rl.tile_win) + tile);
                    const int need = (w.y + kWinBox - 1) / kWinBox * kWinBox;  // whole TMA boxes (wrows is a multiple too)
                    wlo[r] = w.x;
                    wn[r] = need < wrows - used ? need : wrows - used;
                    wb[r] = used * ROWB;
                    used += wn[r];
                }
            }
        }
    }
};

struct WinParams : GatParams {
    alignas(64) CUtensorMap tmap[kWinMaxRel];  // xs of relation r as a [n_src, H*C] fp32 tensor, box = kWinBox rows
};

// one box of kWinBox source rows -> shared memory, completing on `bar` (SASS: UTMALDG)
__device__ __forceinline__ void tma_rows_g2s(uint32_t dst, const CUtensorMap* map, int row0, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(dst), "l"(map), "r"(0), "r"(row0), "r"(bar)
        : "memory");
}
// waits that may last a whole tile: back off so the spinning warp does not take issue slots from the gather warps
__device__ __forceinline__ void mbar_wait_idle(uint32_t bar, uint32_t parity) {
    if (tc::mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    while (!tc::mbar_try_wait(bar, parity)) {
        if (HG_WIN_SLEEP > 0) __nanosleep(HG_WIN_SLEEP);
        if (clock64() - t0 > tc::kWaitCycles) __trap();
    }
}

#ifdef HG_WIN_TRACE
// debug build only (tools/ab_build.sh ... -DHG_WIN_TRACE): clock64 stamps of CTA 0's roles, per tile
constexpr int kTraceTiles = 64, kTraceEv = 12;
__device__ long long g_win_trace[kTraceTiles * kTraceEv];
#define WIN_STAMP(k, ev) do { if (blockIdx.x == 0 && (threadIdx.x & 31) == 0 && (k) < kTraceTiles) g_win_trace[(k) * kTraceEv + (ev)] = clock64(); } while (0)
#else
#define WIN_STAMP(k, ev) do { } while (0)
#endif

template <int HT, int RMAX>
__global__ void __launch_bounds__(kWinThreads, 1)
k_gat_window(const __grid_constant__ WinParams P) {
    extern __shared__ __align__(128) uint8_t s_raw[];
    constexpr int ROWB = HT * kWinC * 4;  // bytes of one window row
    const hg_gat_args& A = P.a;
    const int R = A.n_rel;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t win_bytes = static_cast<uint32_t>(P.wrows) * ROWB;
    const uint32_t est_bytes = win_edge_stage_bytes(P.cap, RMAX);
    uint8_t* const edge0 = s_raw + static_cast<size_t>(kWinStages) * win_bytes;
    uint8_t* const tail = edge0 + static_cast<size_t>(kWinEdgeStages) * est_bytes;
    float* const s_bias = reinterpret_cast<float*>(tail);
    uint64_t* const s_bar = reinterpret_cast<uint64_t*>(tail + kWinC * 4);
    const uint32_t bar_a = tc::smem_u32(s_bar);
    const uint32_t edge_full = bar_a, edge_empty = bar_a + 8 * kWinEdgeStages;
    const uint32_t win_full = bar_a + 16 * kWinEdgeStages, win_empty = win_full + 8 * kWinStages;
    // offsets inside an edge stage
    const uint32_t o_ad = static_cast<uint32_t>(P.cap) * 16;
    const uint32_t o_off = o_ad + RMAX * kMaxTile * 16;
    const uint32_t o_rp = o_off + static_cast<uint32_t>(P.cap) * 4;
    const uint32_t o_ctl = o_rp + (RMAX * (kMaxTile + 1) * 4 + 15) / 16 * 16;

    if (tid < kWinC) {
        float b = 0.f;
        for (int r = 0; r < R; ++r)
            if (A.rel[r].bias) b += __ldg(A.rel[r].bias + tid);
        s_bias[tid] = b;
    }
    if (tid == 0) {
        for (int s = 0; s < kWinEdgeStages; ++s) {
            tc::mbar_init(s_bar + s, 1);                                        // edge_full: lane 0 of the producer warp
            tc::mbar_init(s_bar + kWinEdgeStages + s, kWinGroupWarps);          // edge_empty: one lane per warp of the owning group
        }
        for (int s = 0; s < kWinStages; ++s) {
            tc::mbar_init(s_bar + 2 * kWinEdgeStages + s, 1);                // win_full: one arrive (+ the copies' bytes)
            tc::mbar_init(s_bar + 2 * kWinEdgeStages + kWinStages + s, kWinGroupWarps);  // win_empty
        }
        tc::fence_barrier_init();
    }
    __syncthreads();
    const int64_t n_tiles = ceil_div<int64_t>(A.n_dst, kMaxTile);
    const int n_k = static_cast<int>((n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x);  // tiles of this CTA

    if (warp == kWinConsumerWarps + kWinProducerWarps) {
        // ================================================================== window warp
        int ws = 0;
        uint32_t wph = 0;
        WinTile<ROWB, RMAX> Tn;  // control of the NEXT tile: loaded one iteration ahead, so its DRAM round trip
        {                        // overlaps the wait for a free stage and the copies of the current tile
            const int64_t t0 = static_cast<int64_t>(blockIdx.x) * kMaxTile;
            Tn.load(A, blockIdx.x, t0, (A.n_dst - t0 < kMaxTile) ? static_cast<int>(A.n_dst - t0) : kMaxTile, P.wrows);
        }
        for (int k = 0; k < n_k; ++k) {
            const WinTile<ROWB, RMAX> T = Tn;
            if (k + 1 < n_k) {
                const int64_t tile = blockIdx.x + static_cast<int64_t>(k + 1) * gridDim.x, t0 = tile * kMaxTile;
                Tn.load(A, tile, t0, (A.n_dst - t0 < kMaxTile) ? static_cast<int>(A.n_dst - t0) : kMaxTile, P.wrows);
            }
            WIN_STAMP(k, 0);
            mbar_wait_idle(win_empty + 8 * ws, wph ^ 1u);
            WIN_STAMP(k, 1);
            if (lane == 0) {
                if (T.tot > P.cap || T.used == 0) {  // hub tile / no window: complete the phase without bytes
                    tc::mbar_arrive(win_full + 8 * ws);
                } else {
                    tc::mbar_arrive_expect_tx(win_full + 8 * ws, static_cast<uint32_t>(T.used) * ROWB);
                    const uint32_t win_a = tc::smem_u32(s_raw) + ws * win_bytes;
#pragma unroll
                    for (int r = 0; r < RMAX; ++r) {
                        if (r < R) {
                            for (int row = 0; row < T.wn[r]; row += kWinBox)
                                tma_rows_g2s(win_a + T.wb[r] + row * ROWB, &P.tmap[r], T.wlo[r] + row, win_full + 8 * ws);
                        }
                    }
                }
            }
            __syncwarp();
            if (++ws == kWinStages) ws = 0, wph ^= 1u;
        }
        return;
    }

    if (warp >= kWinConsumerWarps) {
        // ================================================================== edge producers: warp p, tiles k = p, p + 6, ...
        const int p = warp - kWinConsumerWarps;
        auto load_tile = [&](WinTile<ROWB, RMAX>& T, int k) {
            const int64_t tile = blockIdx.x + static_cast<int64_t>(k) * gridDim.x, t0 = tile * kMaxTile;
            T.load(A, tile, t0, (A.n_dst - t0 < kMaxTile) ? static_cast<int>(A.n_dst - t0) : kMaxTile, P.wrows);
        };
        // neighbour ids of pool entries base + u * 32 + lane, u < 8 (-1 past the end): all eight loads in flight at once
        auto load_cols = [&](const WinTile<ROWB, RMAX>& T, int base, int32_t (&j)[8]) {
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int e = base + u * 32 + lane;
                j[u] = -1;
                if (e < T.tot && T.tot <= P.cap) {
                    const int32_t* colp = A.rel[0].col + T.e0[0];
#pragma unroll
                    for (int r = 1; r < RMAX; ++r)
                        if (r < R && e >= T.eb[r]) colp = A.rel[r].col + (T.e0[r] - T.eb[r]);
                    j[u] = __ldg(colp + e);
                }
            }
        };
        WinTile<ROWB, RMAX> Tn;  // control of this warp's NEXT tile, loaded one tile ahead ...
        int32_t jn[8];           // ... and the first 256 neighbour ids of it (one DRAM round trip less per tile)
        if (p < n_k) {
            load_tile(Tn, p);
            load_cols(Tn, 0, jn);
        }
        for (int k = p; k < n_k; k += kWinProducerWarps) {
            const int es = k % kWinEdgeStages;  // the tile's edge stage (a ring longer than the number of producer warps)
            const uint32_t eph = (k / kWinEdgeStages) & 1;
            uint8_t* const st = edge0 + static_cast<size_t>(es) * est_bytes;
            float4* const s_w = reinterpret_cast<float4*>(st);
            float4* const s_ad = reinterpret_cast<float4*>(st + o_ad);
            int32_t* const s_off = reinterpret_cast<int32_t*>(st + o_off);
            int32_t* const s_rp = reinterpret_cast<int32_t*>(st + o_rp);
            int32_t* const s_ctl = reinterpret_cast<int32_t*>(st + o_ctl);
            const int64_t tile = blockIdx.x + static_cast<int64_t>(k) * gridDim.x, t0 = tile * kMaxTile;
            const int rows = (A.n_dst - t0 < kMaxTile) ? static_cast<int>(A.n_dst - t0) : kMaxTile;
            const WinTile<ROWB, RMAX> T = Tn;
            int32_t jc[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) jc[u] = jn[u];
            const bool has_next = k + kWinProducerWarps < n_k;
            if (has_next) load_tile(Tn, k + kWinProducerWarps);
            WIN_STAMP(k, 2);
            mbar_wait_idle(edge_empty + 8 * es, eph ^ 1u);  // stage free (the first pass is on the fresh barrier)
            WIN_STAMP(k, 3);
            if (T.tot > P.cap) {  // hub tile (uniform): the consumers stream it from global memory
                if (lane == 0) {
                    s_ctl[0] = 1, s_ctl[1] = 0, s_ctl[2] = 0;
                    tc::mbar_arrive(edge_full + 8 * es);
                }
                __syncwarp();
                if (has_next) load_cols(Tn, 0, jn);
                continue;
            }
            // ---- stage.  One DRAM round trip per tile on the critical path, every load of the tile in flight at once:
            //   (prefetched during the previous tile of this warp: tile control and the neighbour ids, lanes = pool
            //   entries, 8 per lane)
            //   trip    rowptr slices, destination logits and the source logits of those neighbours.
            constexpr int kRpIt = (RMAX * (kMaxTile + 1) + 31) / 32;
            int32_t rpv[kRpIt];
#pragma unroll
            for (int it = 0; it < kRpIt; ++it) {
                const int idx = it * 32 + lane;
                rpv[it] = 0;
                if (idx < R * (kMaxTile + 1)) {
                    const int r = idx / (kMaxTile + 1), tt = idx - r * (kMaxTile + 1);
                    int d = 0;
#pragma unroll
                    for (int rr = 0; rr < RMAX; ++rr)
                        if (rr == r) d = T.eb[rr] - T.e0[rr];
                    rpv[it] = __ldg(A.rel[r].rowptr + t0 + min(tt, rows)) + d;
                }
            }
            float adv[RMAX][4];
#pragma unroll
            for (int r = 0; r < RMAX; ++r) {
#pragma unroll
                for (int h = 0; h < 4; ++h) adv[r][h] = 0.f;
                if (r < R && lane < rows) {
                    const hg_gat_rel& Rl = A.rel[r];
                    const float* ap = Rl.a_dst + (t0 + lane) * Rl.lda_dst;
#pragma unroll
                    for (int h = 0; h < HT; ++h) adv[r][h] = __ldg(ap + h);
                }
            }
            bool miss = false;
            for (int base = 0; base < T.tot; base += 8 * 32) {
                int32_t j[8];
                int rel_of[8];
                if (base == 0) {
#pragma unroll
                    for (int u = 0; u < 8; ++u) j[u] = jc[u];
                } else {
                    load_cols(T, base, j);
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int e = base + u * 32 + lane;  // pool entry
                    rel_of[u] = 0;
#pragma unroll
                    for (int r = 1; r < RMAX; ++r)
                        if (r < R && e >= T.eb[r]) rel_of[u] = r;
                }
                float v[8][HT];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
#pragma unroll
                    for (int h = 0; h < HT; ++h) v[u][h] = 0.f;
                    if (j[u] >= 0) {
                        const hg_gat_rel& Rl = A.rel[rel_of[u]];
                        const float* src = Rl.a_src + static_cast<int64_t>(j[u]) * Rl.lda_src;
#pragma unroll
                        for (int h = 0; h < HT; ++h) v[u][h] = __ldg(src + h);
                        if (Rl.a_edge) {
                            const int e = base + u * 32 + lane;
                            int d = 0;
#pragma unroll
                            for (int rr = 0; rr < RMAX; ++rr)
                                if (rr == rel_of[u]) d = T.e0[rr] - T.eb[rr];
                            const float* ae = Rl.a_edge + static_cast<int64_t>(e + d) * HT;
#pragma unroll
                            for (int h = 0; h < HT; ++h) v[u][h] += __ldg(ae + h);
                        }
                    }
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    if (j[u] >= 0) {
                        const int e = base + u * 32 + lane;
                        int wlo = 0, wn = 0, wb = 0;
#pragma unroll
                        for (int rr = 0; rr < RMAX; ++rr)
                            if (rr == rel_of[u]) wlo = T.wlo[rr], wn = T.wn[rr], wb = T.wb[rr];
                        const uint32_t d = static_cast<uint32_t>(j[u] - wlo);
                        const bool in = d < static_cast<uint32_t>(wn);
                        miss = miss || !in;
                        s_off[e] = in ? static_cast<int32_t>(wb + d * ROWB) : -1 - j[u];
                        s_w[e] = make_float4(v[u][0], HT > 1 ? v[u][HT > 1 ? 1 : 0] : 0.f, HT > 2 ? v[u][HT > 2 ? 2 : 0] : 0.f,
                                             HT > 3 ? v[u][HT > 3 ? 3 : 0] : 0.f);
                    }
                }
            }
#pragma unroll
            for (int it = 0; it < kRpIt; ++it)
                if (it * 32 + lane < R * (kMaxTile + 1)) s_rp[it * 32 + lane] = rpv[it];
#pragma unroll
            for (int r = 0; r < RMAX; ++r)
                if (r < R && lane < rows) s_ad[r * kMaxTile + lane] = make_float4(adv[r][0], adv[r][1], adv[r][2], adv[r][3]);
            if (has_next) load_cols(Tn, 0, jn);  // in flight during the softmax below
            WIN_STAMP(k, 4);
            const bool any_miss = __any_sync(0xffffffffu, miss);
            __syncwarp();  // the stores above are ordered before the reads below (other lanes' edges)

            // ---- softmax in place: lanes = rows, relation by relation; all heads of an edge are one float4.
            // Each pass handles the row's edges in pairs (both loads first), halving the serialised shared-memory latency.
#pragma unroll
            for (int r = 0; r < RMAX; ++r) {
                if (r < R && lane < rows) {
                    const int beg = s_rp[r * (kMaxTile + 1) + lane], end = s_rp[r * (kMaxTile + 1) + lane + 1];
                    const float4 ad4 = s_ad[r * kMaxTile + lane];
                    const float ad[4] = {ad4.x, ad4.y, ad4.z, ad4.w};
                    float m[HT], l[HT], inv[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int h = 0; h < HT; ++h) m[h] = -INFINITY, l[h] = 0.f;
                    auto sweep = [&](auto&& fn) {  // fn(float (&v)[4]) on every edge of the row, in order, two at a time
                        int e = beg;
                        for (; e + 1 < end; e += 2) {
                            const float4 a4 = s_w[e], b4 = s_w[e + 1];
                            float va[4] = {a4.x, a4.y, a4.z, a4.w}, vb[4] = {b4.x, b4.y, b4.z, b4.w};
                            fn(va);
                            fn(vb);
                            s_w[e] = make_float4(va[0], va[1], va[2], va[3]);
                            s_w[e + 1] = make_float4(vb[0], vb[1], vb[2], vb[3]);
                        }
                        if (e < end) {
                            const float4 a4 = s_w[e];
                            float va[4] = {a4.x, a4.y, a4.z, a4.w};
                            fn(va);
                            s_w[e] = make_float4(va[0], va[1], va[2], va[3]);
                        }
                    };
                    sweep([&](float (&v)[4]) {
#pragma unroll
                        for (int h = 0; h < HT; ++h) {
                            v[h] = leaky(v[h] + ad[h], A.neg_slope);
                            m[h] = fmaxf(m[h], v[h]);
                        }
                    });
                    sweep([&](float (&v)[4]) {
#pragma unroll
                        for (int h = 0; h < HT; ++h) {
                            v[h] = __expf(v[h] - m[h]);
                            l[h] += v[h];
                        }
                    });
#pragma unroll
                    for (int h = 0; h < HT; ++h) inv[h] = 1.f / (l[h] + 1e-16f);
                    sweep([&](float (&v)[4]) {
#pragma unroll
                        for (int h = 0; h < 4; ++h) v[h] *= inv[h];
                    });
                }
            }
            WIN_STAMP(k, 5);
            if (lane == 0) s_ctl[0] = 0, s_ctl[1] = T.used, s_ctl[2] = any_miss ? 1 : 0;
            __syncwarp();  // every lane's stage writes are ordered before lane 0's releasing arrive
            if (lane == 0) tc::mbar_arrive(edge_full + 8 * es);
        }
        return;
    }

    // ====================================================================== consumers
    // Groups of gather warps take alternate tiles, so one group's epilogue / barrier waits overlap the other group's
    // gather and the shared-memory port (the bound of the gather: 768 B per edge) never idles between tiles.
    const int grp = warp / kWinGroupWarps, wi = warp % kWinGroupWarps;
    const int q = lane % kWinTPR, row_in_pass = wi * (32 / kWinTPR) + lane / kWinTPR;  // first chunk / row within a pass
    const float gscale = (A.group == HG_GROUP_MEAN) ? 1.f / static_cast<float>(R) : 1.f;
    const float hscale = 1.f / static_cast<float>(HT);
    for (int k = grp; k < n_k; k += kWinGroups) {
        const int es = k % kWinEdgeStages, ws = k % kWinStages;
        const uint32_t eph = (k / kWinEdgeStages) & 1, wph = (k / kWinStages) & 1;
        const int64_t tile = blockIdx.x + static_cast<int64_t>(k) * gridDim.x, t0 = tile * kMaxTile;
        const int rows = (A.n_dst - t0 < kMaxTile) ? static_cast<int>(A.n_dst - t0) : kMaxTile;
        const uint8_t* const st = edge0 + static_cast<size_t>(es) * est_bytes;
        const float4* const s_w = reinterpret_cast<const float4*>(st);
        const int32_t* const s_off = reinterpret_cast<const int32_t*>(st + o_off);
        const int32_t* const s_rp = reinterpret_cast<const int32_t*>(st + o_rp);
        const int32_t* const s_ctl = reinterpret_cast<const int32_t*>(st + o_ctl);
        const uint32_t b_ef = edge_full + 8 * es, b_ee = edge_empty + 8 * es, b_wf = win_full + 8 * ws, b_we = win_empty + 8 * ws;
        const uint8_t* const wq = s_raw + static_cast<size_t>(ws) * win_bytes + q * 16;

        if (warp == 0) WIN_STAMP(k, 6);
        tc::mbar_wait(b_ef, eph);
        if (warp == 0) WIN_STAMP(k, 7);
        tc::mbar_wait(b_wf, wph);
        if (warp == 0) WIN_STAMP(k, 8);
        const int hub = s_ctl[0], any_miss = s_ctl[2];
        if (hub) {  // uniform: warp per row straight from global memory
            for (int tt = wi; tt < rows; tt += kWinGroupWarps) slow_row<4>(P, t0 + tt, lane, 0, kWinC);
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(b_ee), tc::mbar_arrive(b_we);
            continue;
        }
        for (int pass = 0; pass < kWinPasses; ++pass) {
        const int t = pass * kWinPassRows + row_in_pass;
        float acc[kWinKC][4];
#pragma unroll
        for (int kk = 0; kk < kWinKC; ++kk)
#pragma unroll
            for (int v = 0; v < 4; ++v) acc[kk][v] = 0.f;
        if (t < rows && !any_miss) {
            // Every source of the tile is in the window.  One software-pipelined pass over the row's edges of ALL
            // relations (pool entries; offsets are absolute in the window): the loads of edge i + 1 are issued before the
            // FFMAs of edge i, ping-ponging between two register sets, so a warp exposes one shared-memory latency per
            // row instead of one per edge.  Edge order = relation by relation, CSR order: v6's accumulation order.
            // i-th edge of the row -> pool entry i + delta[r], r = the relation whose cumulative range holds i (selects)
            int cum[RMAX], delta[RMAX], n_edges = 0;
#pragma unroll
            for (int r = 0; r < RMAX; ++r) {
                const int beg = r < R ? s_rp[r * (kMaxTile + 1) + t] : 0, end = r < R ? s_rp[r * (kMaxTile + 1) + t + 1] : 0;
                cum[r] = n_edges;
                delta[r] = beg - n_edges;
                n_edges += end - beg;
            }
            // Three-deep: while the FFMAs of edge i run, the source rows of edge i + 1 are in flight and the
            // offset / weights of edge i + 2 are being read -- no load waits on another load inside an iteration.
            // Loads are predicated on i < n_edges (a finished row = a whole quarter-warp: its LDS.128 requests cost no
            // shared-memory wavefronts); the index math is branch-free.
            auto fetch_ow = [&](int32_t& off, float4& w, int i) {
                int d = delta[0];
#pragma unroll
                for (int r = 1; r < RMAX; ++r)
                    if (i >= cum[r]) d = delta[r];
                if (!HG_WIN_PREDICATE) {
                    const int ic = i < n_edges ? i : n_edges - 1;
                    int dc = delta[0];
#pragma unroll
                    for (int r = 1; r < RMAX; ++r)
                        if (ic >= cum[r]) dc = delta[r];
                    off = s_off[ic + dc];
                    w = s_w[ic + dc];
                } else if (i < n_edges) {
                    off = s_off[i + d];
                    w = s_w[i + d];
                }
            };
            auto fetch_xs = [&](float4 (&val)[HT][kWinKC], int32_t off, bool on) {
                if (on || !HG_WIN_PREDICATE) {
#pragma unroll
                    for (int hh = 0; hh < HT; ++hh) {
#pragma unroll
                        for (int kk = 0; kk < kWinKC; ++kk)
                            val[hh][kk] = *reinterpret_cast<const float4*>(wq + off + hh * (kWinC * 4) + kk * (kWinTPR * 16));
                    }
                }
            };
            auto fma_edge = [&](const float4 (&val)[HT][kWinKC], const float4& w) {
                const float wv[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                for (int hh = 0; hh < HT; ++hh)
#pragma unroll
                    for (int kk = 0; kk < kWinKC; ++kk) {
                        acc[kk][0] = fmaf(wv[hh], val[hh][kk].x, acc[kk][0]);
                        acc[kk][1] = fmaf(wv[hh], val[hh][kk].y, acc[kk][1]);
                        acc[kk][2] = fmaf(wv[hh], val[hh][kk].z, acc[kk][2]);
                        acc[kk][3] = fmaf(wv[hh], val[hh][kk].w, acc[kk][3]);
                    }
            };
            if (n_edges > 0) {
                float4 va[HT][kWinKC], vb[HT][kWinKC], wa, wb, wn = make_float4(0.f, 0.f, 0.f, 0.f);
                int32_t offn = 0;
                fetch_ow(offn, wn, 0);
                fetch_xs(va, offn, true);
                wa = wn;
                fetch_ow(offn, wn, 1);
                for (int i = 0;; i += 2) {
                    fetch_xs(vb, offn, i + 1 < n_edges);
                    wb = wn;
                    fetch_ow(offn, wn, i + 2);
                    fma_edge(va, wa);
                    if (i + 1 >= n_edges) break;
                    fetch_xs(va, offn, i + 2 < n_edges);
                    wa = wn;
                    fetch_ow(offn, wn, i + 3);
                    fma_edge(vb, wb);
                    if (i + 2 >= n_edges) break;
                }
            }
        } else if (t < rows) {
#pragma unroll
            for (int r = 0; r < RMAX; ++r) {
                if (r < R) {
                    const int beg = s_rp[r * (kMaxTile + 1) + t], end = s_rp[r * (kMaxTile + 1) + t + 1];
                    {
                        const char* xg = reinterpret_cast<const char*>(A.rel[r].xs) + q * 16;
                        const uint32_t pitch = static_cast<uint32_t>(A.rel[r].ldxs) * 4u;
                        for (int e = beg; e < end; ++e) {
                            const int32_t off = s_off[e];
                            const float4 wa = s_w[e];
                            const float wv[4] = {wa.x, wa.y, wa.z, wa.w};
                            float4 val[HT][kWinKC];
                            if (off >= 0) {
#pragma unroll
                                for (int hh = 0; hh < HT; ++hh) {
#pragma unroll
                                    for (int kk = 0; kk < kWinKC; ++kk)
                                        val[hh][kk] = *reinterpret_cast<const float4*>(wq + off + hh * (kWinC * 4) + kk * (kWinTPR * 16));
                                }
                            } else {  // source outside the staged window
                                const char* row = xg + static_cast<uint64_t>(static_cast<uint32_t>(-1 - off)) * pitch;
#pragma unroll
                                for (int hh = 0; hh < HT; ++hh) {
#pragma unroll
                                    for (int kk = 0; kk < kWinKC; ++kk)
                                        val[hh][kk] = __ldg(reinterpret_cast<const float4*>(row + hh * (kWinC * 4) + kk * (kWinTPR * 16)));
                                }
                            }
#pragma unroll
                            for (int hh = 0; hh < HT; ++hh)
#pragma unroll
                                for (int kk = 0; kk < kWinKC; ++kk) {
                                    acc[kk][0] = fmaf(wv[hh], val[hh][kk].x, acc[kk][0]);
                                    acc[kk][1] = fmaf(wv[hh], val[hh][kk].y, acc[kk][1]);
                                    acc[kk][2] = fmaf(wv[hh], val[hh][kk].z, acc[kk][2]);
                                    acc[kk][3] = fmaf(wv[hh], val[hh][kk].w, acc[kk][3]);
                                }
                        }
                    }
                }
            }
        }
        if (pass == kWinPasses - 1) {
            __syncwarp();
            if (warp == 0) WIN_STAMP(k, 9);
            if (lane == 0) tc::mbar_arrive(b_ee), tc::mbar_arrive(b_we);  // release: this warp no longer reads the stages
        }
        if (t >= rows) continue;
        // ---- epilogue: head mean, summed biases, group mean, addend, activation, one store per chunk
        // (FMUL then FADD like v6's SASS, so the two kernels round identically)
        const int64_t i = t0 + t;
        const float* __restrict__ addp = A.addend ? A.addend + i * A.ldadd : nullptr;
        float* __restrict__ outp = A.out + i * A.ldout;
#pragma unroll
        for (int kk = 0; kk < kWinKC; ++kk) {
            const int oc = (q + kk * kWinTPR) * 4;
            const float4 b4 = *reinterpret_cast<const float4*>(s_bias + oc);
            const float bv[4] = {b4.x, b4.y, b4.z, b4.w};
            float o[4];
#pragma unroll
            for (int v = 0; v < 4; ++v) o[v] = __fmul_rn(__fadd_rn(__fmul_rn(acc[kk][v], hscale), bv[v]), gscale);
            if (addp) {
                float av[4];
                ldg_vec<4>(addp + oc, av);
#pragma unroll
                for (int v = 0; v < 4; ++v) o[v] = __fadd_rn(o[v], av[v]);
            }
            if (A.act == HG_ACT_RELU) {
#pragma unroll
                for (int v = 0; v < 4; ++v) o[v] = fmaxf(o[v], 0.f);
            } else if (A.act == HG_ACT_ELU) {
                // exp(x) - 1 for x <= 0 as one MUFU.EX2: equal to v6's __expf(x) - 1 (they differ only where exp(x) is
                // denormal, and there both give -1)
#pragma unroll
                for (int v = 0; v < 4; ++v) o[v] = o[v] > 0.f ? o[v] : ex2_ftz(__fmul_rn(o[v], 1.4426950408889634f)) - 1.f;
            }
            st_vec<4>(outp + oc, o);
        }
        }  // pass
        if (warp == 0) WIN_STAMP(k, 10);
    }
}

#ifdef HG_WIN_TRACE
}  // namespace
}  // namespace hg
extern "C" __attribute__((visibility("default"))) int hg_debug_win_trace(long long* host, int n) {
    return static_cast<int>(cudaMemcpyFromSymbol(host, hg::g_win_trace, sizeof(long long) * n));
}
namespace hg {
namespace {
#endif

constexpr int kWinFallback = 1;  // internal: the window kernel does not apply after all, use v6

// cuTensorMapEncodeTiled through the runtime's driver entry point: the library must not NEED libcuda.so.1 at load time
// (it is dlopen'ed on machines without a driver to check its exports).
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled() {
    static const EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess) {
            (void)cudaGetLastError();
            p = nullptr;
        }
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

template <int HT, int RMAX>
int launch_window(GatParams& G, cudaStream_t stream) {
    constexpr int ROWB = HT * kWinC * 4;
    WinParams P;
    static_cast<GatParams&>(P) = G;
    P.tile_rows = kMaxTile;
    P.hs = 4;
    P.n_bias = kWinC;
    P.cap = kWinEdges;
    P.wrows = kWinBytes / ROWB / kWinBox * kWinBox;
    const EncodeTiledFn encode = encode_tiled();
    if (!encode) return kWinFallback;
    for (int r = 0; r < kWinMaxRel; ++r) {
        if (r >= P.a.n_rel) {
            P.tmap[r] = P.tmap[0];
            continue;
        }
        const hg_gat_rel& Rl = P.a.rel[r];
        if (Rl.n_src <= 0 || !Rl.xs || !aligned(Rl.xs, 16) || (Rl.ldxs * 4) % 16 != 0) return kWinFallback;
        const cuuint64_t dims[2] = {static_cast<cuuint64_t>(HT * kWinC), static_cast<cuuint64_t>(Rl.n_src)};
        const cuuint64_t strides[1] = {static_cast<cuuint64_t>(Rl.ldxs) * 4};
        const cuuint32_t box[2] = {static_cast<cuuint32_t>(HT * kWinC), kWinBox};
        const cuuint32_t estr[2] = {1, 1};
        const CUresult cr = encode(&P.tmap[r], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(Rl.xs), dims,
                                                   strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (cr != CUDA_SUCCESS) return kWinFallback;
    }
    const size_t smem = win_smem_bytes(P.wrows, ROWB, P.cap, RMAX);
    HG_CUDA_TRY(cudaFuncSetAttribute(k_gat_window<HT, RMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int64_t need = ceil_div<int64_t>(P.a.n_dst, kMaxTile);
    k_gat_window<HT, RMAX><<<static_cast<unsigned>(need < sms ? need : sms), kWinThreads, smem, stream>>>(P);  // persistent
    return check_launch();
}

template <int RMAX>
int dispatch_window(GatParams& P, cudaStream_t stream) {
    switch (P.a.H) {
        case 1: return launch_window<1, RMAX>(P, stream);
        case 2: return launch_window<2, RMAX>(P, stream);
        case 3: return launch_window<3, RMAX>(P, stream);
        case 4: return launch_window<4, RMAX>(P, stream);
        default: return HG_EUNSUPPORTED;
    }
}

// HG_GAT_WINDOW=0 keeps the v6 kernel (A/B runs)
bool window_enabled() {
    static const bool on = [] {
        const char* e = std::getenv("HG_GAT_WINDOW");
        return !(e && e[0] == '0');
    }();
    return on;
}

// Aggregate-first over wide, 8-byte aligned source rows (the 306-d input features: 153 float2 chunks, 5 per lane).
template <int HT>
int dispatch_wide_shared(GatParams& P, cudaStream_t stream) {
    const int nco = P.f_count / 2;
    if (nco <= 96) return launch<2, 32, 3, 2, HT>(P, stream);
    if (nco <= 128) return launch<2, 32, 4, 2, HT>(P, stream);
    if (nco <= 160) return launch<2, 32, 5, 2, HT>(P, stream);
    return HG_EUNSUPPORTED;
}

// ---------------------------------------------------------------------------------------------------
// Small destination sets (<= 65,536 rows: the rooms side of a scene-graph batch, batch-1 serving) -- head-mean, C = 64,
// H <= 4, inference.  The tile kernels spend a launch like this on pipeline fill (the window kernel: 15 k cycles before
// the first gather) and on 4 tiles per CTA; here 8 threads own a row from rowptr to the store, no shared memory, no
// barrier, every row of the launch resident at once, so the launch costs three dependent L2 round trips.  Each lane
// walks the row's edges itself (the loads are warp-broadcast L1 hits), in v6's order: same bits as the tile kernels.
// A warp that holds a row of more than 64 edges takes the warp-per-row path for its four rows.
// ---------------------------------------------------------------------------------------------------
constexpr int kRowsKernelMax = 65536;

// VEC = 4 covers C <= 64 in multiples of 4 (the hidden layers: C = 64), VEC = 2 even C <= 32 (the output layer: C = 26):
// thread q of a row owns chunks q and q + 8 of the C / VEC vector chunks.
template <int HT, int VEC>
__global__ void __launch_bounds__(256) k_gat_rows(const __grid_constant__ GatParams P) {
    const hg_gat_args& A = P.a;
    const int R = A.n_rel, C = A.C;
    const int tid = threadIdx.x, lane = tid & 31, q = tid & 7;
    const int64_t i = static_cast<int64_t>(blockIdx.x) * 32 + (tid >> 3);
    const bool valid = i < A.n_dst;
    int maxdeg = 0;
    if (valid)
        for (int r = 0; r < R; ++r) {
            const int d = __ldg(A.rel[r].rowptr + i + 1) - __ldg(A.rel[r].rowptr + i);
            maxdeg = d > maxdeg ? d : maxdeg;
        }
    if (__any_sync(0xffffffffu, maxdeg > 64)) {  // warp-uniform
        for (int rr = 0; rr < 4; ++rr) {
            const int64_t ii = static_cast<int64_t>(blockIdx.x) * 32 + (tid >> 5) * 4 + rr;
            if (ii < A.n_dst) slow_row<VEC>(P, ii, lane, 0, C);
        }
        return;
    }
    if (!valid) return;
    const bool live[2] = {q * VEC < C, (q + 8) * VEC < C};
    float acc[2][VEC];
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
        for (int v = 0; v < VEC; ++v) acc[k][v] = 0.f;
    for (int r = 0; r < R; ++r) {
        const hg_gat_rel& Rl = A.rel[r];
        const int32_t beg = __ldg(Rl.rowptr + i), end = __ldg(Rl.rowptr + i + 1);
        if (end <= beg) continue;
        float ad[HT], m[HT], l[HT], inv[HT];
#pragma unroll
        for (int h = 0; h < HT; ++h) ad[h] = __ldg(Rl.a_dst + i * Rl.lda_dst + h), m[h] = -INFINITY, l[h] = 0.f;
        auto score = [&](int32_t e, int32_t j, float (&sc)[HT]) {
#pragma unroll
            for (int h = 0; h < HT; ++h) {
                float v = __ldg(Rl.a_src + static_cast<int64_t>(j) * Rl.lda_src + h);
                if (Rl.a_edge) v += __ldg(Rl.a_edge + static_cast<int64_t>(e) * HT + h);
                sc[h] = leaky(v + ad[h], A.neg_slope);
            }
        };
        for (int32_t e = beg; e < end; ++e) {
            float sc[HT];
            score(e, __ldg(Rl.col + e), sc);
#pragma unroll
            for (int h = 0; h < HT; ++h) m[h] = fmaxf(m[h], sc[h]);
        }
        for (int32_t e = beg; e < end; ++e) {
            float sc[HT];
            score(e, __ldg(Rl.col + e), sc);
#pragma unroll
            for (int h = 0; h < HT; ++h) l[h] += __expf(sc[h] - m[h]);
        }
#pragma unroll
        for (int h = 0; h < HT; ++h) inv[h] = 1.f / (l[h] + 1e-16f);
        const float* __restrict__ xg = Rl.xs + q * VEC;
        const int64_t ld = Rl.ldxs;
#pragma unroll 2
        for (int32_t e = beg; e < end; ++e) {
            const int32_t j = __ldg(Rl.col + e);
            const float* __restrict__ row = xg + static_cast<int64_t>(j) * ld;
            float val[HT][2][VEC];
#pragma unroll
            for (int hh = 0; hh < HT; ++hh)
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    if (live[k]) {
                        ldg_vec<VEC>(row + hh * C + 8 * k * VEC, val[hh][k]);
                    } else {
#pragma unroll
                        for (int v = 0; v < VEC; ++v) val[hh][k][v] = 0.f;
                    }
                }
            float sc[HT], wv[HT];
            score(e, j, sc);
#pragma unroll
            for (int h = 0; h < HT; ++h) wv[h] = __expf(sc[h] - m[h]) * inv[h];
#pragma unroll
            for (int hh = 0; hh < HT; ++hh)
#pragma unroll
                for (int k = 0; k < 2; ++k)
#pragma unroll
                    for (int v = 0; v < VEC; ++v) acc[k][v] = fmaf(wv[hh], val[hh][k][v], acc[k][v]);
        }
    }
    const float gscale = (A.group == HG_GROUP_MEAN) ? 1.f / static_cast<float>(R) : 1.f;
    const float hscale = 1.f / static_cast<float>(HT);
    const float* __restrict__ addp = A.addend ? A.addend + i * A.ldadd : nullptr;
    float* __restrict__ outp = A.out + i * A.ldout;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        if (!live[k]) continue;
        const int oc = (q + 8 * k) * VEC;
        float bs[VEC];  // summed biases of the relations, in relation order (as the tile kernels stage them)
#pragma unroll
        for (int v = 0; v < VEC; ++v) bs[v] = 0.f;
        for (int r = 0; r < R; ++r)
            if (A.rel[r].bias) {
                float b[VEC];
                ldg_vec<VEC>(A.rel[r].bias + oc, b);
#pragma unroll
                for (int v = 0; v < VEC; ++v) bs[v] += b[v];
            }
        float o[VEC];
#pragma unroll
        for (int v = 0; v < VEC; ++v) o[v] = __fmul_rn(__fadd_rn(__fmul_rn(acc[k][v], hscale), bs[v]), gscale);
        if (addp) {
            float av[VEC];
            ldg_vec<VEC>(addp + oc, av);
#pragma unroll
            for (int v = 0; v < VEC; ++v) o[v] = __fadd_rn(o[v], av[v]);
        }
        if (A.act == HG_ACT_RELU) {
#pragma unroll
            for (int v = 0; v < VEC; ++v) o[v] = fmaxf(o[v], 0.f);
        } else if (A.act == HG_ACT_ELU) {
#pragma unroll
            for (int v = 0; v < VEC; ++v) o[v] = o[v] > 0.f ? o[v] : ex2_ftz(__fmul_rn(o[v], 1.4426950408889634f)) - 1.f;
        }
        st_vec<VEC>(outp + oc, o);
    }
}

template <int VEC>
int launch_rows(GatParams& P, cudaStream_t stream) {
    const unsigned grid = static_cast<unsigned>(ceil_div<int64_t>(P.a.n_dst, 32));
    switch (P.a.H) {
        case 1: k_gat_rows<1, VEC><<<grid, 256, 0, stream>>>(P); break;
        case 2: k_gat_rows<2, VEC><<<grid, 256, 0, stream>>>(P); break;
        case 3: k_gat_rows<3, VEC><<<grid, 256, 0, stream>>>(P); break;
        default: k_gat_rows<4, VEC><<<grid, 256, 0, stream>>>(P); break;
    }
    return check_launch();
}

template <int MODE>
int dispatch_heads(GatParams& P, cudaStream_t stream) {
    switch (P.a.H) {
        case 1: return dispatch_special<MODE, 1>(P, stream);
        case 2: return dispatch_special<MODE, 2>(P, stream);
        case 3: return dispatch_special<MODE, 3>(P, stream);
        case 4: return dispatch_special<MODE, 4>(P, stream);
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
    if (a->act < HG_ACT_NONE || a->act > HG_ACT_ELU) return HG_EINVAL;
    if (a->src_shared && !a->concat) return HG_EINVAL;  // aggregate-first output is always [n_dst, H*C]
    if (a->n_dst == 0) return HG_OK;
    if (!a->out) return HG_EINVAL;
    for (int r = 0; r < a->n_rel; ++r) {
        const hg_gat_rel& R = a->rel[r];
        // xs / a_src may be NULL only for an empty source set (every row of that relation is then empty)
        if (!R.a_dst || !R.rowptr || !R.col) return HG_EINVAL;
        if (R.ldxs < 0 || R.ldxs > (INT32_MAX >> 2) || R.lda_src < 0 || R.lda_src > INT32_MAX) return HG_EINVAL;
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
    // small head-mean destination sets: the row kernel (an aggregate-first form of it -- ~12 edges of DRAM-resident source rows
    // per row -- measured 56 us against 29 us for the tile kernel and was dropped)
    if (head_mean && a->H <= 4 && a->n_dst <= kRowsKernelMax && a->dropout_p == 0.f &&
        ((vec == 4 && a->C <= 64) || (vec >= 2 && a->C % 2 == 0 && a->C <= 32))) {
        bool plain = true;
        for (int r = 0; r < a->n_rel; ++r) plain = plain && a->alpha_saved[r] == nullptr;
        if (plain) {
            GatParams P;
            P.a = *a;
            P.f_begin = 0;
            P.f_count = a->C;
            P.wrows = 0;
            return (vec == 4 && a->C % 4 == 0) ? launch_rows<4>(P, stream) : launch_rows<2>(P, stream);
        }
    }
    if (vec == 4 && a->H <= 4 && ((head_mean && a->C <= 128) || (a->src_shared && a->C <= 256))) {
        GatParams P;
        P.a = *a;
        P.f_begin = 0;
        P.f_count = a->C;
        P.wrows = 0;
        if (head_mean && a->C == kWinC && a->n_rel <= kWinMaxRel && a->dropout_p == 0.f && window_enabled()) {
            bool hint = true, plain = true;  // the caller gives windows only where they are short (ops.gat_fused)
            for (int r = 0; r < a->n_rel; ++r) {
                hint = hint && a->rel[r].tile_win != nullptr;
                plain = plain && a->alpha_saved[r] == nullptr;
            }
            if (hint && plain) {
                const int rc = a->n_rel <= 2 ? dispatch_window<2>(P, stream) : dispatch_window<4>(P, stream);
                if (rc != kWinFallback) return rc;
            }
        }
        return head_mean ? dispatch_heads<1>(P, stream) : dispatch_heads<2>(P, stream);
    }
    if (vec == 2 && a->src_shared && a->H <= 4 && a->C > 128 && a->C <= 320) {
        GatParams P;
        P.a = *a;
        P.f_begin = 0;
        P.f_count = a->C;
        P.wrows = 0;
        switch (a->H) {
            case 1: return dispatch_wide_shared<1>(P, stream);
            case 2: return dispatch_wide_shared<2>(P, stream);
            case 3: return dispatch_wide_shared<3>(P, stream);
            default: return dispatch_wide_shared<4>(P, stream);
        }
    }
    const int block = 32 * 8 * vec;  // widest output block one launch covers
    for (int f0 = 0; f0 < width; f0 += block) {
        GatParams P;
        P.a = *a;
        P.wrows = 0;
        P.f_begin = f0;
        P.f_count = (width - f0 < block) ? width - f0 : block;
        const int rc = vec == 4 ? dispatch<4>(P, stream) : (vec == 2 ? dispatch<2>(P, stream) : dispatch<1>(P, stream));
        if (rc != HG_OK) return rc;
    }
    return HG_OK;
}
