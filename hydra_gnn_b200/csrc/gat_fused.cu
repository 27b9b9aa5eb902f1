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
#include "hg_common.cuh"

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
                    for (int v = 0; v < VEC; ++v) o[v] = (acc[h][k][v] * hscale + bp[v]) * gscale;
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
    while (P.tile_rows > HG_GAT_MIN_TILE && ceil_div<int64_t>(A.n_dst, P.tile_rows) < static_cast<int64_t>(sms0) * 6) P.tile_rows >>= 1;
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
    constexpr int kOcc = KC * (MODE == 2 ? HT : 1) <= 4 ? 6 : 4;
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
    if (vec == 4 && a->H <= 4 && ((head_mean && a->C <= 128) || (a->src_shared && a->C <= 256))) {
        GatParams P;
        P.a = *a;
        P.f_begin = 0;
        P.f_count = a->C;
        return head_mean ? dispatch_heads<1>(P, stream) : dispatch_heads<2>(P, stream);
    }
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
