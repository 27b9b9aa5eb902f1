// Pass 1 of the GAT backward (by destination row) as a TILE kernel -- the layout of hg_gat_fused applied to
//     d alpha[e, h] = <g_h[i], xs[col[e], h, :]>,   softmax backward,   d a_dst
// (gat_train.cu::k_bwd_dst_v4 is the verified half-warp-per-row form of the same pass).
//
// STATUS: default where it applies since round 2 (HG_GAT_BWD_TILE=0 disables).  First run on a B200 in round 2:
// tools/check_gat_bwd_tile.py -- every parameter gradient of M-GAT and of the heterogeneous Neural-Tree GAT, dropout 0 and
// 0.4, within 0.05 of the bound against the row-per-half-warp kernel (profiles/r02_bwd_tile_first_run.txt); at 8192 graphs
// 890 vs 942 us (objects layers), 355 vs 413 us (rooms layers), training step 8.03 vs 8.33 ms.  The GPU gradient tests of
// tests/test_gpu_train.py run through it.
//
// A CTA takes 32 destination rows at a time:
//   (1) g = dout * act'(out) * group scale for the tile: threads = (row, column), coalesced; written to global
//       (pass 2 gathers it) and kept in shared memory; bias-gradient column sums stay in registers across tiles;
//   (2) threads = edges of the tile: neighbour id, row of the edge, saved softmax weight, sign of the raw score;
//   (3) threads = (edge, head, quarter of the head's channels): LDG.128 of xs, broadcast LDS.128 of g, two shuffles
//       -> d alpha in shared memory.  Every lane gathers: a row with 5 in-edges no longer idles 11 of 16 lanes;
//   (4) threads = (relation, row, head): t = sum alpha * d alpha;  ds = alpha (d alpha - t) leaky';  dlogit and
//       d a_dst written once.
// A tile with more than `cap` edges in one relation (hub rows) takes the warp-per-row path below.
#include "hg_common.cuh"

namespace hg {
namespace {

constexpr int kTile = 32;
constexpr int kTThreads = 256;
constexpr int kTWarps = kTThreads / 32;
constexpr int kHs = 4;     // heads padded to 4 in the shared-memory layouts
constexpr int kCap = 512;  // edges of one relation staged per tile

__device__ __forceinline__ float keepf(float a_signed, float inv_keep) {
    return (__float_as_uint(a_signed) >> 31) ? 0.f : inv_keep;
}

struct TileSmem {
    float* g;        // [kTile][width]
    int32_t* rp;     // [R][kTile + 1]
    float* ad;       // [R][kTile][kHs]
    int32_t* col;    // [R][cap]
    uint16_t* rowf;  // [R][cap]: row of the edge | (raw score > 0 ? 1 : 0) << (8 + h)
    float* al;       // [R][cap][kHs] saved softmax weights (sign bit = dropped)
    float* dal;      // [R][cap][kHs] d alpha
};

__host__ __device__ inline size_t tile_smem_bytes(int R, int width) {
    return sizeof(float) * kTile * width + sizeof(int32_t) * R * (kTile + 1) + sizeof(float) * R * kTile * kHs +
           sizeof(int32_t) * R * kCap + sizeof(float) * 2 * R * kCap * kHs + sizeof(uint16_t) * R * kCap + 16;
}

__device__ __forceinline__ TileSmem carve(uint8_t* raw, int R, int width) {
    TileSmem S;
    S.g = reinterpret_cast<float*>(raw);                 // 16-byte aligned: width % 4 == 0
    S.al = S.g + kTile * width;
    S.dal = S.al + R * kCap * kHs;
    S.ad = S.dal + R * kCap * kHs;
    S.rp = reinterpret_cast<int32_t*>(S.ad + R * kTile * kHs);
    S.col = S.rp + R * (kTile + 1);
    S.rowf = reinterpret_cast<uint16_t*>(S.col + R * kCap);
    return S;
}

// hub tile: warp per row straight from global memory (same arithmetic as gat_train.cu::k_bwd_dst)
__device__ __noinline__ void slow_row(const hg_gat_train_args& A, const float* sg, int64_t i, int lane, float inv_keep) {
    const int H = A.H, C = A.C;
    const float invH = A.concat ? 1.f : 1.f / static_cast<float>(H);
    for (int r = 0; r < A.n_rel; ++r) {
        const hg_gat_train_rel& R = A.rel[r];
        const int beg = R.rowptr[i], end = R.rowptr[i + 1];
        for (int h = 0; h < H; ++h) {
            const float ad = __ldg(R.a_dst + i * R.lda_dst + h);
            const float* gh = sg + (A.concat ? h * C : 0);
            float t = 0.f;
            for (int p = beg; p < end; ++p) {
                const float* xr = R.xs + static_cast<int64_t>(__ldg(R.col + p)) * R.ldxs + h * C;
                float part = 0.f;
                for (int c = lane; c < C; c += 32) part = fmaf(gh[c], __ldg(xr + c), part);
                const int64_t idx = static_cast<int64_t>(p) * H + h;
                const float a = R.alpha[idx];
                const float da = warp_sum(part) * invH * keepf(a, inv_keep);
                t = fmaf(fabsf(a), da, t);
                if (lane == 0) R.dlogit[idx] = da;
            }
            __syncwarp();
            float dad = 0.f;
            for (int p = beg + lane; p < end; p += 32) {
                const int64_t idx = static_cast<int64_t>(p) * H + h;
                float raw = __ldg(R.a_src + static_cast<int64_t>(__ldg(R.col + p)) * R.lda_src + h) + ad;
                if (R.a_edge) raw += __ldg(R.a_edge + idx);
                const float ds = fabsf(R.alpha[idx]) * (R.dlogit[idx] - t) * (raw > 0.f ? 1.f : A.neg_slope);
                R.dlogit[idx] = ds;
                dad += ds;
            }
            dad = warp_sum(dad);
            if (lane == 0) R.da_dst[i * R.ldda_dst + h] = dad;
            __syncwarp();
        }
    }
}

__global__ void __launch_bounds__(kTThreads, 4) k_bwd_dst_tile(const __grid_constant__ hg_gat_train_args A) {
    extern __shared__ __align__(16) uint8_t s_raw[];
    const int H = A.H, C = A.C, C4 = C / 4, R = A.n_rel;
    const int width = A.concat ? H * C : C;
    const TileSmem S = carve(s_raw, R, width);
    const int tid = threadIdx.x, lane = tid & 31;
    const float inv_keep = A.dropout_p > 0.f ? 1.f / (1.f - A.dropout_p) : 1.f;
    const float gscale = A.group == HG_GROUP_MEAN ? 1.f / static_cast<float>(R) : 1.f;
    const float invH = A.concat ? 1.f : 1.f / static_cast<float>(H);
    const int64_t n_tiles = ceil_div<int64_t>(A.n_dst, kTile);
    float db = 0.f;  // bias-gradient column sum of column `tid` over the tiles of this CTA (width <= 256 threads)

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t t0 = tile * kTile;
        const int rows = (A.n_dst - t0 < kTile) ? static_cast<int>(A.n_dst - t0) : kTile;
        __syncthreads();  // the previous tile is fully consumed

        // ---- (1) g of the tile, rowptr slices, destination scores
        for (int idx = tid; idx < rows * width; idx += kTThreads) {
            const int t = idx / width, c = idx - t * width;
            const float o = A.out[(t0 + t) * A.ldout + c];
            float d = __ldg(A.dout + (t0 + t) * A.lddout + c);
            if (A.act == HG_ACT_ELU) d *= o > 0.f ? 1.f : o + 1.f;
            else if (A.act == HG_ACT_RELU) d = o > 0.f ? d : 0.f;
            d *= gscale;
            A.g[(t0 + t) * A.ldg + c] = d;
            S.g[idx] = d;
        }
        for (int idx = tid; idx < R * (kTile + 1); idx += kTThreads) {
            const int r = idx / (kTile + 1), t = idx - r * (kTile + 1);
            S.rp[idx] = __ldg(A.rel[r].rowptr + t0 + min(t, rows));
        }
        for (int idx = tid; idx < R * kTile * kHs; idx += kTThreads) {
            const int h = idx & (kHs - 1), t = (idx >> 2) & (kTile - 1), r = idx >> 7;
            S.ad[idx] = (t < rows && h < H) ? __ldg(A.rel[r].a_dst + (t0 + t) * A.rel[r].lda_dst + h) : 0.f;
        }
        bool fits = true;  // uniform: straight from global memory, no barrier needed before the branch
        for (int r = 0; r < R; ++r)
            fits = fits && (__ldg(A.rel[r].rowptr + t0 + rows) - __ldg(A.rel[r].rowptr + t0) <= kCap);
        __syncthreads();
        if (A.dbias && tid < width)
            for (int t = 0; t < rows; ++t) db += S.g[t * width + tid];
        if (!fits) {
            for (int t = tid >> 5; t < rows; t += kTWarps) slow_row(A, S.g + t * width, t0 + t, lane, inv_keep);
            continue;
        }

        // ---- (2) threads = edges: neighbour, row of the edge, saved weight, sign of the raw score
        for (int r = 0; r < R; ++r) {
            const hg_gat_train_rel& Rl = A.rel[r];
            const int32_t* rp = S.rp + r * (kTile + 1);
            const int e0 = rp[0], ne = rp[rows] - e0;
            for (int e = tid; e < ne; e += kTThreads) {
                const int32_t j = __ldg(Rl.col + e0 + e);
                int lo = 0, hi = rows;  // row t with rp[t] <= e0 + e < rp[t + 1] (empty rows skipped by the search)
                while (hi - lo > 1) {
                    const int mid = (lo + hi) >> 1;
                    if (rp[mid] <= e0 + e) lo = mid; else hi = mid;
                }
                const float* as = Rl.a_src + static_cast<int64_t>(j) * Rl.lda_src;
                const float* adr = S.ad + ((r * kTile + lo) << 2);
                uint32_t flags = static_cast<uint32_t>(lo);
                float* al = S.al + ((r * kCap + e) << 2);
#pragma unroll
                for (int h = 0; h < kHs; ++h) {
                    float a = 0.f;
                    if (h < H) {
                        const int64_t idx = static_cast<int64_t>(e0 + e) * H + h;
                        a = Rl.alpha[idx];
                        float raw = __ldg(as + h) + adr[h];
                        if (Rl.a_edge) raw += __ldg(Rl.a_edge + idx);
                        if (raw > 0.f) flags |= 1u << (8 + h);
                    }
                    al[h] = a;
                }
                S.col[r * kCap + e] = j;
                S.rowf[r * kCap + e] = static_cast<uint16_t>(flags);
            }
        }
        __syncthreads();

        // ---- (3) threads = (edge, head, quarter): d alpha = <g_h[row], xs[j, h, :]> / H * keep
        for (int r = 0; r < R; ++r) {
            const hg_gat_train_rel& Rl = A.rel[r];
            const int ne = S.rp[r * (kTile + 1) + rows] - S.rp[r * (kTile + 1)];
            const int total = ne * H * 4;
            for (int base = 0; base < total; base += kTThreads) {  // uniform trip count: the shuffles need whole warps
                const int id = base + tid;
                const bool live = id < total;
                const int q = id & 3, eh = id >> 2;
                const int e = live ? eh / H : 0, h = live ? eh - e * H : 0;
                float part = 0.f;
                if (live) {
                    const int32_t j = S.col[r * kCap + e];
                    const int t = S.rowf[r * kCap + e] & 0xff;
                    const float* xr = Rl.xs + static_cast<int64_t>(j) * Rl.ldxs + h * C;
                    const float* gr = S.g + t * width + (A.concat ? h * C : 0);
                    for (int v = q; v < C4; v += 4) {
                        const float4 x4 = __ldg(reinterpret_cast<const float4*>(xr) + v);
                        const float4 g4 = *(reinterpret_cast<const float4*>(gr) + v);
                        part = fmaf(x4.x, g4.x, fmaf(x4.y, g4.y, fmaf(x4.z, g4.z, fmaf(x4.w, g4.w, part))));
                    }
                }
                part += __shfl_xor_sync(0xffffffffu, part, 1);
                part += __shfl_xor_sync(0xffffffffu, part, 2);
                if (live && q == 0) {
                    const int s = ((r * kCap + e) << 2) + h;
                    S.dal[s] = part * invH * keepf(S.al[s], inv_keep);
                }
            }
        }
        __syncthreads();

        // ---- (4) threads = (relation, row, head): softmax backward, dlogit and d a_dst written once
        for (int idx = tid; idx < R * kTile * kHs; idx += kTThreads) {
            const int h = idx & (kHs - 1), t = (idx >> 2) & (kTile - 1), r = idx >> 7;
            if (t >= rows || h >= H) continue;
            const hg_gat_train_rel& Rl = A.rel[r];
            const int32_t* rp = S.rp + r * (kTile + 1);
            const int e0 = rp[0], beg = rp[t] - e0, end = rp[t + 1] - e0;
            const float* al = S.al + ((r * kCap) << 2) + h;
            const float* dal = S.dal + ((r * kCap) << 2) + h;
            const uint16_t* rf = S.rowf + r * kCap;
            float tsum = 0.f;
            for (int e = beg; e < end; ++e) tsum = fmaf(fabsf(al[e << 2]), dal[e << 2], tsum);
            float dad = 0.f;
            for (int e = beg; e < end; ++e) {
                float ds = fabsf(al[e << 2]) * (dal[e << 2] - tsum);
                ds *= ((rf[e] >> (8 + h)) & 1) ? 1.f : A.neg_slope;
                Rl.dlogit[static_cast<int64_t>(e0 + e) * H + h] = ds;
                dad += ds;
            }
            Rl.da_dst[(t0 + t) * Rl.ldda_dst + h] = dad;
        }
    }
    if (A.dbias && tid < width) atomicAdd(A.dbias + tid, db);
}

}  // namespace

// true when the tile kernel applies (float4-aligned xs rows, widths the shared-memory layout holds)
bool gat_bwd_dst_tile_ok(const hg_gat_train_args* a) {
    if (a->C % 4 != 0 || a->H > kHs) return false;
    const int width = a->concat ? a->H * a->C : a->C;
    if (width > kTThreads) return false;
    for (int r = 0; r < a->n_rel; ++r)
        if (!aligned(a->rel[r].xs, 16) || a->rel[r].ldxs % 4 != 0) return false;
    return tile_smem_bytes(a->n_rel, width) <= 200 * 1024;
}

int gat_bwd_dst_tile(const hg_gat_train_args* a, cudaStream_t stream) {
    const int width = a->concat ? a->H * a->C : a->C;
    const size_t smem = tile_smem_bytes(a->n_rel, width);
    // per call, as linear_tc.cu does: the attribute belongs to the current device's context, not to the thread
    HG_CUDA_TRY(cudaFuncSetAttribute(k_bwd_dst_tile, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int64_t tiles = ceil_div<int64_t>(a->n_dst, kTile);
    const int64_t cap = static_cast<int64_t>(sms) * 8;  // grid-stride: the bias gradient is reduced per CTA
    k_bwd_dst_tile<<<static_cast<unsigned>(tiles < cap ? tiles : cap), kTThreads, smem, stream>>>(*a);
    return check_launch();
}

}  // namespace hg
