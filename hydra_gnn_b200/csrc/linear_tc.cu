// K6 tensor-core path: y = act((sum_s x_s W_s^T + bias) * scale + shift) on tcgen05 (sm_100a).
//
// fp32-class accuracy on the TF32 tensor pipe by 3-term error compensation:
//     x = x_hi + x_lo,  W = W_hi + W_lo   (W: hi = rna_tf32(.), lo = rna_tf32(. - hi); x: hi = trunc_tf32, lo = x - hi)
//     x W^T ~= x_hi W_hi^T + x_hi W_lo^T + x_lo W_hi^T          (dropped term ~ 2^-22)
// accumulated in fp32 in TMEM.
//
// Persistent, warp-specialised kernel: one CTA per SM walks the (m-tile, n-tile) list; three
// pipelines run concurrently and never meet at a CTA barrier:
//   warps 2..9    A loaders: x rows are only 8-byte aligned when K = 306, so TMA cannot describe
//                 them; the loaders read the 128 x 32 k-block with vector loads THREE k-blocks ahead
//                 (a register ring, ~48 KB in flight per SM), split hi/lo in registers and store
//                 both tiles in the UMMA canonical K-major SWIZZLE_128B layout;
//   warp 1        W producer: one lane issues 1-D bulk copies (cp.async.bulk, UBLKCP) of the
//                 pre-swizzled W_hi|W_lo image of the k-block (packed once per weight version);
//   warp 0        allocates TMEM; one elected lane issues tcgen05.mma (12 per 32-wide k-block) into
//                 one of TWO accumulator buffers and tcgen05.commit (frees the smem stage / publishes
//                 the accumulator);
//   warps 10..13  epilogue: tcgen05.ld accumulator -> bias/scale/shift/activation -> global, while
//                 the MMA warp already fills the other accumulator buffer with the next tile.
#include "tc_common.cuh"

namespace hg {
namespace {

using namespace tc;

#ifndef HG_TC_EXP
#define HG_TC_EXP 0  // timing experiments only (wrong results): 1 no y stores, 2 no x loads, 4 no W copies, 8 no MMA
#endif

constexpr int BM = 128;
constexpr int BK = 32;                      // fp32 elements per k-block = one 128-byte swizzle row
constexpr int kLoaderWarps = 8;
constexpr int kLoaderThreads = kLoaderWarps * 32;
#ifndef HG_TC_EPI
#define HG_TC_EPI 4
#endif
constexpr int kEpiWarps = HG_TC_EPI;       // 4 or 8: warps sharing a TMEM lane quarter alternate over the 32-column chunks
constexpr int kThreads = 64 + kLoaderThreads + kEpiWarps * 32;  // 448
constexpr int kATileBytes = BM * BK * 4;    // 16 KB
constexpr int kPrefetch = 3;                // k-blocks of A in flight per loader thread

struct TcParams {
    hg_linear_args a;
    const float* wp;      // packed weights: [n_tiles][n_kb][hi|lo][block_n rows x 128 B, swizzled]
    int32_t n_tiles, block_n, n_kb, stages;
    int32_t kb_start[HG_MAX_SEGMENTS + 1];  // first k-block of each segment
    int32_t tmem_cols;
    int32_t vec;          // widest vector all x segments allow (1, 2, 4)
    int32_t cs;           // 1, or 2 = CTA-pair mode (cta_group::2)
    int64_t m_tiles;
};

// ------------------------------------------------------------------------------ weight packing
__global__ void __launch_bounds__(256) k_pack_w(const __grid_constant__ TcParams P, float* __restrict__ wp) {
    const hg_linear_args& A = P.a;
    const int64_t per_stage = static_cast<int64_t>(P.block_n) * BK;  // floats in one hi (or lo) tile
    const int64_t total = static_cast<int64_t>(P.n_tiles) * P.n_kb * per_stage;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total;
         t += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int k_local = static_cast<int>(t % BK);
        const int n_local = static_cast<int>((t / BK) % P.block_n);
        const int kb = static_cast<int>((t / per_stage) % P.n_kb);
        const int nt = static_cast<int>(t / (per_stage * P.n_kb));
        int s = 0;
        while (s + 1 < A.n_seg && kb >= P.kb_start[s + 1]) ++s;
        const int k = (kb - P.kb_start[s]) * BK + k_local;
        const int n = nt * P.block_n + n_local;
        float w = 0.f;
        if (n < A.N && k < A.seg[s].K) w = A.seg[s].w[static_cast<int64_t>(n) * A.seg[s].ldw + k];
        const float hi = tf32_rna(w);
        const float lo = tf32_rna(w - hi);
        float* base = wp + (static_cast<int64_t>(nt) * P.n_kb + kb) * 2 * per_stage;
        const uint32_t off = swz_off(n_local, k_local) / 4;
        base[off] = hi;
        base[per_stage + off] = lo;
    }
}

// ------------------------------------------------------------------------------ main kernel
// Work unit u = (group of CS consecutive m-tiles) x n-tile; cluster c takes u = c, c + n_clusters, ...;
// CTA rank r of the cluster computes m-tile CS * group + r (possibly past the end: all-zero rows, no stores).
// CS = 2 is the CTA-pair mode (tcgen05 cta_group::2): one MMA of M = 256 spans both SMs, each CTA stages its
// own 128 rows of A and HALF of the W rows of the n-tile, and the tensor cores read the other half from the
// peer's shared memory -- W fill and W operand traffic per SM halve (shared-memory bandwidth is the bound of
// the 3-term TF32 scheme).  Rank 0 issues every MMA; rank 1's warp 0 relays "my stage is filled".
// The loaders, the W producer and the MMA warp all walk that sequence (tiles of this CTA in order, k-blocks
// of a tile in order) with incremental cursors: no division in any per-k-block path.
struct Ring {  // pipeline stage + phase parity of the next k-block
    int s = 0;
    uint32_t ph = 0;
    __device__ __forceinline__ void next(int S) {
        if (++s == S) {
            s = 0;
            ph ^= 1u;
        }
    }
};

template <int CS>
struct TilePos {
    int n_tile;
    int64_t m0;
    __device__ __forceinline__ TilePos(const TcParams& P, int it, uint32_t rank) {
        const uint32_t u = blockIdx.x / CS + static_cast<uint32_t>(it) * (gridDim.x / CS);  // < 2^31 (plan())
        const uint32_t grp = u / static_cast<uint32_t>(P.n_tiles);
        n_tile = static_cast<int>(u - grp * P.n_tiles);
        m0 = (static_cast<int64_t>(grp) * CS + rank) * BM;
    }
};

template <int CS>
struct KbCursor {  // k-block g of this CTA: tile `it`, k-block kb = segment seg at column k0
    int it = 0, kb = 0, seg = 0, k0 = 0;
    int n_tile;
    int64_t m0;
    __device__ __forceinline__ KbCursor(const TcParams& P, uint32_t rank) {
        const TilePos<CS> t(P, 0, rank);
        n_tile = t.n_tile;
        m0 = t.m0;
    }
    __device__ __forceinline__ void advance(const TcParams& P, uint32_t rank) {
        ++kb;
        k0 += BK;
        if (kb == P.n_kb) {
            kb = 0; seg = 0; k0 = 0;
            const TilePos<CS> t(P, ++it, rank);
            n_tile = t.n_tile;
            m0 = t.m0;
        } else if (kb == P.kb_start[seg + 1]) {
            ++seg;
            k0 = 0;
        }
    }
};

template <int VEC, int CS>
struct Loader {
    static constexpr int kVecPerRow = BK / VEC;
    static constexpr int kRowStep = kLoaderThreads / kVecPerRow;
    static constexpr int kIters = BM / kRowStep;

    // global -> registers (one k-block slice of this thread: kIters vectors)
    static __device__ __forceinline__ void load(const TcParams& P, const KbCursor<CS>& c, int t, float (&v)[kIters][VEC]) {
        const hg_linear_args& A = P.a;
        const hg_lin_seg& G = A.seg[c.seg];
        const int cv = t % kVecPerRow, rb = t / kVecPerRow;
        const int k = c.k0 + cv * VEC;
        const int64_t step = static_cast<int64_t>(kRowStep) * G.ldx;
        const float* __restrict__ src = G.x + (c.m0 + rb) * G.ldx + k;
#if HG_TC_EXP & 2
        if (true) { for (int i = 0; i < kIters; ++i) for (int e = 0; e < VEC; ++e) v[i][e] = 1.f + t; return; }
#endif
        if (c.m0 + BM <= A.M && c.k0 + BK <= G.K) {  // interior k-block (CTA-uniform): no per-element checks
#pragma unroll
            for (int i = 0; i < kIters; ++i) ldg_vec<VEC>(src + i * step, v[i]);
        } else {
#pragma unroll
            for (int i = 0; i < kIters; ++i) {
                const bool row_ok = c.m0 + rb + i * kRowStep < A.M;
#pragma unroll
                for (int e = 0; e < VEC; ++e) v[i][e] = (row_ok && k + e < G.K) ? __ldg(src + i * step + e) : 0.f;
            }
        }
    }

    // registers -> hi/lo swizzled smem tiles of the stage (shared address `stage`).  hi = x with the 13 low
    // mantissa bits cleared (exactly what the tf32 datapath reads), lo = x - hi (exact; the datapath reads
    // its top 11 bits).
    static __device__ __forceinline__ void store(uint32_t stage, int t, const float (&v)[kIters][VEC]) {
        static_assert(kRowStep % 8 == 0, "rows of one thread must share their swizzle phase");
        const int cv = t % kVecPerRow, rb = t / kVecPerRow;
        const uint32_t dst = stage + swz_off(rb, cv * VEC);  // row rb + i * kRowStep: same phase, (kRowStep / 8) KB further
#pragma unroll
        for (int i = 0; i < kIters; ++i) {
            float hi[VEC], lo[VEC];
#pragma unroll
            for (int e = 0; e < VEC; ++e) {
                hi[e] = __uint_as_float(__float_as_uint(v[i][e]) & 0xFFFFE000u);
                lo[e] = v[i][e] - hi[e];
            }
            sts_vec<VEC>(dst + i * (kRowStep / 8) * 1024, hi);
            sts_vec<VEC>(dst + kATileBytes + i * (kRowStep / 8) * 1024, lo);
        }
    }
};

template <int VEC, int CS>
__global__ void __launch_bounds__(kThreads, 1) k_linear_tc(const __grid_constant__ TcParams P) {
    extern __shared__ uint8_t smem_raw[];
    const hg_linear_args& A = P.a;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int S = P.stages;
    const uint32_t b_tile_bytes = static_cast<uint32_t>(P.block_n / CS) * 128u;  // W rows staged by THIS CTA
    const uint32_t stage_bytes = 2u * kATileBytes + 2u * b_tile_bytes;

    // tiles of this CTA: blockIdx.x, blockIdx.x + gridDim.x, ...
    const uint32_t rank = CS > 1 ? cluster_ctarank() : 0u;
    const int64_t total_units = ceil_div<int64_t>(P.m_tiles, CS) * P.n_tiles;
    const int64_t n_clusters = gridDim.x / CS, cluster_id = blockIdx.x / CS;
    const int my_tiles = static_cast<int>((total_units - cluster_id + n_clusters - 1) / n_clusters);
    const int G = my_tiles * P.n_kb;  // k-blocks this CTA streams (identical for every CTA of the cluster)

    // carve (all as 32-bit shared addresses): [1024-aligned stages][barriers][tmem slot][epilogue affine]
    const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bars = sbase + static_cast<uint32_t>(S) * stage_bytes;
    const uint32_t full_a = bars;                  // [S] loaders -> MMA   (count = loader warps)
    const uint32_t full_b = bars + 8u * S;         // [S] bulk copy -> MMA (count = 1 + tx bytes)
    const uint32_t empty = bars + 16u * S;         // [S] MMA -> loaders / W producer (tcgen05.commit)
    const uint32_t acc_full = bars + 24u * S;      // [2] MMA -> epilogue
    const uint32_t acc_empty = acc_full + 16u;     // [2] epilogue -> MMA (count = epilogue warps of the whole pair)
    const uint32_t peer_full = acc_empty + 16u;    // [S] pair mode, rank 0 only: rank 1's stage is filled (count = 1)
    const uint32_t tmem_slot = peer_full + 8u * S;
    // per-column epilogue affine  y = act(acc * sc[n] + sh[n])  (bias / scale / shift folded once per CTA)
    const uint32_t s_sc = (tmem_slot + 16u + 15u) & ~15u;
    const uint32_t s_sh = s_sc + 4u * static_cast<uint32_t>(P.n_tiles * P.block_n);
    const uint32_t s_xp = s_sh + 4u * static_cast<uint32_t>(P.n_tiles * P.block_n);  // [epilogue warps][4 KB] transpose buffers
    uint8_t* const gen = smem_raw + (sbase - smem_u32(smem_raw));  // generic view of sbase
    for (int n = threadIdx.x; n < P.n_tiles * P.block_n; n += kThreads) {
        float sc = 1.f, sh = 0.f;
        if (n < A.N) {
            if (A.bias) sh = __ldg(A.bias + n);
            if (A.scale) { sc = __ldg(A.scale + n); sh *= sc; }
            if (A.shift) sh += __ldg(A.shift + n);
        }
        reinterpret_cast<float*>(gen + (s_sc - sbase))[n] = sc;
        reinterpret_cast<float*>(gen + (s_sh - sbase))[n] = sh;
    }

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(reinterpret_cast<uint64_t*>(gen + (full_a - sbase)) + s, kLoaderWarps);
            mbar_init(reinterpret_cast<uint64_t*>(gen + (full_b - sbase)) + s, 1);
            mbar_init(reinterpret_cast<uint64_t*>(gen + (empty - sbase)) + s, 1);
            mbar_init(reinterpret_cast<uint64_t*>(gen + (peer_full - sbase)) + s, 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(reinterpret_cast<uint64_t*>(gen + (acc_full - sbase)) + b, 1);
            mbar_init(reinterpret_cast<uint64_t*>(gen + (acc_empty - sbase)) + b, kEpiWarps * CS);
        }
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc<CS>(tmem_slot, static_cast<uint32_t>(P.tmem_cols));
    tc_fence_before();
    __syncthreads();
    if (CS > 1) cluster_sync_all();  // peers' barriers are initialised before anything remote arrives on them
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(gen + (tmem_slot - sbase));

    if (warp == 0 && rank == 0) {
        // ===================== MMA issuer (pair mode: for both CTAs) =====================
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(P.block_n >> 3) << 17) |
                               (static_cast<uint32_t>((BM * CS) >> 4) << 24);
        const uint32_t a_lo_off = kATileBytes >> 4, b_hi_off = (2 * kATileBytes) >> 4;
        const uint32_t b_lo_off = b_hi_off + (b_tile_bytes >> 4);
        Ring r;
        for (int it = 0; it < my_tiles; ++it) {
            const uint32_t ab = static_cast<uint32_t>(it) & 1u;
            if (CS > 1) mbar_wait_cluster(acc_empty + 8u * ab, ((static_cast<uint32_t>(it) >> 1) & 1u) ^ 1u);
            else mbar_wait(acc_empty + 8u * ab, ((static_cast<uint32_t>(it) >> 1) & 1u) ^ 1u);  // epilogue(s) drained this buffer
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + ab * static_cast<uint32_t>(P.block_n);
            for (int kb = 0; kb < P.n_kb; ++kb) {
                mbar_wait(full_a + 8u * r.s, r.ph);
                mbar_wait(full_b + 8u * r.s, r.ph);
                if (CS > 1) mbar_wait_cluster(peer_full + 8u * r.s, r.ph);  // the peer's A rows and W half are in place
                tc_fence_after();
                const uint32_t a_hi = desc_lo(sbase + static_cast<uint32_t>(r.s) * stage_bytes);  // descriptor low words
                if (lane == 0) {
#pragma unroll
                    for (uint32_t ks = 0; ks < ((HG_TC_EXP & 8) ? 0 : BK / 8); ++ks) {  // UMMA_K = 8 tf32 = 32 bytes: start address += 2
                        const uint32_t a = a_hi + 2u * ks;
                        // small terms first, the main term last (one accumulator holds all three)
                        umma_tf32<CS>(d_tmem, a, a + b_lo_off, kDescHi, idesc, ks != 0 ? 1u : static_cast<uint32_t>(kb != 0));
                        umma_tf32<CS>(d_tmem, a + a_lo_off, a + b_hi_off, kDescHi, idesc, 1u);
                        umma_tf32<CS>(d_tmem, a, a + b_hi_off, kDescHi, idesc, 1u);
                    }
                    umma_commit<CS>(empty + 8u * r.s);                           // stage reusable once these MMAs retire
                    if (kb == P.n_kb - 1) umma_commit<CS>(acc_full + 8u * ab);  // accumulator complete
                }
                __syncwarp();
                r.next(S);
            }
        }
    } else if (warp == 0) {
        // ===================== pair mode, rank 1: relay "stage filled" to the issuer =====================
        Ring r;
        for (int g = 0; g < G; ++g) {
            mbar_wait(full_a + 8u * r.s, r.ph);
            mbar_wait(full_b + 8u * r.s, r.ph);
            if (lane == 0) mbar_arrive_remote(peer_full + 8u * r.s, 0u);
            __syncwarp();
            r.next(S);
        }
    } else if (warp == 1) {
        // ===================== W producer =====================
        if (lane == 0) {
            Ring r;
            // packed image of one (n-tile, k-block): [W_hi: block_n rows x 128 B][W_lo: same]; this CTA stages
            // rows [rank * block_n / CS, (rank + 1) * block_n / CS) of both
            const uint32_t img_bytes = 2u * CS * b_tile_bytes, w_bytes = 2u * b_tile_bytes;
            for (int it = 0; it < my_tiles; ++it) {
                const TilePos<CS> tp(P, it, rank);
                const uint8_t* src = reinterpret_cast<const uint8_t*>(P.wp) + static_cast<size_t>(tp.n_tile) * P.n_kb * img_bytes +
                                     rank * b_tile_bytes;
                for (int kb = 0; kb < P.n_kb; ++kb, src += img_bytes) {
                            mbar_wait(empty + 8u * r.s, r.ph ^ 1u);  // the MMAs that read this stage have retired
                    const uint32_t dst = sbase + static_cast<uint32_t>(r.s) * stage_bytes + 2u * kATileBytes;
#if HG_TC_EXP & 4
                    mbar_arrive(full_b + 8u * r.s); r.next(S); continue;
#endif
                    mbar_arrive_expect_tx(full_b + 8u * r.s, w_bytes);
                    if (CS > 1) {
                        bulk_g2s(dst, src, b_tile_bytes, full_b + 8u * r.s);
                        bulk_g2s(dst + b_tile_bytes, src + CS * b_tile_bytes, b_tile_bytes, full_b + 8u * r.s);
                    } else {
                        bulk_g2s(dst, src, w_bytes, full_b + 8u * r.s);
                    }
                    r.next(S);
                }
            }
        }
    } else if (warp < 2 + kLoaderWarps) {
        // ===================== A loaders / converters =====================
        using L = Loader<VEC, CS>;
        const int t = threadIdx.x - 64;
        float v0[L::kIters][VEC], v1[L::kIters][VEC], v2[L::kIters][VEC];
        Ring r;
        KbCursor<CS> cur(P, rank);
        int fetched = 0;
        auto fetch = [&](float (&v)[L::kIters][VEC]) {  // next k-block of the sequence -> registers
            if (fetched < G) {
                L::load(P, cur, t, v);
                cur.advance(P, rank);
            }
            ++fetched;
        };
        auto finish = [&](const float (&v)[L::kIters][VEC]) {  // registers -> stage r.s, publish
            mbar_wait(empty + 8u * r.s, r.ph ^ 1u);  // MMAs of the previous use of this stage have retired
            L::store(sbase + static_cast<uint32_t>(r.s) * stage_bytes, t, v);
            fence_proxy_async();  // generic-proxy stores -> visible to the tensor core (async proxy)
            __syncwarp();
            if (lane == 0) mbar_arrive(full_a + 8u * r.s);
            r.next(S);
        };
        fetch(v0);
        fetch(v1);
        fetch(v2);
        for (int g = 0; g < G; g += kPrefetch) {
            finish(v0);
            fetch(v0);
            if (g + 1 < G) {
                finish(v1);
                fetch(v1);
            }
            if (g + 2 < G) {
                finish(v2);
                fetch(v2);
            }
        }
    } else {
        // ===================== epilogue =====================
        // tcgen05.ld gives lane r the columns of ROW r; a global store wants a warp on consecutive columns.
        // Each warp transposes 32 x 32 chunks through a private 4 KB XOR-swizzled buffer: 8 STS.128 + 8 LDS.128
        // per lane, both conflict-free, then every STG.128 of the warp writes four full 128-byte lines.
        const int q = warp & 3;  // TMEM lane quarter this warp may read
        const bool vec_store = (A.ldy % 4 == 0) && (A.N % 4 == 0) && ((reinterpret_cast<uintptr_t>(A.y) & 15) == 0);
        const int act = A.act;
        const uint32_t xbuf = s_xp + static_cast<uint32_t>(warp - (2 + kLoaderWarps)) * 4096u;
        const uint32_t xrow = xbuf + static_cast<uint32_t>(lane) * 128u;
        const bool has_affine = A.bias || A.scale || A.shift;  // GAT projections have none: skip the pass
        auto affine_act = [&](float (&acc)[32], int n, int cw) {
#pragma unroll
            for (int j4 = 0; j4 < 8; ++j4) {
                if (has_affine && 4 * j4 < cw) {
                    const float4 sc = lds128(s_sc + 4u * (n + 4 * j4)), sh = lds128(s_sh + 4u * (n + 4 * j4));  // broadcast
                    acc[4 * j4 + 0] = fmaf(acc[4 * j4 + 0], sc.x, sh.x);
                    acc[4 * j4 + 1] = fmaf(acc[4 * j4 + 1], sc.y, sh.y);
                    acc[4 * j4 + 2] = fmaf(acc[4 * j4 + 2], sc.z, sh.z);
                    acc[4 * j4 + 3] = fmaf(acc[4 * j4 + 3], sc.w, sh.w);
                }
            }
            if (act == HG_ACT_RELU) {
#pragma unroll
                for (int j = 0; j < 32; ++j) acc[j] = fmaxf(acc[j], 0.f);
            } else if (act == HG_ACT_ELU) {
#pragma unroll
                for (int j = 0; j < 32; ++j) acc[j] = acc[j] > 0.f ? acc[j] : expm1f(acc[j]);
            }
        };
        for (int it = 0; it < my_tiles; ++it) {
            const uint32_t ab = static_cast<uint32_t>(it) & 1u;
            const TilePos<CS> tp(P, it, rank);
            const int n0 = tp.n_tile * P.block_n;
            const int64_t mq = tp.m0 + q * 32;  // first row of this warp's quarter
            mbar_wait(acc_full + 8u * ab, (static_cast<uint32_t>(it) >> 1) & 1u);
            tc_fence_after();
            const uint32_t t_acc = tmem_base + ab * static_cast<uint32_t>(P.block_n) + (static_cast<uint32_t>(q * 32) << 16);
            for (int c0 = 32 * ((warp - (2 + kLoaderWarps)) >> 2); c0 < P.block_n; c0 += 32 * (kEpiWarps / 4)) {
                const int cw = P.block_n - c0 < 32 ? 16 : 32;  // block_n is a multiple of 16
                const int n = n0 + c0;
                float acc[32];
                if (cw == 32) {
                    tmem_ld32(t_acc + static_cast<uint32_t>(c0), acc);
                } else {
                    float lo16[16];
                    tmem_ld16(t_acc + static_cast<uint32_t>(c0), lo16);
#pragma unroll
                    for (int j = 0; j < 16; ++j) acc[j] = lo16[j], acc[16 + j] = 0.f;
                }
                if (n >= A.N) continue;  // warp-uniform
                affine_act(acc, n, cw);
#if HG_TC_EXP & 1
                if (acc[0] != 123.456f) continue;
#endif
                if (vec_store) {
                    __syncwarp();  // the previous chunk has been read back by every lane
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        if (4 * j < cw)
                            sts128(xrow + (static_cast<uint32_t>(j ^ (lane & 7)) << 4), acc[4 * j], acc[4 * j + 1], acc[4 * j + 2], acc[4 * j + 3]);
                    __syncwarp();
                    // read back transposed: all LDS first, then the stores back to back
                    if (cw == 32) {  // lane -> (row 4 i + lane / 8, 16-byte chunk lane % 8)
                        const int ch = lane & 7, r0 = lane >> 3;
                        float4 v[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int row = 4 * i + r0;
                            v[i] = lds128(xbuf + static_cast<uint32_t>(row) * 128u + (static_cast<uint32_t>(ch ^ (row & 7)) << 4));
                        }
                        float* dst = A.y + (mq + r0) * A.ldy + n + 4 * ch;
                        const int64_t step = 4 * A.ldy;
                        if (mq + 32 <= A.M && n + 32 <= A.N) {  // interior chunk (warp-uniform)
#pragma unroll
                            for (int i = 0; i < 8; ++i) *reinterpret_cast<float4*>(dst + i * step) = v[i];
                        } else if (n + 4 * ch < A.N) {
#pragma unroll
                            for (int i = 0; i < 8; ++i)
                                if (mq + r0 + 4 * i < A.M) *reinterpret_cast<float4*>(dst + i * step) = v[i];
                        }
                    } else {  // 16 columns: lane -> (row 8 i + lane / 4, chunk lane % 4)
                        const int ch = lane & 3, r0 = lane >> 2;
                        float4 v[4];
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int row = 8 * i + r0;
                            v[i] = lds128(xbuf + static_cast<uint32_t>(row) * 128u + (static_cast<uint32_t>(ch ^ (row & 7)) << 4));
                        }
                        float* dst = A.y + (mq + r0) * A.ldy + n + 4 * ch;
                        const int64_t step = 8 * A.ldy;
                        if (n + 4 * ch < A.N) {
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                if (mq + r0 + 8 * i < A.M) *reinterpret_cast<float4*>(dst + i * step) = v[i];
                        }
                    }
                } else if (mq + lane < A.M) {
                    float* dst = A.y + (mq + lane) * A.ldy + n;
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (j < cw && n + j < A.N) dst[j] = acc[j];
                }
            }
            tc_fence_before();  // all tcgen05.ld of this buffer done (tmem_ld waits) -> hand it back
            __syncwarp();
            if (lane == 0) {
                if (CS > 1 && rank != 0) mbar_arrive_remote(acc_empty + 8u * ab, 0u);
                else mbar_arrive(acc_empty + 8u * ab);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (CS > 1) cluster_sync_all();  // no CTA leaves while a peer may still multicast into its smem / barriers
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc<CS>(tmem_base, static_cast<uint32_t>(P.tmem_cols));
    }
}

#ifndef HG_TC_PAIR
#define HG_TC_PAIR 1
#endif
constexpr size_t kSmemBudget = 200 * 1024;

bool plan(const hg_linear_args& a, TcParams& P) {
    P.a = a;
    const int n_pad = ceil_div(a.N, 16) * 16;
    P.n_tiles = ceil_div(n_pad, 256);
    P.block_n = ceil_div(ceil_div(n_pad, P.n_tiles), 16) * 16;
    int kb = 0;
    int vec = 4;
    for (int s = 0; s < a.n_seg; ++s) {
        P.kb_start[s] = kb;
        kb += ceil_div(a.seg[s].K, BK);
        while (vec > 1 && (a.seg[s].ldx % vec || !aligned(a.seg[s].x, sizeof(float) * vec))) vec >>= 1;
    }
    for (int s = a.n_seg; s <= HG_MAX_SEGMENTS; ++s) P.kb_start[s] = kb;
    P.n_kb = kb;
    P.vec = vec;
    P.m_tiles = ceil_div<int64_t>(a.M, BM);
    // CTA-pair mode (cta_group::2) when every SM gets at least two m-tiles: halves the W traffic per SM
    const int sms = sm_count() > 0 ? sm_count() : 148;
    // and the k-loop is long enough to amortise the cross-CTA hand-offs (measured: K = 64 is faster unpaired)
    P.cs = (HG_TC_PAIR && P.m_tiles >= 2 * static_cast<int64_t>(sms) && P.n_kb >= 4) ? 2 : 1;
    const size_t stage_bytes = 2 * kATileBytes + 2 * static_cast<size_t>(P.block_n / P.cs) * 128;
    const size_t affine_bytes = 2 * static_cast<size_t>(P.n_tiles) * P.block_n * sizeof(float);
    if (affine_bytes > 32 * 1024) return false;  // N <= 4096
    int stages = static_cast<int>((kSmemBudget - affine_bytes - kEpiWarps * 4096) / stage_bytes);
    stages = stages > 4 ? 4 : stages;
    P.stages = stages;
    int cols = 32;
    while (cols < 2 * P.block_n) cols <<= 1;  // two accumulator buffers
    P.tmem_cols = cols;
    if (ceil_div<int64_t>(P.m_tiles, 1) * P.n_tiles >= (int64_t{1} << 31)) return false;  // 32-bit work-unit ids
    return stages >= 1 && P.n_kb >= 1 && cols <= 512;
}

size_t packed_bytes(const TcParams& P) { return static_cast<size_t>(P.n_tiles) * P.n_kb * 2 * P.block_n * 128; }

int pack_launch(const TcParams& P, float* dst, cudaStream_t stream) {
    const int64_t pack_elems = static_cast<int64_t>(P.n_tiles) * P.n_kb * P.block_n * BK;
    const int64_t want = ceil_div<int64_t>(pack_elems, 256);
    const unsigned pack_grid = static_cast<unsigned>(want < 1184 ? want : 1184);
    k_pack_w<<<pack_grid, 256, 0, stream>>>(P, dst);
    return check_launch();
}

}  // namespace

int linear_tc_supported(const hg_linear_args& a) {
    if (a.M < 1 || a.N < 1) return 0;
    int64_t k_total = 0;
    for (int s = 0; s < a.n_seg; ++s) k_total += a.seg[s].K;
    return k_total >= 1 ? 1 : 0;
}

size_t linear_tc_packed_bytes(const hg_linear_args& a) {
    TcParams P;
    if (!plan(a, P)) return 0;
    return packed_bytes(P);
}

// Split + swizzle the weights of `a` into `dst` (reusable for every call with the same weights / N / K layout).
int linear_tc_pack(const hg_linear_args& a, void* dst, size_t dst_bytes, cudaStream_t stream) {
    TcParams P;
    if (!plan(a, P)) return HG_EUNSUPPORTED;
    if (!dst || dst_bytes < packed_bytes(P)) return HG_EWORKSPACE;
    if (!aligned(dst, 128)) return HG_EALIGN;
    return pack_launch(P, static_cast<float*>(dst), stream);
}

int linear_tc_launch(const hg_linear_args& a, void* ws, size_t ws_bytes, cudaStream_t stream) {
    TcParams P;
    if (!plan(a, P)) return HG_EUNSUPPORTED;
    if (!ws || ws_bytes < packed_bytes(P)) return HG_EWORKSPACE;
    if (!aligned(ws, 128)) return HG_EALIGN;
    P.wp = static_cast<const float*>(ws);
    if (!a.w_prepacked) {
        const int rc = pack_launch(P, static_cast<float*>(ws), stream);
        if (rc != HG_OK) return rc;
    }
    const size_t stage_bytes = 2 * kATileBytes + 2 * static_cast<size_t>(P.block_n / P.cs) * 128;
    const size_t smem = 1024 + P.stages * stage_bytes + (4 * P.stages + 4) * sizeof(uint64_t) + 48 +
                        2 * static_cast<size_t>(P.n_tiles) * P.block_n * sizeof(float) + kEpiWarps * 4096;
    const int sms = sm_count() > 0 ? sm_count() : 148;
    auto go = [&](auto kernel, int cs) -> int {
        HG_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        const int64_t units = ceil_div<int64_t>(P.m_tiles, cs) * P.n_tiles;
        int64_t ctas = units * cs < sms ? units * cs : sms;  // persistent: one CTA per SM
        ctas = ctas / cs * cs;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(static_cast<unsigned>(ctas));
        cfg.blockDim = dim3(kThreads);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = cs;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        HG_CUDA_TRY(cudaLaunchKernelEx(&cfg, kernel, P));
        return check_launch();
    };
    if (P.cs == 2) {
        if (P.vec == 4) return go(k_linear_tc<4, 2>, 2);
        if (P.vec == 2) return go(k_linear_tc<2, 2>, 2);
        return go(k_linear_tc<1, 2>, 2);
    }
    if (P.vec == 4) return go(k_linear_tc<4, 1>, 1);
    if (P.vec == 2) return go(k_linear_tc<2, 1>, 1);
    return go(k_linear_tc<1, 1>, 1);
}

}  // namespace hg
