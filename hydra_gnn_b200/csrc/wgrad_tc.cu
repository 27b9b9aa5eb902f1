// hg_linear_wgrad on tcgen05 (sm_100a):  dW[N, K] += dY[M, N]^T X[M, K], the reduction running over the M rows.
//
// STATUS: default for large weight matrices over many rows (N * K >= 65536, M >= 16384; HG_WGRAD_TC=1 / 0 forces it
// on / off).  First run on a B200 in round 2 (tools/check_wgrad_tc.py, profiles/r02_wgrad_tc_first_run.txt): all 12
// ragged shapes inside the 3-term TF32 bound against fp64 (error 0.02-0.26 of the bound, below the mma.sync kernel's),
// [488,696 x 396 x 306] in 1.35 ms (87.9 TFLOP/s fp32-equivalent) against 2.26 ms; slower than mma.sync on the small
// [396 x 64] / [88 x 64] / [64 x 396] products, which therefore stay on k_wgrad_mma (train_ops.cu).
// tests/test_gpu_train.py::test_linear_wgrad_tcgen05 covers it.
//
// Both operands are row-major with the reduced index m as the ROW, so neither is in the K-major form the tensor
// core reads and neither row pitch is TMA-legal (306 floats = 1224 B).  The loaders therefore read 32-row k-blocks
// with coalesced scalar loads and write them TRANSPOSED into the same canonical K-major SWIZZLE_128B tiles the
// forward uses (tile row = output index, 32 tf32 = 128 B of m per row), split into hi / lo planes for the 3-term
// TF32 scheme (x = hi + lo, hi = rna_tf32(x), lo = x - hi exactly;  P Q ~= P_hi Q_lo + P_lo Q_hi + P_hi Q_hi).
//
//   accumulator  D[128 TMEM lanes = columns p0.. of operand P][bq <= 208 columns = columns q0.. of operand Q], fp32
//   CTA          one (P tile, Q tile, M split); the host picks P = X, Q = dY or the reverse (fewer padded MACs),
//                and as many M splits as fill the SMs once; CTAs of one split are adjacent (L2 reuse of the rows)
//   warp 0       TMEM allocation, MMA issue: per k-block 4 k-steps x 3 terms of M128 x N(bq) x K8, then
//                tcgen05.commit -> the stage's `empty` barrier (and `acc_full` after the last k-block)
//   warps 1..8   loaders: warp w owns rows 4w..4w+3 of the k-block; lane = (column & 7) + 8 (row & 3), which makes
//                every transposing STS.32 of a warp hit 32 distinct banks under the 128-byte swizzle; two k-blocks
//                of loads are in flight in registers ahead of the stores
//   warps 0, 9..11  epilogue (warp 0 after its last commit): tcgen05.ld of TMEM lane quarter warp & 3
//                (lane = P column) -> red.global.add.f32 into dW.  12 warps = 384 threads keeps the register
//                budget at 168 per thread (416 threads would round to 512 and cap it at 128: the loaders spill)
#include "tc_common.cuh"

namespace hg {
namespace {

using namespace tc;

constexpr int kPRows = 128;                  // accumulator rows = columns of P per tile
constexpr int kKb = 32;                      // reduced rows per k-block = one 128-byte swizzle row of tf32
constexpr int kLoadWarps = 8;                // kKb / 4 row groups
constexpr int kThreadsW = 32 * (1 + kLoadWarps + 3);  // 384: warp 0 doubles as the epilogue of lane quarter 0
constexpr int kPTileBytes = kPRows * kKb * 4;                  // 16 KB per plane
constexpr int kMaxBq = 208;
constexpr int kPGroups = kPRows / 8;         // 8-column groups of the P tile
constexpr int kMaxQGroups = kMaxBq / 8;
constexpr size_t kSmemMax = 232448;          // 227 KB opt-in limit per CTA

struct WgParams {
    const float* p; int64_t ldp; int32_t wp;   // P operand [M, wp]: its columns become accumulator rows
    const float* q; int64_t ldq; int32_t wq;   // Q operand [M, wq]: its columns become accumulator columns
    float* out; int64_t sp, sq;                // element (p, q) of the product lives at out[p * sp + q * sq]
    int32_t M, bq, n_ptiles, n_qtiles, kb_per_split, n_kb_total, stages, tmem_cols;
};

struct Ring2 {
    int s = 0;
    uint32_t ph = 0;
    __device__ __forceinline__ void next(int S) {
        if (++s == S) { s = 0; ph ^= 1u; }
    }
};

__global__ void __launch_bounds__(kThreadsW, 1) k_wgrad_tc(const __grid_constant__ WgParams P) {
    extern __shared__ uint8_t smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int S = P.stages;
    const int tiles = P.n_ptiles * P.n_qtiles;
    const int tile = static_cast<int>(blockIdx.x) % tiles, split = static_cast<int>(blockIdx.x) / tiles;
    const int p0 = (tile % P.n_ptiles) * kPRows, q0 = (tile / P.n_ptiles) * P.bq;
    const int kb_beg = split * P.kb_per_split;
    const int n_kb = max(0, min(P.kb_per_split, P.n_kb_total - kb_beg));  // k-blocks this CTA reduces (uniform)
    const uint32_t q_plane = static_cast<uint32_t>(P.bq) * 128u;
    const uint32_t stage_bytes = 2u * kPTileBytes + 2u * q_plane;

    const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bars = sbase + static_cast<uint32_t>(S) * stage_bytes;
    const uint32_t full = bars;               // [S] loaders -> MMA (count = loader warps)
    const uint32_t empty = bars + 8u * S;     // [S] MMA -> loaders (tcgen05.commit)
    const uint32_t acc_full = bars + 16u * S; // MMA -> epilogue
    const uint32_t tmem_slot = acc_full + 8u;
    uint8_t* const gen = smem_raw + (sbase - smem_u32(smem_raw));

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(reinterpret_cast<uint64_t*>(gen + (full - sbase)) + s, kLoadWarps);
            mbar_init(reinterpret_cast<uint64_t*>(gen + (empty - sbase)) + s, 1);
        }
        mbar_init(reinterpret_cast<uint64_t*>(gen + (acc_full - sbase)), 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc<1>(tmem_slot, static_cast<uint32_t>(P.tmem_cols));
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(gen + (tmem_slot - sbase));

    if (n_kb > 0) {
        if (warp == 0) {
            // ===================== MMA issuer =====================
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(P.bq >> 3) << 17) |
                                   (static_cast<uint32_t>(kPRows >> 4) << 24);
            const uint32_t p_lo = kPTileBytes >> 4, q_hi = (2 * kPTileBytes) >> 4, q_lo = q_hi + (q_plane >> 4);
            Ring2 r;
            for (int kb = 0; kb < n_kb; ++kb) {
                mbar_wait(full + 8u * r.s, r.ph);
                tc_fence_after();
                if (lane == 0) {
                    const uint32_t base = desc_lo(sbase + static_cast<uint32_t>(r.s) * stage_bytes);
#pragma unroll
                    for (uint32_t ks = 0; ks < kKb / 8; ++ks) {  // UMMA_K = 8 tf32 = 32 bytes: start address += 2
                        const uint32_t a = base + 2u * ks;
                        umma_tf32<1>(tmem_base, a, a + q_lo, kDescHi, idesc, (kb | static_cast<int>(ks)) != 0 ? 1u : 0u);
                        umma_tf32<1>(tmem_base, a + p_lo, a + q_hi, kDescHi, idesc, 1u);
                        umma_tf32<1>(tmem_base, a, a + q_hi, kDescHi, idesc, 1u);
                    }
                    umma_commit<1>(empty + 8u * r.s);
                    if (kb == n_kb - 1) umma_commit<1>(acc_full);
                }
                __syncwarp();
                r.next(S);
            }
        }
        if (warp >= 1 && warp <= kLoadWarps) {
            // ===================== loaders: global rows -> transposed hi / lo tiles =====================
            const int lw = warp - 1, c8 = lane & 7, r4 = lane >> 3;
            const int mrow = 4 * lw + r4;                       // row of the k-block = tf32 column of the tiles
            const uint32_t t_off = static_cast<uint32_t>(c8 * 128 + (((lw ^ c8) & 7) << 4) + r4 * 4);
            const int nq8 = P.bq >> 3;
            uint32_t pmask = 0, qmask = 0;                      // columns of this lane that exist
#pragma unroll
            for (int g = 0; g < kPGroups; ++g) pmask |= (p0 + 8 * g + c8 < P.wp ? 1u : 0u) << g;
#pragma unroll
            for (int g = 0; g < kMaxQGroups; ++g) qmask |= (g < nq8 && q0 + 8 * g + c8 < P.wq ? 1u : 0u) << g;
            const float* const pcol = P.p + p0 + c8;
            const float* const qcol = P.q + q0 + c8;
            float va[kPGroups + kMaxQGroups], vb[kPGroups + kMaxQGroups];
            int fetched = 0;
            auto fetch = [&](float (&v)[kPGroups + kMaxQGroups]) {
                if (fetched < n_kb) {
                    const int64_t m = static_cast<int64_t>(kb_beg + fetched) * kKb + mrow;
                    const bool ok = m < P.M;
                    const float* pr = pcol + m * P.ldp;
                    const float* qr = qcol + m * P.ldq;
#pragma unroll
                    for (int g = 0; g < kPGroups; ++g) v[g] = (ok && ((pmask >> g) & 1u)) ? __ldg(pr + 8 * g) : 0.f;
#pragma unroll
                    for (int g = 0; g < kMaxQGroups; ++g)
                        if (g < nq8) v[kPGroups + g] = (ok && ((qmask >> g) & 1u)) ? __ldg(qr + 8 * g) : 0.f;
                }
                ++fetched;
            };
            Ring2 r;
            auto finish = [&](const float (&v)[kPGroups + kMaxQGroups]) {
                mbar_wait(empty + 8u * r.s, r.ph ^ 1u);  // the MMAs that read this stage last time have retired
                const uint32_t st = sbase + static_cast<uint32_t>(r.s) * stage_bytes + t_off;
#pragma unroll
                for (int g = 0; g < kPGroups; ++g) {
                    const float hi = tf32_rna(v[g]);
                    const float h1[1] = {hi}, l1[1] = {v[g] - hi};
                    sts_vec<1>(st + g * 1024u, h1);
                    sts_vec<1>(st + kPTileBytes + g * 1024u, l1);
                }
#pragma unroll
                for (int g = 0; g < kMaxQGroups; ++g) {
                    if (g < nq8) {
                        const float hi = tf32_rna(v[kPGroups + g]);
                        const float h1[1] = {hi}, l1[1] = {v[kPGroups + g] - hi};
                        sts_vec<1>(st + 2u * kPTileBytes + g * 1024u, h1);
                        sts_vec<1>(st + 2u * kPTileBytes + q_plane + g * 1024u, l1);
                    }
                }
                fence_proxy_async();  // generic-proxy stores -> visible to the tensor core (async proxy)
                __syncwarp();
                if (lane == 0) mbar_arrive(full + 8u * r.s);
                r.next(S);
            };
            fetch(va);
            fetch(vb);
            for (int g = 0; g < n_kb; g += 2) {
                finish(va);
                fetch(va);
                if (g + 1 < n_kb) {
                    finish(vb);
                    fetch(vb);
                }
            }
        } else {
            // ===================== epilogue: TMEM -> red.global.add (warps 0, 9, 10, 11) =====================
            const int quarter = warp & 3;  // TMEM lane quarter this warp may read
            const int p = p0 + quarter * 32 + lane;
            mbar_wait(acc_full, 0u);
            tc_fence_after();
            const uint32_t t_acc = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
            float* const orow = P.out + static_cast<int64_t>(p) * P.sp;
            for (int c0 = 0; c0 < P.bq; c0 += 32) {
                const int cw = P.bq - c0 < 32 ? 16 : 32;  // bq is a multiple of 16
                float acc[32];
                if (cw == 32) {
                    tmem_ld32(t_acc + static_cast<uint32_t>(c0), acc);
                } else {
                    float lo16[16];
                    tmem_ld16(t_acc + static_cast<uint32_t>(c0), lo16);
#pragma unroll
                    for (int j = 0; j < 16; ++j) acc[j] = lo16[j], acc[16 + j] = 0.f;
                }
                if (p < P.wp) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int qc = q0 + c0 + j;
                        if (j < cw && qc < P.wq) atomicAdd(orow + static_cast<int64_t>(qc) * P.sq, acc[j]);
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc<1>(tmem_base, static_cast<uint32_t>(P.tmem_cols));
    }
}

// padded multiply-accumulates + tile fill per reduced row, the quantity the orientation is chosen by
struct Shape {
    int n_pt, n_qt, bq;
    int64_t cost;
};
Shape shape_for(int wp, int wq) {
    Shape s;
    s.n_pt = ceil_div(wp, kPRows);
    const int wq16 = ceil_div(wq, 16) * 16;
    s.n_qt = ceil_div(wq16, kMaxBq);
    s.bq = ceil_div(ceil_div(wq16, s.n_qt), 16) * 16;
    s.cost = static_cast<int64_t>(s.n_pt) * s.n_qt * (kPRows + s.bq);
    return s;
}

}  // namespace

// dW[N, K] += dY^T X (the caller has zeroed dW when it does not accumulate).  Returns HG_OK or an error code.
int wgrad_tc(const float* dy, int64_t lddy, const float* x, int64_t ldx, int32_t M, int32_t N, int32_t K, float* dw,
             int64_t lddw, cudaStream_t stream) {
    if (M <= 0 || N <= 0 || K <= 0) return HG_OK;
    WgParams P{};
    const Shape a = shape_for(K, N), b = shape_for(N, K);  // a: P = X (rows of D = k); b: P = dY (rows of D = n)
    const bool x_is_p = a.cost <= b.cost;
    const Shape& s = x_is_p ? a : b;
    if (x_is_p) {
        P.p = x; P.ldp = ldx; P.wp = K; P.q = dy; P.ldq = lddy; P.wq = N; P.sp = 1; P.sq = lddw;
    } else {
        P.p = dy; P.ldp = lddy; P.wp = N; P.q = x; P.ldq = ldx; P.wq = K; P.sp = lddw; P.sq = 1;
    }
    P.out = dw;
    P.M = M;
    P.bq = s.bq;
    P.n_ptiles = s.n_pt;
    P.n_qtiles = s.n_qt;
    const size_t stage_bytes = 2 * static_cast<size_t>(kPTileBytes) + 2 * static_cast<size_t>(s.bq) * 128;
    const size_t fixed = 1024 + 256;  // alignment slack + barriers + TMEM slot
    int stages = static_cast<int>((kSmemMax - fixed) / stage_bytes);
    stages = stages > 4 ? 4 : stages;
    if (stages < 2) return HG_EUNSUPPORTED;
    P.stages = stages;
    P.tmem_cols = s.bq <= 32 ? 32 : (s.bq <= 64 ? 64 : (s.bq <= 128 ? 128 : 256));
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int tiles = s.n_pt * s.n_qt;
    P.n_kb_total = ceil_div(M, kKb);
    int splits = sms / tiles;
    splits = splits < 1 ? 1 : (splits > P.n_kb_total ? P.n_kb_total : splits);
    P.kb_per_split = ceil_div(P.n_kb_total, splits);
    splits = ceil_div(P.n_kb_total, P.kb_per_split);
    const size_t smem = stages * stage_bytes + fixed;
    // per call, as linear_tc.cu does: the attribute belongs to the current device's context, not to the thread
    HG_CUDA_TRY(cudaFuncSetAttribute(k_wgrad_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    k_wgrad_tc<<<static_cast<unsigned>(tiles * splits), kThreadsW, smem, stream>>>(P);
    return check_launch();
}

}  // namespace hg
