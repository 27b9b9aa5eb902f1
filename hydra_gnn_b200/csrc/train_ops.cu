// Weight gradient of the projection (dW = dY^T X) and the Adam update (include/hydra_gnn_b200.h, "Training step").
//
// hg_linear_wgrad: fp32 FMA SGEMM, 64 x 64 output tiles, 4 x 4 micro-tiles, the M (row) dimension split across
// CTAs so that the grid covers the 148 SMs several times; partial tiles are added with red.global.add.f32.
// Both operands are read in their natural row-major layout (rows of dY give the N index, rows of X the K index),
// so every global load is coalesced and no transpose is materialised.
#include "hg_common.cuh"

#include <cstdlib>

namespace hg {

// wgrad_tc.cu: the tcgen05 variant (default for N * K >= 65536 and M >= 16384; HG_WGRAD_TC=1 / 0 forces it on / off)
int wgrad_tc(const float* dy, int64_t lddy, const float* x, int64_t ldx, int32_t M, int32_t N, int32_t K, float* dw,
             int64_t lddw, cudaStream_t stream);

namespace {

constexpr int kTN = 64, kTK = 64, kTM = 16, kWgThreads = 256;

__global__ void __launch_bounds__(kWgThreads) k_wgrad(const float* __restrict__ dy, int64_t lddy,
                                                      const float* __restrict__ x, int64_t ldx, int M, int N, int K,
                                                      float* __restrict__ dw, int64_t lddw, int rows_per_split) {
    __shared__ __align__(16) float sA[2][kTM][kTN];
    __shared__ __align__(16) float sB[2][kTM][kTK];
    const int tid = threadIdx.x;
    const int n0 = blockIdx.x * kTN, k0 = blockIdx.y * kTK;
    const int m_beg = blockIdx.z * rows_per_split;
    const int m_end = min(M, m_beg + rows_per_split);
    if (m_beg >= m_end) return;
    const int lrow = tid >> 4, lcol = (tid & 15) * 4;  // loader: row of the m-tile, first of 4 columns
    const int tn = (tid >> 4) * 4, tk = (tid & 15) * 4;  // micro-tile origin
    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
    float ra[4], rb[4];
    auto load = [&](int m0) {
        const int m = m0 + lrow;
        const bool ok = m < m_end;
        const float* pa = dy + static_cast<int64_t>(m) * lddy + n0 + lcol;
        const float* pb = x + static_cast<int64_t>(m) * ldx + k0 + lcol;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            ra[t] = (ok && n0 + lcol + t < N) ? __ldg(pa + t) : 0.f;
            rb[t] = (ok && k0 + lcol + t < K) ? __ldg(pb + t) : 0.f;
        }
    };
    load(m_beg);
    int buf = 0;
    for (int m0 = m_beg; m0 < m_end; m0 += kTM) {
        *reinterpret_cast<float4*>(&sA[buf][lrow][lcol]) = make_float4(ra[0], ra[1], ra[2], ra[3]);
        *reinterpret_cast<float4*>(&sB[buf][lrow][lcol]) = make_float4(rb[0], rb[1], rb[2], rb[3]);
        __syncthreads();
        if (m0 + kTM < m_end) load(m0 + kTM);
#pragma unroll
        for (int mm = 0; mm < kTM; ++mm) {
            const float4 a4 = *reinterpret_cast<const float4*>(&sA[buf][mm][tn]);
            const float4 b4 = *reinterpret_cast<const float4*>(&sB[buf][mm][tk]);
            const float av[4] = {a4.x, a4.y, a4.z, a4.w};
            const float bv[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(av[a], bv[b], acc[a][b]);
        }
        buf ^= 1;  // the next iteration writes the other buffer: one barrier per m-tile
    }
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        const int n = n0 + tn + a;
        if (n >= N) continue;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int k = k0 + tk + b;
            if (k < K) atomicAdd(dw + static_cast<int64_t>(n) * lddw + k, acc[a][b]);
        }
    }
}


// ---- tensor-pipe variant: mma.sync m16n8k8 TF32 with the 3-term split (x = hi + lo, hi = mantissa truncated to
// 10 bits, lo = x - hi exact; products lo*hi + hi*lo + hi*hi, fp32 accumulate) -- fp32-class accuracy, the same
// scheme as the forward projection.  Roles: mma-M = n (rows of dW), mma-N = k (columns of dW), mma-K = m (the
// reduced row index), so A(n, m) = dY[m][n] and B(m, k) = X[m][k] are both read from the m-major tiles as loaded.
constexpr int kMN = 64, kMK = 64, kMM = 32, kMThreads = 128, kMLd = 72;  // 72: fragment reads hit 32 distinct banks

__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
    hi = __float_as_uint(x) & 0xffffe000u;
    lo = __float_as_uint(x - __uint_as_float(hi));
}

__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <bool VEC2>
__global__ void __launch_bounds__(kMThreads, 4) k_wgrad_mma(const float* __restrict__ dy, int64_t lddy,
                                                         const float* __restrict__ x, int64_t ldx, int M, int N, int K,
                                                         float* __restrict__ dw, int64_t lddw, int rows_per_split) {
    __shared__ __align__(16) float sA[2][kMM][kMLd];
    __shared__ __align__(16) float sB[2][kMM][kMLd];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int gid = lane >> 2, tig = lane & 3;
    const int n0 = blockIdx.x * kMN, k0 = blockIdx.y * kMK;
    const int m_beg = blockIdx.z * rows_per_split;
    const int m_end = min(M, m_beg + rows_per_split);
    if (m_beg >= m_end) return;
    const int wn = (w >> 1) * 32, wk = (w & 1) * 32;
    const int lcol = (tid & 31) * 2, lrow = tid >> 5;  // loader: 2 columns, rows lrow + 4 i
    float acc[2][4][4];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int t = 0; t < 4; ++t) acc[i][j][t] = 0.f;
    float2 ra[8], rb[8];
    // Column clamp instead of predicates: dW[n][k] depends on column n of dY and column k of X only, so a clamped
    // (in-bounds, wrong) column only feeds accumulators that are never stored.  Rows are predicated in the one partial
    // stage at the end of the row range.
    const int ca = VEC2 ? min(n0 + lcol, (N - 2) & ~1) : min(n0 + lcol, N - 1);
    const int cb = VEC2 ? min(k0 + lcol, (K - 2) & ~1) : min(k0 + lcol, K - 1);
    const int ca1 = min(n0 + lcol + 1, N - 1), cb1 = min(k0 + lcol + 1, K - 1);  // scalar variant
    const float* pa = dy + static_cast<int64_t>(m_beg + lrow) * lddy;
    const float* pb = x + static_cast<int64_t>(m_beg + lrow) * ldx;
    auto load = [&](int m0) {
        if (m0 + kMM <= m_end) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (VEC2) {
                    ra[i] = __ldg(reinterpret_cast<const float2*>(pa + static_cast<int64_t>(4 * i) * lddy + ca));
                    rb[i] = __ldg(reinterpret_cast<const float2*>(pb + static_cast<int64_t>(4 * i) * ldx + cb));
                } else {
                    ra[i] = make_float2(__ldg(pa + static_cast<int64_t>(4 * i) * lddy + ca),
                                        __ldg(pa + static_cast<int64_t>(4 * i) * lddy + ca1));
                    rb[i] = make_float2(__ldg(pb + static_cast<int64_t>(4 * i) * ldx + cb),
                                        __ldg(pb + static_cast<int64_t>(4 * i) * ldx + cb1));
                }
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const bool ok = m0 + lrow + 4 * i < m_end;
                ra[i] = rb[i] = make_float2(0.f, 0.f);
                if (ok) {
                    if (VEC2) {
                        ra[i] = __ldg(reinterpret_cast<const float2*>(pa + static_cast<int64_t>(4 * i) * lddy + ca));
                        rb[i] = __ldg(reinterpret_cast<const float2*>(pb + static_cast<int64_t>(4 * i) * ldx + cb));
                    } else {
                        ra[i] = make_float2(__ldg(pa + static_cast<int64_t>(4 * i) * lddy + ca),
                                            __ldg(pa + static_cast<int64_t>(4 * i) * lddy + ca1));
                        rb[i] = make_float2(__ldg(pb + static_cast<int64_t>(4 * i) * ldx + cb),
                                            __ldg(pb + static_cast<int64_t>(4 * i) * ldx + cb1));
                    }
                }
            }
        }
        pa += static_cast<int64_t>(kMM) * lddy;
        pb += static_cast<int64_t>(kMM) * ldx;
    };
    load(m_beg);
    int buf = 0;
    for (int m0 = m_beg; m0 < m_end; m0 += kMM) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            *reinterpret_cast<float2*>(&sA[buf][lrow + 4 * i][lcol]) = ra[i];
            *reinterpret_cast<float2*>(&sB[buf][lrow + 4 * i][lcol]) = rb[i];
        }
        __syncthreads();
        if (m0 + kMM < m_end) load(m0 + kMM);
#pragma unroll
        for (int ks = 0; ks < kMM / 8; ++ks) {
            const int r0 = ks * 8 + tig, r1 = r0 + 4;
            uint32_t ah[2][4], al[2][4], bh[4][2], bl[4][2];
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int c = wn + i * 16 + gid;
                split_tf32(sA[buf][r0][c], ah[i][0], al[i][0]);
                split_tf32(sA[buf][r0][c + 8], ah[i][1], al[i][1]);
                split_tf32(sA[buf][r1][c], ah[i][2], al[i][2]);
                split_tf32(sA[buf][r1][c + 8], ah[i][3], al[i][3]);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int c = wk + j * 8 + gid;
                split_tf32(sB[buf][r0][c], bh[j][0], bl[j][0]);
                split_tf32(sB[buf][r1][c], bh[j][1], bl[j][1]);
            }
            // term-major order: eight independent accumulators between two MMAs on the same one
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) mma_tf32(acc[i][j], al[i], bh[j]);
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) mma_tf32(acc[i][j], ah[i], bl[j]);
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) mma_tf32(acc[i][j], ah[i], bh[j]);
        }
        buf ^= 1;
    }
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const int n = n0 + wn + i * 16 + gid + (t >> 1) * 8;
                const int k = k0 + wk + j * 8 + 2 * tig + (t & 1);
                if (n < N && k < K) atomicAdd(dw + static_cast<int64_t>(n) * lddw + k, acc[i][j][t]);
            }
}

// g = dy * act'(y), dbias += column sums of g.  Block = 32 columns x 8 rows, grid-stride over rows.
__global__ void __launch_bounds__(256) k_act_bwd(const float* __restrict__ dy, int64_t lddy, const float* __restrict__ y,
                                                 int64_t ldy, int64_t M, int N, int act, float* __restrict__ g,
                                                 int64_t ldg, float* __restrict__ dbias) {
    __shared__ float s_sum[8][33];
    const int c = blockIdx.y * 32 + threadIdx.x;
    float acc = 0.f;
    if (c < N) {
        for (int64_t m = static_cast<int64_t>(blockIdx.x) * 8 + threadIdx.y; m < M; m += static_cast<int64_t>(gridDim.x) * 8) {
            float d = dy[m * lddy + c];
            if (act != HG_ACT_NONE) {
                const float o = __ldg(y + m * ldy + c);
                if (act == HG_ACT_ELU) d *= o > 0.f ? 1.f : o + 1.f;
                else d = o > 0.f ? d : 0.f;
            }
            if (g) g[m * ldg + c] = d;
            acc += d;
        }
    }
    if (!dbias) return;
    s_sum[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && c < N) {
        float t = 0.f;
#pragma unroll
        for (int r = 0; r < 8; ++r) t += s_sum[r][threadIdx.x];
        atomicAdd(dbias + c, t);
    }
}

__global__ void k_zero_rows(float* __restrict__ p, int64_t ld, int N, int K) {
    const int64_t total = static_cast<int64_t>(N) * K;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total;
         t += static_cast<int64_t>(gridDim.x) * blockDim.x)
        p[(t / K) * ld + t % K] = 0.f;
}

__global__ void __launch_bounds__(256) k_adam(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                              float* __restrict__ v, int64_t n, float lr, float b1, float b2, float eps,
                                              float wd, float gscale, int step, const int32_t* __restrict__ step_dev) {
    // torch.optim.Adam (single-tensor path): step_size = lr / (1 - b1^t); denom = sqrt(v) / sqrt(1 - b2^t) + eps
    const int t_step = step_dev ? *step_dev : step;
    const double bc1 = 1.0 - pow(static_cast<double>(b1), static_cast<double>(t_step));
    const double bc2 = 1.0 - pow(static_cast<double>(b2), static_cast<double>(t_step));
    const float step_size = static_cast<float>(static_cast<double>(lr) / bc1);
    const float inv_sqrt_bc2 = static_cast<float>(1.0 / sqrt(bc2));
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < n;
         t += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const float pt = p[t];
        const float gt = fmaf(wd, pt, g[t] * gscale);
        const float mt = b1 * m[t] + (1.f - b1) * gt;
        const float vt = b2 * v[t] + (1.f - b2) * gt * gt;
        m[t] = mt;
        v[t] = vt;
        p[t] = pt - step_size * mt / (sqrtf(vt) * inv_sqrt_bc2 + eps);
    }
}

}  // namespace
}  // namespace hg

extern "C" int hg_linear_wgrad(const float* dy, int64_t lddy, const float* x, int64_t ldx, int32_t M, int32_t N, int32_t K,
                               float* dw, int64_t lddw, int32_t accumulate, void* stream_) {
    using namespace hg;
    if (M < 0 || N < 0 || K < 0 || lddy < N || ldx < K || lddw < K) return HG_EINVAL;
    if (N == 0 || K == 0) return HG_OK;
    if (!dw || (M > 0 && (!dy || !x))) return HG_EINVAL;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int sms = sm_count() > 0 ? sm_count() : 148;
    if (!accumulate) {
        const int64_t total = static_cast<int64_t>(N) * K;
        const int64_t g = ceil_div<int64_t>(total, 256);
        k_zero_rows<<<static_cast<unsigned>(g < sms * 8 ? g : sms * 8), 256, 0, stream>>>(dw, lddw, N, K);
    }
    if (M > 0) {
        // tcgen05 variant (wgrad_tc.cu): measured on a B200 at M = 488,696 -- [396 x 306] 1.35 ms against 2.26 ms for the
        // mma.sync kernel below, but 454 / 180 / 465 us against 436 / 134 / 435 us at [396 x 64], [88 x 64], [64 x 396]
        // (profiles/r02_wgrad_tc_first_run.txt) -- so it is the default only for large weight matrices over many rows.
        // HG_WGRAD_TC=1 / 0 forces it on / off (read per call so one process can A/B).
        const char* tc_env = getenv("HG_WGRAD_TC");
        const bool tc_auto = static_cast<int64_t>(N) * K >= 65536 && M >= 16384;
        if (tc_env ? tc_env[0] == '1' : tc_auto) {
            const int rc = check_launch();  // the zeroing launch
            return rc != HG_OK ? rc : wgrad_tc(dy, lddy, x, ldx, M, N, K, dw, lddw, stream);
        }
        static const bool use_fma = getenv("HG_WGRAD_FMA") != nullptr;  // A/B: the CUDA-core variant
        const int tn = ceil_div(N, kMN), tk = ceil_div(K, kMK);
        // one wave: as many row splits as fit the resident CTA slots (a 2nd, nearly empty wave doubles the time)
        static thread_local int occ_mma = 0, occ_fma = 0;
        if (!occ_mma) {
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_mma, k_wgrad_mma<true>, kMThreads, 0);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_fma, k_wgrad, kWgThreads, 0);
            occ_mma = occ_mma > 0 ? occ_mma : 1;
            occ_fma = occ_fma > 0 ? occ_fma : 1;
        }
        int splits = sms * (use_fma ? occ_fma : occ_mma) / (tn * tk);
        const int max_splits = ceil_div(M, 4 * kMM);
        splits = splits < 1 ? 1 : (splits > max_splits ? max_splits : splits);
        int rows = ceil_div(M, splits);
        rows = ceil_div(rows, kMM) * kMM;
        splits = ceil_div(M, rows);
        const dim3 grid(tn, tk, splits);
        if (use_fma) {
            k_wgrad<<<grid, kWgThreads, 0, stream>>>(dy, lddy, x, ldx, M, N, K, dw, lddw, rows);
        } else if (N % 2 == 0 && K % 2 == 0 && lddy % 2 == 0 && ldx % 2 == 0 && aligned(dy, 8) && aligned(x, 8)) {
            // float2 loads: with an odd width the last column would share its pair with the clamp and load the wrong one
            k_wgrad_mma<true><<<grid, kMThreads, 0, stream>>>(dy, lddy, x, ldx, M, N, K, dw, lddw, rows);
        } else {
            k_wgrad_mma<false><<<grid, kMThreads, 0, stream>>>(dy, lddy, x, ldx, M, N, K, dw, lddw, rows);
        }
    }
    return check_launch();
}

extern "C" int hg_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                            float beta1, float beta2, float eps, float weight_decay, float grad_scale, int32_t step,
                            const int32_t* step_dev, void* stream_) {
    using namespace hg;
    if (n < 0 || (!step_dev && step < 1)) return HG_EINVAL;
    if (n == 0) return HG_OK;
    if (!param || !grad || !exp_avg || !exp_avg_sq) return HG_EINVAL;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int64_t g = ceil_div<int64_t>(n, 256);
    k_adam<<<static_cast<unsigned>(g < sms * 8 ? g : sms * 8), 256, 0, stream>>>(
        param, grad, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay, grad_scale, step, step_dev);
    return check_launch();
}

extern "C" int hg_act_bwd(const float* dy, int64_t lddy, const float* y, int64_t ldy, int64_t M, int32_t N, int32_t act,
                          float* g, int64_t ldg, float* dbias, void* stream_) {
    using namespace hg;
    if (M < 0 || N < 0) return HG_EINVAL;
    if (M == 0 || N == 0 || (!g && !dbias)) return HG_OK;
    if (!dy || (act != HG_ACT_NONE && !y)) return HG_EINVAL;
    if (act == HG_ACT_NONE && !dbias && g == dy) return HG_OK;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int sms = sm_count() > 0 ? sm_count() : 148;
    const int cols = ceil_div(N, 32);
    int64_t gx = ceil_div<int64_t>(M, 8);
    const int64_t cap = ceil_div<int64_t>(static_cast<int64_t>(sms) * 8, cols);
    gx = gx < cap ? gx : cap;
    k_act_bwd<<<dim3(static_cast<unsigned>(gx > 0 ? gx : 1), cols), dim3(32, 8), 0, stream>>>(dy, lddy, y, ldy, M, N, act, g,
                                                                                          ldg, dbias);
    return check_launch();
}
