// K6 -- fused projection  y = act((sum_s x_s W_s^T + bias) * scale + shift).
//
// This file holds the dispatcher and the fp32 CUDA-core path (HG_PREC_FP32): a classic
// 128x64x16 shared-memory tiled SGEMM with 8x4 register micro-tiles, used for small / odd
// shapes and as the on-device cross-check of the tensor-core path.  The tcgen05 path
// (HG_PREC_TF32X3) lives in linear_tc.cu.
#include "hg_common.cuh"

namespace hg {

int linear_tc_supported(const hg_linear_args& a);
size_t linear_tc_packed_bytes(const hg_linear_args& a);
int linear_tc_pack(const hg_linear_args& a, void* dst, size_t dst_bytes, cudaStream_t stream);
int linear_tc_launch(const hg_linear_args& a, void* ws, size_t ws_bytes, cudaStream_t stream);

namespace {

constexpr int BM = 128, BN = 64, BK = 16;
constexpr int TM = 8, TN = 4;
constexpr int kThreads = (BM / TM) * (BN / TN);  // 256

struct LinParams {
    hg_linear_args a;
};

__global__ void __launch_bounds__(kThreads) k_linear_fp32(const __grid_constant__ LinParams P) {
    const hg_linear_args& A = P.a;
    __shared__ float sx[BK][BM + 4];
    __shared__ float sw[BK][BN + 4];
    const int tid = threadIdx.x;
    const int tx = tid % (BN / TN);  // column group
    const int ty = tid / (BN / TN);  // row group
    const int64_t m0 = static_cast<int64_t>(blockIdx.x) * BM;
    const int n0 = blockIdx.y * BN;

    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

    for (int s = 0; s < A.n_seg; ++s) {
        const hg_lin_seg& S = A.seg[s];
        for (int k0 = 0; k0 < S.K; k0 += BK) {
            // x tile: BM x BK, thread t loads elements (row = t / BK + 16*i..., col = t % BK)
#pragma unroll
            for (int it = 0; it < (BM * BK) / kThreads; ++it) {
                const int idx = it * kThreads + tid;
                const int r = idx / BK, c = idx % BK;
                const int64_t m = m0 + r;
                const int k = k0 + c;
                sx[c][r] = (m < A.M && k < S.K) ? __ldg(S.x + m * S.ldx + k) : 0.f;
            }
#pragma unroll
            for (int it = 0; it < (BN * BK) / kThreads; ++it) {
                const int idx = it * kThreads + tid;
                const int r = idx / BK, c = idx % BK;
                const int n = n0 + r;
                const int k = k0 + c;
                sw[c][r] = (n < A.N && k < S.K) ? __ldg(S.w + static_cast<int64_t>(n) * S.ldw + k) : 0.f;
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < BK; ++kk) {
                float xr[TM], wr[TN];
#pragma unroll
                for (int i = 0; i < TM; ++i) xr[i] = sx[kk][ty * TM + i];
#pragma unroll
                for (int j = 0; j < TN; ++j) wr[j] = sw[kk][tx * TN + j];
#pragma unroll
                for (int i = 0; i < TM; ++i)
#pragma unroll
                    for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(xr[i], wr[j], acc[i][j]);
            }
            __syncthreads();
        }
    }

#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int64_t m = m0 + ty * TM + i;
        if (m >= A.M) continue;
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            const int n = n0 + tx * TN + j;
            if (n >= A.N) continue;
            float v = acc[i][j];
            if (A.bias) v += __ldg(A.bias + n);
            if (A.scale) v *= __ldg(A.scale + n);
            if (A.shift) v += __ldg(A.shift + n);
            A.y[m * A.ldy + n] = apply_act(v, A.act);
        }
    }
}

int validate(const hg_linear_args* a) {
    if (!a) return HG_EINVAL;
    if (a->n_seg < 1 || a->n_seg > HG_MAX_SEGMENTS || a->M < 0 || a->N < 0) return HG_EINVAL;
    if (a->act < HG_ACT_NONE || a->act > HG_ACT_ELU) return HG_EINVAL;
    if (a->M == 0 || a->N == 0) return HG_OK;
    if (!a->y) return HG_EINVAL;
    for (int s = 0; s < a->n_seg; ++s) {
        if (a->seg[s].K < 0) return HG_EINVAL;
        if (a->seg[s].K > 0 && (!a->seg[s].x || !a->seg[s].w)) return HG_EINVAL;
    }
    return HG_OK;
}

}  // namespace
}  // namespace hg

extern "C" size_t hg_linear_workspace_bytes(const hg_linear_args* a) {
    if (!a || hg::validate(a) != HG_OK) return 0;
    if (a->precision == HG_PREC_TF32X3 && hg::linear_tc_supported(*a)) return hg::linear_tc_packed_bytes(*a);
    return 0;
}

extern "C" int hg_linear_pack_weights(const hg_linear_args* a, void* dst, size_t dst_bytes, void* stream_) {
    using namespace hg;
    if (!a || a->n_seg < 1 || a->n_seg > HG_MAX_SEGMENTS || a->N < 1) return HG_EINVAL;
    for (int s = 0; s < a->n_seg; ++s)
        if (a->seg[s].K < 0 || (a->seg[s].K > 0 && !a->seg[s].w)) return HG_EINVAL;
    if (a->precision != HG_PREC_TF32X3) return HG_OK;  // the fp32 path reads W as it is
    hg_linear_args b = *a;
    if (b.M < 1) b.M = 1;
    if (!linear_tc_supported(b)) return HG_EUNSUPPORTED;
    return linear_tc_pack(b, dst, dst_bytes, static_cast<cudaStream_t>(stream_));
}

extern "C" int hg_linear(const hg_linear_args* a, void* workspace, size_t workspace_bytes, void* stream_) {
    using namespace hg;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    int rc = validate(a);
    if (rc != HG_OK) return rc;
    if (a->M == 0 || a->N == 0) return HG_OK;
    if (a->precision == HG_PREC_TF32X3) {
        if (!linear_tc_supported(*a)) return HG_EUNSUPPORTED;
        return linear_tc_launch(*a, workspace, workspace_bytes, stream);
    }
    if (a->precision != HG_PREC_FP32) return HG_EINVAL;
    LinParams P;
    P.a = *a;
    const dim3 grid(static_cast<unsigned>(ceil_div<int64_t>(a->M, BM)), static_cast<unsigned>(ceil_div(a->N, BN)));
    k_linear_fp32<<<grid, kThreads, 0, stream>>>(P);
    return check_launch();
}
