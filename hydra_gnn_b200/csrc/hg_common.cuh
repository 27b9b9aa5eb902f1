// Shared device/host helpers for the hydra_gnn_b200 kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "hydra_gnn_b200.h"

namespace hg {

constexpr int kWarp = 32;

int record_cuda(cudaError_t e);  // stores e thread-locally; returns HG_ECUDA or HG_OK
int check_launch();              // cudaGetLastError() -> code

#define HG_CUDA_TRY(expr)                                   \
    do {                                                    \
        cudaError_t _e = (expr);                            \
        if (_e != cudaSuccess) return ::hg::record_cuda(_e); \
    } while (0)

template <typename T>
__host__ __device__ inline T ceil_div(T a, T b) { return (a + b - 1) / b; }

inline bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

// Grid sizing: B200 has 148 SMs; grids are rounded to whole waves of (SMs x resident CTAs).
int sm_count();

template <int VEC> struct VecT;
template <> struct VecT<1> { using type = float; };
template <> struct VecT<2> { using type = float2; };
template <> struct VecT<4> { using type = float4; };

// Read-only-path vector loads (ld.global.nc) / plain stores; VEC in {1,2,4}.
template <int VEC>
__device__ __forceinline__ void ldg_vec(const float* __restrict__ p, float (&v)[VEC]) {
    if constexpr (VEC == 4) {
        float4 t = __ldg(reinterpret_cast<const float4*>(p));
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else if constexpr (VEC == 2) {
        float2 t = __ldg(reinterpret_cast<const float2*>(p));
        v[0] = t.x; v[1] = t.y;
    } else {
        v[0] = __ldg(p);
    }
}

template <int VEC>
__device__ __forceinline__ void st_vec(float* __restrict__ p, const float (&v)[VEC]) {
    if constexpr (VEC == 4) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    } else if constexpr (VEC == 2) {
        *reinterpret_cast<float2*>(p) = make_float2(v[0], v[1]);
    } else {
        *p = v[0];
    }
}

__device__ __forceinline__ float apply_act(float v, int act) {
    if (act == HG_ACT_RELU) return v > 0.f ? v : 0.f;
    if (act == HG_ACT_ELU) return v > 0.f ? v : expm1f(v);  // F.elu, alpha = 1
    return v;
}


// Attention-dropout decision of (relation slot r, flat index idx = CSR position * H + head): splitmix64 of the seed
// and the counter.  Returns 1/(1-p) when kept, 0 when dropped.
__device__ __forceinline__ float dropout_keep_scale(uint64_t seed, int r, int64_t idx, float p, float inv_keep) {
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (static_cast<uint64_t>(idx) * HG_MAX_RELATIONS + r + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    const float u = static_cast<float>(z >> 40) * (1.0f / 16777216.0f);
    return u >= p ? inv_keep : 0.f;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace hg
