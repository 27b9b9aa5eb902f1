// Library-level entry points: error text, ABI version, device probe.
#include "hg_common.cuh"

namespace hg {
namespace {
thread_local int t_last_cuda = 0;
}

int record_cuda(cudaError_t e) {
    t_last_cuda = static_cast<int>(e);
    return e == cudaSuccess ? HG_OK : HG_ECUDA;
}

int check_launch() { return record_cuda(cudaGetLastError()); }

int sm_count() {
    static thread_local int cached_dev = -1;
    static thread_local int cached = 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (dev != cached_dev) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
        cached = n;
        cached_dev = dev;
    }
    return cached;
}
}  // namespace hg

extern "C" const char* hg_strerror(int code) {
    switch (code) {
        case HG_OK: return "ok";
        case HG_EINVAL: return "invalid argument";
        case HG_ERANGE: return "size or index out of int32 range";
        case HG_EWORKSPACE: return "workspace too small";
        case HG_ECUDA: return "CUDA runtime error (see hg_last_cuda_error)";
        case HG_EALIGN: return "pointer or leading dimension misaligned";
        case HG_EUNSUPPORTED: return "unsupported shape for this kernel";
        case HG_ECAPACITY: return "problem exceeds the fixed scratch capacity of this *_static entry point";
        default: return "unknown error";
    }
}

extern "C" int hg_abi_version(void) { return HG_ABI_VERSION; }

extern "C" int hg_last_cuda_error(void) { return hg::t_last_cuda; }

extern "C" int hg_device_sm_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
        (void)cudaGetLastError();
        return 0;
    }
    return hg::sm_count();
}

// sizeof() of every ABI struct, so a binding can verify its layout at load time.
extern "C" int hg_sizeof(int which) {
    switch (which) {
        case 0: return static_cast<int>(sizeof(hg_plan_rel));
        case 1: return static_cast<int>(sizeof(hg_seg_args));
        case 2: return static_cast<int>(sizeof(hg_gat_rel));
        case 3: return static_cast<int>(sizeof(hg_gat_args));
        case 4: return static_cast<int>(sizeof(hg_lin_seg));
        case 5: return static_cast<int>(sizeof(hg_linear_args));
        case 6: return static_cast<int>(sizeof(hg_gat_train_rel));
        case 7: return static_cast<int>(sizeof(hg_gat_train_args));
        case 8: return static_cast<int>(sizeof(hg_htree_split_args));
        case 9: return static_cast<int>(sizeof(hg_synth_args));
        default: return -1;
    }
}
