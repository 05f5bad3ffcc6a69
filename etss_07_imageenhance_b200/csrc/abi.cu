// Version / error-text / device-check entry points of the C ABI (include/b200dce.h).
#include "common.cuh"

extern "C" int b200_version(void) { return 100; }

extern "C" const char* b200_strerror(int code) {
    switch (code) {
        case B200_OK: return "ok";
        case B200_ERR_NULL: return "a required pointer is NULL";
        case B200_ERR_BAD_SHAPE: return "bad shape (non-positive dims, or H/W not divisible by scale_factor)";
        case B200_ERR_MISALIGNED: return "tensor base pointer is not 16-byte aligned";
        case B200_ERR_UNSUPPORTED: return "unsupported hyper-parameter for the sm_100a kernels";
        case B200_ERR_WORKSPACE: return "workspace too small";
        case B200_ERR_LAUNCH: return "CUDA launch failed";
        case B200_ERR_NO_DEVICE: return "no sm_100 (B200) CUDA device is current";
        case B200_ERR_TOO_LARGE: return "problem too large for the built-in scratch array of this *_static entry point (use the variant that takes a workspace)";
        default: return "unknown b200dce error";
    }
}

extern "C" int b200_check_device(void) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        (void)cudaGetLastError();
        return B200_ERR_NO_DEVICE;
    }
    // the library carries sm_100a SASS only (arch-specific target: it does not load on sm_103 or any other 10.x part)
    int major = 0, minor = -1;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev) != cudaSuccess) {
        (void)cudaGetLastError();
        return B200_ERR_NO_DEVICE;
    }
    return major == 10 && minor == 0 ? B200_OK : B200_ERR_NO_DEVICE;
}
