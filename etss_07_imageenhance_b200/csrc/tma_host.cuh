// Host-side access to cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time libcuda
// dependency).  Shared by the Zero-DCE tensor-core path (dce_tc.cu) and the Zero-DCE++ low-resolution net (dcepp.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

namespace b200 {

typedef CUresult (*tma_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline tma_encode_fn tma_encoder() {
    static tma_encode_fn fn = []() -> tma_encode_fn {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess) {
            (void)cudaGetLastError();
            return nullptr;
        }
        return reinterpret_cast<tma_encode_fn>(p);
    }();
    return fn;
}

}  // namespace b200
