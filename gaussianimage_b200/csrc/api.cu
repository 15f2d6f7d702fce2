// api.cu -- version / status helpers of the C ABI declared in include/gi2d.h.
#include "gi2d_common.cuh"

extern "C" {

int gi2d_abi_version(void) { return 1; }

const char *gi2d_status_string(int status) {
    if (status == GI2D_OK) return "ok";
    if (status == GI2D_ERR_INVALID) return "invalid argument";
    if (status == GI2D_ERR_WORKSPACE) return "workspace too small";
    if (status >= GI2D_ERR_CUDA_BASE) return cudaGetErrorString((cudaError_t)(status - GI2D_ERR_CUDA_BASE));
    return "unknown status";
}

}  // extern "C"
