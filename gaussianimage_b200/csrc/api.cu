// api.cu -- version / status helpers of the C ABI declared in include/gi2d.h.
#include "gi2d_common.cuh"

#include <mutex>

namespace gi2d {

unsigned int *sync_slot(cudaStream_t stream) {
    constexpr int EAGER = 4096, TOTAL = 16384, STRIDE = 8;      // 32-byte pairs: one L2 sector each
    static unsigned int *pool[16] = {nullptr};
    static unsigned int eager_next[16] = {0}, capture_next[16] = {0};
    static std::mutex lock;                                  // ctypes / the host glue release the GIL: host threads may race here
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return nullptr;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream, &cs) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    std::lock_guard<std::mutex> guard(lock);
    if (!pool[dev]) {
        if (cs != cudaStreamCaptureStatusNone) return nullptr;                       // no allocation / memset inside a capture
        unsigned int *p = nullptr;
        if (cudaMalloc(&p, sizeof(unsigned int) * STRIDE * TOTAL) != cudaSuccess ||
            cudaMemset(p, 0, sizeof(unsigned int) * STRIDE * TOTAL) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        pool[dev] = p;
    }
    unsigned int slot;
    if (cs == cudaStreamCaptureStatusNone) {
        slot = eager_next[dev]++ % EAGER;
    } else {
        if (capture_next[dev] >= (unsigned)(TOTAL - EAGER)) return nullptr;
        slot = EAGER + capture_next[dev]++;
    }
    return pool[dev] + (size_t)STRIDE * slot;
}

}  // namespace gi2d

extern "C" {

int gi2d_abi_version(void) { return 2; }

const char *gi2d_status_string(int status) {
    if (status == GI2D_OK) return "ok";
    if (status == GI2D_ERR_INVALID) return "invalid argument";
    if (status == GI2D_ERR_WORKSPACE) return "workspace too small";
    if (status == GI2D_ERR_STATIC_LIMIT) return "size exceeds the library's static scratch (see the limit in include/gi2d.h)";
    if (status >= GI2D_ERR_CUDA_BASE) return cudaGetErrorString((cudaError_t)(status - GI2D_ERR_CUDA_BASE));
    return "unknown status";
}

}  // extern "C"
