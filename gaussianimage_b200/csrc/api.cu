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

// Fill kernel that stands in for a 1-D memset node of a captured graph (gi2d_graph_fold_memsets).
__global__ void graph_fill_kernel(unsigned char *dst, unsigned int value, unsigned int elem, size_t width) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= width) return;
    if (elem == 4) reinterpret_cast<unsigned int *>(dst)[i] = value;
    else if (elem == 2) reinterpret_cast<unsigned short *>(dst)[i] = (unsigned short)value;
    else dst[i] = (unsigned char)value;
}

}  // namespace gi2d

extern "C" {

int gi2d_abi_version(void) { return 2; }

// See include/gi2d.h.  Small 1-D memset nodes only (what cudaMemsetAsync of a semaphore / counter buffer records); anything
// else is left alone.
int gi2d_graph_fold_memsets(void *graph_, size_t max_bytes, int *n_replaced) {
    cudaGraph_t graph = reinterpret_cast<cudaGraph_t>(graph_);
    if (!graph) return GI2D_ERR_INVALID;
    if (n_replaced) *n_replaced = 0;
    size_t n = 0;
    cudaError_t e = cudaGraphGetNodes(graph, nullptr, &n);
    if (e != cudaSuccess) return GI2D_ERR_CUDA_BASE + (int)e;
    if (n == 0) return GI2D_OK;
    cudaGraphNode_t *nodes = new cudaGraphNode_t[n];
    e = cudaGraphGetNodes(graph, nodes, &n);
    int rc = GI2D_OK, done = 0;
    for (size_t k = 0; k < n && e == cudaSuccess; ++k) {
        cudaGraphNodeType type;
        if ((e = cudaGraphNodeGetType(nodes[k], &type)) != cudaSuccess) break;
        if (type != cudaGraphNodeTypeMemset) continue;
        cudaMemsetParams mp;
        if ((e = cudaGraphMemsetNodeGetParams(nodes[k], &mp)) != cudaSuccess) break;
        if (mp.height != 1 || mp.width == 0 || (mp.elementSize != 1 && mp.elementSize != 2 && mp.elementSize != 4) ||
            mp.width * mp.elementSize > max_bytes)
            continue;
        size_t nd = 0, nt = 0;
        if ((e = cudaGraphNodeGetDependencies(nodes[k], nullptr, &nd)) != cudaSuccess) break;
        if ((e = cudaGraphNodeGetDependentNodes(nodes[k], nullptr, &nt)) != cudaSuccess) break;
        cudaGraphNode_t *deps = new cudaGraphNode_t[nd + 1], *outs = new cudaGraphNode_t[nt + 1], *froms = new cudaGraphNode_t[nt + 1];
        if (nd) e = cudaGraphNodeGetDependencies(nodes[k], deps, &nd);
        if (e == cudaSuccess && nt) e = cudaGraphNodeGetDependentNodes(nodes[k], outs, &nt);
        if (e == cudaSuccess) {
            unsigned char *dst = static_cast<unsigned char *>(mp.dst);
            unsigned int value = mp.value, elem = mp.elementSize;
            size_t width = mp.width;
            void *args[4] = {&dst, &value, &elem, &width};
            cudaKernelNodeParams kp = {};
            kp.func = reinterpret_cast<void *>(gi2d::graph_fill_kernel);
            kp.blockDim = dim3(128);
            kp.gridDim = dim3((unsigned)((width + 127) / 128));
            kp.sharedMemBytes = 0;
            kp.kernelParams = args;
            kp.extra = nullptr;
            cudaGraphNode_t fill;
            e = cudaGraphAddKernelNode(&fill, graph, deps, nd, &kp);
            if (e == cudaSuccess && nt) {
                for (size_t j = 0; j < nt; ++j) froms[j] = fill;
                e = cudaGraphAddDependencies(graph, froms, outs, nt);
            }
            if (e == cudaSuccess) e = cudaGraphDestroyNode(nodes[k]);
            if (e == cudaSuccess) ++done;
        }
        delete[] deps; delete[] outs; delete[] froms;
    }
    delete[] nodes;
    if (e != cudaSuccess) rc = GI2D_ERR_CUDA_BASE + (int)e;
    if (n_replaced) *n_replaced = done;
    return rc;
}

const char *gi2d_status_string(int status) {
    if (status == GI2D_OK) return "ok";
    if (status == GI2D_ERR_INVALID) return "invalid argument";
    if (status == GI2D_ERR_WORKSPACE) return "workspace too small";
    if (status == GI2D_ERR_STATIC_LIMIT) return "size exceeds the library's static scratch (see the limit in include/gi2d.h)";
    if (status >= GI2D_ERR_CUDA_BASE) return cudaGetErrorString((cudaError_t)(status - GI2D_ERR_CUDA_BASE));
    return "unknown status";
}

}  // extern "C"
