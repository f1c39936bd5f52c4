// Device-memory and IPC helpers of the peer-memory runtime (host side).
//
// Buffers that other processes map (receive slots, flag words, gradient arenas) come from
// cudaMalloc directly so a cudaIpcMemHandle can be taken for them; handles travel through the
// torch.distributed store (setup path only).
#include <cuda.h>
#include <string.h>

#include <map>
#include <mutex>
#include <string>

#include "common.cuh"

extern "C" int lwc_malloc(void** out_dev, size_t bytes) {
    if (!out_dev) return LWC_EINVAL;
    *out_dev = nullptr;
    LWC_CUDA_CHECK(cudaMalloc(out_dev, bytes ? bytes : 256));
    return 0;
}

extern "C" int lwc_free(void* dev) {
    if (!dev) return 0;
    LWC_CUDA_CHECK(cudaFree(dev));
    return 0;
}

extern "C" int lwc_memset_async(void* dev, int value, size_t bytes, void* stream) {
    LWC_CUDA_CHECK(cudaMemsetAsync(dev, value, bytes, (cudaStream_t)stream));
    return 0;
}

// A cudaIpcMemHandle names the driver's whole underlying allocation: cudaMalloc may carve several small requests
// out of one block, and cudaIpcOpenMemHandle then returns the BLOCK's base in the importing process.  So the
// exporter also publishes the offset of its pointer inside the block (lwc_ipc_base_offset) and the importer adds
// it to what lwc_ipc_open_handle returns.  Opens are cached and reference-counted per handle: two heaps that share
// a block map it once.
namespace {
struct OpenedIpc { void* base; int refs; };
std::mutex g_ipc_mu;
std::map<std::string, OpenedIpc> g_ipc_by_handle;
std::map<void*, std::string> g_ipc_by_base;
}  // namespace

extern "C" int lwc_ipc_get_handle(void* dev, void* handle_out_64B) {
    if (!dev || !handle_out_64B) return LWC_EINVAL;
    cudaIpcMemHandle_t h;
    LWC_CUDA_CHECK(cudaIpcGetMemHandle(&h, dev));
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle_out_64B, &h, 64);
    return 0;
}

extern "C" int lwc_ipc_base_offset(void* dev, uint64_t* offset_out) {
    if (!dev || !offset_out) return LWC_EINVAL;
    CUdeviceptr base = 0;
    size_t size = 0;
    if (cuMemGetAddressRange(&base, &size, (CUdeviceptr)dev) != CUDA_SUCCESS) return LWC_ESTATE;
    *offset_out = (uint64_t)((CUdeviceptr)dev - base);
    return 0;
}

extern "C" int lwc_ipc_open_handle(const void* handle_64B, void** out_dev) {
    if (!handle_64B || !out_dev) return LWC_EINVAL;
    const std::string key((const char*)handle_64B, 64);
    std::lock_guard<std::mutex> lock(g_ipc_mu);
    auto it = g_ipc_by_handle.find(key);
    if (it != g_ipc_by_handle.end()) {
        it->second.refs++;
        *out_dev = it->second.base;
        return 0;
    }
    cudaIpcMemHandle_t h;
    memcpy(&h, handle_64B, 64);
    LWC_CUDA_CHECK(cudaIpcOpenMemHandle(out_dev, h, cudaIpcMemLazyEnablePeerAccess));
    g_ipc_by_handle[key] = OpenedIpc{*out_dev, 1};
    g_ipc_by_base[*out_dev] = key;
    return 0;
}

extern "C" int lwc_ipc_close_handle(void* dev) {
    if (!dev) return 0;
    std::lock_guard<std::mutex> lock(g_ipc_mu);
    auto it = g_ipc_by_base.find(dev);
    if (it == g_ipc_by_base.end()) return LWC_EINVAL;
    auto h = g_ipc_by_handle.find(it->second);
    if (--h->second.refs > 0) return 0;
    g_ipc_by_handle.erase(h);
    g_ipc_by_base.erase(it);
    LWC_CUDA_CHECK(cudaIpcCloseMemHandle(dev));
    return 0;
}

extern "C" int lwc_enable_peer_access(int peer_device) {
    int cur = 0;
    LWC_CUDA_CHECK(cudaGetDevice(&cur));
    if (cur == peer_device) return 0;
    int can = 0;
    LWC_CUDA_CHECK(cudaDeviceCanAccessPeer(&can, cur, peer_device));
    if (!can) return LWC_ESTATE;
    cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return 0; }
    return (int)e;
}

extern "C" const char* lwc_error_string(int code) {
    switch (code) {
        case 0: return "success";
        case LWC_EINVAL: return "lwc: invalid argument";
        case LWC_ENOMEM: return "lwc: out of memory";
        case LWC_ETOOBIG: return "lwc: size beyond the supported maximum";
        case LWC_ESTATE: return "lwc: object not in a usable state";
        default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "lwc: unknown error";
    }
}
