// Device-memory and IPC helpers of the peer-memory runtime (host side).
//
// Buffers that other processes map (receive slots, flag words, gradient arenas) come from
// cudaMalloc directly so a cudaIpcMemHandle can be taken for them; handles travel through the
// torch.distributed store (setup path only).
#include <string.h>

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

extern "C" int lwc_ipc_get_handle(void* dev, void* handle_out_64B) {
    if (!dev || !handle_out_64B) return LWC_EINVAL;
    cudaIpcMemHandle_t h;
    LWC_CUDA_CHECK(cudaIpcGetMemHandle(&h, dev));
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle_out_64B, &h, 64);
    return 0;
}

extern "C" int lwc_ipc_open_handle(const void* handle_64B, void** out_dev) {
    if (!handle_64B || !out_dev) return LWC_EINVAL;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle_64B, 64);
    LWC_CUDA_CHECK(cudaIpcOpenMemHandle(out_dev, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}

extern "C" int lwc_ipc_close_handle(void* dev) {
    if (!dev) return 0;
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
