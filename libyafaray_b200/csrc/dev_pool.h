// Scene and BVH-builder buffers come from the device's stream-ordered memory pool (release threshold raised in yafgpu_create, so freed
// blocks stay in the pool): re-uploading a scene re-uses the previous scene's memory instead of paying cudaFree's device-wide
// synchronisation and the map/unmap of cudaMalloc on every call.  All allocations and frees are ordered on the context's stream.
#pragma once
#include <cuda_runtime.h>

namespace yg {
template<class T> static inline cudaError_t poolAlloc(T **p, size_t bytes, cudaStream_t s) { return cudaMallocAsync((void **)p, bytes ? bytes : 16, s); }
static inline void poolFree(void *p, cudaStream_t s) { if(p) cudaFreeAsync(p, s); }
}
