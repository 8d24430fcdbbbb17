// Peer-mapped buffers for the final frame gather ("write straight into a peer-mapped slot of the
// consumer rank's buffer").  One process per GPU: the consumer rank allocates the frame-ordered gather buffers with
// cudaMalloc and exports them as CUDA IPC handles; every producer rank maps them and moves its finished frames with
// ONE cudaMemcpyAsync per step - a copy-engine transfer over NVLink that occupies no SM on either GPU (an NCCL
// send/recv pair runs a kernel on both ends, which takes SMs away from the persistent conv kernels of the next
// step).  Completion is signalled through 32-bit flags in the consumer's memory: the producer's stream writes the step
// number behind its copy, the consumer's stream waits for it with cuStreamWaitValue32 (no spinning kernel).
// No counterpart in the reference (it has no multi-GPU inference path).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cudaTypedefs.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/ghost_b200.h"
#include "crop.cuh"  // ghost::set_last_error

namespace {

int fail(const char* what, cudaError_t e) {
  char msg[256];
  snprintf(msg, sizeof(msg), "%s: %s", what, cudaGetErrorString(e));
  ghost::set_last_error(msg);
  cudaGetLastError();
  return GHOST_ERR_CUDA;
}

// after the copy of a step: publish the step number in the consumer's flag word (a store over NVLink)
__global__ void peer_flag_kernel(volatile uint32_t* flag, uint32_t value) {
  __threadfence_system();
  *flag = value;
  __threadfence_system();
}

typedef CUresult (*WaitValue32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
WaitValue32Fn wait_value_fn() {
  static WaitValue32Fn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<WaitValue32Fn>(p);
  }
  return fn;
}

}  // namespace

extern "C" {

int ghost_peer_alloc(size_t bytes, void** dev_ptr, unsigned char* handle64) {
  if (!dev_ptr || !handle64 || bytes == 0) {
    ghost::set_last_error("ghost_peer_alloc: bad argument");
    return GHOST_ERR_INVALID;
  }
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e != cudaSuccess) return fail("ghost_peer_alloc: cudaMalloc", e);
  e = cudaMemset(p, 0, bytes);
  if (e != cudaSuccess) return fail("ghost_peer_alloc: cudaMemset", e);
  cudaIpcMemHandle_t h;
  e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return fail("ghost_peer_alloc: cudaIpcGetMemHandle", e);
  }
  static_assert(sizeof(h) == 64, "CUDA IPC handles are 64 bytes");
  memcpy(handle64, &h, 64);
  *dev_ptr = p;
  return GHOST_OK;
}

int ghost_peer_open(const unsigned char* handle64, void** dev_ptr) {
  if (!dev_ptr || !handle64) {
    ghost::set_last_error("ghost_peer_open: bad argument");
    return GHOST_ERR_INVALID;
  }
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return fail("ghost_peer_open: cudaIpcOpenMemHandle", e);
  *dev_ptr = p;
  return GHOST_OK;
}

int ghost_peer_close(void* dev_ptr) {
  cudaError_t e = cudaIpcCloseMemHandle(dev_ptr);
  return e == cudaSuccess ? GHOST_OK : fail("ghost_peer_close", e);
}

int ghost_peer_free(void* dev_ptr) {
  cudaError_t e = cudaFree(dev_ptr);
  return e == cudaSuccess ? GHOST_OK : fail("ghost_peer_free", e);
}

int ghost_peer_copy(void* dst, const void* src, size_t bytes, void* stream) {
  cudaError_t e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, reinterpret_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? GHOST_OK : fail("ghost_peer_copy", e);
}

int ghost_peer_signal(uint32_t* flag, uint32_t value, void* stream) {
  peer_flag_kernel<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream)>>>(flag, value);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? GHOST_OK : fail("ghost_peer_signal", e);
}

int ghost_peer_wait(uint32_t* flag, uint32_t value, void* stream) {
  WaitValue32Fn fn = wait_value_fn();
  if (!fn) {
    ghost::set_last_error("ghost_peer_wait: cuStreamWaitValue32 is not available");
    return GHOST_ERR_CUDA;
  }
  CUresult r = fn(reinterpret_cast<CUstream>(stream), reinterpret_cast<CUdeviceptr>(flag), value, CU_STREAM_WAIT_VALUE_GEQ);
  if (r != CUDA_SUCCESS) {
    char msg[96];
    snprintf(msg, sizeof(msg), "ghost_peer_wait: cuStreamWaitValue32 failed (%d)", static_cast<int>(r));
    ghost::set_last_error(msg);
    return GHOST_ERR_CUDA;
  }
  return GHOST_OK;
}

}  // extern "C"
