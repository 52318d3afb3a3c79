// Symmetric (peer-mapped) device memory for the in-kernel collectives.  No counterpart in the
// reference: its only inter-process data path is the TensorPipe RPC of CPU tensors
// (/root/reference/simple_distributed.py:47-49, :80 `.cpu()`), it has no collectives at all.
//
// Every rank cudaMalloc's a buffer of the same size, exports it as a CUDA IPC handle,
// the handles travel through the torch.distributed store, and every rank maps its
// peers' buffers into its own address space.  After that a kernel can ld/st/red any
// peer's buffer directly over NVLink / NVSwitch -- no NCCL call on the data path.
// (One process per GPU, so cudaIpc* is the mapping mechanism; peer access is enabled
// lazily by cudaIpcOpenMemHandle.)
#include "sdml_host.h"

#include <string.h>

using namespace sdml;

extern "C" int sdml_symm_alloc(long long bytes, void** out_ptr) {
  if (bytes <= 0 || !out_ptr) return set_error("sdml_symm_alloc: bad arguments");
  void* p = nullptr;
  SDML_CUDA_OK(cudaMalloc(&p, static_cast<size_t>(bytes)));
  SDML_CUDA_OK(cudaMemset(p, 0, static_cast<size_t>(bytes)));
  SDML_CUDA_OK(cudaDeviceSynchronize());
  *out_ptr = p;
  return 0;
}

extern "C" int sdml_symm_free(void* ptr) {
  if (ptr) SDML_CUDA_OK(cudaFree(ptr));
  return 0;
}

// handle_out: 64 bytes (CUDA_IPC_HANDLE_SIZE)
extern "C" int sdml_symm_export(void* ptr, unsigned char* handle_out) {
  cudaIpcMemHandle_t h;
  SDML_CUDA_OK(cudaIpcGetMemHandle(&h, ptr));
  static_assert(sizeof(h) == 64, "IPC handle size");
  memcpy(handle_out, &h, sizeof(h));
  return 0;
}

extern "C" int sdml_symm_import(const unsigned char* handle, void** out_ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  SDML_CUDA_OK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *out_ptr = p;
  return 0;
}

extern "C" int sdml_symm_unimport(void* ptr) {
  if (ptr) SDML_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

// Strided device-to-device copy on the copy engine (`height` rows of `width` bytes, pitches in
// bytes); dst or src may be a peer-mapped pointer.  Used to push column ranges of transposed
// operands into a tensor-parallel peer's buffer without spending SMs on it.
extern "C" int sdml_memcpy2d_async(void* dst, long long dpitch, const void* src, long long spitch,
                                   long long width, long long height, cudaStream_t stream) {
  if (!dst || !src || width <= 0 || height <= 0 || dpitch < width || spitch < width)
    return set_error("sdml_memcpy2d_async: bad arguments");
  SDML_CUDA_OK(cudaMemcpy2DAsync(dst, static_cast<size_t>(dpitch), src, static_cast<size_t>(spitch),
                                 static_cast<size_t>(width), static_cast<size_t>(height),
                                 cudaMemcpyDeviceToDevice, stream));
  return 0;
}
