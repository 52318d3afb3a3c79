// (No counterpart in the reference, which has no native code: /root/reference has two files,
// simple_distributed.py and readme.md.)
// Host-side helpers shared by every translation unit of libsdml_b200.so:
// error reporting through the C ABI, cached device attributes, TMA tensor-map
// encoding (driver entry point fetched at run time, no link against libcuda),
// and the launch counter the benchmark harness reads.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace sdml {

// Record a message retrievable through sdml_last_error(); returns -1.
int set_error(const char* msg);
int set_cuda_error(cudaError_t e, const char* expr, const char* file, int line);

#define SDML_CUDA_OK(expr)                                                   \
  do {                                                                       \
    cudaError_t _e = (expr);                                                 \
    if (_e != cudaSuccess) return ::sdml::set_cuda_error(_e, #expr, __FILE__, __LINE__); \
  } while (0)

int num_sms(int device);

// Deadline + error record handed to every kernel that waits for another GPU
// (set once per process with sdml_set_wait_timeout; default: unbounded, no record).
struct HostWaitGuard {
  unsigned long long timeout_ns;
  uint32_t* err;
};
HostWaitGuard wait_guard();
void count_launch(int n = 1);

enum TmapDtype { TMAP_BF16 = 0, TMAP_U8 = 1, TMAP_F32 = 2 };
enum TmapSwizzle { TMAP_SWIZZLE_NONE = 0, TMAP_SWIZZLE_128B = 3 };

// 2-D row-major tensor [outer, inner] with leading dimension `ld_elems`;
// box = {box_inner, box_outer} elements.  Out-of-bounds reads are zero-filled.
int make_tmap_2d(CUtensorMap* map, const void* ptr, TmapDtype dt, uint64_t inner, uint64_t outer,
                 uint64_t ld_elems, uint32_t box_inner, uint32_t box_outer, TmapSwizzle swz);

}  // namespace sdml
