// Host-side common code for libsdml_b200.so (see sdml_host.h).  Infrastructure only: the reference
// (/root/reference/simple_distributed.py) has no native code, so nothing here mirrors it.
#include "sdml_host.h"

#include <atomic>
#include <mutex>
#include <stdio.h>
#include <string.h>

namespace sdml {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

int set_error(const char* msg) {
  snprintf(g_err, sizeof(g_err), "%s", msg);
  return -1;
}

int set_cuda_error(cudaError_t e, const char* expr, const char* file, int line) {
  snprintf(g_err, sizeof(g_err), "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e),
           file, line, expr);
  return (int)e;
}

int num_sms(int device) {
  static int cache[64] = {};
  if (device < 0 || device >= 64) device = 0;
  if (cache[device] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0)
      n = 148;
    cache[device] = n;
  }
  return cache[device];
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

static std::atomic<unsigned long long> g_wait_ns{0};
static std::atomic<uint32_t*> g_wait_err{nullptr};
HostWaitGuard wait_guard() { return HostWaitGuard{g_wait_ns.load(), g_wait_err.load()}; }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int make_tmap_2d(CUtensorMap* map, const void* ptr, TmapDtype dt, uint64_t inner, uint64_t outer,
                 uint64_t ld_elems, uint32_t box_inner, uint32_t box_outer, TmapSwizzle swz) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return set_error("cuTensorMapEncodeTiled driver entry point not available");
  CUtensorMapDataType cdt;
  uint64_t esz;
  switch (dt) {
    case TMAP_BF16: cdt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16; esz = 2; break;
    case TMAP_U8:   cdt = CU_TENSOR_MAP_DATA_TYPE_UINT8;    esz = 1; break;
    default:        cdt = CU_TENSOR_MAP_DATA_TYPE_FLOAT32;  esz = 4; break;
  }
  cuuint64_t gdim[2] = {inner, outer};
  cuuint64_t gstride[1] = {ld_elems * esz};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, cdt, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swz == TMAP_SWIZZLE_128B ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[256];
    snprintf(buf, sizeof(buf),
             "cuTensorMapEncodeTiled failed (%d): inner=%llu outer=%llu ld=%llu box={%u,%u} ptr=%p",
             (int)r, (unsigned long long)inner, (unsigned long long)outer,
             (unsigned long long)ld_elems, box_inner, box_outer, ptr);
    return set_error(buf);
  }
  return 0;
}

}  // namespace sdml

extern "C" const char* sdml_last_error() { return sdml::g_err; }
extern "C" long long sdml_launch_count() { return sdml::g_launches.load(); }
extern "C" int sdml_abi_version() { return 1; }

// Bound every in-kernel wait for another GPU launched by this process from now on: after
// timeout_ns (0 = unbounded) a waiting kernel writes {code, seen, wanted, 0} into err_host
// (4 x uint32 of pinned host memory that outlives the kernels; may be NULL) and carries on.
// Codes: 11 all-gather flag, 12 reduce-scatter landing slots, 13 weight-gradient landing slots,
// 14 refreshed-weights counter, 15 staged DP all-reduce flag.
extern "C" int sdml_set_wait_timeout(unsigned long long timeout_ns, void* err_host) {
  sdml::g_wait_ns.store(timeout_ns);
  sdml::g_wait_err.store(static_cast<uint32_t*>(err_host));
  return 0;
}
