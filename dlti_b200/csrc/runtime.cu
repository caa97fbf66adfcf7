// Host runtime shared by all kernels: launch accounting, error strings, device
// properties and TMA tensor-map encoding (driver entry point fetched at run time so
// the library links without libcuda on the build box).
// No reference counterpart (the reference has no native code).
#include "common.cuh"

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>

#include <cudaTypedefs.h>

static std::atomic<long long> g_launches{0};
static char g_error[512] = {0};

void dlti_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
void dlti_set_error(const char* msg) {
  std::strncpy(g_error, msg, sizeof(g_error) - 1);
  g_error[sizeof(g_error) - 1] = 0;
}

DLTI_API long long dlti_launch_count() { return g_launches.load(std::memory_order_relaxed); }
DLTI_API void dlti_reset_launch_count() { g_launches.store(0); }
DLTI_API const char* dlti_last_error() { return g_error; }
DLTI_API const char* dlti_error_string(int code) {
  switch (code) {
    case 0: return "success";
    case DLTI_ERR_BAD_ARG: return "dlti: bad argument";
    case DLTI_ERR_UNSUPPORTED: return "dlti: unsupported configuration";
    case DLTI_ERR_DRIVER: return "dlti: CUDA driver entry point unavailable";
    case DLTI_ERR_TOO_LARGE: return "dlti: size exceeds the compiled-in maximum";
    default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "dlti: unknown error";
  }
}
DLTI_API int dlti_version() { return 100; }

int dlti_num_sms() {
  static int sms[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (sms[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0)
      v = 148;
    sms[dev] = v;
  }
  return sms[dev];
}
DLTI_API int dlti_sm_count() { return dlti_num_sms(); }

static std::atomic<int> g_pdl{-1};
int dlti_pdl_enabled() {
  int v = g_pdl.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = std::getenv("DLTI_PDL");
    v = (e && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 1;
    g_pdl.store(v, std::memory_order_relaxed);
  }
  return v;
}
DLTI_API void dlti_set_pdl(int level) { g_pdl.store(level < 0 ? 0 : (level > 2 ? 2 : level), std::memory_order_relaxed); }
DLTI_API int dlti_get_pdl() { return dlti_pdl_enabled(); }

static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;
static std::once_flag g_encode_once;

static void load_encode() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess &&
      q == cudaDriverEntryPointSuccess)
    g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
}

int dlti_make_tmap_2d(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer,
                      uint64_t ld_elems, uint32_t box_inner, uint32_t box_outer, int elem_bytes,
                      int swizzle_bytes) {
  std::call_once(g_encode_once, load_encode);
  if (!g_encode) {
    dlti_set_error("cuTensorMapEncodeTiled not available");
    return DLTI_ERR_DRIVER;
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15) || ((ld_elems * elem_bytes) & 15)) {
    dlti_set_error("TMA operand must be 16-byte aligned with a 16-byte-multiple row pitch");
    return DLTI_ERR_BAD_ARG;
  }
  CUtensorMapDataType dt = elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                           : elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                             : CU_TENSOR_MAP_DATA_TYPE_UINT8;
  cuuint64_t gdim[2] = {inner, outer};
  cuuint64_t gstride[1] = {ld_elems * (uint64_t)elem_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMapSwizzle sw = swizzle_bytes == 128  ? CU_TENSOR_MAP_SWIZZLE_128B
                          : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                          : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                                : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = g_encode(out, dt, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[256];
    std::snprintf(buf, sizeof(buf),
                  "cuTensorMapEncodeTiled failed (%d): base=%p inner=%llu outer=%llu ld=%llu box=%ux%u",
                  (int)r, base, (unsigned long long)inner, (unsigned long long)outer,
                  (unsigned long long)ld_elems, box_inner, box_outer);
    dlti_set_error(buf);
    return DLTI_ERR_DRIVER;
  }
  return 0;
}

// Cached 16-bit / 128B-swizzle tensor maps: the training step encodes the same (pointer, shape, box) a few
// thousand times per step (weights, recycled activation buffers); a map is a pure function of its key, so a
// hit is always valid. cuTensorMapEncodeTiled costs ~1-2 us of host time per call.
namespace {
struct TmapKey {
  const void* base;
  uint64_t inner, outer, ld;
  uint32_t bi, bo;
  bool operator==(const TmapKey& o) const {
    return base == o.base && inner == o.inner && outer == o.outer && ld == o.ld && bi == o.bi && bo == o.bo;
  }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    size_t h = reinterpret_cast<size_t>(k.base);
    auto mix = [&](uint64_t v) { h ^= v + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2); };
    mix(k.inner); mix(k.outer); mix(k.ld); mix(((uint64_t)k.bi << 32) | k.bo);
    return h;
  }
};
std::mutex g_tm_mu;
std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> g_tm_cache;
}  // namespace

int dlti_get_tmap(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t bi,
                  uint32_t bo) {
  TmapKey key{base, inner, outer, ld, bi, bo};
  {
    std::lock_guard<std::mutex> g(g_tm_mu);
    auto it = g_tm_cache.find(key);
    if (it != g_tm_cache.end()) {
      *out = it->second;
      return 0;
    }
  }
  int rc = dlti_make_tmap_2d(out, base, inner, outer, ld, bi, bo, 2, 128);
  if (rc) return rc;
  std::lock_guard<std::mutex> g(g_tm_mu);
  if (g_tm_cache.size() > 65536) g_tm_cache.clear();
  g_tm_cache.emplace(key, *out);
  return 0;
}

// MN-major operand [K, N] (N contiguous) as ONE box per stage: a 3D view {64 n-elements, K rows, N/64 chunks}
// whose box {64, box_k, chunks} lands in shared memory chunk after chunk — the layout the MN-major UMMA
// descriptor expects (successive 64-element MN chunks box_k * 128 bytes apart) — so the producer issues one TMA
// instead of one per chunk. Requires N % 64 == 0. Cached like dlti_get_tmap.
int dlti_get_tmap_mn3d(CUtensorMap* out, const void* base, uint64_t N, uint64_t K, uint64_t ld, uint32_t box_k,
                       uint32_t chunks) {
  if (N % 64) return DLTI_ERR_UNSUPPORTED;
  TmapKey key{base, N, K, ld, 0x80000000u | box_k, chunks};
  {
    std::lock_guard<std::mutex> g(g_tm_mu);
    auto it = g_tm_cache.find(key);
    if (it != g_tm_cache.end()) {
      *out = it->second;
      return 0;
    }
  }
  std::call_once(g_encode_once, load_encode);
  if (!g_encode) return DLTI_ERR_DRIVER;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || ((ld * 2) & 15)) return DLTI_ERR_BAD_ARG;
  cuuint64_t gdim[3] = {64, K, N / 64};
  cuuint64_t gstride[2] = {ld * 2, 128};
  cuuint32_t box[3] = {64, box_k, chunks};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), gdim, gstride, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    dlti_set_error("cuTensorMapEncodeTiled (3D MN-major view) failed");
    return DLTI_ERR_DRIVER;
  }
  std::lock_guard<std::mutex> g(g_tm_mu);
  g_tm_cache.emplace(key, *out);
  return 0;
}

// ---- CUDA IPC helpers for the symmetric arena (fallback when torch symmetric memory is
// unavailable): open a peer's cudaIpcMemHandle_t (64 bytes) in this process. P2P access over
// NVLink is enabled lazily by the driver. -------------------------------------------------------
DLTI_API int dlti_ipc_get_handle(const void* dev_ptr, void* handle_out_64) {
  cudaIpcMemHandle_t h;
  DLTI_CUDA_CHECK(cudaIpcGetMemHandle(&h, const_cast<void*>(dev_ptr)));
  std::memcpy(handle_out_64, &h, sizeof(h));
  return 0;
}
DLTI_API int dlti_ipc_open(const void* handle_64, void** out_ptr) {
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle_64, sizeof(h));
  DLTI_CUDA_CHECK(cudaIpcOpenMemHandle(out_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
DLTI_API int dlti_ipc_close(void* ptr) {
  DLTI_CUDA_CHECK(cudaIpcCloseMemHandle(ptr));
  return 0;
}
DLTI_API int dlti_mem_base(const void* dev_ptr, void** base_out, unsigned long long* size_out) {
  CUdeviceptr base = 0;
  size_t size = 0;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn)
    return DLTI_ERR_DRIVER;
  typedef CUresult (*fn_t)(CUdeviceptr*, size_t*, CUdeviceptr);
  if (reinterpret_cast<fn_t>(fn)(&base, &size, (CUdeviceptr)dev_ptr) != CUDA_SUCCESS) return DLTI_ERR_DRIVER;
  *base_out = (void*)base;
  *size_out = size;
  return 0;
}
