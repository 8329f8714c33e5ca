// tensormap.cu — host-side TMA descriptor encoding through the driver entry point
// (resolved at run time with cudaGetDriverEntryPoint, so the library has no link-time
// dependency on libcuda and can be built on a machine without a GPU driver).
#include "sxb_common.cuh"

namespace sxb {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn resolve_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

int encode_tensor_map(CUtensorMap* m, int dtype_code, int rank, const void* ptr,
                      const unsigned long long* dims, const unsigned long long* strides_bytes,
                      const unsigned int* box, int swizzle_bytes) {
  EncodeTiledFn enc = resolve_encode();
  if (!enc) return SXB_ERR_DRIVER;
  CUtensorMapDataType dt;
  switch (dtype_code) {
    case kBF16: dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16; break;
    case kF16: dt = CU_TENSOR_MAP_DATA_TYPE_FLOAT16; break;
    case kF32: dt = CU_TENSOR_MAP_DATA_TYPE_FLOAT32; break;
    case kFP8E4M3: case kU8: case kI8: dt = CU_TENSOR_MAP_DATA_TYPE_UINT8; break;
    default: return SXB_ERR_BAD_ARG;
  }
  CUtensorMapSwizzle sw = CU_TENSOR_MAP_SWIZZLE_NONE;
  if (swizzle_bytes == 128) sw = CU_TENSOR_MAP_SWIZZLE_128B;
  else if (swizzle_bytes == 64) sw = CU_TENSOR_MAP_SWIZZLE_64B;
  else if (swizzle_bytes == 32) sw = CU_TENSOR_MAP_SWIZZLE_32B;
  cuuint64_t d[5];
  cuuint64_t st[4];
  cuuint32_t b[5], es[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) st[i] = strides_bytes[i];
  CUresult r = enc(m, dt, (cuuint32_t)rank, const_cast<void*>(ptr), d, st, b, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? SXB_OK : SXB_ERR_DRIVER;
}

// ---- local-ownership test for anything a TMA multicast tensor map will cover ---------------------
// TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another
// GPU's memory hung the GPU.  A range counts as local only if (1) the runtime reports an ordinary device
// allocation of the current device, (2) the driver's allocation range contains the WHOLE extent and (3) the
// extent touches none of the ranges registered as foreign (peer / multicast mappings of symmetric memory,
// registered by comm/symm.py when a buffer is rendezvoused).  Anything else selects the non-multicast kernels.
namespace {
struct ForeignRange { uintptr_t lo, hi; };
ForeignRange g_foreign[4096];
int g_n_foreign = 0;
typedef CUresult (*AddrRangeFn)(CUdeviceptr*, size_t*, CUdeviceptr);
AddrRangeFn resolve_addr_range() {
  static AddrRangeFn fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<AddrRangeFn>(ptr);
  }
  return fn;
}
}  // namespace

bool range_is_local(const void* p, long long bytes) {
  if (p == nullptr || bytes <= 0) return false;
  const uintptr_t lo = reinterpret_cast<uintptr_t>(p), hi = lo + (uintptr_t)bytes;
  for (int i = 0; i < g_n_foreign; ++i)
    if (lo < g_foreign[i].hi && g_foreign[i].lo < hi) return false;
  cudaPointerAttributes a{};
  int dev = -1;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  if (cudaGetDevice(&dev) != cudaSuccess) return false;
  if (a.type != cudaMemoryTypeDevice || a.device != dev) return false;
  AddrRangeFn ar = resolve_addr_range();
  if (!ar) return false;
  CUdeviceptr base = 0;
  size_t size = 0;
  if (ar(&base, &size, (CUdeviceptr)lo) != CUDA_SUCCESS) return false;
  return (uintptr_t)base <= lo && hi <= (uintptr_t)base + size;
}

}  // namespace sxb

// Declare [p, p+bytes) as NOT this GPU's memory (a peer's buffer or a multicast window mapped into this
// process).  Append-only; returns SXB_ERR_TOO_LARGE when the table is full (callers then must not rely on it).
SXB_API int sxb_register_foreign_range(const void* p, long long bytes) {
  if (p == nullptr || bytes <= 0) return SXB_ERR_BAD_ARG;
  const uintptr_t lo = reinterpret_cast<uintptr_t>(p), hi = lo + (uintptr_t)bytes;
  for (int i = 0; i < sxb::g_n_foreign; ++i)
    if (sxb::g_foreign[i].lo == lo && sxb::g_foreign[i].hi >= hi) return SXB_OK;
  if (sxb::g_n_foreign >= 4096) return SXB_ERR_TOO_LARGE;
  sxb::g_foreign[sxb::g_n_foreign].lo = lo;
  sxb::g_foreign[sxb::g_n_foreign].hi = hi;
  ++sxb::g_n_foreign;
  return SXB_OK;
}
// 1 when [p, p+bytes) passes the local-ownership test above, else 0.
SXB_API int sxb_range_is_local(const void* p, long long bytes) { return sxb::range_is_local(p, bytes) ? 1 : 0; }

// Library/version probe used by Python to verify that the loaded .so matches the sources.
SXB_API int sxb_abi_version() { return 3; }

SXB_API int sxb_device_info(int* sm_count, int* cc_major, int* cc_minor, long long* total_mem) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  cudaDeviceProp p;
  e = cudaGetDeviceProperties(&p, dev);
  if (e != cudaSuccess) return (int)e;
  if (sm_count) *sm_count = p.multiProcessorCount;
  if (cc_major) *cc_major = p.major;
  if (cc_minor) *cc_minor = p.minor;
  if (total_mem) *total_mem = (long long)p.totalGlobalMem;
  return SXB_OK;
}
