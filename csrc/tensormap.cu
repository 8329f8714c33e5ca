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

}  // namespace sxb

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
