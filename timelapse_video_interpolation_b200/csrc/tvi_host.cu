// libtvi host plumbing: thread-local error string, version, device query and the
// cuTensorMapEncodeTiled wrapper (resolved at run time through the runtime's driver entry
// point lookup, so the library has no link-time dependency on libcuda and can be dlopen'ed on
// a machine without a GPU — the "does it load / export every symbol" CPU test relies on that).
#include <stdarg.h>
#include <string.h>

#include <mutex>

#include "tvi_common.cuh"

namespace tvi {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int make_tensor_map(CUtensorMap* map, const void* base, int elem_bytes, int rank, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box, Swizzle swz) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return fail(TVI_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
  // cuTensorMapEncodeTiled is a DRIVER call: it needs a context current on the calling thread.  A thread that
  // has only been handed a device by cudaSetDevice-less frameworks (the autograd engine's worker thread calls
  // into libtvi with the backward GEMM as its very first CUDA work) may not have one bound yet
  // (CUDA_ERROR_INVALID_CONTEXT); cudaFree(0) binds the runtime's primary context.  Once per thread.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    TVI_CUDA_OK(cudaFree(nullptr));
    ctx_bound = true;
  }
  CUtensorMapDataType dt;
  switch (elem_bytes) {
    case 2: dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16; break;
    case 4: dt = CU_TENSOR_MAP_DATA_TYPE_FLOAT32; break;
    default: return fail(TVI_ERR_INVALID, "make_tensor_map: unsupported element size %d", elem_bytes);
  }
  cuuint64_t gdims[5];
  cuuint64_t gstrides[4];
  cuuint32_t gbox[5];
  cuuint32_t estr[5];
  for (int i = 0; i < rank; ++i) {
    gdims[i] = dims[i];
    gbox[i] = box[i];
    estr[i] = 1;
    if (i + 1 < rank) gstrides[i] = strides_bytes[i];
  }
  CUtensorMapSwizzle s = CU_TENSOR_MAP_SWIZZLE_NONE;
  switch (swz) {
    case Swizzle::None: s = CU_TENSOR_MAP_SWIZZLE_NONE; break;
    case Swizzle::B32: s = CU_TENSOR_MAP_SWIZZLE_32B; break;
    case Swizzle::B64: s = CU_TENSOR_MAP_SWIZZLE_64B; break;
    case Swizzle::B128: s = CU_TENSOR_MAP_SWIZZLE_128B; break;
  }
  CUresult r = fn(map, dt, (cuuint32_t)rank, const_cast<void*>(base), gdims, gstrides, gbox, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, s, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(TVI_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (rank %d, dims %llu x %llu, box %u x %u)",
                (int)r, rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 1), box[0],
                rank > 1 ? box[1] : 1);
  return TVI_OK;
}

}  // namespace tvi

extern "C" {

const char* tvi_version(void) { return "libtvi 0.1 (sm_100a)"; }

const char* tvi_last_error(void) { return tvi::g_err; }

int tvi_device_info(int* sm_count, int* cc_major, int* cc_minor) {
  int dev = 0;
  TVI_CUDA_OK(cudaGetDevice(&dev));
  int n = 0, ma = 0, mi = 0;
  TVI_CUDA_OK(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
  TVI_CUDA_OK(cudaDeviceGetAttribute(&ma, cudaDevAttrComputeCapabilityMajor, dev));
  TVI_CUDA_OK(cudaDeviceGetAttribute(&mi, cudaDevAttrComputeCapabilityMinor, dev));
  if (sm_count) *sm_count = n;
  if (cc_major) *cc_major = ma;
  if (cc_minor) *cc_minor = mi;
  return TVI_OK;
}

}  // extern "C"
