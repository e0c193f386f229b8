// Library-wide host plumbing: error strings, device queries, tensor-map encoding.
#include <stdarg.h>

#include <atomic>

#include "common.cuh"


static thread_local char tl_error[512] = "";
std::atomic<int64_t> g_sdb_launches{0};

void sdb_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(tl_error, sizeof(tl_error), fmt, ap);
  va_end(ap);
}

void sdb_count_launch() { g_sdb_launches.fetch_add(1, std::memory_order_relaxed); }

extern "C" int sdb_version(void) { return 100; }
extern "C" const char* sdb_last_error(void) { return tl_error; }
extern "C" int64_t sdb_launch_count(void) { return g_sdb_launches.load(); }

// one reader per translation unit (each holds its own copy of the device flag, see common.cuh); filled by static
// initialisers while the library loads, so no locking is needed afterwards
static SdbFlagReader g_flag_readers[32];
static int g_n_flag_readers = 0;
void sdb_register_timeout_flag(SdbFlagReader fn) {
  if (g_n_flag_readers < 32) g_flag_readers[g_n_flag_readers++] = fn;
}

extern "C" int sdb_debug_timeout_flag(void) {
  if (cudaDeviceSynchronize() != cudaSuccess) return -1;
  int any = 0;
  for (int i = 0; i < g_n_flag_readers; ++i) {
    const int v = g_flag_readers[i]();
    if (v < 0) return -1;
    any |= v;
  }
  return any;
}

int sdb_num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess || !p)
    return nullptr;
  fn = reinterpret_cast<PFN_encodeTiled>(p);
  return fn;
}

int sdb_make_tensor_map(CUtensorMap* out, const void* base, int elem_bytes, int rank,
                        const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box,
                        int swizzle_bytes) {
  PFN_encodeTiled fn = get_encode_fn();
  if (!fn) {
    sdb_set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver / device?)");
    return SDB_ECUDA;
  }
  cuuint64_t gdims[5], gstrides[4];
  cuuint32_t gbox[5], estr[5];
  for (int i = 0; i < rank; ++i) {
    gdims[i] = dims[i];
    gbox[i] = box[i];
    estr[i] = 1;
    if (i > 0) gstrides[i - 1] = strides_bytes[i];
  }
  CUtensorMapSwizzle sw = CU_TENSOR_MAP_SWIZZLE_NONE;
  if (swizzle_bytes == 32) sw = CU_TENSOR_MAP_SWIZZLE_32B;
  else if (swizzle_bytes == 64) sw = CU_TENSOR_MAP_SWIZZLE_64B;
  else if (swizzle_bytes == 128) sw = CU_TENSOR_MAP_SWIZZLE_128B;
  CUtensorMapDataType dt = elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                           : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  CUresult r = fn(out, dt, (cuuint32_t)rank, const_cast<void*>(base), gdims, gstrides, gbox, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    sdb_set_error("cuTensorMapEncodeTiled failed: CUresult %d (rank %d, dims %llu %llu %llu %llu, "
                  "box %u %u %u %u, base %p)",
                  (int)r, rank, (unsigned long long)dims[0],
                  (unsigned long long)(rank > 1 ? dims[1] : 0),
                  (unsigned long long)(rank > 2 ? dims[2] : 0),
                  (unsigned long long)(rank > 3 ? dims[3] : 0), box[0], rank > 1 ? box[1] : 0,
                  rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0, base);
    return SDB_ECUDA;
  }
  return SDB_OK;
}
