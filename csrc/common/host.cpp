// b200flow — host utilities: last-error slot for the C ABI, TMA descriptor
// encoding via cudaGetDriverEntryPoint, per-device SM count cache.
#include "host.h"

#include <mutex>

namespace b2f {

static thread_local std::string g_last_error;

void set_error(const std::string& msg) { g_last_error = msg; }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled needs a current driver context; threads that have only
// ever been handed work by another thread (e.g. autograd workers) may not have
// one yet, so bind the primary context once per thread.
static void ensure_context() {
  static thread_local bool ready = false;
  if (!ready) {
    cudaFree(nullptr);
    ready = true;
  }
}

static EncodeTiledFn get_encode() {
  ensure_context();
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

static int dtype_info(TmaDtype dt, CUtensorMapDataType* cu, uint32_t* esize) {
  switch (dt) {
    case TmaDtype::BF16: *cu = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16; *esize = 2; return 0;
    case TmaDtype::F16:  *cu = CU_TENSOR_MAP_DATA_TYPE_FLOAT16;  *esize = 2; return 0;
    case TmaDtype::U8:   *cu = CU_TENSOR_MAP_DATA_TYPE_UINT8;    *esize = 1; return 0;
    case TmaDtype::F32:  *cu = CU_TENSOR_MAP_DATA_TYPE_FLOAT32;  *esize = 4; return 0;
  }
  return -1;
}

int make_tma_3d(CUtensorMap* out, const void* base, TmaDtype dt, uint64_t batch, uint64_t rows,
                uint64_t cols, uint64_t batch_stride, uint64_t ld, uint32_t box_rows,
                uint32_t box_cols, int swizzle_bytes) {
  EncodeTiledFn enc = get_encode();
  if (!enc) {
    set_error("b200flow: cuTensorMapEncodeTiled driver entry point unavailable (no CUDA driver?)");
    return -1;
  }
  CUtensorMapDataType cu;
  uint32_t es;
  if (dtype_info(dt, &cu, &es)) return -1;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || ((ld * es) & 15) || ((batch_stride * es) & 15)) {
    set_error("b200flow: TMA operand must be 16-byte aligned with 16-byte-multiple strides");
    return -2;
  }
  if (swizzle_bytes == 128 && box_cols * es != 128) {
    set_error("b200flow: 128B-swizzled TMA box must be exactly 128 bytes wide");
    return -2;
  }
  cuuint64_t dims[3] = {cols, rows, batch};
  cuuint64_t strides[2] = {ld * es, batch_stride * es};
  cuuint32_t box[3] = {box_cols, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUtensorMapSwizzle sw = swizzle_bytes == 128  ? CU_TENSOR_MAP_SWIZZLE_128B
                          : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                          : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                                : CU_TENSOR_MAP_SWIZZLE_NONE;
  uint32_t rank = batch > 1 || batch_stride ? 3 : 2;
  CUresult r = enc(out, cu, rank, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("b200flow: cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r) +
              " (rows=" + std::to_string(rows) + " cols=" + std::to_string(cols) +
              " ld=" + std::to_string(ld) + ")");
    return -1;
  }
  return 0;
}

int make_tma_2d(CUtensorMap* out, const void* base, TmaDtype dt, uint64_t rows, uint64_t cols,
                uint64_t ld, uint32_t box_rows, uint32_t box_cols, int swizzle_bytes) {
  return make_tma_3d(out, base, dt, 1, rows, cols, 0, ld, box_rows, box_cols, swizzle_bytes);
}

int make_tma_4d(CUtensorMap* out, const void* base, TmaDtype dt, uint64_t d3, uint64_t d2,
                uint64_t rows, uint64_t cols, uint64_t s3, uint64_t s2, uint64_t s1,
                uint32_t box_rows, uint32_t box_cols, int swizzle_bytes) {
  EncodeTiledFn enc = get_encode();
  if (!enc) {
    set_error("b200flow: cuTensorMapEncodeTiled driver entry point unavailable (no CUDA driver?)");
    return -1;
  }
  CUtensorMapDataType cu;
  uint32_t es;
  if (dtype_info(dt, &cu, &es)) return -1;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || ((s1 * es) & 15) || ((s2 * es) & 15) ||
      ((s3 * es) & 15)) {
    set_error("b200flow: TMA operand must be 16-byte aligned with 16-byte-multiple strides");
    return -2;
  }
  // TMA wants dims ordered by increasing stride after the contiguous one; the
  // attention layouts ([b, s, h, d] with rows = s, d2 = h) have s2 < s1, so the
  // caller-visible coordinate order is fixed here as (col, d2, row, d3).
  cuuint64_t dims[4] = {cols, d2, rows, d3};
  cuuint64_t strides[3] = {s2 * es, s1 * es, s3 * es};
  cuuint32_t box[4] = {box_cols, 1, box_rows, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                               : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = enc(out, cu, 4, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("b200flow: cuTensorMapEncodeTiled(4d) failed with CUresult " + std::to_string((int)r));
    return -1;
  }
  return 0;
}

int sm_count() {
  static int cache[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (!cache[dev]) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cache[dev] = n;
  }
  return cache[dev];
}

}  // namespace b2f

extern "C" const char* b2f_last_error() { return b2f::g_last_error.c_str(); }
