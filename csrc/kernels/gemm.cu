// b200flow K0 — plain (single-GPU) tcgen05 GEMM entry points. See gemm_core.cuh.
#include "gemm_core.cuh"
#include "../common/host.h"
#include "../../include/b200flow.h"

using namespace b2f;
using namespace b2f::gemm;

__global__ void __launch_bounds__(NUM_THREADS, 1)
b2f_gemm_bf16_kernel(const __grid_constant__ CUtensorMap tma_a,
                     const __grid_constant__ CUtensorMap tma_b, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  PlainPolicy pol;
  gemm_body(&tma_a, &tma_b, p, pol, blockIdx.x, gridDim.x, smem_raw);
}

namespace b2f {
namespace gemm {

// Build the operand tensor map for one side of the product.
//   K-major : memory is [rows = M or N, cols = K], box = (tile_rows x 64)
//   MN-major: memory is [rows = K, cols = M or N], box = (64 x 64)
int make_operand_map(CUtensorMap* map, const void* ptr, int mn, int k, long long ld,
                     int batch, long long batch_stride, int mn_major, int tile_mn) {
  if (!mn_major)
    return make_tma_3d(map, ptr, TmaDtype::BF16, batch, mn, k, batch > 1 ? batch_stride : (long long)mn * ld,
                       ld, tile_mn, BK, 128);
  return make_tma_3d(map, ptr, TmaDtype::BF16, batch, k, mn, batch > 1 ? batch_stride : (long long)k * ld,
                     ld, BK, 64, 128);
}

int check_params(const Params& p) {
  B2F_REQUIRE(p.M > 0 && p.N > 0 && p.K > 0 && p.batch > 0, "GEMM dims must be positive");
  B2F_REQUIRE((reinterpret_cast<uintptr_t>(p.C) & 15) == 0, "C must be 16-byte aligned");
  B2F_REQUIRE(!(p.flags & EPI_ACCUM) || (p.flags & EPI_OUT_F32), "accumulate needs fp32 C");
  return 0;
}

}  // namespace gemm
}  // namespace b2f

extern "C" int b2f_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K,
                             long long lda, long long ldb, long long ldc, int a_mn_major,
                             int b_mn_major, int batch, long long strideA, long long strideB,
                             long long strideC, const void* bias, const void* residual,
                             long long ldr, long long strideR, void* aux, long long ldaux,
                             long long strideAux, float alpha, int flags, void* stream) {
  Params p;
  p.M = M; p.N = N; p.K = K; p.batch = batch;
  p.a_mn_major = a_mn_major; p.b_mn_major = b_mn_major;
  p.C = C; p.ldc = ldc; p.strideC = strideC;
  p.bias = static_cast<const __nv_bfloat16*>(bias);
  p.residual = static_cast<const __nv_bfloat16*>(residual);
  p.ldr = ldr; p.strideR = strideR;
  p.aux = static_cast<__nv_bfloat16*>(aux);
  p.ldaux = ldaux; p.strideAux = strideAux;
  p.alpha = alpha; p.flags = flags;
  if (int rc = check_params(p)) return rc;

  CUtensorMap ma, mb;
  if (int rc = make_operand_map(&ma, A, M, K, lda, batch, strideA, a_mn_major, BM)) return rc;
  if (int rc = make_operand_map(&mb, B, N, K, ldb, batch, strideB, b_mn_major, BN)) return rc;

  static bool attr_set[64] = {false};
  int dev = 0;
  B2F_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    B2F_CHECK_CUDA(cudaFuncSetAttribute(b2f_gemm_bf16_kernel,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    attr_set[dev & 63] = true;
  }
  const int num_tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN) * batch;
  int grid = num_tiles < sm_count() ? num_tiles : sm_count();
  b2f_gemm_bf16_kernel<<<grid, NUM_THREADS, SMEM_BYTES, static_cast<cudaStream_t>(stream)>>>(
      ma, mb, p);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
