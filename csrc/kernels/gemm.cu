// b200flow K0 — plain (single-GPU) tcgen05 GEMM entry points. See gemm_core.cuh.
#include <cuda_fp8.h>

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

__global__ void __launch_bounds__(NUM_THREADS, 1)
b2f_gemm_fp8_kernel(const __grid_constant__ CUtensorMap tma_a,
                    const __grid_constant__ CUtensorMap tma_b, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  PlainPolicy pol;
  gemm_body<PlainPolicy, 1>(&tma_a, &tma_b, p, pol, blockIdx.x, gridDim.x, smem_raw);
}

// Row-wise e4m3 quantisation: q[r, :] = x[r, :] / s_r, s_r = amax_r / 448 (one warp per row).
__global__ void __launch_bounds__(256)
quantize_rows_e4m3_kernel(const __nv_bfloat16* __restrict__ x, uint8_t* __restrict__ q,
                          float* __restrict__ scale, long long rows, int K, long long ldx) {
  const int lane = threadIdx.x & 31;
  const int kv = K >> 3;
  for (long long r = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); r < rows; r += (long long)gridDim.x * 8) {
    const uint4* xr = reinterpret_cast<const uint4*>(x + r * ldx);
    float amax = 0.f;
    for (int c = lane; c < kv; c += 32) {
      const uint4 u = ld_nc_v4(xr + c);
      const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), cc = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
      amax = fmaxf(amax, fmaxf(fmaxf(fabsf(a.x), fabsf(a.y)), fmaxf(fabsf(b.x), fabsf(b.y))));
      amax = fmaxf(amax, fmaxf(fmaxf(fabsf(cc.x), fabsf(cc.y)), fmaxf(fabsf(d.x), fabsf(d.y))));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    const float s = amax > 0.f ? amax / 448.f : 1.f;
    const float inv = 1.f / s;
    if (lane == 0) scale[r] = s;
    uint2* qr = reinterpret_cast<uint2*>(q + r * (long long)K);
    for (int c = lane; c < kv; c += 32) {
      const uint4 u = ld_nc_v4(xr + c);
      const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), cc = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
      auto cv = [&](float lo, float hi) -> uint32_t {
        return (uint32_t)__nv_cvt_float2_to_fp8x2(make_float2(lo * inv, hi * inv), __NV_SATFINITE, __NV_E4M3);
      };
      uint2 o;
      o.x = cv(a.x, a.y) | (cv(b.x, b.y) << 16);
      o.y = cv(cc.x, cc.y) | (cv(d.x, d.y) << 16);
      qr[c] = o;
    }
  }
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
  p.row_scale = nullptr; p.col_scale = nullptr;
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

// C = (Aq * Bq^T) * row_scale[m] * col_scale[n] (+bias, +gelu ...): e4m3 operands, both K-major:
// Aq [M, K] u8, Bq [N, K] u8.  K must be a multiple of 16.
extern "C" int b2f_gemm_fp8(const void* Aq, const void* Bq, void* C, int M, int N, int K, long long lda,
                            long long ldb, long long ldc, const float* row_scale, const float* col_scale,
                            const void* bias, const void* residual, long long ldr, void* aux,
                            long long ldaux, int flags, void* stream) {
  Params p{};
  p.M = M; p.N = N; p.K = K; p.batch = 1;
  p.C = C; p.ldc = ldc;
  p.bias = static_cast<const __nv_bfloat16*>(bias);
  p.residual = static_cast<const __nv_bfloat16*>(residual);
  p.ldr = ldr;
  p.aux = static_cast<__nv_bfloat16*>(aux);
  p.ldaux = ldaux;
  p.alpha = 1.f; p.flags = flags;
  p.row_scale = row_scale; p.col_scale = col_scale;
  if (int rc = check_params(p)) return rc;
  B2F_REQUIRE(K % 16 == 0, "fp8 gemm: K must be a multiple of 16");
  CUtensorMap ma, mb;
  if (int rc = make_tma_3d(&ma, Aq, TmaDtype::U8, 1, M, K, (long long)M * lda, lda, BM, 128, 128)) return rc;
  if (int rc = make_tma_3d(&mb, Bq, TmaDtype::U8, 1, N, K, (long long)N * ldb, ldb, BN, 128, 128)) return rc;
  static bool attr_set[64] = {false};
  int dev = 0;
  B2F_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    B2F_CHECK_CUDA(cudaFuncSetAttribute(b2f_gemm_fp8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        SMEM_BYTES));
    attr_set[dev & 63] = true;
  }
  const int tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
  const int grid = tiles < sm_count() ? tiles : sm_count();
  b2f_gemm_fp8_kernel<<<grid, NUM_THREADS, SMEM_BYTES, static_cast<cudaStream_t>(stream)>>>(ma, mb, p);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_quantize_rows_e4m3(const void* x, void* q, float* scale, long long rows, int K,
                                      long long ldx, void* stream) {
  B2F_REQUIRE(K % 8 == 0 && rows > 0 && ldx % 8 == 0, "quantize: K and ld must be multiples of 8");
  long long blocks = (rows + 7) / 8;
  const long long cap = (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  quantize_rows_e4m3_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<uint8_t*>(q), scale, rows, K, ldx);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
