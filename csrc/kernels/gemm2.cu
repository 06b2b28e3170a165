// b200flow K0 (2-CTA variant) — tcgen05 GEMM with cta_group::2: a CTA pair on one TPC shares a
// 256x256 output tile.  Each CTA stages its own 128 rows of A and one 128-row half of B per
// k-block (32 KB instead of the 48 KB of the 1-CTA 128x256 kernel -> one third less L2->SM
// operand traffic and room for 6 pipeline stages), the leader CTA's single MMA thread issues
// `tcgen05.mma.cta_group::2` for both SMs, and every CTA drains its own 128 accumulator rows
// from its own TMEM.
//
// Barrier protocol (all mbarriers live at identical offsets in both CTAs):
//   full[s]       leader's: 2 arrivals (one producer per CTA) + the bytes of both CTAs' TMA
//                 loads (the loads use the .cta_group::2 form so they may signal the peer)
//   empty[s]      per CTA: 1 arrival from a multicast tcgen05.commit (mask 0b11)
//   tmem_full[a]  per CTA: 1 arrival from a multicast tcgen05.commit
//   tmem_empty[a] leader's: 8 arrivals = 4 epilogue warps x 2 CTAs (remote mbarrier.arrive)
//
// Same operand-majorness options and epilogue as the 1-CTA core (gemm_core.cuh).
#include <stdlib.h>

#include "gemm2_core.cuh"
#include "../common/host.h"
#include "../../include/b200flow.h"

using namespace b2f;
using namespace b2f::gemm;
using namespace b2f::gemm2;

namespace b2f {
namespace gemm {
int make_operand_map(CUtensorMap* map, const void* ptr, int mn, int k, long long ld, int batch,
                     long long batch_stride, int mn_major, int tile_mn);
int check_params(const Params& p);
}  // namespace gemm
}  // namespace b2f

namespace {

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
b2f_gemm2_bf16_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                      const Params p) {
  extern __shared__ uint8_t smem_raw[];
  Plain2 pol;
  gemm2_body<Plain2, S2>(&tma_a, &tma_b, p, pol, blockIdx.x >> 1, gridDim.x >> 1, smem_raw);
}

// bf16 output through the shared-memory staging buffer (full 6-deep ring + 32 KB staging)
constexpr int STAGED_STAGES = S2;
constexpr int STAGED_SMEM = smem2_bytes(STAGED_STAGES, STAGE_BYTES_2);
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
b2f_gemm2_bf16_staged_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                             const __grid_constant__ CUtensorMap tma_c, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  Staged2 pol;
  pol.staging = smem + STAGED_STAGES * ST2_BYTES + 1024;
  pol.map_c = &tma_c;
  gemm2_body<Staged2, STAGED_STAGES>(&tma_a, &tma_b, p, pol, blockIdx.x >> 1, gridDim.x >> 1, smem_raw);
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
b2f_gemm2_f32_staged_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                            const __grid_constant__ CUtensorMap tma_c, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  StagedF32 pol;
  pol.staging = smem + STAGED_STAGES * ST2_BYTES + 1024;
  pol.map_c = &tma_c;
  gemm2_body<StagedF32, STAGED_STAGES>(&tma_a, &tma_b, p, pol, blockIdx.x >> 1, gridDim.x >> 1, smem_raw);
}

}  // namespace

extern "C" int b2f_gemm_bf16_2cta(const void* A, const void* B, void* C, int M, int N, int K,
                                  long long lda, long long ldb, long long ldc, int a_mn_major,
                                  int b_mn_major, int batch, long long strideA, long long strideB,
                                  long long strideC, const void* bias, const void* residual,
                                  long long ldr, long long strideR, void* aux, long long ldaux,
                                  long long strideAux, float alpha, int flags, void* stream) {
  Params p{};
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
  // per-CTA boxes: 128 rows of A, 128 rows of B (K-major) / 64x64 chunks (MN-major)
  if (int rc = make_operand_map(&ma, A, M, K, lda, batch, strideA, a_mn_major, HALF)) return rc;
  if (int rc = make_operand_map(&mb, B, N, K, ldb, batch, strideB, b_mn_major, HALF)) return rc;
  static bool attr_set[64] = {false};
  int dev = 0;
  B2F_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    B2F_CHECK_CUDA(cudaFuncSetAttribute(b2f_gemm2_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        SMEM2_BYTES));
    attr_set[dev & 63] = true;
  }
  const int tiles = ((M + T2_BM - 1) / T2_BM) * ((N + T2_BN - 1) / T2_BN) * batch;
  int pairs = sm_count() / 2;
  if (pairs > tiles) pairs = tiles;
  static const bool direct = [] { const char* e = getenv("B2F_GEMM2_EPI"); return e && e[0] == 'd'; }();
  const bool can_stage = !(flags & EPI_OUT_F32) && (ldc % 8 == 0) && (N % 8 == 0) && (strideC % 8 == 0);
  const bool can_stage_f32 = (flags & EPI_OUT_F32) && (ldc % 4 == 0) && (N % 4 == 0) && (strideC % 4 == 0);
  if (can_stage_f32 && !direct) {
    static bool f32_attr[64] = {false};
    if (!f32_attr[dev & 63]) {
      B2F_CHECK_CUDA(cudaFuncSetAttribute(b2f_gemm2_f32_staged_kernel,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, STAGED_SMEM));
      f32_attr[dev & 63] = true;
    }
    CUtensorMap mc;
    if (int rc = make_tma_3d(&mc, C, TmaDtype::F32, batch, M, N, batch > 1 ? strideC : (long long)M * ldc, ldc,
                             32, 32, 128)) return rc;
    b2f_gemm2_f32_staged_kernel<<<pairs * 2, NUM_THREADS, STAGED_SMEM, static_cast<cudaStream_t>(stream)>>>(ma, mb, mc, p);
  } else if (can_stage && !direct) {
    static bool staged_attr[64] = {false};
    if (!staged_attr[dev & 63]) {
      B2F_CHECK_CUDA(cudaFuncSetAttribute(b2f_gemm2_bf16_staged_kernel,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, STAGED_SMEM));
      staged_attr[dev & 63] = true;
    }
    CUtensorMap mc;
    if (int rc = make_tma_3d(&mc, C, TmaDtype::BF16, batch, M, N, batch > 1 ? strideC : (long long)M * ldc, ldc,
                             32, 64, 128)) return rc;
    b2f_gemm2_bf16_staged_kernel<<<pairs * 2, NUM_THREADS, STAGED_SMEM, static_cast<cudaStream_t>(stream)>>>(ma, mb, mc, p);
  } else {
    b2f_gemm2_bf16_kernel<<<pairs * 2, NUM_THREADS, SMEM2_BYTES, static_cast<cudaStream_t>(stream)>>>(ma, mb, p);
  }
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

#ifdef B2F_GEMM2_DEBUG
extern "C" int b2f_gemm2_debug(long long* out8) {
  B2F_CHECK_CUDA(cudaDeviceSynchronize());
  B2F_CHECK_CUDA(cudaMemcpyFromSymbol(out8, g_gemm2_dbg, sizeof(long long) * 8));
  return 0;
}
#endif
