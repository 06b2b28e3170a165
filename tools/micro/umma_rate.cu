// Micro-benchmark: issue rate of tcgen05.mma (M=128, K=16, bf16) for SS / TS operand forms and
// several N, all SMs busy.  Operands are whatever is in shared memory (values do not matter).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -o umma_rate umma_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../csrc/common/ptx.cuh"
using namespace b2f;
constexpr uint32_t DESC_HI = (1024u >> 4) | (1u << 14) | (2u << 29);
__device__ __forceinline__ uint64_t mk_desc(uint32_t lo) {
  uint64_t d; asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(DESC_HI)); return d;
}
extern __shared__ uint8_t sm_raw[];
// MODE 0: SS K-major A,B.  MODE 1: TS (A in TMEM), B MN-major.  MODE 2: SS with A K-major, B MN-major
// MODE 3: SS with A and B MN-major (the wgrad form)
template <int MODE, int N>
__global__ void __launch_bounds__(128, 1) rate_kernel(int iters, long long* out) {
  __shared__ uint64_t bar;
  __shared__ uint32_t tbase;
  uint8_t* sm = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(sm_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (warp == 0) { tmem_alloc(&tbase, 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  if (warp == 0) {
    const uint32_t a = ((smem_u32(sm) & 0x3FFFF) >> 4) | (1u << 16);
    const uint32_t b = ((smem_u32(sm + 65536) & 0x3FFFF) >> 4) | (1u << 16);
    const uint32_t bmn = ((smem_u32(sm + 65536) & 0x3FFFF) >> 4) | (1024u << 16);
    const uint32_t idesc_kk = make_idesc(1, 1, 128, N, 0, 0), idesc_kmn = make_idesc(1, 1, 128, N, 0, 1);
    const uint32_t idesc_mnmn = make_idesc(1, 1, 128, N, 1, 1);
    const uint32_t amn = ((smem_u32(sm) & 0x3FFFF) >> 4) | (512u << 16);   // LBO = 8 KB chunk stride
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          if (MODE == 0) umma_f16_ss(256 * (it & 1), mk_desc(a + (k >> 2) * 1024 + (k & 3) * 2), mk_desc(b + (k >> 2) * 1024 * (N / 128 > 0 ? N / 128 : 1) + (k & 3) * 2), idesc_kk, k > 0);
          if (MODE == 1) umma_f16_ts(256, 0 + k * 8, mk_desc(bmn + k * 128), idesc_kmn, k > 0);
          if (MODE == 2) umma_f16_ss(256 * (it & 1), mk_desc(a + (k >> 2) * 1024 + (k & 3) * 2), mk_desc(bmn + k * 128), idesc_kmn, k > 0);
          if (MODE == 3) umma_f16_ss(256 * (it & 1), mk_desc(amn + (k & 3) * 128), mk_desc(bmn + (k & 3) * 128), idesc_mnmn, k > 0);
        }
      }
      __syncwarp();
    }
    if (elect_one()) umma_commit(&bar);
    __syncwarp();
    mbar_wait(&bar, 0);
    long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(0, 512); }
}
template <int MODE, int N>
void run(const char* name, long long* d_out) {
  const int smem = 200 * 1024, iters = 2000;
  cudaFuncSetAttribute(rate_kernel<MODE, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  rate_kernel<MODE, N><<<148, 128, smem>>>(100, d_out);
  cudaDeviceSynchronize();
  rate_kernel<MODE, N><<<148, 128, smem>>>(iters, d_out);
  cudaError_t e = cudaDeviceSynchronize();
  long long cyc = 0;
  cudaMemcpy(&cyc, d_out, 8, cudaMemcpyDeviceToHost);
  printf("%-28s N=%3d  %7.1f cycles / MMA (ideal %d)  err=%s\n", name, N, (double)cyc / (iters * 8.0), N / 2, cudaGetErrorString(e));
}
int main() {
  long long* d_out; cudaMalloc(&d_out, 8);
  run<0, 64>("SS  A,B K-major", d_out);
  run<0, 128>("SS  A,B K-major", d_out);
  run<0, 256>("SS  A,B K-major", d_out);
  run<2, 128>("SS  A K-major, B MN-major", d_out);
  run<2, 256>("SS  A K-major, B MN-major", d_out);
  run<3, 128>("SS  A,B MN-major", d_out);
  run<3, 256>("SS  A,B MN-major", d_out);
  run<1, 64>("TS  A in TMEM, B MN-major", d_out);
  run<1, 128>("TS  A in TMEM, B MN-major", d_out);
  run<1, 256>("TS  A in TMEM, B MN-major", d_out);
  return 0;
}
