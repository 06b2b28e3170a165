// b200flow K0 (block-scaled fp8) — MXFP8 quantiser / dequantiser and the plain block-scaled GEMM entry point.
// See gemm_mx_core.cuh for the kernel body and the scale-factor layout.  Capability parity: the expert GEMMs of the
// DeepSpeed-MoE / Megatron-DeepSpeed stack the reference pins (containers/megatron-deepspeed/Dockerfile:12-13), which
// BASELINE.json config 5 runs in block-scaled fp8.
#include "gemm_mx_core.cuh"
#include "../common/host.h"
#include "../../include/b200flow.h"

using namespace b2f;
using namespace b2f::gemm_mx;

__global__ void __launch_bounds__(NUM_THREADS, 1)
b2f_gemm_mx_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                   const gemm::Params p, const MxScales sc) {
  extern __shared__ uint8_t smem_raw[];
  PlainMxPolicy pol;
  gemm_mx_body(&tma_a, &tma_b, p, sc, pol, blockIdx.x, gridDim.x, smem_raw);
}

// bf16 [batch][R, K] -> e4m3 bytes [batch][R, K] + UE8M0 scales in the atom layout.  One lane per 32-element
// group: 64 B in, 32 B + 1 scale byte out; a warp covers 1024 consecutive elements of a row.
__global__ void __launch_bounds__(256)
quantize_mx_kernel(const __nv_bfloat16* __restrict__ x, uint8_t* __restrict__ q, uint8_t* __restrict__ sf,
                   int R, int K, long long ldx, int atoms, int batch, long long x_batch, long long q_batch,
                   long long sf_batch) {
  const int groups_per_row = K >> 5;
  const long long total = (long long)batch * R * groups_per_row;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % groups_per_row);
    const long long rb = i / groups_per_row;
    const int r = (int)(rb % R);
    const int b = (int)(rb / R);
    const uint4* src = reinterpret_cast<const uint4*>(x + b * x_batch + (long long)r * ldx + g * 32);
    uint4 u[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) u[j] = ld_nc_v4(src + j);
    float v[32];
    float amax = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t w[4] = {u[j].x, u[j].y, u[j].z, u[j].w};
#pragma unroll
      for (int h = 0; h < 4; ++h) {
        const float2 f = unpack_bf16x2(w[h]);
        v[j * 8 + h * 2] = f.x;
        v[j * 8 + h * 2 + 1] = f.y;
        amax = fmaxf(amax, fmaxf(fabsf(f.x), fabsf(f.y)));
      }
    }
    float inv;
    const uint32_t e = ue8m0_scale(amax, inv);
    uint32_t o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j)
      o[j] = cvt_e4m3x2(v[j * 4] * inv, v[j * 4 + 1] * inv) | (cvt_e4m3x2(v[j * 4 + 2] * inv, v[j * 4 + 3] * inv) << 16);
    uint4* dst = reinterpret_cast<uint4*>(q + b * q_batch + (long long)r * K + g * 32);
    dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
    dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
    sf[b * sf_batch + sf_offset(r, g >> 2, g & 3, atoms)] = (uint8_t)e;
  }
}

// MX copy of the TRANSPOSE of x: x is [batch][N, K] (K contiguous), the output has rows k and contracts over n -
// q [batch][K, N] bytes, scales for (K rows, N deep).  The backward GEMMs of the expert FFNs multiply by W^T; this
// replaces a strided transpose copy followed by the row quantiser.  One CTA owns a 128 (n) x 128 (k) tile: it goes
// through shared memory once, every thread then owns output row k = tid % 128 and two of the four 32-wide n groups.
constexpr int QT_TILE = 128;
constexpr int QT_PITCH = QT_TILE + 8;      // bf16 elements; 272 B rows keep the 16-byte stores aligned
__global__ void __launch_bounds__(256)
quantize_mx_t_kernel(const __nv_bfloat16* __restrict__ x, uint8_t* __restrict__ q, uint8_t* __restrict__ sf,
                     int N, int K, long long ldx, int atoms, long long x_batch, long long q_batch,
                     long long sf_batch) {
  __shared__ __align__(16) __nv_bfloat16 tile[QT_TILE * QT_PITCH];
  const int k0 = blockIdx.x * QT_TILE, n0 = blockIdx.y * QT_TILE, b = blockIdx.z;
  const __nv_bfloat16* xb = x + b * x_batch;
  const int c8 = (threadIdx.x & 15) * 8;                      // 8 consecutive k per 16-byte load
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int n = (threadIdx.x >> 4) + 16 * i;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (k0 + c8 < K) v = ld_nc_v4(reinterpret_cast<const uint4*>(xb + (long long)(n0 + n) * ldx + k0 + c8));
    *reinterpret_cast<uint4*>(&tile[n * QT_PITCH + c8]) = v;
  }
  __syncthreads();
  const int r = threadIdx.x & 127;
  if (k0 + r >= K) return;
#pragma unroll
  for (int gg = 0; gg < 2; ++gg) {
    const int g = (threadIdx.x >> 7) * 2 + gg;
    float v[32];
    float amax = 0.f;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      v[j] = __bfloat162float(tile[(g * 32 + j) * QT_PITCH + r]);
      amax = fmaxf(amax, fabsf(v[j]));
    }
    float inv;
    const uint32_t e = ue8m0_scale(amax, inv);
    uint32_t o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j)
      o[j] = cvt_e4m3x2(v[j * 4] * inv, v[j * 4 + 1] * inv) | (cvt_e4m3x2(v[j * 4 + 2] * inv, v[j * 4 + 3] * inv) << 16);
    uint4* dst = reinterpret_cast<uint4*>(q + b * q_batch + (long long)(k0 + r) * N + n0 + g * 32);
    dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
    dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
    sf[b * sf_batch + sf_offset(k0 + r, n0 >> 7, g, atoms)] = (uint8_t)e;
  }
}

// e4m3 bytes + atom-layout scales -> bf16 (the bf16 wgrad GEMMs and the tests read quantised operands back).
__global__ void __launch_bounds__(256)
dequantize_mx_kernel(const uint8_t* __restrict__ q, const uint8_t* __restrict__ sf, __nv_bfloat16* __restrict__ y,
                     int R, int K, long long ldy, int atoms, int batch, long long q_batch, long long sf_batch,
                     long long y_batch) {
  const int groups_per_row = K >> 5;
  const long long total = (long long)batch * R * groups_per_row;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i % groups_per_row);
    const long long rb = i / groups_per_row;
    const int r = (int)(rb % R);
    const int b = (int)(rb / R);
    const uint32_t e = sf[b * sf_batch + sf_offset(r, g >> 2, g & 3, atoms)];
    const float s = __uint_as_float(e << 23);
    const uint4* src = reinterpret_cast<const uint4*>(q + b * q_batch + (long long)r * K + g * 32);
    uint4* dst = reinterpret_cast<uint4*>(y + b * y_batch + (long long)r * ldy + g * 32);
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const uint4 u = src[j];
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
      uint32_t o[8];
#pragma unroll
      for (int h = 0; h < 4; ++h) {
        float f[4];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t pair;
          const uint16_t two = (uint16_t)(w[h] >> (16 * c));
          asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(pair) : "h"(two));
          const __half2 hv = *reinterpret_cast<const __half2*>(&pair);
          f[c * 2] = __low2float(hv);
          f[c * 2 + 1] = __high2float(hv);
        }
        o[h * 2] = pack_bf16x2(f[0] * s, f[1] * s);
        o[h * 2 + 1] = pack_bf16x2(f[2] * s, f[3] * s);
      }
      dst[j * 2] = make_uint4(o[0], o[1], o[2], o[3]);
      dst[j * 2 + 1] = make_uint4(o[4], o[5], o[6], o[7]);
    }
  }
}

extern "C" int b2f_mx_sf_atoms(int rows) { return (rows + 127) / 128 + 1; }

extern "C" long long b2f_mx_sf_bytes(int rows, int K) {
  return (long long)(K / 128) * b2f_mx_sf_atoms(rows) * 512;
}

extern "C" int b2f_quantize_mx(const void* x, void* q, void* sf, int rows, int K, long long ldx, int batch,
                               long long x_batch, void* stream) {
  B2F_REQUIRE(rows > 0 && K > 0 && K % 128 == 0 && ldx % 8 == 0 && batch > 0,
              "quantize_mx: K must be a multiple of 128, the row pitch of 8");
  const int atoms = b2f_mx_sf_atoms(rows);
  const long long groups = (long long)batch * rows * (K / 32);
  long long blocks = (groups + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  quantize_mx_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<uint8_t*>(q), static_cast<uint8_t*>(sf), rows, K, ldx,
      atoms, batch, x_batch, (long long)rows * K, b2f_mx_sf_bytes(rows, K));
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// x [batch][N, K] -> MX of x^T: q [batch][K, N], scales of (K rows, N deep).  N % 128 == 0, K % 8 == 0.
extern "C" int b2f_quantize_mx_t(const void* x, void* q, void* sf, int N, int K, long long ldx, int batch,
                                 long long x_batch, void* stream) {
  B2F_REQUIRE(N > 0 && K > 0 && N % 128 == 0 && K % 8 == 0 && ldx % 8 == 0 && batch > 0 && batch < 65536,
              "quantize_mx_t: N must be a multiple of 128, K and the row pitch of 8");
  const int atoms = b2f_mx_sf_atoms(K);
  dim3 grid((K + QT_TILE - 1) / QT_TILE, N / QT_TILE, batch);
  quantize_mx_t_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<uint8_t*>(q), static_cast<uint8_t*>(sf), N, K, ldx, atoms,
      x_batch, (long long)N * K, b2f_mx_sf_bytes(K, N));
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_dequantize_mx(const void* q, const void* sf, void* y, int rows, int K, long long ldy,
                                 int batch, long long y_batch, void* stream) {
  B2F_REQUIRE(rows > 0 && K > 0 && K % 128 == 0 && ldy % 8 == 0 && batch > 0,
              "dequantize_mx: K must be a multiple of 128, the row pitch of 8");
  const int atoms = b2f_mx_sf_atoms(rows);
  const long long groups = (long long)batch * rows * (K / 32);
  long long blocks = (groups + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  dequantize_mx_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint8_t*>(q), static_cast<const uint8_t*>(sf), static_cast<__nv_bfloat16*>(y), rows, K,
      ldy, atoms, batch, (long long)rows * K, b2f_mx_sf_bytes(rows, K), y_batch);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

namespace b2f {
namespace gemm {
int check_params(const Params& p);   // gemm.cu
}
namespace gemm_mx {

int make_mx_maps(CUtensorMap* ma, CUtensorMap* mb, const void* Aq, const void* Bq, int M, int N, int K, int batch) {
  if (int rc = make_tma_3d(ma, Aq, TmaDtype::U8, batch, M, K, (long long)M * K, K, BM, BK, 128)) return rc;
  return make_tma_3d(mb, Bq, TmaDtype::U8, batch, N, K, (long long)N * K, K, BN, BK, 128);
}

}  // namespace gemm_mx
}  // namespace b2f

// C[b] = Aq[b] (x) Bq[b]^T with MX block scales (+bias) (gelu, aux pre-activation) (+residual); Aq [batch][M, K],
// Bq [batch][N, K] dense e4m3 bytes, scales as written by b2f_quantize_mx for (M, K) and (N, K).
extern "C" int b2f_gemm_mx(const void* Aq, const void* sfa, const void* Bq, const void* sfb, void* C, int M, int N,
                           int K, int batch, long long ldc, long long strideC, const void* bias,
                           long long bias_batch, const void* residual, long long ldr, long long strideR,
                           void* aux, long long ldaux, long long strideAux, int flags, void* stream) {
  gemm::Params p{};
  p.M = M; p.N = N; p.K = K; p.batch = batch;
  p.C = C; p.ldc = ldc; p.strideC = strideC;
  p.bias = static_cast<const __nv_bfloat16*>(bias);
  p.residual = static_cast<const __nv_bfloat16*>(residual);
  p.ldr = ldr; p.strideR = strideR;
  p.aux = static_cast<__nv_bfloat16*>(aux);
  p.ldaux = ldaux; p.strideAux = strideAux;
  p.alpha = 1.f; p.flags = flags;
  if (int rc = gemm::check_params(p)) return rc;
  B2F_REQUIRE(K % 128 == 0, "gemm_mx: K must be a multiple of 128");
  B2F_REQUIRE(batch == 1 || bias == nullptr || bias_batch == 0, "gemm_mx: per-batch bias goes through the MoE entry points");
  MxScales sc;
  sc.sfa = static_cast<const uint8_t*>(sfa);
  sc.sfb = static_cast<const uint8_t*>(sfb);
  sc.a_atoms = b2f_mx_sf_atoms(M);
  sc.b_atoms = b2f_mx_sf_atoms(N);
  sc.sfa_batch = b2f_mx_sf_bytes(M, K);
  sc.sfb_batch = b2f_mx_sf_bytes(N, K);
  CUtensorMap ma, mb;
  if (int rc = make_mx_maps(&ma, &mb, Aq, Bq, M, N, K, batch)) return rc;
  static bool attr_set[64] = {false};
  int dev = 0;
  B2F_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    B2F_CHECK_CUDA(cudaFuncSetAttribute(b2f_gemm_mx_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    attr_set[dev & 63] = true;
  }
  const int tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN) * batch;
  const int grid = tiles < sm_count() ? tiles : sm_count();
  b2f_gemm_mx_kernel<<<grid, NUM_THREADS, SMEM_BYTES, static_cast<cudaStream_t>(stream)>>>(ma, mb, p, sc);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
