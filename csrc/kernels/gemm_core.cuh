// b200flow K0 — persistent warp-specialised tcgen05 GEMM core for sm_100a.
//
//   C[b][M,N] = epilogue( alpha * A[b] * B[b] )      bf16 in, fp32 accumulate in TMEM
//
// Roles (192 threads): warp 0 = TMA producer (one lane), warp 1 = MMA issuer
// (one lane) + TMEM owner, warps 2..5 = epilogue (one accumulator row per
// thread, TMEM -> registers -> global).  128x256x64 tiles, 4-stage smem ring,
// two 256-column TMEM accumulator stages so the epilogue of tile i overlaps the
// main loop of tile i+1.  Operands may be K-major or MN-major (selected at run
// time through the UMMA descriptors), which gives forward (NT), dgrad (NN) and
// wgrad (TN) GEMMs from one kernel without any transposes.
//
// The kernel body is a device-function template over a `Policy` so that the
// fused compute+collective kernels (all-gather->GEMM, GEMM->reduce-scatter,
// MoE dispatch/combine) reuse the exact same pipeline and only override tile
// order, the "operand tile is ready" wait, and the epilogue destination.
//
// Capability parity: cuBLAS GEMMs + apex fused_dense / fused_weight_gradient_mlp
//  that the reference pulls in through
// containers/megatron-deepspeed/Dockerfile:6-13.
#pragma once
#include "../common/ptx.cuh"

namespace b2f {
namespace gemm {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;  // 64 bf16 = 128 B = one swizzle-128B row
constexpr int STAGES = 4;
constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KB
constexpr int B_STAGE_BYTES = BN * BK * 2;  // 32 KB
constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
constexpr int CHUNK_BYTES = 64 * BK * 2;  // one 64(MN) x 64(K) MN-major box = 8 KB
constexpr int NUM_THREADS = 192;
constexpr int TMEM_COLS = 512;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;

enum EpiFlags : int {
  EPI_GELU = 1,        // v = gelu(v) (tanh approximation, as Megatron's bias_gelu)
  EPI_OUT_F32 = 2,     // C is fp32
  EPI_ACCUM = 4,       // C += v (fp32 C only; wgrad main_grad accumulate)
  EPI_RELU = 8,        // v = max(v, 0)
  EPI_BIAS_ROW = 16,   // bias indexed by row (m) instead of column (n)
  EPI_DGELU = 32,      // `residual` holds a saved pre-activation z: v *= gelu'(z) instead of v += z
};

struct Params {
  int M, N, K;
  int batch;
  int a_mn_major, b_mn_major;
  // epilogue
  void* C;
  long long ldc, strideC;
  const __nv_bfloat16* bias;       // [N] (or [M] with EPI_BIAS_ROW) or nullptr
  const __nv_bfloat16* residual;   // [M, N] bf16 or nullptr
  long long ldr, strideR;
  __nv_bfloat16* aux;              // optional pre-activation copy, [M, N]
  long long ldaux, strideAux;
  float alpha;
  int flags;
  // fp8 path: dequantisation scales, v = acc * row_scale[m] * col_scale[n]
  const float* row_scale;
  const float* col_scale;
};

struct SharedStorage {
  uint64_t full[STAGES];
  uint64_t empty[STAGES];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
};

__device__ __forceinline__ float gelu_tanh(float x) {
  // 0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3)))
  float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  return 0.5f * x * (1.0f + t);
}

__device__ __forceinline__ float gelu_tanh_grad(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  const float du = 0.7978845608028654f * (1.f + 3.f * 0.044715f * x * x);
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * du;
}

// Default policy: plain GEMM, row-major tile order (m fastest), no cross-GPU hooks.
struct PlainPolicy {
  static constexpr int GROUP_M = 16;  // 2048 rows of A stay in L2 while B streams past (see gemm2_core.cuh)
  __device__ __forceinline__ void map_tile(int t, int num_m, int num_n, int& b, int& mb,
                                           int& nb) const {
    const int per_batch = num_m * num_n;
    b = t / per_batch;
    const int r = t - b * per_batch;
    const int per_group = GROUP_M * num_n;
    const int g = r / per_group;
    const int m0 = g * GROUP_M;
    const int gm = num_m - m0 < GROUP_M ? num_m - m0 : GROUP_M;
    const int rr = r - g * per_group;
    nb = rr / gm;
    mb = m0 + (rr - nb * gm);
  }
  // Called by the producer lane before the first TMA of a tile.
  __device__ __forceinline__ void wait_operands(int, int, int) const {}
  // Called by every epilogue thread once the accumulator tile is complete in TMEM.
  __device__ __forceinline__ void before_epilogue(int, int, int) const {}
  // Convert/store 32 columns of one accumulator row (default: the standard epilogue).
  __device__ __forceinline__ void store_chunk(const Params& p, int b, int mb, int nb, int m, int n0,
                                              const uint32_t (&acc)[32]) const;
  // Called by every epilogue thread after its rows of the tile are stored.
  __device__ __forceinline__ void tile_done(int, int, int) const {}
  __device__ __forceinline__ void finish() const {}
};

// Epilogue arithmetic of one 32-column chunk of one accumulator row (everything but the C store).
__device__ __forceinline__ void epilogue_values(const Params& p, int b, int m, int n0,
                                                const uint32_t (&acc)[32], float (&v)[32]) {
  const int flags = p.flags;
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(acc[i]) * p.alpha;
  const bool full = (n0 + 32 <= p.N);
  if (p.row_scale) {
    const float rs = p.row_scale[m];
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] *= rs;
  }
  if (p.col_scale) {
    for (int i = 0; i < 32; ++i)
      if (n0 + i < p.N) v[i] *= p.col_scale[n0 + i];
  }
  if (p.bias) {
    if (flags & EPI_BIAS_ROW) {
      float bv = h2f(p.bias[m]);
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] += bv;
    } else if (full) {
      const uint4* bp = reinterpret_cast<const uint4*>(p.bias + n0);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 u = __ldg(bp + q);
        float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y), f2 = unpack_bf16x2(u.z),
               f3 = unpack_bf16x2(u.w);
        v[q * 8 + 0] += f0.x; v[q * 8 + 1] += f0.y; v[q * 8 + 2] += f1.x; v[q * 8 + 3] += f1.y;
        v[q * 8 + 4] += f2.x; v[q * 8 + 5] += f2.y; v[q * 8 + 6] += f3.x; v[q * 8 + 7] += f3.y;
      }
    } else {
      for (int i = 0; i < 32; ++i)
        if (n0 + i < p.N) v[i] += h2f(p.bias[n0 + i]);
    }
  }
  if (p.aux) {
    __nv_bfloat16* ap = p.aux + (long long)b * p.strideAux + (long long)m * p.ldaux + n0;
    if (full && ((p.ldaux & 7) == 0)) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 u;
        u.x = pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]); u.y = pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]);
        u.z = pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]); u.w = pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7]);
        *reinterpret_cast<uint4*>(ap + q * 8) = u;
      }
    } else {
      for (int i = 0; i < 32; ++i)
        if (n0 + i < p.N) ap[i] = f2h(v[i]);
    }
  }
  if (flags & EPI_GELU) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = gelu_tanh(v[i]);
  }
  if (flags & EPI_RELU) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.0f);
  }
  if (p.residual) {
    const __nv_bfloat16* rp =
        p.residual + (long long)b * p.strideR + (long long)m * p.ldr + n0;
    float r[32];
    if (full && ((p.ldr & 7) == 0)) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 u = *reinterpret_cast<const uint4*>(rp + q * 8);
        float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y), f2 = unpack_bf16x2(u.z),
               f3 = unpack_bf16x2(u.w);
        r[q * 8 + 0] = f0.x; r[q * 8 + 1] = f0.y; r[q * 8 + 2] = f1.x; r[q * 8 + 3] = f1.y;
        r[q * 8 + 4] = f2.x; r[q * 8 + 5] = f2.y; r[q * 8 + 6] = f3.x; r[q * 8 + 7] = f3.y;
      }
    } else {
      for (int i = 0; i < 32; ++i) r[i] = (n0 + i < p.N) ? h2f(rp[i]) : 0.f;
    }
    if (flags & EPI_DGELU) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] *= gelu_tanh_grad(r[i]);
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] += r[i];
    }
  }
}

// C store of one chunk (bf16 or fp32, optional fp32 accumulate).
__device__ __forceinline__ void epilogue_store(const Params& p, int b, int m, int n0, const float (&v)[32]) {
  const int flags = p.flags;
  const bool full = (n0 + 32 <= p.N);
  if (flags & EPI_OUT_F32) {
    float* cp = reinterpret_cast<float*>(p.C) + (long long)b * p.strideC + (long long)m * p.ldc + n0;
    if (full && ((p.ldc & 3) == 0)) {
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        float4 o = make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
        if (flags & EPI_ACCUM) {
          float4 c = *reinterpret_cast<const float4*>(cp + q * 4);
          o.x += c.x; o.y += c.y; o.z += c.z; o.w += c.w;
        }
        *reinterpret_cast<float4*>(cp + q * 4) = o;
      }
    } else {
      for (int i = 0; i < 32; ++i)
        if (n0 + i < p.N) cp[i] = (flags & EPI_ACCUM) ? cp[i] + v[i] : v[i];
    }
  } else {
    __nv_bfloat16* cp = reinterpret_cast<__nv_bfloat16*>(p.C) + (long long)b * p.strideC +
                        (long long)m * p.ldc + n0;
    if (full && ((p.ldc & 7) == 0)) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 u;
        u.x = pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]); u.y = pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]);
        u.z = pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]); u.w = pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7]);
        *reinterpret_cast<uint4*>(cp + q * 8) = u;
      }
    } else {
      for (int i = 0; i < 32; ++i)
        if (n0 + i < p.N) cp[i] = f2h(v[i]);
    }
  }
}

// Store/convert one 32-column chunk of one accumulator row.
__device__ __forceinline__ void epilogue_chunk(const Params& p, int b, int m, int n0,
                                               const uint32_t (&acc)[32]) {
  if (m >= p.M) return;
  float v[32];
  epilogue_values(p, b, m, n0, acc, v);
  epilogue_store(p, b, m, n0, v);
}

__device__ __forceinline__ void PlainPolicy::store_chunk(const Params& p, int b, int, int, int m,
                                                         int n0, const uint32_t (&acc)[32]) const {
  epilogue_chunk(p, b, m, n0, acc);
}

// The whole persistent GEMM. `first_tile`/`tile_stride` let a fused kernel give
// only a subset of its CTAs to the GEMM role (the rest move data between GPUs).
// KIND: 0 = bf16 operands (kind::f16, 64 elements per 128 B row), 1 = fp8 e4m3 operands
// (kind::f8f6f4, 128 elements per row, K-major only).
// NSTAGES: depth of the operand ring (a fused kernel may trade a stage for epilogue staging space;
// the shared-memory layout is [NSTAGES x 48 KB ring][SharedStorage, 1 KB][policy-owned tail]).
template <class Policy, int KIND = 0, int NSTAGES = STAGES>
__device__ __forceinline__ void gemm_body(const CUtensorMap* tma_a, const CUtensorMap* tma_b,
                                          const Params& p, const Policy& pol, int first_tile,
                                          int tile_stride, uint8_t* smem_raw) {
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  SharedStorage* ss = reinterpret_cast<SharedStorage*>(smem + NSTAGES * STAGE_BYTES);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_m = (p.M + BM - 1) / BM;
  const int num_n = (p.N + BN - 1) / BN;
  const int num_tiles = num_m * num_n * p.batch;
  constexpr int BKE = KIND == 1 ? 2 * BK : BK;      // elements per k-block (one 128 B row)
  const int num_kb = (p.K + BKE - 1) / BKE;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(tma_a);
    tma_prefetch_desc(tma_b);
#pragma unroll
    for (int s = 0; s < NSTAGES; ++s) {
      mbar_init(&ss->full[s], 1);
      mbar_init(&ss->empty[s], 1);
    }
    mbar_init(&ss->tmem_full[0], 1);
    mbar_init(&ss->tmem_full[1], 1);
    mbar_init(&ss->tmem_empty[0], 4);
    mbar_init(&ss->tmem_empty[1], 4);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(&ss->tmem_base, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = ss->tmem_base;

  if (warp == 0) {
    if (lane == 0) {
      // ---------------------------------------------------------- TMA producer
      int stage = 0;
      uint32_t phase = 0;
      for (int t = first_tile; t < num_tiles; t += tile_stride) {
        int b, mb, nb;
        pol.map_tile(t, num_m, num_n, b, mb, nb);
        pol.wait_operands(b, mb, nb);
        const int m0 = mb * BM, n0 = nb * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&ss->empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          uint8_t* sb = sa + A_STAGE_BYTES;
          mbar_arrive_expect_tx(&ss->full[stage], STAGE_BYTES);
          const int k0 = kb * BKE;
          if (!p.a_mn_major) {
            tma_load_3d(sa, tma_a, &ss->full[stage], k0, m0, b);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              tma_load_3d(sa + j * CHUNK_BYTES, tma_a, &ss->full[stage], m0 + j * 64, k0, b);
          }
          if (!p.b_mn_major) {
            tma_load_3d(sb, tma_b, &ss->full[stage], k0, n0, b);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_3d(sb + j * CHUNK_BYTES, tma_b, &ss->full[stage], n0 + j * 64, k0, b);
          }
          if (++stage == NSTAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    {
      // ------------------------------------------------------------ MMA issuer: the whole warp
      // walks the loop, one elected lane issues (keeps descriptors in uniform registers)
      const uint32_t idesc = KIND == 1 ? make_idesc(0, 0, BM, BN, 0, 0)
                                       : make_idesc(B2F_MMA_FMT, B2F_MMA_FMT, BM, BN, p.a_mn_major, p.b_mn_major);
      // K-major: +32 B per UMMA_K inside the swizzle atom; MN-major: +16 rows * 128 B.
      const uint32_t a_kstep = p.a_mn_major ? 2048u : 32u;
      const uint32_t b_kstep = p.b_mn_major ? 2048u : 32u;
      const uint32_t a_lbo = p.a_mn_major ? (uint32_t)CHUNK_BYTES : 16u;
      const uint32_t b_lbo = p.b_mn_major ? (uint32_t)CHUNK_BYTES : 16u;
      constexpr uint32_t DHI = desc_hi_sw128(1024);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int t = first_tile; t < num_tiles; t += tile_stride, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&ss->tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&ss->full[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
          const uint32_t la = desc_lo_sw128(sa, a_lbo), lb = desc_lo_sw128(sa + A_STAGE_BYTES, b_lbo);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              const uint64_t da = desc_from(la + k * (a_kstep >> 4), DHI);
              const uint64_t db = desc_from(lb + k * (b_kstep >> 4), DHI);
              if (KIND == 1) umma_f8_ss(d_tmem, da, db, idesc, (kb | k) != 0);
              else umma_f16_ss(d_tmem, da, db, idesc, (kb | k) != 0);
            }
            umma_commit(&ss->empty[stage]);  // frees the smem slot when these MMAs retire
            if (kb == num_kb - 1) umma_commit(&ss->tmem_full[as]);
          }
          __syncwarp();
          if (++stage == NSTAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else {
    // ---------------------------------------------------------------- epilogue
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    int it = 0;
    for (int t = first_tile; t < num_tiles; t += tile_stride, ++it) {
      int b, mb, nb;
      pol.map_tile(t, num_m, num_n, b, mb, nb);
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      mbar_wait(&ss->tmem_full[as], aphase);
      tc_fence_after();
      pol.before_epilogue(b, mb, nb);
      const int m = mb * BM + q * 32 + lane;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        const int n0 = nb * BN + c * 32;
        if (n0 >= p.N) break;  // warp-uniform
        uint32_t acc[32];
        tmem_ld_32x32b_x32(taddr + c * 32, acc);
        tmem_ld_wait();
        pol.store_chunk(p, b, mb, nb, m, n0, acc);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&ss->tmem_empty[as]);
      pol.tile_done(b, mb, nb);
    }
    pol.finish();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

}  // namespace gemm
}  // namespace b2f
