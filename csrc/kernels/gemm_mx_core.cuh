// b200flow K0 (block-scaled fp8) — persistent tcgen05 GEMM with MX scale factors in tensor memory.
//
//   C[b][M,N] = epilogue( sum_kb  (A[b] * 2^sfa) * (B[b] * 2^sfb) )
//
// A [M, K] and B [N, K] hold e4m3 bytes, K-major; every 32 consecutive K elements of a row share one
// UE8M0 scale (value 2^(byte - 127)) — the OCP MXFP8 format that `tcgen05.mma.kind::mxf8f6f4.block_scale`
// consumes natively.  The scales never pass through registers: a k-block's scale bytes travel
// global -> shared with one bulk copy per operand and shared -> tensor memory with `tcgen05.cp`, and the
// MMA reads them from TMEM next to the accumulator.
//
// Scale-factor layout in global memory ("atoms", written by the quantiser in this file's .cu): for one
// operand with R rows and K columns,  sf[kb][a][512 B]  with kb = k / 128, a = row / 128, and inside an atom
// byte (row % 32) * 16 + ((row % 128) / 32) * 4 + (k % 128) / 32 — exactly the 32 x 128-bit block that
// `tcgen05.cp.32x128b.warpx4` replicates over the four lane quarters of TMEM, so that row r finds the scales of
// its four 32-element K groups in lane r, column r / 32 (mod 4).  An operand is allocated with one spare atom
// per k-block so that a tile which starts in the middle of an atom can always fetch two whole atoms.
//
// Tile 128 x 192 x 128 (K in elements = bytes): two 192-column accumulator stages (384 TMEM columns) leave
// 128 columns for the scale factors of the 5-stage operand ring (4 columns for A, 8 for B per stage), so the
// epilogue of tile i still overlaps the main loop of tile i+1.  A 192-wide tile starts on a 64-column boundary:
// odd tiles begin two TMEM columns into their first B atom, which the MMA's scale address expresses directly.
//
// Roles as in gemm_core.cuh (192 threads: TMA warp, MMA warp, four epilogue warps) and the same Policy hooks, so
// the MoE kernels (csrc/comm/fused_moe.cu) reuse the pipeline: a tile's operand rows may arrive from other GPUs
// (`wait_operands`), and the epilogue may quantise its output for the next GEMM or store gate-scaled rows into
// the GPU that owns the token (`store_chunk`).
//
// Capability parity: DeepSpeed-MoE expert FFNs on cuBLAS / Transformer-Engine fp8.
#pragma once
#include "gemm_core.cuh"

namespace b2f {
namespace gemm_mx {

using gemm::Params;

constexpr int BM = 128;
constexpr int BN = 192;
constexpr int BK = 128;                       // fp8 elements = bytes: one 128 B swizzle row
constexpr int STAGES = 5;
constexpr int A_BYTES = BM * BK;              // 16 KB
constexpr int B_BYTES = BN * BK;              // 24 KB
constexpr int SFA_BYTES = 512;                // one atom: 128 rows x 4 K groups
constexpr int SFB_BYTES = 1024;               // two atoms: up to 256 columns
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int SF_STAGE_BYTES = SFA_BYTES + SFB_BYTES;
constexpr int NUM_THREADS = 192;
constexpr int TMEM_COLS = 512;
constexpr int ACC_COLS = BN;                  // per accumulator stage
constexpr int SF_TMEM_BASE = 2 * ACC_COLS;    // 384
constexpr int SF_TMEM_STAGE = 12;             // 4 (A) + 8 (B) columns per ring stage
constexpr int SMEM_BYTES = STAGES * (STAGE_BYTES + SF_STAGE_BYTES) + 1024 /*align slack*/ + 256 /*barriers*/;
static_assert(SF_TMEM_BASE + STAGES * SF_TMEM_STAGE <= TMEM_COLS, "scale factors do not fit in tensor memory");
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");

// Scale-factor operands of one launch.
struct MxScales {
  const uint8_t* sfa;        // [batch][K/128][a_atoms][512]
  const uint8_t* sfb;        // [batch][K/128][b_atoms][512]
  int a_atoms, b_atoms;      // atoms allocated per k-block (>= ceil(rows / 128) + 1)
  long long sfa_batch, sfb_batch;   // bytes between batches
};

struct SharedStorage {
  uint64_t full[STAGES];
  uint64_t empty[STAGES];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
};

// Instruction descriptor of kind::mxf8f6f4.block_scale: e4m3 x e4m3, UE8M0 scales, fp32 accumulate, both K-major.
// Bits: [4,6) B scale id, [7,10) A format, [10,13) B format, [17,23) N >> 3, [23] scale format (1 = UE8M0),
// [24,29) M >> 4, [29,31) A scale id.  The scale ids select which of the four bytes of a 32-bit TMEM scale
// column the instruction uses, i.e. the 32-element K group inside the 128-element k-block.
__host__ __device__ constexpr uint32_t make_idesc_mx(uint32_t m, uint32_t n) {
  return ((n >> 3) << 17) | (1u << 23) | ((m >> 4) << 24);
}

__device__ __forceinline__ void umma_mxf8_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                             uint32_t sfa_tmem, uint32_t sfb_tmem, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf8f6f4.block_scale [%0], %1, %2, %3, [%4], [%5], p;\n\t"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(sfa_tmem), "r"(sfb_tmem), "r"(accumulate)
      : "memory");
}

// shared -> tensor memory: one 32 x 128-bit block (an atom of 512 B), replicated to all four lane quarters.
__device__ __forceinline__ void tmem_cp_32x128b_warpx4(uint32_t taddr, uint64_t sdesc) {
  asm volatile("tcgen05.cp.cta_group::1.32x128b.warpx4 [%0], %1;" ::"r"(taddr), "l"(sdesc) : "memory");
}

// Shared-memory descriptor of an un-swizzled K-major block of 8-row x 16-byte core matrices stacked along rows
// (SBO = 128 B); one core matrix wide, so LBO is not used.
__device__ __forceinline__ uint64_t make_sf_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>((128u >> 4) & 0x3FFF) << 32;
  d |= 1ull << 46;
  return d;
}

// e4m3 pair from two floats (round to nearest, saturating).
__device__ __forceinline__ uint32_t cvt_e4m3x2(float lo, float hi) {
  uint16_t r;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %2, %1;" : "=h"(r) : "f"(lo), "f"(hi));
  return r;
}

// UE8M0 scale of a 32-element group with absolute maximum `amax`: the smallest power of two 2^e with
// amax / 2^e <= 448 (the largest e4m3 value).  Returns the biased byte (e + 127); `inv` = 2^-e.
__device__ __forceinline__ uint32_t ue8m0_scale(float amax, float& inv) {
  const float t = __fdiv_rn(amax, 448.0f);
  const uint32_t bits = __float_as_uint(t);
  uint32_t e = (bits >> 23) & 0xFF;
  if (bits & 0x7FFFFF) e += 1;                 // round the exponent up: the quantised values must not overflow
  e = e < 1u ? 1u : (e > 253u ? 253u : e);
  inv = __uint_as_float((254u - e) << 23);
  return e;
}

// Byte offset of the scale of (row, 32-element K group g of k-block kb) inside an operand's scale array.
__device__ __forceinline__ long long sf_offset(int row, int kb, int g, int atoms) {
  return ((long long)kb * atoms + (row >> 7)) * 512 + (row & 31) * 16 + ((row & 127) >> 5) * 4 + g;
}

struct PlainMxPolicy {
  static constexpr int GROUP_M = 16;
  __device__ __forceinline__ void map_tile(int t, int num_m, int num_n, int& b, int& mb, int& nb) const {
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
  // false: the tile has no rows to compute (all three roles skip it consistently)
  __device__ __forceinline__ bool tile_active(int, int, int) const { return true; }
  __device__ __forceinline__ void wait_operands(int, int, int) const {}
  __device__ __forceinline__ void before_epilogue(int, int, int) const {}
  __device__ __forceinline__ void store_chunk(const Params& p, int b, int, int, int m, int n0,
                                              const uint32_t (&acc)[32]) const {
    gemm::epilogue_chunk(p, b, m, n0, acc);
  }
  __device__ __forceinline__ void tile_done(int, int, int) const {}
  __device__ __forceinline__ void finish() const {}
};

// NST: depth of the operand ring (a fused kernel may trade a stage for epilogue staging space; the shared-memory
// layout is [NST x 40 KB operands][NST x 1.5 KB scales][SharedStorage][policy-owned tail]).
template <class Policy, int NST = STAGES>
__device__ __forceinline__ void gemm_mx_body(const CUtensorMap* tma_a, const CUtensorMap* tma_b, const Params& p,
                                             const MxScales& sc, const Policy& pol, int first_tile,
                                             int tile_stride, uint8_t* smem_raw) {
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* sf_smem = smem + NST * STAGE_BYTES;
  SharedStorage* ss = reinterpret_cast<SharedStorage*>(sf_smem + NST * SF_STAGE_BYTES);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_m = (p.M + BM - 1) / BM;
  const int num_n = (p.N + BN - 1) / BN;
  const int num_tiles = num_m * num_n * p.batch;
  const int num_kb = p.K / BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(tma_a);
    tma_prefetch_desc(tma_b);
#pragma unroll
    for (int s = 0; s < NST; ++s) {
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
        if (!pol.tile_active(b, mb, nb)) continue;
        pol.wait_operands(b, mb, nb);
        const int m0 = mb * BM, n0 = nb * BN;
        const uint8_t* sfa = sc.sfa + (long long)b * sc.sfa_batch + (long long)mb * 512;
        const uint8_t* sfb = sc.sfb + (long long)b * sc.sfb_batch + (long long)(n0 >> 7) * 512;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&ss->empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          uint8_t* sfs = sf_smem + stage * SF_STAGE_BYTES;
          mbar_arrive_expect_tx(&ss->full[stage], STAGE_BYTES + SF_STAGE_BYTES);
          tma_load_3d(sa, tma_a, &ss->full[stage], kb * BK, m0, b);
          tma_load_3d(sa + A_BYTES, tma_b, &ss->full[stage], kb * BK, n0, b);
          bulk_load_1d(sfs, sfa + (long long)kb * sc.a_atoms * 512, SFA_BYTES, &ss->full[stage]);
          bulk_load_1d(sfs + SFA_BYTES, sfb + (long long)kb * sc.b_atoms * 512, SFB_BYTES, &ss->full[stage]);
          if (++stage == NST) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (whole warp walks, one lane issues)
    const uint32_t idesc = make_idesc_mx(BM, BN);
    constexpr uint32_t DHI = desc_hi_sw128(1024);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int t = first_tile; t < num_tiles; t += tile_stride) {
      int b, mb, nb;
      pol.map_tile(t, num_m, num_n, b, mb, nb);
      if (!pol.tile_active(b, mb, nb)) continue;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      ++it;
      mbar_wait(&ss->tmem_empty[as], aphase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * ACC_COLS;
      // a 192-wide tile starts 0 or 64 columns into its first B atom: 0 or 2 TMEM columns
      const uint32_t sfb_col0 = ((nb * BN) & 127) >> 5;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&ss->full[stage], phase);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
        const uint32_t sfs = smem_u32(sf_smem + stage * SF_STAGE_BYTES);
        const uint32_t la = desc_lo_sw128(sa, 16), lb = desc_lo_sw128(sa + A_BYTES, 16);
        const uint32_t sf_t = tmem_base + SF_TMEM_BASE + stage * SF_TMEM_STAGE;
        if (elect_one()) {
          tmem_cp_32x128b_warpx4(sf_t, make_sf_desc(sfs));
          tmem_cp_32x128b_warpx4(sf_t + 4, make_sf_desc(sfs + SFA_BYTES));
          tmem_cp_32x128b_warpx4(sf_t + 8, make_sf_desc(sfs + SFA_BYTES + 512));
#pragma unroll
          for (int k = 0; k < BK / 32; ++k) {
            const uint64_t da = desc_from(la + k * 2, DHI);          // +32 B inside the swizzle atom
            const uint64_t db = desc_from(lb + k * 2, DHI);
            const uint32_t id = idesc | ((uint32_t)k << 29) | ((uint32_t)k << 4);
            umma_mxf8_ss(d_tmem, da, db, id, sf_t, sf_t + 4 + sfb_col0, (kb | k) != 0);
          }
          umma_commit(&ss->empty[stage]);
          if (kb == num_kb - 1) umma_commit(&ss->tmem_full[as]);
        }
        __syncwarp();
        if (++stage == NST) { stage = 0; phase ^= 1; }
      }
    }
    __syncwarp();
  } else {
    // ---------------------------------------------------------------- epilogue
    const int q = warp & 3;
    int it = 0;
    for (int t = first_tile; t < num_tiles; t += tile_stride) {
      int b, mb, nb;
      pol.map_tile(t, num_m, num_n, b, mb, nb);
      if (!pol.tile_active(b, mb, nb)) continue;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      ++it;
      mbar_wait(&ss->tmem_full[as], aphase);
      tc_fence_after();
      pol.before_epilogue(b, mb, nb);
      const int m = mb * BM + q * 32 + lane;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * ACC_COLS;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        const int n0 = nb * BN + c * 32;
        if (n0 >= p.N) break;
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

}  // namespace gemm_mx
}  // namespace b2f
