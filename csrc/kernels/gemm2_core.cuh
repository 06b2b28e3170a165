// b200flow K0 (2-CTA variant) — device body of the cta_group::2 GEMM, shared by the plain kernel
// (gemm2.cu) and the fused compute+collective kernels (comm/fused_gemm.cu).  See gemm2.cu for the
// design notes and the barrier protocol.
#pragma once
#include "gemm_core.cuh"

namespace b2f {
namespace gemm2 {
using namespace b2f::gemm;

constexpr int T2_BM = 256, T2_BN = 256;     // tile of the CTA pair
constexpr int HALF = 128;                   // rows of A / rows of B staged per CTA
constexpr int S2 = 6;                       // stages
constexpr int A2_BYTES = HALF * BK * 2;     // 16 KB
constexpr int B2_BYTES = HALF * BK * 2;     // 16 KB
constexpr int ST2_BYTES = A2_BYTES + B2_BYTES;
constexpr int SMEM2_BYTES = S2 * ST2_BYTES + 1024 + 256;
constexpr int smem2_bytes(int nst, int tail) { return nst * ST2_BYTES + 1024 /*barriers*/ + tail + 1024 /*align*/; }

struct Shared2 {
  uint64_t full[S2];
  uint64_t empty[S2];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local` in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  // default semantics (no cluster-scope membar): a .release.cluster arrive made the producer wait for
  // its previous TMA round trip every k-block and halved the kernel's throughput
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tma_load_3d_2sm(void* smem_dst, const CUtensorMap* m, uint32_t bar_cluster_addr,
                                                int32_t c0, int32_t c1, int32_t c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster_addr), "r"(c0), "r"(c1),
      "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_f16_ss_2sm(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                                uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this smem offset in BOTH CTAs once all prior MMAs have retired
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar) {
  const uint16_t mask = 0x3;
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"(mask)
      : "memory");
}

// Cycle accounting of the leader's MMA thread in pair 0 (development aid, read with
// b2f_gemm2_debug; compiled in with -DB2F_GEMM2_DEBUG): [0] total loop, [1] waiting for operands, [2] waiting for a free accumulator,
// [3] k-blocks issued.
#ifdef B2F_GEMM2_DEBUG
static __device__ long long g_gemm2_dbg[8];
#endif

// The whole persistent CTA-pair GEMM as a device function.  `first_pair` / `npairs` let a fused
// kernel keep some clusters for data movement; NST is the ring depth (shared-memory layout:
// [NST x 32 KB ring][Shared2, 1 KB][policy-owned tail]).  Policy hooks (all indices in 128-row
// blocks of the CTA that calls them, mb128 = 2 * pair-tile row + cluster rank):
//   map_tile(t, num_m, num_n, b, mb, nb)            tile order over 256x256 pair tiles
//   wait_operands(b, mb128, nb)                     producer: rows of A for this block are ready
//   select_a(def, mb128, m0) -> map                 producer: tensor map / row origin for A
//   before_epilogue / store_chunk / tile_done       epilogue, as in the 1-CTA core
// tile_done may take the Params (policies that store C themselves) or not
template <class Policy>
__device__ __forceinline__ auto tile_done_hook(const Policy& pol, const Params& p, int b, int mb128, int nb)
    -> decltype(pol.tile_done(p, b, mb128, nb)) {
  return pol.tile_done(p, b, mb128, nb);
}
template <class Policy>
__device__ __forceinline__ auto tile_done_hook(const Policy& pol, const Params&, int b, int mb128, int nb)
    -> decltype(pol.tile_done(b, mb128, nb)) {
  return pol.tile_done(b, mb128, nb);
}

template <class Policy, int NST>
__device__ __forceinline__ void gemm2_body(const CUtensorMap* tma_a_p, const CUtensorMap* tma_b_p,
                                           const Params& p, const Policy& pol, int first_pair,
                                           int npairs, uint8_t* smem_raw, int tile_begin = 0,
                                           int tile_end = -1) {
  const CUtensorMap& tma_a = *tma_a_p;
  const CUtensorMap& tma_b = *tma_b_p;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  Shared2* ss = reinterpret_cast<Shared2*>(smem + NST * ST2_BYTES);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();
  const bool leader = cta == 0;
  const int num_m = (p.M + T2_BM - 1) / T2_BM;
  const int num_n = (p.N + T2_BN - 1) / T2_BN;
  // [tile_begin, tile_end): the slice of the tile order this group of `npairs` clusters walks (default: all of it)
  const int num_tiles = tile_end < 0 ? num_m * num_n * p.batch : tile_end;
  const int num_kb = (p.K + BK - 1) / BK;
  const int pair = tile_begin + first_pair;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    for (int s = 0; s < NST; ++s) {
      mbar_init(&ss->full[s], 2);
      mbar_init(&ss->empty[s], 1);
    }
    mbar_init(&ss->tmem_full[0], 1);
    mbar_init(&ss->tmem_full[1], 1);
    mbar_init(&ss->tmem_empty[0], 8);
    mbar_init(&ss->tmem_empty[1], 8);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc_2sm(&ss->tmem_base, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();            // peer's barriers are initialised before anyone signals them
  tc_fence_after();
  const uint32_t tmem_base = ss->tmem_base;

  auto map_tile = [&](int t, int& b, int& mb, int& nb) { pol.map_tile(t, num_m, num_n, b, mb, nb); };

  if (warp == 0) {
    if (lane == 0) {
      // ------------------------------------------------ TMA producer (one per CTA)
      int stage = 0;
      uint32_t phase = 0;
      for (int t = pair; t < num_tiles; t += npairs) {
        int b, mb, nb;
        map_tile(t, b, mb, nb);
        int m0 = mb * T2_BM + (int)cta * HALF;           // this CTA's accumulator rows
        const int n0 = nb * T2_BN + (int)cta * HALF;     // this CTA's half of the B tile
        pol.wait_operands(b, 2 * mb + (int)cta, nb);
        const CUtensorMap* map_a = pol.select_a(&tma_a, 2 * mb + (int)cta, m0);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&ss->empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * ST2_BYTES;
          uint8_t* sb = sa + A2_BYTES;
          const uint32_t full_leader = map_to_cta(smem_u32(&ss->full[stage]), 0);
          if (leader) mbar_arrive_expect_tx(&ss->full[stage], 2 * ST2_BYTES);
          else mbar_arrive_cluster(full_leader);
          const int k0 = kb * BK;
          if (!p.a_mn_major) {
            tma_load_3d_2sm(sa, map_a, full_leader, k0, m0, b);
          } else {
            tma_load_3d_2sm(sa, map_a, full_leader, m0, k0, b);
            tma_load_3d_2sm(sa + CHUNK_BYTES, map_a, full_leader, m0 + 64, k0, b);
          }
          if (!p.b_mn_major) {
            tma_load_3d_2sm(sb, &tma_b, full_leader, k0, n0, b);
          } else {
            tma_load_3d_2sm(sb, &tma_b, full_leader, n0, k0, b);
            tma_load_3d_2sm(sb + CHUNK_BYTES, &tma_b, full_leader, n0 + 64, k0, b);
          }
          if (++stage == NST) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (leader) {
      // ------------------------------------------- MMA issuer (leader CTA only).  The whole warp
      // walks the loop (uniform control flow, operands stay in uniform registers) and one elected
      // lane issues; descriptors are a precomputed low word plus a constant per K step.
      const uint32_t idesc = make_idesc(B2F_MMA_FMT, B2F_MMA_FMT, T2_BM, T2_BN, p.a_mn_major, p.b_mn_major);
      const uint32_t a_kstep = p.a_mn_major ? 2048u : 32u;
      const uint32_t b_kstep = p.b_mn_major ? 2048u : 32u;
      const uint32_t a_lbo = p.a_mn_major ? (uint32_t)CHUNK_BYTES : 16u;
      const uint32_t b_lbo = p.b_mn_major ? (uint32_t)CHUNK_BYTES : 16u;
      constexpr uint32_t DHI = desc_hi_sw128(1024);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
#ifdef B2F_GEMM2_DEBUG
      long long t_full = 0, t_acc = 0, nkb = 0;
      const long long t_begin = clock64();
#endif
      for (int t = pair; t < num_tiles; t += npairs, ++it) {
        const int as = it & 1;
#ifdef B2F_GEMM2_DEBUG
        long long c0 = clock64();
#endif
        mbar_wait(&ss->tmem_empty[as], ((it >> 1) & 1) ^ 1);
#ifdef B2F_GEMM2_DEBUG
        t_acc += clock64() - c0;
#endif
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * T2_BN;
        for (int kb = 0; kb < num_kb; ++kb) {
#ifdef B2F_GEMM2_DEBUG
          c0 = clock64();
#endif
          mbar_wait(&ss->full[stage], phase);
#ifdef B2F_GEMM2_DEBUG
          t_full += clock64() - c0;
          ++nkb;
#endif
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * ST2_BYTES);
          const uint32_t la = desc_lo_sw128(sa, a_lbo), lb = desc_lo_sw128(sa + A2_BYTES, b_lbo);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              umma_f16_ss_2sm(d_tmem, desc_from(la + k * (a_kstep >> 4), DHI),
                              desc_from(lb + k * (b_kstep >> 4), DHI), idesc, (kb | k) != 0);
            umma_commit_2sm(&ss->empty[stage]);
            if (kb == num_kb - 1) umma_commit_2sm(&ss->tmem_full[as]);
          }
          __syncwarp();
          if (++stage == NST) { stage = 0; phase ^= 1; }
        }
      }
#ifdef B2F_GEMM2_DEBUG
      if (pair == 0 && lane == 0) {
        g_gemm2_dbg[0] = clock64() - t_begin;
        g_gemm2_dbg[1] = t_full;
        g_gemm2_dbg[2] = t_acc;
        g_gemm2_dbg[3] = nkb;
      }
#endif
    }
    __syncwarp();
  } else {
    // -------------------------------------------- epilogue (both CTAs, own rows)
    const int q = warp & 3;
    int it = 0;
    for (int t = pair; t < num_tiles; t += npairs, ++it) {
      int b, mb, nb;
      map_tile(t, b, mb, nb);
      const int as = it & 1;
      mbar_wait(&ss->tmem_full[as], (it >> 1) & 1);
      tc_fence_after();
      const int mb128 = 2 * mb + (int)cta;
      pol.before_epilogue(b, mb128, nb);
      const int m = mb * T2_BM + (int)cta * HALF + q * 32 + lane;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * T2_BN;
#pragma unroll 1
      for (int c = 0; c < T2_BN / 32; ++c) {
        const int n0 = nb * T2_BN + c * 32;
        if (n0 >= p.N) break;
        uint32_t acc[32];
        tmem_ld_32x32b_x32(taddr + c * 32, acc);
        tmem_ld_wait();
        pol.store_chunk(p, b, mb128, nb, m, n0, acc);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(map_to_cta(smem_u32(&ss->tmem_empty[as]), 0));
      tile_done_hook(pol, p, b, mb128, nb);
    }
    pol.finish();
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();            // neither CTA frees TMEM / exits while the peer still uses the pair
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_2sm(tmem_base, TMEM_COLS);
  }
}

// Default policy: operands always ready, ordinary epilogue.  Tile order: m-fastest inside groups
// of GROUP_M pair-tile rows, n next, groups last — the A rows of a group (GROUP_M x 256 x K) stay in
// L2 while B streams past once per group, so neither operand is re-read from HBM when M is large.
struct Plain2 {
  static constexpr int GROUP_M = 8;
  __device__ __forceinline__ void map_tile(int t, int num_m, int num_n, int& b, int& mb, int& nb) const {
    const int per_batch = num_m * num_n;
    b = t / per_batch;
    const int r = t - b * per_batch;
    const int per_group = GROUP_M * num_n;
    const int g = r / per_group;
    const int m0 = g * GROUP_M;
    const int gm = num_m - m0 < GROUP_M ? num_m - m0 : GROUP_M;   // rows in this (maybe last) group
    const int rr = r - g * per_group;
    nb = rr / gm;
    mb = m0 + (rr - nb * gm);
  }
  __device__ __forceinline__ void wait_operands(int, int, int) const {}
  __device__ __forceinline__ const CUtensorMap* select_a(const CUtensorMap* def, int, int&) const { return def; }
  __device__ __forceinline__ void before_epilogue(int, int, int) const {}
  __device__ __forceinline__ void store_chunk(const Params& p, int b, int, int, int m, int n0,
                                              const uint32_t (&acc)[32]) const {
    epilogue_chunk(p, b, m, n0, acc);
  }
  __device__ __forceinline__ void tile_done(int, int, int) const {}
  __device__ __forceinline__ void finish() const {}
};

// Same as Plain2, but a bf16 C tile leaves through shared memory and the TMA store path.  The
// staging buffer is two 64-column blocks of [128 rows x 128 B] in the 128B-swizzle pattern
// (conflict-free 16 B writes).  Each epilogue warp works on its own 32 rows: two 32-column chunks
// fill a block, lane 0 issues one `cp.async.bulk.tensor` store for the warp's 32 x 64 box, and the
// block is reused two chunks later once that store has read it (`wait_group.read 1`).  L2 then sees
// whole-line writes instead of eight 16-byte partial-sector stores per row and chunk (ncu: half the
// write sectors, an eighth of the write requests); partial tiles are clipped by the tensor map; and
// 32 KB of staging leaves room for the full 6-stage operand ring.
constexpr int STAGE_BLOCK = HALF * 128;           // one 64-column block: 128 rows x 128 B = 16 KB
constexpr int STAGE_BYTES_2 = 2 * STAGE_BLOCK;    // 32 KB

struct Staged2 : Plain2 {
  uint8_t* staging;             // 1024-byte aligned
  const CUtensorMap* map_c;     // [batch, M, N] bf16, box 64 cols x 32 rows, SWIZZLE_128B
  __device__ __forceinline__ void issue_block(const Params& p, int b, int mb128, int n0_block, int blk) const {
    asm volatile("fence.proxy.async;" ::: "memory");   // generic-proxy staging writes -> TMA reads
    __syncwarp();
    if ((threadIdx.x & 31) == 0) {
      const int row0 = ((threadIdx.x >> 5) & 3) * 32;            // this warp's TMEM quarter
      const int m0 = mb128 * HALF + row0;
      if (m0 < p.M) tma_store_3d(map_c, staging + blk * STAGE_BLOCK + row0 * 128, n0_block, m0, b);
      tma_store_commit();
    }
  }
  __device__ __forceinline__ void before_epilogue(int, int, int) const {
    if ((threadIdx.x & 31) == 0) tma_store_wait_read<0>();  // previous tile's blocks have left the buffer
    __syncwarp();
  }
  __device__ __forceinline__ void store_chunk(const Params& p, int b, int mb128, int nb, int m, int n0,
                                              const uint32_t (&acc)[32]) const {
    const int c = (n0 - nb * T2_BN) >> 5;                       // 32-column chunk inside the tile
    const int blk = (c >> 1) & 1;
    if (c >= 4 && (c & 1) == 0) {
      // this block was shipped two chunks ago: only the newer store may still be reading
      if ((threadIdx.x & 31) == 0) tma_store_wait_read<1>();
      __syncwarp();
    }
    if (m < p.M) {   // rows past M: whatever is in the staging buffer is clipped by the store
      float v[32];
      epilogue_values(p, b, m, n0, acc, v);
      const int row = m - mb128 * HALF;
      uint8_t* base = staging + blk * STAGE_BLOCK + row * 128;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 u;
        u.x = pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]); u.y = pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]);
        u.z = pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]); u.w = pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7]);
        const int j = ((c & 1) * 4 + q) ^ (row & 7);            // 128B swizzle: 16 B chunk ^= row % 8
        *reinterpret_cast<uint4*>(base + j * 16) = u;
      }
    }
    // second half of a block (or the last chunk of a tile whose width is an odd number of chunks)
    if ((c & 1) == 1 || n0 + 32 >= p.N) issue_block(p, b, mb128, nb * T2_BN + (c >> 1) * 64, blk);
  }
  __device__ __forceinline__ void tile_done(int, int, int) const {}
  __device__ __forceinline__ void finish() const {
    if ((threadIdx.x & 31) == 0) tma_store_wait<0>();
  }
};

// fp32 C through the same staging buffer: one 32-column chunk (128 B of fp32 per row) is one
// staging block, shipped per warp as a 32 x 32 box with a plain tensor store or — for the wgrad
// `C +=` — a reducing tensor store (`cp.reduce.async.bulk.tensor ... .add`), so main_grad is never
// read by the SM and L2 sees whole-line updates.
struct StagedF32 : Plain2 {
  uint8_t* staging;             // 1024-byte aligned, 2 blocks of [128 rows x 128 B]
  const CUtensorMap* map_c;     // [batch, M, N] fp32, box 32 cols x 32 rows, SWIZZLE_128B
  __device__ __forceinline__ void before_epilogue(int, int, int) const {
    if ((threadIdx.x & 31) == 0) tma_store_wait_read<0>();
    __syncwarp();
  }
  __device__ __forceinline__ void store_chunk(const Params& p, int b, int mb128, int nb, int m, int n0,
                                              const uint32_t (&acc)[32]) const {
    const int c = (n0 - nb * T2_BN) >> 5;
    const int blk = c & 1;
    if (c >= 2) {   // this block was shipped two chunks ago: only the newer store may still be reading
      if ((threadIdx.x & 31) == 0) tma_store_wait_read<1>();
      __syncwarp();
    }
    if (m < p.M) {
      float v[32];
      epilogue_values(p, b, m, n0, acc, v);
      const int row = m - mb128 * HALF;
      uint8_t* base = staging + blk * STAGE_BLOCK + row * 128;
#pragma unroll
      for (int q = 0; q < 8; ++q)
        *reinterpret_cast<float4*>(base + ((q ^ (row & 7)) * 16)) = make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
    }
    asm volatile("fence.proxy.async;" ::: "memory");
    __syncwarp();
    if ((threadIdx.x & 31) == 0) {
      const int row0 = ((threadIdx.x >> 5) & 3) * 32;
      const int m0 = mb128 * HALF + row0;
      if (m0 < p.M) {
        if (p.flags & EPI_ACCUM) tma_reduce_add_3d(map_c, staging + blk * STAGE_BLOCK + row0 * 128, n0, m0, b);
        else tma_store_3d(map_c, staging + blk * STAGE_BLOCK + row0 * 128, n0, m0, b);
      }
      tma_store_commit();
    }
  }
  __device__ __forceinline__ void tile_done(int, int, int) const {}
  __device__ __forceinline__ void finish() const {
    if ((threadIdx.x & 31) == 0) tma_store_wait<0>();
  }
};

}  // namespace gemm2
}  // namespace b2f
