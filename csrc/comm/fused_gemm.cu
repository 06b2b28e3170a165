// b200flow K3 / K4 — tensor-parallel GEMMs fused with their sequence-parallel
// collective over NVLink peer memory (sm_100a).  Both reuse the tcgen05 GEMM core
// (kernels/gemm_core.cuh) through its Policy hooks.
//
// K3  all-gather -> GEMM      Y[T, N] = AG_rows(X_shard)[T, K] * B
//     One launch: the first `comm_ctas` CTAs pull the peers' row shards (and copy the
//     local one) into the gathered operand buffer 128-row block by block and publish
//     a per-block flag; the remaining CTAs run the persistent GEMM, local shard's
//     blocks first, and their TMA producer waits on the flag of the block it is
//     about to load.  The transfer hides under the math tile by tile, and the
//     gathered buffer is kept for the wgrad GEMM of the backward pass.
//
// K4  GEMM -> reduce-scatter  Y_shard[T/w, N] = sum_ranks(X * B)[own rows] (+bias, +residual)
//     Tiles whose rows belong to a peer are computed first and their bf16 partials
//     stored straight into that peer's receive buffer from the epilogue registers
//     (st.global on the mapped peer address) followed by a release flag; the rank's
//     own tiles come last and their epilogue waits for the peers' flags, adds the
//     received partials in fp32, then bias/residual, and writes the final shard.
//
// Cross-call reuse of the symmetric buffers is safe with two alternating slots
// (see DESIGN.md §fused-TP for the argument).
//
// Replaces NCCL all-gather + cuBLAS and cuBLAS + NCCL reduce-scatter of Megatron's
// sequence-parallel Column/RowParallelLinear.
#include <stdlib.h>

#include "../kernels/gemm_core.cuh"
#include "../kernels/gemm2_core.cuh"
#include "../common/host.h"
#include "../../include/b200flow.h"

using namespace b2f;
using namespace b2f::gemm;

namespace b2f {
namespace gemm {
int make_operand_map(CUtensorMap* map, const void* ptr, int mn, int k, long long ld, int batch,
                     long long batch_stride, int mn_major, int tile_mn);
int check_params(const Params& p);
}  // namespace gemm
}  // namespace b2f

namespace {

constexpr int MAX_WORLD = 8;

struct PeerPtrs {
  void* ptr[MAX_WORLD];
};

__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void wait_ge_gpu(const uint32_t* p, uint32_t target) {
  long long t0 = clock64();
  while (static_cast<int32_t>(ld_acquire_gpu(p) - target) < 0) {
    if (clock64() - t0 > B2F_SPIN_LIMIT_CYCLES) {
      printf("b200flow: local tile flag wait timed out (block %d)\n", blockIdx.x);
      __trap();
    }
  }
}
__device__ __forceinline__ uint4 ld_cv_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.cv.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}

// ------------------------------------------------------------------ K3
struct AGArgs {
  PeerPtrs x_shards;        // symmetric: each rank's [Ts, K] bf16 shard
  __nv_bfloat16* xg;        // local gathered [T, K]
  PeerPtrs ready;           // symmetric u32[world] "my shard is written" (slot base)
  uint32_t* tile_flags;     // local u32[num_m]
  int rank, world;
  int rows_per_rank, K;
  int comm_ctas;
  uint32_t epoch;
  int own_last;
  int tail_begin;           // CTA-pair kernel: tiles [tail_begin, end) of the tile order belong to the gather clusters,
                            // which join the GEMM once their share of the gather is done (< 0: they just exit)
};

struct AGPolicy {
  const uint32_t* tile_flags;
  uint32_t epoch;
  int rank, blocks_per_rank;
  __device__ __forceinline__ void map_tile(int t, int num_m, int num_n, int& b, int& mb,
                                           int& nb) const {
    // n fastest inside a row block, row blocks starting at this rank's own shard: the
    // first 1/world of the tiles needs no remote data, which is the time the pull has
    b = 0;
    const int mi = t / num_n;
    nb = t - mi * num_n;
    mb = (mi + rank * blocks_per_rank) % num_m;
  }
  __device__ __forceinline__ void wait_operands(int, int mb, int) const {
    wait_ge_gpu(tile_flags + mb, epoch);
    asm volatile("fence.proxy.async;" ::: "memory");  // generic-proxy copy -> TMA (async proxy) read
  }
  __device__ __forceinline__ void before_epilogue(int, int, int) const {}
  __device__ __forceinline__ void store_chunk(const Params& p, int b, int, int, int m, int n0,
                                              const uint32_t (&acc)[32]) const {
    epilogue_chunk(p, b, m, n0, acc);
  }
  __device__ __forceinline__ void tile_done(int, int, int) const {}
  __device__ __forceinline__ void finish() const {}
};

// Gather role: one thread per comm CTA drives the copy engine.  A 128-row block of a shard
// (128 x K bf16, contiguous) is moved in 32 KB chunks: bulk load (peer or local global -> this
// CTA's shared memory, completion on an mbarrier), then bulk store (shared -> local gathered
// buffer).  Six 32 KB slots, five loads in flight per CTA, so a handful of CTAs keeps enough
// bytes in flight to cover NVLink latency; no registers or LSU instructions touch the data.
constexpr int AG_CHUNK = 32 * 1024;
constexpr int AG_SLOTS = 6;
constexpr int AG_AHEAD = 5;

__device__ void ag_comm_role(const AGArgs& a, uint8_t* smem_raw) {
  if (threadIdx.x != 0) return;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + AG_SLOTS * AG_CHUNK);
  for (int i = 0; i < AG_SLOTS; ++i) mbar_init(&full[i], 1);
  fence_mbar_init();
  const int blocks_per_rank = a.rows_per_rank / BM;
  const int units = blocks_per_rank * a.world;
  const long long unit_bytes = (long long)BM * a.K * 2;
  const int chunks_per_unit = (int)((unit_bytes + AG_CHUNK - 1) / AG_CHUNK);
  if (blockIdx.x == 0) {
    // announce: this rank's shard for `epoch` is complete (stream order guarantees it)
    __threadfence_system();
    for (int r = 0; r < a.world; ++r)
      st_release_sys(static_cast<uint32_t*>(a.ready.ptr[r]) + a.rank, a.epoch);
  }
  // this CTA's chunk sequence: units blockIdx.x, +comm_ctas, ... each split into chunks
  int my_units = 0;
  for (int u = blockIdx.x; u < units; u += a.comm_ctas) ++my_units;
  const long long total = (long long)my_units * chunks_per_unit;
  auto locate = [&](long long c, const uint8_t*& src, uint8_t*& dst, uint32_t& bytes, int& unit_idx) {
    const int ui = (int)(c / chunks_per_unit), ci = (int)(c % chunks_per_unit);
    const int u = blockIdx.x + ui * a.comm_ctas;
    // ring order; `own_last` (CTA-pair kernel: the GEMM reads own rows from the shard) moves the
    // local copy behind the remote pulls
    const int srank = (a.rank + u / blocks_per_rank + (a.own_last ? 1 : 0)) % a.world;
    const int blk = u % blocks_per_rank;
    const long long off = (long long)ci * AG_CHUNK;
    const long long rem = unit_bytes - off;
    bytes = (uint32_t)(rem < AG_CHUNK ? rem : AG_CHUNK);
    src = static_cast<const uint8_t*>(a.x_shards.ptr[srank]) + (long long)blk * unit_bytes + off;
    dst = reinterpret_cast<uint8_t*>(a.xg) + ((long long)srank * blocks_per_rank + blk) * unit_bytes + off;
    unit_idx = srank * blocks_per_rank + blk;
    return srank;
  };
  int ready_seen = -1;  // ranks are visited in ring order, so one "last checked" value suffices
  auto issue_load = [&](long long c) {
    const uint8_t* src; uint8_t* dst; uint32_t bytes; int ui;
    const int srank = locate(c, src, dst, bytes, ui);
    if (srank != a.rank && srank != ready_seen) {
      wait_flag_ge(static_cast<const uint32_t*>(a.ready.ptr[a.rank]) + srank, a.epoch);
      ready_seen = srank;
    }
    const int slot = (int)(c % AG_SLOTS);
    if (c % chunks_per_unit == 0) race_delay();
    mbar_arrive_expect_tx(&full[slot], bytes);
    bulk_load_1d(smem + slot * AG_CHUNK, src, bytes, &full[slot]);
  };
  long long next_load = 0;
  for (; next_load < total && next_load < AG_AHEAD; ++next_load) issue_load(next_load);
  int pending_flag = -1;        // unit whose stores are all issued but not yet known complete
  int since_pending = 0;        // store groups committed after that unit's last one
  for (long long c = 0; c < total; ++c) {
    const uint8_t* src; uint8_t* dst; uint32_t bytes; int ui;
    locate(c, src, dst, bytes, ui);
    const int slot = (int)(c % AG_SLOTS);
    mbar_wait(&full[slot], (uint32_t)((c / AG_SLOTS) & 1));
    bulk_store_1d(dst, smem + slot * AG_CHUNK, bytes);
    tma_store_commit();
    if (pending_flag >= 0) ++since_pending;
    // slot of chunk c+AG_AHEAD was last read by the store of chunk c-1: allow only the newest
    // group (this one) to still be reading
    tma_store_wait_read<1>();
    if (next_load < total) issue_load(next_load++);
    if (pending_flag >= 0 && since_pending == 2) {
      tma_store_wait<2>();      // everything older than the two newest groups has landed
      asm volatile("fence.proxy.async;" ::: "memory");
      __threadfence();
      st_release_gpu(a.tile_flags + pending_flag, a.epoch);
      pending_flag = -1;
    }
    if ((c + 1) % chunks_per_unit == 0) {
      if (pending_flag >= 0 || chunks_per_unit < 3 || c + 1 == total) {
        // short units or the tail: drain and publish directly
        tma_store_wait<0>();
        asm volatile("fence.proxy.async;" ::: "memory");
        __threadfence();
        if (pending_flag >= 0) st_release_gpu(a.tile_flags + pending_flag, a.epoch);
        st_release_gpu(a.tile_flags + ui, a.epoch);
        pending_flag = -1;
      } else {
        pending_flag = ui;
        since_pending = 0;
      }
    }
  }
  tma_store_wait<0>();
  // this shared memory becomes the GEMM's operand ring and barrier block next
  for (int i = 0; i < AG_SLOTS; ++i)
    asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(&full[i])) : "memory");
}

__global__ void __launch_bounds__(NUM_THREADS, 1)
b2f_ag_gemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                   const Params p, const AGArgs a) {
  extern __shared__ uint8_t smem_raw[];
  if ((int)blockIdx.x < a.comm_ctas) {
    ag_comm_role(a, smem_raw);
    return;
  }
  AGPolicy pol{a.tile_flags, a.epoch, a.rank, a.rows_per_rank / BM};
  gemm_body(&tma_a, &tma_b, p, pol, (int)blockIdx.x - a.comm_ctas, (int)gridDim.x - a.comm_ctas,
            smem_raw);
}

// ---- K3 on the CTA-pair core: 256x256 pair tiles, each CTA of a pair waits for (and loads) its own
// 128-row block.  Rows of this rank's own shard are read straight from the shard (second tensor
// map), so the GEMM starts at once and the gather role copies the own shard last.
struct AG2Policy {
  const uint32_t* tile_flags;
  const CUtensorMap* own_map;   // tensor map over this rank's [Ts, K] shard
  uint32_t epoch;
  int rank, blocks_per_rank;    // 128-row blocks per rank
  __device__ __forceinline__ void map_tile(int t, int num_m, int num_n, int& b, int& mb, int& nb) const {
    b = 0;
    const int mi = t / num_n;
    nb = t - mi * num_n;
    mb = (mi + rank * (blocks_per_rank >> 1)) % num_m;   // pair-tile rows, own shard first
  }
  __device__ __forceinline__ void wait_operands(int, int mb128, int) const {
    if (mb128 / blocks_per_rank == rank) return;           // own rows come from the shard itself
    wait_ge_gpu(tile_flags + mb128, epoch);
    asm volatile("fence.proxy.async;" ::: "memory");
  }
  __device__ __forceinline__ const CUtensorMap* select_a(const CUtensorMap* def, int mb128, int& m0) const {
    if (mb128 / blocks_per_rank != rank) return def;
    m0 -= rank * blocks_per_rank * BM;
    return own_map;
  }
  // C leaves through the swizzled staging buffer and TMA tensor stores (see gemm2::Staged2)
  gemm2::Staged2 out;
  __device__ __forceinline__ void before_epilogue(int b, int mb128, int nb) const { out.before_epilogue(b, mb128, nb); }
  __device__ __forceinline__ void store_chunk(const Params& p, int b, int mb128, int nb, int m, int n0,
                                              const uint32_t (&acc)[32]) const {
    out.store_chunk(p, b, mb128, nb, m, n0, acc);
  }
  __device__ __forceinline__ void tile_done(int, int, int) const {}
  __device__ __forceinline__ void finish() const { out.finish(); }
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
b2f_ag_gemm2_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_own,
                    const __grid_constant__ CUtensorMap tma_b, const __grid_constant__ CUtensorMap tma_c,
                    const Params p, const AGArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  AG2Policy pol{a.tile_flags, &tma_own, a.epoch, a.rank, a.rows_per_rank / BM};
  pol.out.staging = smem + gemm2::S2 * gemm2::ST2_BYTES + 1024;
  pol.out.map_c = &tma_c;
  if ((int)blockIdx.x < a.comm_ctas) {   // comm_ctas is even: whole clusters
    // gather role first (one thread per CTA drives the copies; the others wait for it at the GEMM's first barrier),
    // then the cluster takes its slice of the tail of the tile order: no SM sits idle once the gather is done
    ag_comm_role(a, smem_raw);
    if (a.tail_begin < 0) return;
    gemm2::gemm2_body<AG2Policy, gemm2::S2>(&tma_a, &tma_b, p, pol, (int)blockIdx.x >> 1, a.comm_ctas >> 1, smem_raw,
                                            a.tail_begin, -1);
    return;
  }
  const int num_tiles = ((p.M + 255) / 256) * ((p.N + 255) / 256);
  gemm2::gemm2_body<AG2Policy, gemm2::S2>(&tma_a, &tma_b, p, pol, ((int)blockIdx.x - a.comm_ctas) >> 1,
                                          ((int)gridDim.x - a.comm_ctas) >> 1, smem_raw, 0,
                                          a.tail_begin < 0 ? num_tiles : a.tail_begin);
}

// ------------------------------------------------------------------ K4
struct RSArgs {
  PeerPtrs recv;            // symmetric: each rank's receive buffer [world][Ts, N] bf16
  PeerPtrs flags;           // symmetric u32 [world][tiles_per_shard]
  int rank, world;
  int rows_per_rank;        // Ts
  int N;
  uint32_t epoch;
  int dbg;                  // timing experiments only (B2F_RS_DEBUG): 1 = no sends, no receives; 2 = sends only
  int interleave;           // tile order: 1 = rounds of (one tile per peer, then the own tile with the same index),
                            // flags published as soon as a tile's rows have landed; 0 = all peer tiles first
  // K4' (GEMM -> all-reduce): finished rows also go to every peer's full output, then one
  // "all my rows have landed" flag per peer closes the call
  int allreduce;
  PeerPtrs bcast;           // symmetric: each rank's full output [T, N] bf16, row pitch ld_bcast
  long long ld_bcast;
  PeerPtrs done;            // symmetric u32[world]
  uint32_t* cta_counter;    // local u32, zero between calls
};

// Partial tiles bound for a peer are staged in shared memory as bf16 rows (528 B pitch: 512 B of
// data + 16 B so the one-row-per-thread 16 B writes spread over the banks) and leave as one bulk
// copy per row, i.e. 512 B NVLink writes instead of scattered 16 B stores.
constexpr int RS_STAGES = 3;
constexpr int RS_PITCH = BN * 2 + 16;
constexpr int RS_SMEM_BYTES = RS_STAGES * STAGE_BYTES + 1024 /*barriers*/ + BM * RS_PITCH + 16 /*rx barrier*/ + 1024 /*align*/;

struct RSPolicy {
  RSArgs a;
  int blocks_per_rank, num_n;
  uint8_t* staging;
  __nv_bfloat16* out;   // local output shard [Ts, N] and its row pitch (elements)
  long long ldo;
  int pitch;            // bytes between staged rows (the CTA-pair kernel trades the bank-spreading pad for a ring stage)
  // the flag of a shipped tile is published one tile later (its bulk copies have long landed by
  // then), so the epilogue never sits out an NVLink round trip
  mutable int pend_owner = -1, pend_mb_local = 0, pend_nb = 0;
  mutable uint32_t rx_phase = 0;
  __device__ __forceinline__ uint64_t* rx_bar() const {
    return reinterpret_cast<uint64_t*>(staging + BM * pitch);
  }
  // kernel prologue (one thread, before the CTA-wide sync): barrier for the partial-row prefetch
  __device__ __forceinline__ void init_smem() const { mbar_init(rx_bar(), 128); fence_mbar_init(); }
  __device__ __forceinline__ void flush() const {
    if (pend_owner < 0) return;
    tma_store_wait<0>();  // this thread's row of the pending tile is in the peer's memory
    asm volatile("fence.proxy.async;" ::: "memory");
    asm volatile("bar.sync 1, 128;" ::: "memory");  // ... and so is everyone else's
    if (threadIdx.x == 64) {
      __threadfence_system();
      st_release_sys(flag_at(pend_owner, a.rank, pend_mb_local, pend_nb), a.epoch);
    }
    pend_owner = -1;
  }
  __device__ __forceinline__ void finish() const {
    flush();
    tma_store_wait<0>();
    if (!a.allreduce) return;
    // K4': this CTA's rows are in every peer's output.  The last CTA of the grid to get here tells the
    // peers and then holds the kernel open until every peer has said the same, so that whatever follows
    // on this stream reads a complete matrix.
    asm volatile("fence.proxy.async;" ::: "memory");
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (threadIdx.x == 64) {
      __threadfence_system();
      const unsigned arrived = atomicAdd(a.cta_counter, 1u);
      if (arrived == gridDim.x - 1) {
        atomicExch(a.cta_counter, 0u);
        __threadfence_system();
        for (int s = 0; s < a.world; ++s)
          if (s != a.rank) st_release_sys(static_cast<uint32_t*>(a.done.ptr[s]) + a.rank, a.epoch);
        for (int s = 0; s < a.world; ++s)
          if (s != a.rank) wait_flag_ge(static_cast<uint32_t*>(a.done.ptr[a.rank]) + s, a.epoch);
      }
    }
  }
  __device__ __forceinline__ int owner_of(int mb) const { return mb / blocks_per_rank; }
  __device__ __forceinline__ void map_tile(int t, int num_m, int nn, int& b, int& mb, int& nb) const {
    b = 0;
    const int per_owner = blocks_per_rank * nn;
    int owner, r;
    if (a.interleave) {
      // round k: tile k of every peer (ring order), then own tile k.  The sends spread over the whole kernel instead
      // of its first (world-1)/world, and an own tile's partials were produced by the peers in the same round.
      const int k = t / a.world, s = t - k * a.world;
      owner = (a.rank + 1 + s) % a.world;
      r = k;
    } else {
      // peers' row blocks first (ring order after this rank), own blocks last
      const int oi = t / per_owner;
      owner = (a.rank + 1 + oi) % a.world;
      r = t - oi * per_owner;
    }
    nb = r / blocks_per_rank;
    mb = owner * blocks_per_rank + (r - nb * blocks_per_rank);
  }
  __device__ __forceinline__ void wait_operands(int, int, int) const {}
  __device__ __forceinline__ uint32_t* flag_at(int on_rank, int src, int mb_local, int nb) const {
    return static_cast<uint32_t*>(a.flags.ptr[on_rank]) +
           ((size_t)src * blocks_per_rank + mb_local) * num_n + nb;
  }
  __device__ __forceinline__ void before_epilogue(int, int mb, int nb) const {
    if ((threadIdx.x & 31) == 0) race_delay();
    __syncwarp();
    tma_store_wait_read<0>();    // the staging rows of the previous tile have been read out
    if (owner_of(mb) != a.rank) return;
    flush();  // publish before waiting on the peers, or two ranks could wait on each other
    if (a.dbg) return;
    const int mb_local = mb - a.rank * blocks_per_rank;
    if ((threadIdx.x & 31) == 0) {
      for (int s = 0; s < a.world; ++s)
        if (s != a.rank) wait_flag_ge(flag_at(a.rank, s, mb_local, nb), a.epoch);
    }
    __syncwarp();
    if (a.world == 2) {
      // one bulk copy per row pulls the peer's 512 B partial row into the (idle) staging buffer
      const int r = (int)threadIdx.x - 64;
      const int row_in_tile = (((r >> 5) + 2) & 3) * 32 + (r & 31);
      const int n0 = nb * BN;
      const int cols = a.N - n0 < BN ? a.N - n0 : BN;
      const __nv_bfloat16* src = static_cast<const __nv_bfloat16*>(a.recv.ptr[a.rank]) +
                                 ((size_t)(a.rank ^ 1) * a.rows_per_rank + (size_t)mb_local * BM + row_in_tile) * a.N + n0;
      asm volatile("fence.proxy.async;" ::: "memory");
      mbar_arrive_expect_tx(rx_bar(), (uint32_t)cols * 2);
      bulk_load_1d(staging + (size_t)row_in_tile * pitch, src, (uint32_t)cols * 2, rx_bar());
      mbar_wait(rx_bar(), rx_phase);
      rx_phase ^= 1;
    }
  }
  __device__ __forceinline__ void store_chunk(const Params& p, int b, int mb, int nb, int m, int n0,
                                              const uint32_t (&acc)[32]) const {
    const int owner = owner_of(mb);
    const int row_local = m - owner * a.rows_per_rank;
    if (m >= p.M || n0 >= p.N) return;
    if (owner != a.rank) {
      // partial tile -> bf16 row in the staging buffer; tile_done ships the rows
      uint8_t* dst = staging + (size_t)(m - mb * BM) * pitch + (size_t)(n0 - nb * BN) * 2;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 u;
        u.x = pack_bf16x2(__uint_as_float(acc[q * 8 + 0]), __uint_as_float(acc[q * 8 + 1]));
        u.y = pack_bf16x2(__uint_as_float(acc[q * 8 + 2]), __uint_as_float(acc[q * 8 + 3]));
        u.z = pack_bf16x2(__uint_as_float(acc[q * 8 + 4]), __uint_as_float(acc[q * 8 + 5]));
        u.w = pack_bf16x2(__uint_as_float(acc[q * 8 + 6]), __uint_as_float(acc[q * 8 + 7]));
        *reinterpret_cast<uint4*>(dst + q * 16) = u;
      }
      return;
    }
    // own rows: add every peer's partial, then the ordinary epilogue on the local row index
    uint32_t sum[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) sum[i] = acc[i];
    for (int s = 0; s < a.world; ++s) {
      if (s == a.rank || a.dbg) continue;
      const __nv_bfloat16* src = static_cast<const __nv_bfloat16*>(a.recv.ptr[a.rank]) +
                                 ((size_t)s * a.rows_per_rank + row_local) * a.N + n0;
      const uint8_t* staged = staging + (size_t)(m - mb * BM) * pitch + (size_t)(n0 - nb * BN) * 2;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint4 u = a.world == 2 ? *reinterpret_cast<const uint4*>(staged + q * 16) : ld_cv_v4(src + q * 8);
        const float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y), f2 = unpack_bf16x2(u.z),
                     f3 = unpack_bf16x2(u.w);
        sum[q * 8 + 0] = __float_as_uint(__uint_as_float(sum[q * 8 + 0]) + f0.x);
        sum[q * 8 + 1] = __float_as_uint(__uint_as_float(sum[q * 8 + 1]) + f0.y);
        sum[q * 8 + 2] = __float_as_uint(__uint_as_float(sum[q * 8 + 2]) + f1.x);
        sum[q * 8 + 3] = __float_as_uint(__uint_as_float(sum[q * 8 + 3]) + f1.y);
        sum[q * 8 + 4] = __float_as_uint(__uint_as_float(sum[q * 8 + 4]) + f2.x);
        sum[q * 8 + 5] = __float_as_uint(__uint_as_float(sum[q * 8 + 5]) + f2.y);
        sum[q * 8 + 6] = __float_as_uint(__uint_as_float(sum[q * 8 + 6]) + f3.x);
        sum[q * 8 + 7] = __float_as_uint(__uint_as_float(sum[q * 8 + 7]) + f3.y);
      }
    }
    // Params describe the *shard* output: C/residual are [Ts, N], indexed by the local row.  The
    // finished bf16 row chunk goes back into this thread's staging row (over the partial it just
    // consumed) and leaves with one bulk copy per row in tile_done.
    Params q = p;
    q.M = a.rows_per_rank;
    float v[32];
    epilogue_values(q, b, row_local, n0, sum, v);
    uint8_t* dst = staging + (size_t)(m - mb * BM) * pitch + (size_t)(n0 - nb * BN) * 2;
#pragma unroll
    for (int qq = 0; qq < 4; ++qq) {
      uint4 u;
      u.x = pack_bf16x2(v[qq * 8 + 0], v[qq * 8 + 1]); u.y = pack_bf16x2(v[qq * 8 + 2], v[qq * 8 + 3]);
      u.z = pack_bf16x2(v[qq * 8 + 4], v[qq * 8 + 5]); u.w = pack_bf16x2(v[qq * 8 + 6], v[qq * 8 + 7]);
      *reinterpret_cast<uint4*>(dst + qq * 16) = u;
    }
  }
  __device__ __forceinline__ void tile_done(int, int mb, int nb) const {
    const int owner = owner_of(mb);
    if (owner == a.rank) {   // (before_epilogue already flushed) finished rows -> the local output shard
      const int r = (int)threadIdx.x - 64;
      const int row_in_tile = (((r >> 5) + 2) & 3) * 32 + (r & 31);
      const int row_local = mb * BM + row_in_tile - owner * a.rows_per_rank;
      const int n0 = nb * BN;
      const int cols = a.N - n0 < BN ? a.N - n0 : BN;
      asm volatile("fence.proxy.async;" ::: "memory");
      bulk_store_1d(out + (size_t)row_local * ldo + n0, staging + (size_t)row_in_tile * pitch, (uint32_t)cols * 2);
      if (a.allreduce) {   // K4': the same finished row into every peer's full output
        const size_t row_global = (size_t)mb * BM + row_in_tile;
        for (int s = 0; s < a.world; ++s)
          if (s != a.rank)
            bulk_store_1d(static_cast<__nv_bfloat16*>(a.bcast.ptr[s]) + row_global * a.ld_bcast + n0,
                          staging + (size_t)row_in_tile * pitch, (uint32_t)cols * 2);
      }
      tma_store_commit();
      return;
    }
    // this thread's row: staging -> owner's receive slot for this source rank, one bulk copy
    const int r = (int)threadIdx.x - 64;  // epilogue threads are 64..191, TMEM quarter = warp & 3
    const int row_in_tile = (((r >> 5) + 2) & 3) * 32 + (r & 31);
    const int row_local = mb * BM + row_in_tile - owner * a.rows_per_rank;
    const int n0 = nb * BN;
    const int cols = a.N - n0 < BN ? a.N - n0 : BN;
    __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(a.recv.ptr[owner]) +
                         ((size_t)a.rank * a.rows_per_rank + row_local) * a.N + n0;
    flush();
    if (a.dbg == 1) return;
    asm volatile("fence.proxy.async;" ::: "memory");  // generic-proxy staging writes -> bulk copy reads
    bulk_store_1d(dst, staging + (size_t)row_in_tile * pitch, (uint32_t)cols * 2);
    tma_store_commit();
    pend_owner = owner;
    pend_mb_local = mb - owner * blocks_per_rank;
    pend_nb = nb;
    // interleaved order: the owner is about to need this tile, so sit out the NVLink round trip here (the MMA warp
    // is already on the next tile) instead of publishing one tile later
    if (a.interleave) flush();
  }
};

__global__ void __launch_bounds__(NUM_THREADS, 1)
b2f_gemm_rs_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                   const Params p, const RSArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  RSPolicy pol{a, a.rows_per_rank / BM, (p.N + BN - 1) / BN, smem + RS_STAGES * STAGE_BYTES + 1024,
               reinterpret_cast<__nv_bfloat16*>(p.C), p.ldc, RS_PITCH};
  if (threadIdx.x == 0) pol.init_smem();
  gemm_body<RSPolicy, 0, RS_STAGES>(&tma_a, &tma_b, p, pol, blockIdx.x, gridDim.x, smem_raw);
}

// ---- K4 on the CTA-pair core: same protocol, each CTA of a pair handles its own 128-row block.
// Five ring stages (the plain kernel has six; with four the main loop ran ~10 % slower in single-GPU loopback,
// profiles/r2_call6_k3_k4_loopback.txt) fit next to the 128-row staging block once its rows are packed at 512 B: the
// one-row-per-thread 16 B staging writes then collide on the banks, ~2 us per tile inside an epilogue that overlaps a
// 15+ us main loop.
constexpr int RS2_STAGES = 5;
constexpr int RS2_PITCH = BN * 2;
constexpr int RS2_SMEM_BYTES = gemm2::smem2_bytes(RS2_STAGES, BM * RS2_PITCH + 16);
static_assert(RS2_SMEM_BYTES <= 227 * 1024, "K4 CTA-pair kernel: shared memory budget");

struct RS2Policy {
  RSPolicy in;   // 128-row-block logic is identical; only the tile order differs
  __device__ __forceinline__ void map_tile(int t, int num_m, int nn, int& b, int& mb, int& nb) const {
    b = 0;
    const int pairs_per_rank = in.blocks_per_rank >> 1;
    const int per_owner = pairs_per_rank * nn;
    int owner, r;
    if (in.a.interleave) {
      const int k = t / in.a.world, s = t - k * in.a.world;
      owner = (in.a.rank + 1 + s) % in.a.world;
      r = k;
    } else {
      const int oi = t / per_owner;
      owner = (in.a.rank + 1 + oi) % in.a.world;
      r = t - oi * per_owner;
    }
    nb = r / pairs_per_rank;
    mb = owner * pairs_per_rank + (r - nb * pairs_per_rank);
  }
  __device__ __forceinline__ void wait_operands(int, int, int) const {}
  __device__ __forceinline__ const CUtensorMap* select_a(const CUtensorMap* def, int, int&) const { return def; }
  __device__ __forceinline__ void before_epilogue(int b, int mb128, int nb) const { in.before_epilogue(b, mb128, nb); }
  __device__ __forceinline__ void store_chunk(const Params& p, int b, int mb128, int nb, int m, int n0,
                                              const uint32_t (&acc)[32]) const {
    in.store_chunk(p, b, mb128, nb, m, n0, acc);
  }
  __device__ __forceinline__ void tile_done(int b, int mb128, int nb) const { in.tile_done(b, mb128, nb); }
  __device__ __forceinline__ void finish() const { in.finish(); }
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
b2f_gemm_rs2_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                    const Params p, const RSArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  RS2Policy pol{RSPolicy{a, a.rows_per_rank / BM, (p.N + BN - 1) / BN,
                         smem + RS2_STAGES * gemm2::ST2_BYTES + 1024, reinterpret_cast<__nv_bfloat16*>(p.C), p.ldc,
                         RS2_PITCH}};
  if (threadIdx.x == 0) pol.in.init_smem();
  gemm2::gemm2_body<RS2Policy, RS2_STAGES>(&tma_a, &tma_b, p, pol, blockIdx.x >> 1, gridDim.x >> 1, smem_raw);
}

int fill(PeerPtrs* t, const long long* ptrs, int world) {
  B2F_REQUIRE(world >= 2 && world <= MAX_WORLD, "fused TP GEMM: group size must be 2..8");
  for (int i = 0; i < MAX_WORLD; ++i) t->ptr[i] = i < world ? reinterpret_cast<void*>(ptrs[i]) : nullptr;
  return 0;
}

template <typename Kern>
int ensure_smem(Kern k) {
  B2F_CHECK_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
  return 0;
}

}  // namespace

// Y[T, N] = AG(X)[T, K] * B (+bias).  x_shard_ptrs: each rank's [T/world, K] bf16 shard in
// symmetric memory; xg: local [T, K] gathered output (kept by the caller for wgrad);
// ready_ptrs: symmetric u32[world]; tile_flags: local u32[T/128].
extern "C" int b2f_ag_gemm_bf16(const long long* x_shard_ptrs, void* xg, const void* B, void* Y,
                                int T, int N, int K, long long ldb, long long ldy, int b_mn_major,
                                const void* bias, const long long* ready_ptrs, void* tile_flags,
                                int rank, int world, int comm_ctas, unsigned epoch, void* stream) {
  B2F_REQUIRE(T % (world * BM) == 0, "ag_gemm: rows per rank must be a multiple of 128");
  B2F_REQUIRE(K % 8 == 0, "ag_gemm: K must be a multiple of 8");
  Params p{};
  p.M = T; p.N = N; p.K = K; p.batch = 1;
  p.a_mn_major = 0; p.b_mn_major = b_mn_major;
  p.C = Y; p.ldc = ldy; p.bias = static_cast<const __nv_bfloat16*>(bias);
  p.alpha = 1.f; p.flags = 0;
  if (int rc = check_params(p)) return rc;
  AGArgs a{};
  if (int rc = fill(&a.x_shards, x_shard_ptrs, world)) return rc;
  if (int rc = fill(&a.ready, ready_ptrs, world)) return rc;
  a.xg = static_cast<__nv_bfloat16*>(xg);
  a.tile_flags = static_cast<uint32_t*>(tile_flags);
  a.rank = rank; a.world = world; a.rows_per_rank = T / world; a.K = K; a.epoch = epoch;
  const int sms = sm_count();
  if (comm_ctas < 1) comm_ctas = 8;
  if (comm_ctas > sms / 2) comm_ctas = sms / 2;
  a.comm_ctas = comm_ctas;
  static const bool force_1cta = [] { const char* e = getenv("B2F_FUSED_TP_CORE"); return e && e[0] == '1'; }();
  if (!force_1cta && (T / world) % 256 == 0 && N >= 256 && N % 8 == 0 && ldy % 8 == 0) {
    // CTA-pair core: per-CTA boxes of 128 rows of A / 128 rows of B
    a.own_last = 1;
    if (a.comm_ctas & 1) ++a.comm_ctas;
    CUtensorMap ma2, mown, mb2;
    if (int rc = make_operand_map(&ma2, xg, T, K, K, 1, 0, 0, BM)) return rc;
    if (int rc = make_operand_map(&mown, reinterpret_cast<const void*>(x_shard_ptrs[rank]), T / world, K, K, 1, 0, 0, BM)) return rc;
    if (int rc = make_operand_map(&mb2, B, N, K, ldb, 1, 0, b_mn_major, BM)) return rc;
    CUtensorMap mc;
    if (int rc = make_tma_3d(&mc, Y, TmaDtype::BF16, 1, T, N, (long long)T * ldy, ldy, 32, 64, 128)) return rc;
    constexpr int AG2_SMEM = gemm2::smem2_bytes(gemm2::S2, gemm2::STAGE_BYTES_2);
    B2F_CHECK_CUDA(cudaFuncSetAttribute(b2f_ag_gemm2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AG2_SMEM));
    const int tiles2 = (T / 256) * ((N + 255) / 256);
    int pairs = (sms - a.comm_ctas) / 2;
    if (pairs > tiles2) pairs = tiles2;
    // The gather clusters rejoin the GEMM when their copies are done.  Their slice of the tile order is sized so that
    // everyone finishes together: a cluster pair of SMs computes at ~1/74 of the sustained GEMM rate, a gather CTA
    // moves ~15 GB/s (peer -> shared -> local), measured on this box (profiles/r2_tp_perf*.txt).
    static const double cta_gbs = [] { const char* e = getenv("B2F_AG_CTA_GBS"); return e ? atof(e) : 15.0; }();
    static const bool rejoin = [] { const char* e = getenv("B2F_AG_REJOIN"); return !(e && e[0] == '0'); }();
    a.tail_begin = -1;
    if (rejoin && pairs > 0 && tiles2 > 2 * pairs) {
      const double t_gemm = 2.0 * T * (double)N * K / 1.30e15;                  // whole GEMM on every SM, seconds
      const double gather_bytes = (double)T * K * 2;                            // peers' shards + own copy
      const double t_gather = gather_bytes / (a.comm_ctas * cta_gbs * 1e9);     // time the gather clusters are busy
      const int nc = a.comm_ctas / 2;
      const double all = nc + pairs;
      // work (in cluster-seconds) = t_gemm * all; gather clusters contribute (T_end - t_gather) each
      const double t_end = (t_gemm * all + nc * t_gather) / all;
      double share = t_end > t_gather ? nc * (t_end - t_gather) / (t_gemm * all) : 0.0;
      int tail = (int)(share * tiles2);
      if (tail > 0) a.tail_begin = tiles2 - tail;
    }
    b2f_ag_gemm2_kernel<<<a.comm_ctas + 2 * pairs, NUM_THREADS, AG2_SMEM, static_cast<cudaStream_t>(stream)>>>(
        ma2, mown, mb2, mc, p, a);
    B2F_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  CUtensorMap ma, mb;
  if (int rc = make_operand_map(&ma, xg, T, K, K, 1, 0, 0, BM)) return rc;
  if (int rc = make_operand_map(&mb, B, N, K, ldb, 1, 0, b_mn_major, BN)) return rc;
  if (int rc = ensure_smem(b2f_ag_gemm_kernel)) return rc;
  const int tiles = (T / BM) * ((N + BN - 1) / BN);
  int gemm_ctas = sms - comm_ctas;
  if (gemm_ctas > tiles) gemm_ctas = tiles;
  b2f_ag_gemm_kernel<<<comm_ctas + gemm_ctas, NUM_THREADS, SMEM_BYTES, static_cast<cudaStream_t>(stream)>>>(
      ma, mb, p, a);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Y_shard[T/world, N] = RS(X[T, K] * B) (+bias) (+residual_shard).
// recv_ptrs: each rank's symmetric receive buffer [world][T/world, N] bf16;
// flag_ptrs: symmetric u32 [world][(T/world/128) * ceil(N/256)].
struct ARExtra {
  const long long* out_ptrs;   // each rank's full output [T, N]
  const long long* done_ptrs;
  void* cta_counter;
};

static int gemm_rs_launch(const void* X, const void* B, void* Y_shard, int T, int N, int K,
                          long long ldx, long long ldb, long long ldy, int b_mn_major,
                          const void* bias, const void* residual, long long ldr,
                          const long long* recv_ptrs, const long long* flag_ptrs, int rank,
                          int world, unsigned epoch, void* stream, const ARExtra* ar) {
  B2F_REQUIRE(T % (world * BM) == 0, "gemm_rs: rows per rank must be a multiple of 128");
  B2F_REQUIRE(N % 32 == 0, "gemm_rs: N must be a multiple of 32");
  B2F_REQUIRE(ldy % 8 == 0 && (reinterpret_cast<uintptr_t>(Y_shard) & 15) == 0, "gemm_rs: output rows must be 16-byte aligned");
  Params p{};
  p.M = T; p.N = N; p.K = K; p.batch = 1;
  p.a_mn_major = 0; p.b_mn_major = b_mn_major;
  p.C = Y_shard; p.ldc = ldy;
  p.bias = static_cast<const __nv_bfloat16*>(bias);
  p.residual = static_cast<const __nv_bfloat16*>(residual);
  p.ldr = ldr;
  p.alpha = 1.f; p.flags = 0;
  if (int rc = check_params(p)) return rc;
  RSArgs a{};
  if (int rc = fill(&a.recv, recv_ptrs, world)) return rc;
  if (int rc = fill(&a.flags, flag_ptrs, world)) return rc;
  a.rank = rank; a.world = world; a.rows_per_rank = T / world; a.N = N; a.epoch = epoch;
  if (const char* e = getenv("B2F_RS_DEBUG")) a.dbg = atoi(e);
  // measured at the Llama TP2 shapes (profiles/r2_call7_*): the interleaved order is 0-7 % slower than "peers first",
  // so it stays an experiment switch
  static const int interleave = [] { const char* e = getenv("B2F_RS_INTERLEAVE"); return e ? atoi(e) : 0; }();
  a.interleave = interleave;
  if (ar) {
    a.allreduce = 1;
    a.dbg = 0;
    if (int rc = fill(&a.bcast, ar->out_ptrs, world)) return rc;
    if (int rc = fill(&a.done, ar->done_ptrs, world)) return rc;
    a.ld_bcast = ldy;
    a.cta_counter = static_cast<uint32_t*>(ar->cta_counter);
  }
  static const bool force_1cta = [] { const char* e = getenv("B2F_FUSED_TP_CORE"); return e && e[0] == '1'; }();
  if (!force_1cta && (T / world) % 256 == 0 && N >= 256) {
    CUtensorMap ma2, mb2;
    if (int rc = make_operand_map(&ma2, X, T, K, ldx, 1, 0, 0, BM)) return rc;
    if (int rc = make_operand_map(&mb2, B, N, K, ldb, 1, 0, b_mn_major, BM)) return rc;
    B2F_CHECK_CUDA(cudaFuncSetAttribute(b2f_gemm_rs2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        RS2_SMEM_BYTES));
    const int tiles2 = (T / 256) * ((N + 255) / 256);
    int pairs = sm_count() / 2;
    if (pairs > tiles2) pairs = tiles2;
    b2f_gemm_rs2_kernel<<<2 * pairs, NUM_THREADS, RS2_SMEM_BYTES, static_cast<cudaStream_t>(stream)>>>(ma2, mb2, p, a);
    B2F_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  CUtensorMap ma, mb;
  if (int rc = make_operand_map(&ma, X, T, K, ldx, 1, 0, 0, BM)) return rc;
  if (int rc = make_operand_map(&mb, B, N, K, ldb, 1, 0, b_mn_major, BN)) return rc;
  B2F_CHECK_CUDA(cudaFuncSetAttribute(b2f_gemm_rs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      RS_SMEM_BYTES));
  const int tiles = (T / BM) * ((N + BN - 1) / BN);
  const int sms = sm_count();
  b2f_gemm_rs_kernel<<<tiles < sms ? tiles : sms, NUM_THREADS, RS_SMEM_BYTES, static_cast<cudaStream_t>(stream)>>>(
      ma, mb, p, a);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_gemm_rs_bf16(const void* X, const void* B, void* Y_shard, int T, int N, int K,
                                long long ldx, long long ldb, long long ldy, int b_mn_major,
                                const void* bias, const void* residual, long long ldr,
                                const long long* recv_ptrs, const long long* flag_ptrs, int rank,
                                int world, unsigned epoch, void* stream) {
  return gemm_rs_launch(X, B, Y_shard, T, N, K, ldx, ldb, ldy, b_mn_major, bias, residual, ldr, recv_ptrs,
                        flag_ptrs, rank, world, epoch, stream, nullptr);
}

// K4': Y[T, N] = all-reduce(X[T, K] * B) (+bias) (+residual) on every rank — Megatron's row-parallel
// linear (and the dgrad of a column-parallel one) without sequence parallelism.
// The reduce-scatter half is K4 unchanged; a rank's finished rows then go into its own output and,
// with one 512-byte bulk copy per row and peer, into every peer's, and the kernel ends only when every
// peer has reported that all of its rows have landed.
// out_ptrs: each rank's [T, N] bf16 output in symmetric memory (row pitch ldy); residual: this rank's
// own rows [T/world, N]; done_ptrs: symmetric u32[world]; cta_counter: local u32, zero at first use.
extern "C" int b2f_gemm_ar_bf16(const void* X, const void* B, const long long* out_ptrs, int T, int N, int K,
                                long long ldx, long long ldb, long long ldy, int b_mn_major,
                                const void* bias, const void* residual, long long ldr,
                                const long long* recv_ptrs, const long long* flag_ptrs,
                                const long long* done_ptrs, void* cta_counter, int rank, int world,
                                unsigned epoch, void* stream) {
  B2F_REQUIRE(out_ptrs && done_ptrs && cta_counter, "gemm_ar: output, done-flag and counter buffers required");
  B2F_REQUIRE(world >= 2 && world <= MAX_WORLD && rank >= 0 && rank < world, "gemm_ar: bad rank / group size");
  B2F_REQUIRE(T % world == 0, "gemm_ar: rows must divide by the group size");
  ARExtra ar{out_ptrs, done_ptrs, cta_counter};
  void* own_rows = reinterpret_cast<__nv_bfloat16*>(out_ptrs[rank]) + (size_t)rank * (T / world) * ldy;
  return gemm_rs_launch(X, B, own_rows, T, N, K, ldx, ldb, ldy, b_mn_major, bias, residual, ldr, recv_ptrs,
                        flag_ptrs, rank, world, epoch, stream, &ar);
}
