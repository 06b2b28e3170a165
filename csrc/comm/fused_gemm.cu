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
#include "../kernels/gemm_core.cuh"
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
};

__device__ void ag_comm_role(const AGArgs& a) {
  const int blocks_per_rank = a.rows_per_rank / BM;
  const int units = blocks_per_rank * a.world;
  const long long row_vecs = (long long)a.K / 8;          // uint4 per row
  const long long unit_vecs = row_vecs * BM;
  if (blockIdx.x == 0 && (int)threadIdx.x < a.world) {
    // announce: this rank's shard for `epoch` is complete (stream order guarantees it)
    __threadfence_system();
    st_release_sys(static_cast<uint32_t*>(a.ready.ptr[threadIdx.x]) + a.rank, a.epoch);
  }
  for (int u = blockIdx.x; u < units; u += a.comm_ctas) {
    // own shard first, then peers in ring order
    const int src = (a.rank + u / blocks_per_rank) % a.world;
    const int blk = u % blocks_per_rank;
    if (src != a.rank) {
      if (threadIdx.x == 0)
        wait_flag_ge(static_cast<const uint32_t*>(a.ready.ptr[a.rank]) + src, a.epoch);
      __syncthreads();
    }
    const uint4* s = reinterpret_cast<const uint4*>(
        static_cast<const __nv_bfloat16*>(a.x_shards.ptr[src]) + (long long)blk * BM * a.K);
    uint4* d = reinterpret_cast<uint4*>(a.xg + ((long long)src * a.rows_per_rank + (long long)blk * BM) * a.K);
    long long i = threadIdx.x;
    const long long step = blockDim.x;
    for (; i + 7 * step < unit_vecs; i += 8 * step) {   // 8 independent 16 B loads in flight
      uint4 v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = ld_v4(s + i + j * step);
#pragma unroll
      for (int j = 0; j < 8; ++j) d[i + j * step] = v[j];
    }
    for (; i < unit_vecs; i += step) d[i] = ld_v4(s + i);
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      asm volatile("fence.proxy.async;" ::: "memory");
      st_release_gpu(a.tile_flags + src * blocks_per_rank + blk, a.epoch);
    }
  }
}

__global__ void __launch_bounds__(NUM_THREADS, 1)
b2f_ag_gemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                   const Params p, const AGArgs a) {
  extern __shared__ uint8_t smem_raw[];
  if ((int)blockIdx.x < a.comm_ctas) {
    ag_comm_role(a);
    return;
  }
  AGPolicy pol{a.tile_flags, a.epoch, a.rank, a.rows_per_rank / BM};
  gemm_body(&tma_a, &tma_b, p, pol, (int)blockIdx.x - a.comm_ctas, (int)gridDim.x - a.comm_ctas,
            smem_raw);
}

// ------------------------------------------------------------------ K4
struct RSArgs {
  PeerPtrs recv;            // symmetric: each rank's receive buffer [world][Ts, N] bf16
  PeerPtrs flags;           // symmetric u32 [world][tiles_per_shard]
  int rank, world;
  int rows_per_rank;        // Ts
  int N;
  uint32_t epoch;
};

struct RSPolicy {
  RSArgs a;
  int blocks_per_rank, num_n;
  __device__ __forceinline__ int owner_of(int mb) const { return mb / blocks_per_rank; }
  __device__ __forceinline__ void map_tile(int t, int num_m, int nn, int& b, int& mb, int& nb) const {
    // peers' row blocks first (ring order after this rank), own blocks last
    b = 0;
    const int per_owner = blocks_per_rank * nn;
    const int oi = t / per_owner;
    const int owner = (a.rank + 1 + oi) % a.world;
    const int r = t - oi * per_owner;
    nb = r / blocks_per_rank;
    mb = owner * blocks_per_rank + (r - nb * blocks_per_rank);
  }
  __device__ __forceinline__ void wait_operands(int, int, int) const {}
  __device__ __forceinline__ uint32_t* flag_at(int on_rank, int src, int mb_local, int nb) const {
    return static_cast<uint32_t*>(a.flags.ptr[on_rank]) +
           ((size_t)src * blocks_per_rank + mb_local) * num_n + nb;
  }
  __device__ __forceinline__ void before_epilogue(int, int mb, int nb) const {
    if (owner_of(mb) != a.rank) return;
    const int mb_local = mb - a.rank * blocks_per_rank;
    if ((threadIdx.x & 31) == 0) {
      for (int s = 0; s < a.world; ++s)
        if (s != a.rank) wait_flag_ge(flag_at(a.rank, s, mb_local, nb), a.epoch);
    }
    __syncwarp();
  }
  __device__ __forceinline__ void store_chunk(const Params& p, int b, int mb, int nb, int m, int n0,
                                              const uint32_t (&acc)[32]) const {
    const int owner = owner_of(mb);
    const int row_local = m - owner * a.rows_per_rank;
    if (m >= p.M || n0 >= p.N) return;
    if (owner != a.rank) {
      // partial tile -> owner's receive slot for this source rank
      __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(a.recv.ptr[owner]) +
                           ((size_t)a.rank * a.rows_per_rank + row_local) * a.N + n0;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 u;
        u.x = pack_bf16x2(__uint_as_float(acc[q * 8 + 0]), __uint_as_float(acc[q * 8 + 1]));
        u.y = pack_bf16x2(__uint_as_float(acc[q * 8 + 2]), __uint_as_float(acc[q * 8 + 3]));
        u.z = pack_bf16x2(__uint_as_float(acc[q * 8 + 4]), __uint_as_float(acc[q * 8 + 5]));
        u.w = pack_bf16x2(__uint_as_float(acc[q * 8 + 6]), __uint_as_float(acc[q * 8 + 7]));
        st_v4(dst + q * 8, u);
      }
      return;
    }
    // own rows: add every peer's partial, then the ordinary epilogue on the local row index
    uint32_t sum[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) sum[i] = acc[i];
    for (int s = 0; s < a.world; ++s) {
      if (s == a.rank) continue;
      const __nv_bfloat16* src = static_cast<const __nv_bfloat16*>(a.recv.ptr[a.rank]) +
                                 ((size_t)s * a.rows_per_rank + row_local) * a.N + n0;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint4 u = ld_cv_v4(src + q * 8);
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
    // Params describe the *shard* output: C/residual are [Ts, N], indexed by the local row
    Params q = p;
    q.M = a.rows_per_rank;
    epilogue_chunk(q, b, row_local, n0, sum);
  }
  __device__ __forceinline__ void tile_done(int, int mb, int nb) const {
    const int owner = owner_of(mb);
    if (owner == a.rank) return;
    // all 128 epilogue threads' peer stores must be out before the flag
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (threadIdx.x == 64) {  // first epilogue thread
      __threadfence_system();
      st_release_sys(flag_at(owner, a.rank, mb - owner * blocks_per_rank, nb), a.epoch);
    }
  }
};

__global__ void __launch_bounds__(NUM_THREADS, 1)
b2f_gemm_rs_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                   const Params p, const RSArgs a) {
  extern __shared__ uint8_t smem_raw[];
  RSPolicy pol{a, a.rows_per_rank / BM, (p.N + BN - 1) / BN};
  gemm_body(&tma_a, &tma_b, p, pol, blockIdx.x, gridDim.x, smem_raw);
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
  if (comm_ctas < 1) comm_ctas = 16;
  if (comm_ctas > sms / 2) comm_ctas = sms / 2;
  a.comm_ctas = comm_ctas;
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
extern "C" int b2f_gemm_rs_bf16(const void* X, const void* B, void* Y_shard, int T, int N, int K,
                                long long ldx, long long ldb, long long ldy, int b_mn_major,
                                const void* bias, const void* residual, long long ldr,
                                const long long* recv_ptrs, const long long* flag_ptrs, int rank,
                                int world, unsigned epoch, void* stream) {
  B2F_REQUIRE(T % (world * BM) == 0, "gemm_rs: rows per rank must be a multiple of 128");
  B2F_REQUIRE(N % 32 == 0, "gemm_rs: N must be a multiple of 32");
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
  CUtensorMap ma, mb;
  if (int rc = make_operand_map(&ma, X, T, K, ldx, 1, 0, 0, BM)) return rc;
  if (int rc = make_operand_map(&mb, B, N, K, ldb, 1, 0, b_mn_major, BN)) return rc;
  if (int rc = ensure_smem(b2f_gemm_rs_kernel)) return rc;
  const int tiles = (T / BM) * ((N + BN - 1) / BN);
  const int sms = sm_count();
  b2f_gemm_rs_kernel<<<tiles < sms ? tiles : sms, NUM_THREADS, SMEM_BYTES, static_cast<cudaStream_t>(stream)>>>(
      ma, mb, p, a);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
