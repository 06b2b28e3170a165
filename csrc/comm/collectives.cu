// b200flow K1 / K2 — gradient collectives fused with their elementwise work, over
// NVLink 5 / NVSwitch peer memory (sm_100a).
//
// All buffers named "symmetric" are the same-sized allocation on every rank of the
// group, mapped into every peer's address space (b200flow/comm/arena.py); kernels get
// a device table of the `world` base pointers.  Cross-GPU ordering uses per-block
// epoch flags in symmetric memory: block b of every rank signals block b of every
// peer (st.release.sys after a system fence) and waits for all of them
// (ld.acquire.sys).  Epochs only grow, so flags never need clearing.  All ranks must
// launch the same grid, and the grid is kept <= the SM count so every block is
// resident (no block can wait on one that has not been scheduled).
//
//  K1  b2f_allreduce_f32      in-place two-shot all-reduce of an fp32 gradient bucket with
//                             the 1/N averaging (and an optional bf16 cast-out) fused in;
//                             P2P pull/push, or in-switch NVLS reduction through the
//                             multicast mapping (multimem.ld_reduce / multimem.st).
//  K2a b2f_zero1_reduce       reduce-scatter by peer pull of the owner's shard + sum of
//                             squares of the reduced shard (for the global clip norm).
//  K2b b2f_zero1_adam_gather  clip + AdamW on the fp32 master shard, bf16 working copy pushed
//                             straight into every rank's flat parameter buffer (all-gather).
//
// Replaces: torch DDP Reducer + ncclAllReduce, Horovod fused-buffer allreduce, DeepSpeed
// ZeRO-1 reduce-scatter / FusedAdam / all_gather phases.
#include "../common/ptx.cuh"
#include "../common/host.h"
#include "../../include/b200flow.h"

using namespace b2f;

namespace {

constexpr int MAX_WORLD = 8;
constexpr int COMM_THREADS = 512;

struct PeerTable {
  void* ptr[MAX_WORLD];
};

// Block-level barrier across ranks. flags: symmetric u32 array, layout [slot][world].
__device__ __forceinline__ void cross_rank_block_barrier(const PeerTable& flags, int rank,
                                                         int world, int slot, uint32_t epoch) {
  if (threadIdx.x == 0) race_delay();
  __syncthreads();
  if ((int)threadIdx.x < world) {
    const int p = threadIdx.x;
    __threadfence_system();
    uint32_t* remote = static_cast<uint32_t*>(flags.ptr[p]) + (size_t)slot * world + rank;
    st_release_sys(remote, epoch);
    const uint32_t* mine = static_cast<const uint32_t*>(flags.ptr[rank]) + (size_t)slot * world + p;
    wait_flag_ge(mine, epoch);
  }
  __syncthreads();
}

__device__ __forceinline__ float4 ld_f4(const float* p) {
  float4 v;
  asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_f4(float* p, const float4& v) {
  asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

// ---------------------------------------------------------------------- K1
// grads: symmetric fp32 buffers; the bucket is elements [off, off + n) (n % 4 == 0).
// Each block owns a contiguous chunk; inside it rank r reduces sub-slice r and
// writes the averaged values back into every rank's buffer (in place: an index is
// read from all peers and written to all peers only by its owner).
template <bool MULTIMEM>
__global__ void __launch_bounds__(COMM_THREADS)
allreduce_f32_kernel(PeerTable grads, float* mc_base, PeerTable flags, int rank, int world,
                     long long off, long long n, float scale, __nv_bfloat16* bf16_out,
                     uint32_t epoch) {
  const long long nv = n >> 2;                       // float4 count
  const long long per_block = (nv + gridDim.x - 1) / gridDim.x;
  const long long b0 = (long long)blockIdx.x * per_block;
  const long long b1 = b0 + per_block < nv ? b0 + per_block : nv;
  const long long len = b1 > b0 ? b1 - b0 : 0;
  const long long per_rank = (len + world - 1) / world;
  const long long r0 = b0 + (long long)rank * per_rank;
  const long long r1 = r0 + per_rank < b1 ? r0 + per_rank : b1;

  cross_rank_block_barrier(flags, rank, world, blockIdx.x, epoch);  // all grads final

  if (MULTIMEM) {
    float* mc = mc_base + off;
    long long i = r0 + threadIdx.x;
    for (; i + 3 * COMM_THREADS < r1; i += 4 * COMM_THREADS) {  // 4 reductions in flight
      float4 s[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) s[j] = multimem_ld_reduce_f4(mc + (i + j * COMM_THREADS) * 4);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        s[j].x *= scale; s[j].y *= scale; s[j].z *= scale; s[j].w *= scale;
        multimem_st_f4(mc + (i + j * COMM_THREADS) * 4, s[j]);
      }
    }
    for (; i < r1; i += COMM_THREADS) {
      float4 s = multimem_ld_reduce_f4(mc + i * 4);
      s.x *= scale; s.y *= scale; s.z *= scale; s.w *= scale;
      multimem_st_f4(mc + i * 4, s);
    }
  } else {
    for (long long i0 = r0; i0 < r1; i0 += 2 * COMM_THREADS) {  // 2 x world loads in flight
      float4 s[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const long long i = i0 + j * COMM_THREADS + threadIdx.x;
        s[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < r1) {
#pragma unroll
          for (int p = 0; p < MAX_WORLD; ++p) {
            if (p < world) {
              const float4 v = ld_f4(static_cast<const float*>(grads.ptr[p]) + off + i * 4);
              s[j].x += v.x; s[j].y += v.y; s[j].z += v.z; s[j].w += v.w;
            }
          }
        }
      }
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const long long i = i0 + j * COMM_THREADS + threadIdx.x;
        if (i < r1) {
          s[j].x *= scale; s[j].y *= scale; s[j].z *= scale; s[j].w *= scale;
#pragma unroll
          for (int p = 0; p < MAX_WORLD; ++p)
            if (p < world) st_f4(static_cast<float*>(grads.ptr[p]) + off + i * 4, s[j]);
        }
      }
    }
  }

  cross_rank_block_barrier(flags, rank, world, blockIdx.x, epoch + 1);  // all slices landed

  if (bf16_out) {  // fused cast of the averaged bucket for a bf16 consumer
    const float* mine = static_cast<const float*>(grads.ptr[rank]) + off;
    for (long long i = b0 + threadIdx.x; i < b1; i += COMM_THREADS) {
      const float4 v = ld_f4(mine + i * 4);
      uint2 o;
      o.x = pack_bf16x2(v.x, v.y);
      o.y = pack_bf16x2(v.z, v.w);
      reinterpret_cast<uint2*>(bf16_out + off)[i] = o;
    }
  }
}

// ---------------------------------------------------------------- K1, bf16 on the wire
// The same averaged all-reduce with half the NVLink bytes: gradients stay fp32 in the flat buffer, but what crosses
// the links is bf16 (the precision DDP's bf16 compression hook ships).  Three phases inside one kernel, per block:
//   A  cast this rank's chunk fp32 -> bf16 into its symmetric staging buffer (local pass)
//   B  after a cross-rank barrier, rank r reduces sub-slice r of the chunk over every rank's staging in fp32
//      (in-switch `multimem.ld_reduce ... acc::f32 ... bf16x2` when a multicast mapping exists, peer loads otherwise),
//      scales by 1/N and writes the bf16 result into every rank's staging (`multimem.st` / peer stores)
//   C  after a second barrier, staging bf16 -> fp32 back into the gradient buffer (local pass)
// A and C are HBM passes (12 bytes per element at ~6.5 TB/s) next to 4 (N-1)/N bytes per element of NVLink traffic
// at ~0.77 TB/s: the wire time halves and the local passes add about a fifth of what was saved.
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4." B2F_HALF_PTX " {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_bf16x8(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4." B2F_HALF_PTX " [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ uint4 scale_bf16x8(const uint4& v, float s) {
  const float2 a = unpack_bf16x2(v.x), b = unpack_bf16x2(v.y), c = unpack_bf16x2(v.z), d = unpack_bf16x2(v.w);
  uint4 o;
  o.x = pack_bf16x2(a.x * s, a.y * s); o.y = pack_bf16x2(b.x * s, b.y * s);
  o.z = pack_bf16x2(c.x * s, c.y * s); o.w = pack_bf16x2(d.x * s, d.y * s);
  return o;
}

// grads: this rank's fp32 flat buffer (local); stage: symmetric bf16 staging buffers, element i of the bucket at
// stage[i]; n % 8 == 0.  stage_mc: multicast address of the staging buffers or null.
template <bool MULTIMEM>
__global__ void __launch_bounds__(COMM_THREADS)
allreduce_bf16wire_kernel(float* grads, PeerTable stage, __nv_bfloat16* stage_mc, PeerTable flags, int rank, int world,
                          long long off, long long n, float scale, uint32_t epoch) {
  const long long nv = n >> 3;                       // 8-element (16-byte bf16) vectors
  const long long per_block = (nv + gridDim.x - 1) / gridDim.x;
  const long long b0 = (long long)blockIdx.x * per_block;
  const long long b1 = b0 + per_block < nv ? b0 + per_block : nv;
  const long long len = b1 > b0 ? b1 - b0 : 0;
  const long long per_rank = (len + world - 1) / world;
  const long long r0 = b0 + (long long)rank * per_rank;
  const long long r1 = r0 + per_rank < b1 ? r0 + per_rank : b1;
  float* g = grads + off;
  uint4* mine = reinterpret_cast<uint4*>(stage.ptr[rank]);

  // A: fp32 -> bf16 staging
  for (long long i = b0 + threadIdx.x; i < b1; i += COMM_THREADS) {
    const float4 lo = ld_f4(g + i * 8), hi = ld_f4(g + i * 8 + 4);
    uint4 o;
    o.x = pack_bf16x2(lo.x, lo.y); o.y = pack_bf16x2(lo.z, lo.w);
    o.z = pack_bf16x2(hi.x, hi.y); o.w = pack_bf16x2(hi.z, hi.w);
    mine[i] = o;
  }
  cross_rank_block_barrier(flags, rank, world, blockIdx.x, epoch);      // every rank's chunk is staged

  // B: reduce the owned sub-slice across ranks, average, hand the result to everyone
  if (MULTIMEM) {
    uint4* mc = reinterpret_cast<uint4*>(stage_mc);
    long long i = r0 + threadIdx.x;
    for (; i + 3 * COMM_THREADS < r1; i += 4 * COMM_THREADS) {
      uint4 v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = multimem_ld_reduce_bf16x8(mc + i + j * COMM_THREADS);
#pragma unroll
      for (int j = 0; j < 4; ++j) multimem_st_bf16x8(mc + i + j * COMM_THREADS, scale_bf16x8(v[j], scale));
    }
    for (; i < r1; i += COMM_THREADS) multimem_st_bf16x8(mc + i, scale_bf16x8(multimem_ld_reduce_bf16x8(mc + i), scale));
  } else {
    for (long long i = r0 + threadIdx.x; i < r1; i += COMM_THREADS) {
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int p = 0; p < MAX_WORLD; ++p) {
        if (p < world) {
          const uint4 v = ld_v4(reinterpret_cast<const uint4*>(stage.ptr[p]) + i);
          const float2 a = unpack_bf16x2(v.x), b = unpack_bf16x2(v.y), c = unpack_bf16x2(v.z), d = unpack_bf16x2(v.w);
          acc[0] += a.x; acc[1] += a.y; acc[2] += b.x; acc[3] += b.y;
          acc[4] += c.x; acc[5] += c.y; acc[6] += d.x; acc[7] += d.y;
        }
      }
      uint4 o;
      o.x = pack_bf16x2(acc[0] * scale, acc[1] * scale); o.y = pack_bf16x2(acc[2] * scale, acc[3] * scale);
      o.z = pack_bf16x2(acc[4] * scale, acc[5] * scale); o.w = pack_bf16x2(acc[6] * scale, acc[7] * scale);
#pragma unroll
      for (int p = 0; p < MAX_WORLD; ++p)
        if (p < world) st_v4(reinterpret_cast<uint4*>(stage.ptr[p]) + i, o);
    }
  }
  cross_rank_block_barrier(flags, rank, world, blockIdx.x, epoch + 1);  // every sub-slice has landed everywhere

  // C: averaged bf16 -> fp32 gradient buffer
  for (long long i = b0 + threadIdx.x; i < b1; i += COMM_THREADS) {
    const uint4 v = mine[i];
    const float2 a = unpack_bf16x2(v.x), b = unpack_bf16x2(v.y), c = unpack_bf16x2(v.z), d = unpack_bf16x2(v.w);
    st_f4(g + i * 8, make_float4(a.x, a.y, b.x, b.y));
    st_f4(g + i * 8 + 4, make_float4(c.x, c.y, d.x, d.y));
  }
}

// --------------------------------------------------------------------- K2a
// reduced[i] = sum_p grads_p[lo + i], i in [0, n); partial[block] = sum reduced^2.
// MULTIMEM: one `multimem.ld_reduce` per vector sums the shard across every GPU inside the switch, so a rank
// receives n x 4 bytes instead of (world - 1) x that.
template <bool MULTIMEM>
__global__ void __launch_bounds__(COMM_THREADS)
zero1_reduce_kernel(PeerTable grads, const float* grads_mc, PeerTable flags, int rank, int world, long long lo,
                    long long n, float* __restrict__ reduced, float* __restrict__ partial,
                    uint32_t epoch) {
  cross_rank_block_barrier(flags, rank, world, blockIdx.x, epoch);  // every rank's backward is done
  const long long nv = n >> 2;
  float acc = 0.f;
  // four vectors per thread and trip: 4 (x world) independent loads in flight cover the NVLink round trip
  const long long stride = (long long)gridDim.x * COMM_THREADS;
  for (long long i0 = (long long)blockIdx.x * COMM_THREADS + threadIdx.x; i0 < nv; i0 += 4 * stride) {
    float4 s[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long i = i0 + j * stride;
      s[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i >= nv) continue;
      if (MULTIMEM) {
        s[j] = multimem_ld_reduce_f4(grads_mc + lo + i * 4);
      } else {
#pragma unroll
        for (int p = 0; p < MAX_WORLD; ++p) {
          if (p < world) {
            const float4 v = ld_f4(static_cast<const float*>(grads.ptr[p]) + lo + i * 4);
            s[j].x += v.x; s[j].y += v.y; s[j].z += v.z; s[j].w += v.w;
          }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long i = i0 + j * stride;
      if (i >= nv) continue;
      reinterpret_cast<float4*>(reduced)[i] = s[j];
      acc += s[j].x * s[j].x + s[j].y * s[j].y + s[j].z * s[j].z + s[j].w * s[j].w;
    }
  }
  __shared__ float red[COMM_THREADS / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < COMM_THREADS / 32; ++w) s += red[w];
    partial[blockIdx.x] = s;
  }
}

__global__ void sum_partials_kernel(const float* __restrict__ partial, int n, float* __restrict__ out) {
  __shared__ float red[32];
  float acc = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) acc += partial[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w];
    out[0] = s;
  }
}

// --------------------------------------------------------------------- K2b
struct AdamHyper {
  float lr, beta1, beta2, eps, weight_decay, bc1, bc2;
};

// Eight elements per thread and iteration: the bf16 result leaves as one 16-byte store per destination (or one
// `multimem.st` that the switch replicates to every GPU when a multicast mapping of the parameters exists).
template <bool MULTIMEM>
__global__ void __launch_bounds__(COMM_THREADS)
zero1_adam_gather_kernel(const float* __restrict__ reduced, float* __restrict__ master,
                         float* __restrict__ m, float* __restrict__ v, PeerTable params, __nv_bfloat16* params_mc,
                         PeerTable flags, int rank, int world, long long lo, long long n,
                         AdamHyper h, const float* __restrict__ grad_scale,
                         const int* __restrict__ skip, uint32_t epoch) {
  const bool do_step = !(skip && *skip);
  const float gs = grad_scale ? *grad_scale : 1.f;
  const long long nv8 = n >> 3;
  for (long long i = (long long)blockIdx.x * COMM_THREADS + threadIdx.x; i < nv8;
       i += (long long)gridDim.x * COMM_THREADS) {
    uint32_t packed[4];
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const long long j = i * 2 + half;
      float4 p4 = reinterpret_cast<float4*>(master)[j];
      if (do_step) {
        const float4 g4 = reinterpret_cast<const float4*>(reduced)[j];
        float4 m4 = reinterpret_cast<float4*>(m)[j];
        float4 v4 = reinterpret_cast<float4*>(v)[j];
        float* pp = &p4.x; float* mm = &m4.x; float* vv = &v4.x; const float* gg = &g4.x;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float g = gg[k] * gs;
          mm[k] = h.beta1 * mm[k] + (1.f - h.beta1) * g;
          vv[k] = h.beta2 * vv[k] + (1.f - h.beta2) * g * g;
          pp[k] = pp[k] - h.lr * ((mm[k] / h.bc1) / (sqrtf(vv[k] / h.bc2) + h.eps) + h.weight_decay * pp[k]);
        }
        reinterpret_cast<float4*>(master)[j] = p4;
        reinterpret_cast<float4*>(m)[j] = m4;
        reinterpret_cast<float4*>(v)[j] = v4;
      }
      packed[half * 2] = pack_bf16x2(p4.x, p4.y);
      packed[half * 2 + 1] = pack_bf16x2(p4.z, p4.w);
    }
    const uint4 o = make_uint4(packed[0], packed[1], packed[2], packed[3]);
    if (MULTIMEM) {
      multimem_st_bf16x8(reinterpret_cast<uint4*>(params_mc + lo) + i, o);
    } else {
#pragma unroll
      for (int p = 0; p < MAX_WORLD; ++p)
        if (p < world) st_v4(reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(params.ptr[p]) + lo) + i, o);
    }
  }
  cross_rank_block_barrier(flags, rank, world, blockIdx.x, epoch);  // every shard has landed everywhere
}

// Peer-memory self test: out[i] = sum over ranks of buf_p[i] (pull), after a barrier.
__global__ void __launch_bounds__(COMM_THREADS)
peer_sum_test_kernel(PeerTable bufs, PeerTable flags, int rank, int world, long long n,
                     float* __restrict__ out, uint32_t epoch) {
  cross_rank_block_barrier(flags, rank, world, blockIdx.x, epoch);
  for (long long i = (long long)blockIdx.x * COMM_THREADS + threadIdx.x; i < n;
       i += (long long)gridDim.x * COMM_THREADS) {
    float s = 0.f;
    for (int p = 0; p < world; ++p) s += static_cast<const float*>(bufs.ptr[p])[i];
    out[i] = s;
  }
}

__global__ void peer_barrier_kernel(PeerTable flags, int rank, int world, uint32_t epoch) {
  cross_rank_block_barrier(flags, rank, world, 0, epoch);
}

int fill_table(PeerTable* t, const long long* ptrs, int world) {
  B2F_REQUIRE(world >= 1 && world <= MAX_WORLD, "peer group size must be 1..8");
  for (int i = 0; i < MAX_WORLD; ++i) t->ptr[i] = i < world ? reinterpret_cast<void*>(ptrs[i]) : nullptr;
  return 0;
}

}  // namespace

extern "C" int b2f_comm_max_blocks() {
  const int n = b2f::sm_count();
  return n < 128 ? n : 128;
}

// grads_ptrs / flags_ptrs: host arrays of `world` device addresses (one per rank).
// flags need >= blocks * world u32 per rank, zero-initialised once.
extern "C" int b2f_allreduce_f32(const long long* grads_ptrs, long long multicast_ptr,
                                 const long long* flags_ptrs, int rank, int world, long long off,
                                 long long n, float scale, void* bf16_out, int blocks,
                                 unsigned epoch, void* stream) {
  B2F_REQUIRE(n > 0 && n % 4 == 0 && off % 4 == 0, "allreduce: offset and length must be multiples of 4");
  B2F_REQUIRE(blocks >= 1 && blocks <= b2f_comm_max_blocks(), "allreduce: block count out of range");
  PeerTable g, f;
  if (int rc = fill_table(&g, grads_ptrs, world)) return rc;
  if (int rc = fill_table(&f, flags_ptrs, world)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (multicast_ptr)
    allreduce_f32_kernel<true><<<blocks, COMM_THREADS, 0, st>>>(
        g, reinterpret_cast<float*>(multicast_ptr), f, rank, world, off, n, scale,
        static_cast<__nv_bfloat16*>(bf16_out), epoch);
  else
    allreduce_f32_kernel<false><<<blocks, COMM_THREADS, 0, st>>>(
        g, nullptr, f, rank, world, off, n, scale, static_cast<__nv_bfloat16*>(bf16_out), epoch);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// partial: fp32 scratch [blocks]; sumsq_out: fp32 [1].
extern "C" int b2f_zero1_reduce(const long long* grads_ptrs, const long long* flags_ptrs, int rank,
                                int world, long long lo, long long n, float* reduced,
                                float* partial, float* sumsq_out, int blocks, unsigned epoch,
                                void* stream) {
  B2F_REQUIRE(n > 0 && n % 4 == 0 && lo % 4 == 0, "zero1: shard bounds must be multiples of 4");
  B2F_REQUIRE(blocks >= 1 && blocks <= b2f_comm_max_blocks(), "zero1: block count out of range");
  PeerTable g, f;
  if (int rc = fill_table(&g, grads_ptrs, world)) return rc;
  if (int rc = fill_table(&f, flags_ptrs, world)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  zero1_reduce_kernel<false><<<blocks, COMM_THREADS, 0, st>>>(g, nullptr, f, rank, world, lo, n, reduced, partial,
                                                              epoch);
  B2F_CHECK_CUDA(cudaGetLastError());
  sum_partials_kernel<<<1, 256, 0, st>>>(partial, blocks, sumsq_out);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// The same with the gradient buffers' NVLS multicast address: the shard is summed inside the switch.
extern "C" int b2f_zero1_reduce_mc(const long long* grads_ptrs, long long grads_multicast_ptr,
                                   const long long* flags_ptrs, int rank, int world, long long lo, long long n,
                                   float* reduced, float* partial, float* sumsq_out, int blocks, unsigned epoch,
                                   void* stream) {
  B2F_REQUIRE(n > 0 && n % 4 == 0 && lo % 4 == 0, "zero1: shard bounds must be multiples of 4");
  B2F_REQUIRE(blocks >= 1 && blocks <= b2f_comm_max_blocks(), "zero1: block count out of range");
  B2F_REQUIRE(grads_multicast_ptr != 0, "zero1_reduce_mc: needs the multicast mapping of the gradient buffers");
  PeerTable g, f;
  if (int rc = fill_table(&g, grads_ptrs, world)) return rc;
  if (int rc = fill_table(&f, flags_ptrs, world)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  zero1_reduce_kernel<true><<<blocks, COMM_THREADS, 0, st>>>(
      g, reinterpret_cast<const float*>(grads_multicast_ptr), f, rank, world, lo, n, reduced, partial, epoch);
  B2F_CHECK_CUDA(cudaGetLastError());
  sum_partials_kernel<<<1, 256, 0, st>>>(partial, blocks, sumsq_out);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// K1 with bf16 on the wire: grads = this rank's fp32 flat buffer, stage_ptrs = symmetric bf16 staging buffers of at
// least n elements (stage_multicast_ptr: their NVLS multicast address or 0); n % 8 == 0, off % 4 == 0.
extern "C" int b2f_allreduce_bf16wire(void* grads, const long long* stage_ptrs, long long stage_multicast_ptr,
                                      const long long* flags_ptrs, int rank, int world, long long off, long long n,
                                      float scale, int blocks, unsigned epoch, void* stream) {
  B2F_REQUIRE(n > 0 && n % 8 == 0 && off % 4 == 0, "allreduce_bf16wire: length % 8 and offset % 4 must be 0");
  B2F_REQUIRE(blocks >= 1 && blocks <= b2f_comm_max_blocks(), "allreduce: block count out of range");
  PeerTable sg, f;
  if (int rc = fill_table(&sg, stage_ptrs, world)) return rc;
  if (int rc = fill_table(&f, flags_ptrs, world)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (stage_multicast_ptr)
    allreduce_bf16wire_kernel<true><<<blocks, COMM_THREADS, 0, st>>>(
        static_cast<float*>(grads), sg, reinterpret_cast<__nv_bfloat16*>(stage_multicast_ptr), f, rank, world, off, n,
        scale, epoch);
  else
    allreduce_bf16wire_kernel<false><<<blocks, COMM_THREADS, 0, st>>>(static_cast<float*>(grads), sg, nullptr, f, rank,
                                                                      world, off, n, scale, epoch);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_zero1_adam_gather_mc(const float* reduced, float* master, float* m, float* v,
                                        const long long* params_ptrs, long long params_multicast_ptr,
                                        const long long* flags_ptrs, int rank, int world, long long lo, long long n,
                                        float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                                        const float* grad_scale, const int* skip, int blocks, unsigned epoch,
                                        void* stream);

extern "C" int b2f_zero1_adam_gather(const float* reduced, float* master, float* m, float* v,
                                     const long long* params_ptrs, const long long* flags_ptrs,
                                     int rank, int world, long long lo, long long n, float lr,
                                     float beta1, float beta2, float eps, float weight_decay,
                                     int step, const float* grad_scale, const int* skip,
                                     int blocks, unsigned epoch, void* stream) {
  return b2f_zero1_adam_gather_mc(reduced, master, m, v, params_ptrs, 0, flags_ptrs, rank, world, lo, n, lr, beta1,
                                 beta2, eps, weight_decay, step, grad_scale, skip, blocks, epoch, stream);
}

// params_multicast_ptr != 0: the updated bf16 shard leaves with `multimem.st` (one store, replicated by the switch).
extern "C" int b2f_zero1_adam_gather_mc(const float* reduced, float* master, float* m, float* v,
                                        const long long* params_ptrs, long long params_multicast_ptr,
                                        const long long* flags_ptrs, int rank, int world, long long lo, long long n,
                                        float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                                        const float* grad_scale, const int* skip, int blocks, unsigned epoch,
                                        void* stream) {
  B2F_REQUIRE(n > 0 && n % 8 == 0 && lo % 8 == 0 && step >= 1, "zero1: shard bounds must be multiples of 8, step >= 1");
  B2F_REQUIRE(blocks >= 1 && blocks <= b2f_comm_max_blocks(), "zero1: block count out of range");
  PeerTable p, f;
  if (int rc = fill_table(&p, params_ptrs, world)) return rc;
  if (int rc = fill_table(&f, flags_ptrs, world)) return rc;
  AdamHyper h{lr, beta1, beta2, eps, weight_decay, 1.f - powf(beta1, (float)step),
              1.f - powf(beta2, (float)step)};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (params_multicast_ptr)
    zero1_adam_gather_kernel<true><<<blocks, COMM_THREADS, 0, st>>>(
        reduced, master, m, v, p, reinterpret_cast<__nv_bfloat16*>(params_multicast_ptr), f, rank, world, lo, n, h,
        grad_scale, skip, epoch);
  else
    zero1_adam_gather_kernel<false><<<blocks, COMM_THREADS, 0, st>>>(reduced, master, m, v, p, nullptr, f, rank, world,
                                                                     lo, n, h, grad_scale, skip, epoch);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Stream-ordered barrier across the ranks of a peer group: everything every rank enqueued
// before it (including stores into peers' memory) is complete and visible after it.
extern "C" int b2f_peer_barrier(const long long* flags_ptrs, int rank, int world, unsigned epoch,
                                void* stream) {
  PeerTable f;
  if (int rc = fill_table(&f, flags_ptrs, world)) return rc;
  peer_barrier_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(f, rank, world, epoch);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_peer_sum_test(const long long* buf_ptrs, const long long* flags_ptrs, int rank,
                                 int world, long long n, float* out, int blocks, unsigned epoch,
                                 void* stream) {
  PeerTable b, f;
  if (int rc = fill_table(&b, buf_ptrs, world)) return rc;
  if (int rc = fill_table(&f, flags_ptrs, world)) return rc;
  peer_sum_test_kernel<<<blocks, COMM_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(
      b, f, rank, world, n, out, epoch);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
