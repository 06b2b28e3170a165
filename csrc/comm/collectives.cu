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

// --------------------------------------------------------------------- K2a
// reduced[i] = sum_p grads_p[lo + i], i in [0, n); partial[block] = sum reduced^2.
__global__ void __launch_bounds__(COMM_THREADS)
zero1_reduce_kernel(PeerTable grads, PeerTable flags, int rank, int world, long long lo,
                    long long n, float* __restrict__ reduced, float* __restrict__ partial,
                    uint32_t epoch) {
  cross_rank_block_barrier(flags, rank, world, blockIdx.x, epoch);  // every rank's backward is done
  const long long nv = n >> 2;
  float acc = 0.f;
  for (long long i = (long long)blockIdx.x * COMM_THREADS + threadIdx.x; i < nv;
       i += (long long)gridDim.x * COMM_THREADS) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int p = 0; p < MAX_WORLD; ++p) {
      if (p < world) {
        const float4 v = ld_f4(static_cast<const float*>(grads.ptr[p]) + lo + i * 4);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
    }
    reinterpret_cast<float4*>(reduced)[i] = s;
    acc += s.x * s.x + s.y * s.y + s.z * s.z + s.w * s.w;
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

__global__ void __launch_bounds__(COMM_THREADS)
zero1_adam_gather_kernel(const float* __restrict__ reduced, float* __restrict__ master,
                         float* __restrict__ m, float* __restrict__ v, PeerTable params,
                         PeerTable flags, int rank, int world, long long lo, long long n,
                         AdamHyper h, const float* __restrict__ grad_scale,
                         const int* __restrict__ skip, uint32_t epoch) {
  const bool do_step = !(skip && *skip);
  const float gs = grad_scale ? *grad_scale : 1.f;
  const long long nv = n >> 2;
  for (long long i = (long long)blockIdx.x * COMM_THREADS + threadIdx.x; i < nv;
       i += (long long)gridDim.x * COMM_THREADS) {
    float4 p4 = reinterpret_cast<float4*>(master)[i];
    if (do_step) {
      const float4 g4 = reinterpret_cast<const float4*>(reduced)[i];
      float4 m4 = reinterpret_cast<float4*>(m)[i];
      float4 v4 = reinterpret_cast<float4*>(v)[i];
      float* pp = &p4.x; float* mm = &m4.x; float* vv = &v4.x; const float* gg = &g4.x;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float g = gg[k] * gs;
        mm[k] = h.beta1 * mm[k] + (1.f - h.beta1) * g;
        vv[k] = h.beta2 * vv[k] + (1.f - h.beta2) * g * g;
        pp[k] = pp[k] - h.lr * ((mm[k] / h.bc1) / (sqrtf(vv[k] / h.bc2) + h.eps) + h.weight_decay * pp[k]);
      }
      reinterpret_cast<float4*>(master)[i] = p4;
      reinterpret_cast<float4*>(m)[i] = m4;
      reinterpret_cast<float4*>(v)[i] = v4;
    }
    uint2 o;
    o.x = pack_bf16x2(p4.x, p4.y);
    o.y = pack_bf16x2(p4.z, p4.w);
#pragma unroll
    for (int p = 0; p < MAX_WORLD; ++p)
      if (p < world)
        reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(params.ptr[p]) + lo)[i] = o;
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
  zero1_reduce_kernel<<<blocks, COMM_THREADS, 0, st>>>(g, f, rank, world, lo, n, reduced, partial,
                                                       epoch);
  B2F_CHECK_CUDA(cudaGetLastError());
  sum_partials_kernel<<<1, 256, 0, st>>>(partial, blocks, sumsq_out);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_zero1_adam_gather(const float* reduced, float* master, float* m, float* v,
                                     const long long* params_ptrs, const long long* flags_ptrs,
                                     int rank, int world, long long lo, long long n, float lr,
                                     float beta1, float beta2, float eps, float weight_decay,
                                     int step, const float* grad_scale, const int* skip,
                                     int blocks, unsigned epoch, void* stream) {
  B2F_REQUIRE(n > 0 && n % 4 == 0 && lo % 4 == 0 && step >= 1, "zero1: bad shard bounds / step");
  B2F_REQUIRE(blocks >= 1 && blocks <= b2f_comm_max_blocks(), "zero1: block count out of range");
  PeerTable p, f;
  if (int rc = fill_table(&p, params_ptrs, world)) return rc;
  if (int rc = fill_table(&f, flags_ptrs, world)) return rc;
  AdamHyper h{lr, beta1, beta2, eps, weight_decay, 1.f - powf(beta1, (float)step),
              1.f - powf(beta2, (float)step)};
  zero1_adam_gather_kernel<<<blocks, COMM_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(
      reduced, master, m, v, p, f, rank, world, lo, n, h, grad_scale, skip, epoch);
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
