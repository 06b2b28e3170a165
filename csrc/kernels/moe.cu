// b200flow K5a + MoE data movement — fused top-1 router and token permutation kernels (sm_100a).
//
// K5a router: one launch turns gate logits [T, E] into everything the expert-parallel layer
// needs — softmax, arg-max expert, gate probability, the token's slot inside its expert
// (an ordered prefix count, so routing is deterministic), capacity dropping, per-expert
// token counts and the load-balancing auxiliary loss terms — emitting a *permutation*
// (token -> row of the dispatch buffer) instead of DeepSpeed's dense one-hot
// [tokens, experts, capacity] einsum operands.
//
// Row movers: gather tokens into the [E, C, H] dispatch layout (optionally writing straight
// into peer memory of the expert's owner = the all-to-all) and the gate-weighted combine.
#include "../common/ptx.cuh"
#include "../common/host.h"
#include "../../include/b200flow.h"

using namespace b2f;

namespace {

constexpr int MAX_E = 64;
constexpr int RT = 512;    // tokens (= threads) per router block

// The router runs as two kernels over ceil(T / RT) blocks, with a small caller-provided scratch
// ([blocks][E] ints + [blocks][E] floats) between them:
//   pass 1  softmax, arg-max, gate probability; per-block token count and probability mass per expert
//   pass 2  a token's slot = tokens of its expert in earlier blocks (summed from the scratch)
//           + earlier tokens of its expert in this block (ballot ranks); capacity dropping;
//           block 0 also reduces the per-block partials into counts / me / ce in a fixed order.
// Slots follow token order, so routing is deterministic.

// logits fp32 [T, E]. Outputs of the pair: expert[T] (-1 = dropped), slot[T], gate[T], probs [T, E]
// (saved for backward), counts[E] (kept tokens), me[E] = sum_t p[t, e], ce[E] = #tokens whose
// arg-max is e.
__global__ void __launch_bounds__(RT)
router_pass1_kernel(const float* __restrict__ logits, int T, int E, int* __restrict__ expert,
                    float* __restrict__ gate, float* __restrict__ probs, int* __restrict__ blk_cnt,
                    float* __restrict__ blk_me) {
  __shared__ float me_w[RT / 32][MAX_E];   // per-warp partial sums, added in warp order (deterministic)
  __shared__ int ce_s[MAX_E];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int e = threadIdx.x; e < E; e += RT) ce_s[e] = 0;
  __syncthreads();
  const int t = blockIdx.x * RT + threadIdx.x;
  int best = -1;
  float bp = 0.f;
  const float* z = logits + (long long)t * E;
  float mx = -INFINITY, inv = 0.f;
  if (t < T) {
    for (int e = 0; e < E; ++e) mx = fmaxf(mx, z[e]);
    float sum = 0.f;
    for (int e = 0; e < E; ++e) sum += __expf(z[e] - mx);
    inv = 1.f / sum;
  }
  for (int e = 0; e < E; ++e) {
    float p = 0.f;
    if (t < T) {
      p = __expf(z[e] - mx) * inv;
      probs[(long long)t * E + e] = p;
      if (p > bp) { bp = p; best = e; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) p += __shfl_xor_sync(0xffffffffu, p, o);
    if (lane == 0) me_w[warp][e] = p;
  }
  for (int e = 0; e < E; ++e) {
    const unsigned m = __ballot_sync(0xffffffffu, best == e);
    if (lane == 0 && m) atomicAdd(&ce_s[e], __popc(m));
  }
  if (t < T) {
    expert[t] = best;   // provisional: pass 2 applies the capacity
    gate[t] = bp;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < E; e += RT) {
    float m = 0.f;
    for (int w = 0; w < RT / 32; ++w) m += me_w[w][e];
    blk_cnt[blockIdx.x * E + e] = ce_s[e];
    blk_me[blockIdx.x * E + e] = m;
  }
}

__global__ void __launch_bounds__(RT)
router_pass2_kernel(int T, int E, int capacity, int nblk, int* __restrict__ expert, int* __restrict__ slot,
                    float* __restrict__ gate, const int* __restrict__ blk_cnt,
                    const float* __restrict__ blk_me, int* __restrict__ counts, float* __restrict__ me,
                    float* __restrict__ ce) {
  __shared__ int base[MAX_E];
  __shared__ int warp_cnt[RT / 32][MAX_E];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // tokens of each expert in earlier blocks
  for (int e = threadIdx.x; e < E; e += RT) {
    int acc = 0;
    for (int bI = 0; bI < (int)blockIdx.x; ++bI) acc += blk_cnt[bI * E + e];
    base[e] = acc;
  }
  if (blockIdx.x == 0) {
    for (int e = threadIdx.x; e < E; e += RT) {
      int tot = 0;
      float m = 0.f;
      for (int bI = 0; bI < nblk; ++bI) { tot += blk_cnt[bI * E + e]; m += blk_me[bI * E + e]; }
      counts[e] = tot < capacity ? tot : capacity;
      ce[e] = (float)tot;
      me[e] = m;
    }
  }
  const int t = blockIdx.x * RT + threadIdx.x;
  const int best = t < T ? expert[t] : -1;
  int rank_in_warp = 0;
  for (int e = 0; e < E; ++e) {
    const unsigned m = __ballot_sync(0xffffffffu, best == e);
    if (lane == 0) warp_cnt[warp][e] = __popc(m);
    if (best == e) rank_in_warp = __popc(m & ((1u << lane) - 1));
  }
  __syncthreads();
  if (t < T) {
    int my_slot = base[best] + rank_in_warp;
    for (int w = 0; w < warp; ++w) my_slot += warp_cnt[w][best];
    const bool keep = my_slot < capacity;
    expert[t] = keep ? best : -1;
    slot[t] = keep ? my_slot : -1;
    if (!keep) gate[t] = 0.f;
  }
}

// dst_row(t) = expert[t] * capacity + slot[t]. dst: [E * capacity, H] rows. One warp per token.
// With peer pointers, expert e's rows live on rank e / experts_per_rank at
// [src_rank][local_expert][capacity] (the dispatch all-to-all done by stores).
struct PeerRows {
  void* ptr[8];
};

__global__ void __launch_bounds__(256)
moe_dispatch_kernel(const __nv_bfloat16* __restrict__ x, const int* __restrict__ expert,
                    const int* __restrict__ slot, int T, int H, int capacity,
                    __nv_bfloat16* __restrict__ dst_local, PeerRows peers, int world,
                    int experts_per_rank, int src_rank) {
  const int warps_per_block = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  const int hv = H >> 3;
  if (lane == 0) race_delay();
  __syncwarp();
  for (int t = blockIdx.x * warps_per_block + (threadIdx.x >> 5); t < T; t += gridDim.x * warps_per_block) {
    const int e = expert[t];
    if (e < 0) continue;
    __nv_bfloat16* row;
    if (world > 1) {
      const int owner = e / experts_per_rank, le = e - owner * experts_per_rank;
      row = static_cast<__nv_bfloat16*>(peers.ptr[owner]) +
            (((long long)src_rank * experts_per_rank + le) * capacity + slot[t]) * H;
    } else {
      row = dst_local + ((long long)e * capacity + slot[t]) * H;
    }
    const uint4* s = reinterpret_cast<const uint4*>(x + (long long)t * H);
    uint4* d = reinterpret_cast<uint4*>(row);
    for (int c = lane; c < hv; c += 32) st_v4(d + c, ld_nc_v4(s + c));
  }
}

// y[t] = gate[t] * src_row(t) (0 for dropped tokens); src: the combine buffer after the return
// all-to-all, or peer memory of the experts' owners (pull).
__global__ void __launch_bounds__(256)
moe_combine_kernel(const __nv_bfloat16* __restrict__ src_local, PeerRows peers, int world,
                   int experts_per_rank, int src_rank, const int* __restrict__ expert,
                   const int* __restrict__ slot, const float* __restrict__ gate, int T, int H,
                   int capacity, __nv_bfloat16* __restrict__ y) {
  const int warps_per_block = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  const int hv = H >> 3;
  for (int t = blockIdx.x * warps_per_block + (threadIdx.x >> 5); t < T; t += gridDim.x * warps_per_block) {
    const int e = expert[t];
    uint4* d = reinterpret_cast<uint4*>(y + (long long)t * H);
    if (e < 0) {
      for (int c = lane; c < hv; c += 32) d[c] = make_uint4(0, 0, 0, 0);
      continue;
    }
    const __nv_bfloat16* row;
    if (world > 1) {
      const int owner = e / experts_per_rank, le = e - owner * experts_per_rank;
      row = static_cast<const __nv_bfloat16*>(peers.ptr[owner]) +
            (((long long)src_rank * experts_per_rank + le) * capacity + slot[t]) * H;
    } else {
      row = src_local + ((long long)e * capacity + slot[t]) * H;
    }
    const float g = gate[t];
    const uint4* s = reinterpret_cast<const uint4*>(row);
    for (int c = lane; c < hv; c += 32) {
      const uint4 u = ld_v4(s + c);
      float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), cc = unpack_bf16x2(u.z), dd = unpack_bf16x2(u.w);
      uint4 o;
      o.x = pack_bf16x2(a.x * g, a.y * g); o.y = pack_bf16x2(b.x * g, b.y * g);
      o.z = pack_bf16x2(cc.x * g, cc.y * g); o.w = pack_bf16x2(dd.x * g, dd.y * g);
      d[c] = o;
    }
  }
}

// Backward of the combine w.r.t. the gate: dgate[t] = <dy[t], out_row(t)>.
__global__ void __launch_bounds__(256)
moe_gate_grad_kernel(const __nv_bfloat16* __restrict__ rows, const int* __restrict__ expert,
                     const int* __restrict__ slot, const __nv_bfloat16* __restrict__ dy, int T, int H,
                     int capacity, float* __restrict__ dgate) {
  const int warps_per_block = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  const int hv = H >> 3;
  for (int t = blockIdx.x * warps_per_block + (threadIdx.x >> 5); t < T; t += gridDim.x * warps_per_block) {
    const int e = expert[t];
    float acc = 0.f;
    if (e >= 0) {
      const uint4* a = reinterpret_cast<const uint4*>(rows + ((long long)e * capacity + slot[t]) * H);
      const uint4* b = reinterpret_cast<const uint4*>(dy + (long long)t * H);
      for (int c = lane; c < hv; c += 32) {
        const uint4 u = ld_nc_v4(a + c), v = ld_nc_v4(b + c);
        float2 u0 = unpack_bf16x2(u.x), u1 = unpack_bf16x2(u.y), u2 = unpack_bf16x2(u.z), u3 = unpack_bf16x2(u.w);
        float2 v0 = unpack_bf16x2(v.x), v1 = unpack_bf16x2(v.y), v2 = unpack_bf16x2(v.z), v3 = unpack_bf16x2(v.w);
        acc += u0.x * v0.x + u0.y * v0.y + u1.x * v1.x + u1.y * v1.y + u2.x * v2.x + u2.y * v2.y +
               u3.x * v3.x + u3.y * v3.y;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) dgate[t] = acc;
  }
}

int fill_peers(PeerRows* p, const long long* ptrs, int world) {
  B2F_REQUIRE(world >= 1 && world <= 8, "moe: expert-parallel group size must be 1..8");
  for (int i = 0; i < 8; ++i) p->ptr[i] = (ptrs && i < world) ? reinterpret_cast<void*>(ptrs[i]) : nullptr;
  return 0;
}

inline int row_grid(int T) {
  int b = (T + 7) / 8;
  const int cap = b2f::sm_count() * 8;
  return b > cap ? cap : (b < 1 ? 1 : b);
}

}  // namespace

extern "C" long long b2f_moe_router_scratch_bytes(int T, int E) {
  const long long nblk = (T + RT - 1) / RT;
  return nblk * E * 8;
}

// scratch: b2f_moe_router_scratch_bytes(T, E) bytes of device memory (contents irrelevant).
extern "C" int b2f_moe_router_top1(const float* logits, int T, int E, int capacity, int* expert,
                                   int* slot, float* gate, float* probs, int* counts, float* me,
                                   float* ce, void* scratch, void* stream) {
  B2F_REQUIRE(T > 0 && E >= 1 && E <= MAX_E && capacity > 0, "router: need 1 <= E <= 64, capacity > 0");
  B2F_REQUIRE(scratch != nullptr, "router: scratch buffer required (b2f_moe_router_scratch_bytes)");
  const int nblk = (T + RT - 1) / RT;
  int* blk_cnt = static_cast<int*>(scratch);
  float* blk_me = reinterpret_cast<float*>(blk_cnt + (long long)nblk * E);
  router_pass1_kernel<<<nblk, RT, 0, (cudaStream_t)stream>>>(logits, T, E, expert, gate, probs, blk_cnt, blk_me);
  B2F_CHECK_CUDA(cudaGetLastError());
  router_pass2_kernel<<<nblk, RT, 0, (cudaStream_t)stream>>>(T, E, capacity, nblk, expert, slot, gate, blk_cnt,
                                                             blk_me, counts, me, ce);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// peer_ptrs == NULL (world 1): rows go to dst_local [E * capacity, H].  Otherwise they are stored
// into the owners' buffers laid out [world(src), experts_per_rank, capacity, H].
extern "C" int b2f_moe_dispatch(const void* x, const int* expert, const int* slot, int T, int H,
                                int capacity, void* dst_local, const long long* peer_ptrs, int world,
                                int experts_per_rank, int src_rank, void* stream) {
  B2F_REQUIRE(H % 8 == 0 && T > 0, "moe dispatch: H must be a multiple of 8");
  PeerRows p;
  if (int rc = fill_peers(&p, peer_ptrs, peer_ptrs ? world : 1)) return rc;
  moe_dispatch_kernel<<<row_grid(T), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, expert, slot, T, H, capacity, (__nv_bfloat16*)dst_local, p,
      peer_ptrs ? world : 1, experts_per_rank, src_rank);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_moe_combine(const void* src_local, const long long* peer_ptrs, int world,
                               int experts_per_rank, int src_rank, const int* expert, const int* slot,
                               const float* gate, int T, int H, int capacity, void* y, void* stream) {
  B2F_REQUIRE(H % 8 == 0 && T > 0, "moe combine: H must be a multiple of 8");
  PeerRows p;
  if (int rc = fill_peers(&p, peer_ptrs, peer_ptrs ? world : 1)) return rc;
  moe_combine_kernel<<<row_grid(T), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)src_local, p, peer_ptrs ? world : 1, experts_per_rank, src_rank, expert, slot,
      gate, T, H, capacity, (__nv_bfloat16*)y);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_moe_gate_grad(const void* rows, const int* expert, const int* slot, const void* dy,
                                 int T, int H, int capacity, float* dgate, void* stream) {
  B2F_REQUIRE(H % 8 == 0 && T > 0, "moe gate grad: H must be a multiple of 8");
  moe_gate_grad_kernel<<<row_grid(T), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)rows, expert, slot, (const __nv_bfloat16*)dy, T, H, capacity, dgate);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
