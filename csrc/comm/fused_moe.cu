// b200flow K5 — expert-parallel MoE FFN fused with its two all-to-alls over NVLink peer memory (sm_100a),
// block-scaled fp8 (MXFP8) on the wire and in the expert GEMMs.
//
// What DeepSpeed-MoE does as  einsum-dispatch -> all_to_all_single -> expert FFN on cuBLAS -> all_to_all_single ->
// einsum-combine  (SURVEY.md §2F "EP", §2H C13-C15) becomes two kernels and no barrier between them:
//
//   UP    every CTA first takes its share of the *dispatch*: for a 128-slot block of one expert's queue it gathers
//         the tokens, quantises them to e4m3 with one UE8M0 scale per 32 hidden elements, and ships token bytes
//         and scale atoms straight into the buffers of the GPU that owns the expert (512-byte bulk stores from a
//         shared-memory staging block), then publishes that block's flag there.  The same CTAs then run the expert
//         GEMM (tcgen05 kind::mxf8f6f4.block_scale, scales in tensor memory): the TMA producer of an M-tile waits
//         for exactly the flag of the 128 rows it is about to load, so tiles start as their tokens land and tiles
//         beyond an expert's token count are never computed.  The epilogue applies bias + GELU (forward; the
//         pre-activation is kept for backward) or multiplies by gelu'(pre) (backward), keeps the bf16 result the
//         weight-gradient GEMMs need, and quantises its 32-column chunk for the second GEMM.
//   DOWN  the second GEMM (again block-scaled fp8) whose epilogue is the *combine*: finished bf16 rows leave as one
//         bulk store per row into the [expert, slot, hidden] buffer of the GPU the token came from; a per-source
//         counter says when all rows are home.  The source then applies the gate and un-permutes locally.
//
// Per call and per direction, 1 byte + 1/32 byte per element crosses NVLink on the way out (bf16 before: 2 bytes)
// and 2 bytes on the way back.  Cross-call reuse: token buffers and flags are epoch-tagged or double-buffered by
// call parity (see parallel/moe.py for the argument that two parities suffice).
//
// Requires capacity % 128 == 0 and hidden % 512 == 0 (a 128-row tile then has a single source GPU and a staging
// block is 128 rows x 512 bytes); the layer falls back to the unfused path otherwise.
#include "../kernels/gemm_mx_core.cuh"
#include "../common/host.h"
#include "../../include/b200flow.h"

using namespace b2f;
using namespace b2f::gemm_mx;

namespace b2f {
namespace gemm_mx {
int make_mx_maps(CUtensorMap* ma, CUtensorMap* mb, const void* Aq, const void* Bq, int M, int N, int K, int batch);
}
}  // namespace b2f

extern "C" int b2f_mx_sf_atoms(int rows);
extern "C" long long b2f_mx_sf_bytes(int rows, int K);

namespace {

constexpr int MAX_WORLD = 8;
constexpr int MAX_E = 64;
constexpr int SLAB = 512;                       // hidden columns (= fp8 bytes) staged per step
constexpr int STG_BYTES = 128 * SLAB + 4 * 512; // 128 rows + 4 scale atoms
constexpr int UP_SMEM = gemm_mx::SMEM_BYTES;    // the dispatch phase reuses the operand ring as staging
static_assert(2 * STG_BYTES + 1024 <= UP_SMEM, "staging must fit in the GEMM's shared memory");

struct PeerPtrs {
  void* ptr[MAX_WORLD];
};

struct MoeGeom {
  int world, rank, El, C, R;     // R = world * C rows per local expert
  int tiles_per_src;             // C / 128
  uint32_t epoch;
};

__device__ __forceinline__ uint32_t cnt_tag(uint32_t epoch) { return (epoch & 0xFFFu) << 20; }

// Kept-token count that source `src` sent for local expert `le` (waits for this call's value to arrive).
__device__ __forceinline__ int wait_count(const uint32_t* cnt, const MoeGeom& g, int src, int le) {
  const uint32_t* p = cnt + src * g.El + le;
  const uint32_t tag = cnt_tag(g.epoch);
  long long t0 = clock64();
  uint32_t v;
  while (((v = ld_acquire_sys(p)) & 0xFFF00000u) != tag) {
    if (clock64() - t0 > B2F_SPIN_LIMIT_CYCLES) {
      printf("b200flow: MoE count wait timed out (block %d)\n", blockIdx.x);
      __trap();
    }
  }
  return (int)(v & 0xFFFFFu);
}

// ------------------------------------------------------------------------------------------------ UP
struct UpArgs {
  // dispatch side (this GPU's tokens)
  const __nv_bfloat16* x;        // [T, H]
  const int* token_of;           // [E][C] token index of a slot, -1 = empty
  const int* counts;             // [E] kept tokens per (global) expert
  PeerPtrs xq, xsf, flags, cnt;  // owners' buffers: u8 [El][R][H], u8 [El][H/128][atoms][512], u32 [El][R/128], u32 [world][El]
  int H, E;
  int atoms_x;                   // scale atoms per k-block of the token buffer
  long long xsf_batch;           // bytes per local expert
  // GEMM side (this GPU's experts)
  const uint32_t* my_flags;
  const uint32_t* my_cnt;
  const __nv_bfloat16* bias;     // [El][N] or nullptr
  __nv_bfloat16* keep;           // [El][R][N] bf16: forward = pre-activation, backward = d(pre-activation)
  const __nv_bfloat16* pre;      // backward: saved pre-activation [El][R][N]
  uint8_t* hq;                   // [El][R][N] e4m3 for the second GEMM
  uint8_t* hsf;                  // its scale atoms [El][N/128][atoms_h][512]
  int atoms_h;
  long long hsf_batch;
  int backward;                  // 0: h = gelu(acc + bias); 1: h = acc * gelu'(pre)
  MoeGeom g;
};

struct UpPolicy {
  UpArgs a;
  int num_n;
  __device__ __forceinline__ void map_tile(int t, int num_m, int nn, int& b, int& mb, int& nb) const {
    // all N tiles of one 128-row block back to back (the block's tokens stay in L2); row blocks in arrival order:
    // this GPU's own tokens first, then the peers' in ring order, expert by expert inside
    const int mi = t / nn;
    nb = t - mi * nn;
    const int per_src = a.g.El * a.g.tiles_per_src;
    const int si = mi / per_src;
    const int src = (a.g.rank + si) % a.g.world;
    const int r = mi - si * per_src;
    b = r / a.g.tiles_per_src;
    mb = src * a.g.tiles_per_src + (r - b * a.g.tiles_per_src);
  }
  __device__ __forceinline__ bool tile_active(int b, int mb, int) const {
    const int src = mb / a.g.tiles_per_src, j = mb - src * a.g.tiles_per_src;
    return j * 128 < wait_count(a.my_cnt, a.g, src, b);
  }
  __device__ __forceinline__ void wait_operands(int b, int mb, int) const {
    wait_flag_ge(a.my_flags + b * (a.g.R >> 7) + mb, a.g.epoch);
    asm volatile("fence.proxy.async;" ::: "memory");   // peers' stores -> this GPU's TMA (async proxy) reads
  }
  __device__ __forceinline__ void before_epilogue(int, int, int) const {}
  __device__ __forceinline__ void store_chunk(const Params& p, int b, int, int, int m, int n0,
                                              const uint32_t (&acc)[32]) const {
    if (m >= p.M) return;
    const long long row = (long long)b * p.M + m;
    float v[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(acc[i]);
    __nv_bfloat16* kp = a.keep + row * p.N + n0;
    if (!a.backward) {
      if (a.bias) {
        const uint4* bp = reinterpret_cast<const uint4*>(a.bias + (long long)b * p.N + n0);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint4 u = __ldg(bp + q);
          const float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y), f2 = unpack_bf16x2(u.z), f3 = unpack_bf16x2(u.w);
          v[q * 8 + 0] += f0.x; v[q * 8 + 1] += f0.y; v[q * 8 + 2] += f1.x; v[q * 8 + 3] += f1.y;
          v[q * 8 + 4] += f2.x; v[q * 8 + 5] += f2.y; v[q * 8 + 6] += f3.x; v[q * 8 + 7] += f3.y;
        }
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {      // keep the pre-activation (rounded to bf16, as the unfused path does)
        uint4 u;
        u.x = pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]); u.y = pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]);
        u.z = pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]); u.w = pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7]);
        *reinterpret_cast<uint4*>(kp + q * 8) = u;
      }
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = gemm::gelu_tanh(v[i]);
    } else {
      const uint4* zp = reinterpret_cast<const uint4*>(a.pre + row * p.N + n0);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint4 u = zp[q];
        const float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y), f2 = unpack_bf16x2(u.z), f3 = unpack_bf16x2(u.w);
        const float z[8] = {f0.x, f0.y, f1.x, f1.y, f2.x, f2.y, f3.x, f3.y};
#pragma unroll
        for (int i = 0; i < 8; ++i) v[q * 8 + i] *= gemm::gelu_tanh_grad(z[i]);
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {      // d(pre-activation) in bf16 for the weight and bias gradients
        uint4 u;
        u.x = pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]); u.y = pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]);
        u.z = pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]); u.w = pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7]);
        *reinterpret_cast<uint4*>(kp + q * 8) = u;
      }
    }
    // quantise the 32-column chunk: it is exactly one scale group of the second GEMM's K dimension
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) amax = fmaxf(amax, fabsf(v[i]));
    float inv;
    const uint32_t e = ue8m0_scale(amax, inv);
    uint32_t o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j)
      o[j] = cvt_e4m3x2(v[j * 4] * inv, v[j * 4 + 1] * inv) | (cvt_e4m3x2(v[j * 4 + 2] * inv, v[j * 4 + 3] * inv) << 16);
    uint4* hp = reinterpret_cast<uint4*>(a.hq + row * p.N + n0);
    hp[0] = make_uint4(o[0], o[1], o[2], o[3]);
    hp[1] = make_uint4(o[4], o[5], o[6], o[7]);
    a.hsf[(long long)b * a.hsf_batch + sf_offset(m, n0 >> 7, (n0 & 127) >> 5, a.atoms_h)] = (uint8_t)e;
  }
  __device__ __forceinline__ void tile_done(int, int, int) const {}
  __device__ __forceinline__ void finish() const {}
};

// Dispatch phase of one CTA: 128-slot blocks of the experts' queues, `first`, `first + stride`, ...
__device__ __forceinline__ void dispatch_phase(const UpArgs& a, uint8_t* smem, int first, int stride) {
  const MoeGeom& g = a.g;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (blockIdx.x == 0 && tid < a.E) {
    // every owner learns how many of this GPU's tokens each of its experts gets (tiles beyond are never computed)
    const int e = tid, dest = e / g.El, le = e - dest * g.El;
    st_release_sys(static_cast<uint32_t*>(a.cnt.ptr[dest]) + g.rank * g.El + le,
                   cnt_tag(g.epoch) | (uint32_t)a.counts[e]);
  }
  const int slabs = a.H / SLAB;
  int n = 0;            // staging-buffer use counter of this CTA (one bulk group per thread per use)
  int u = 0;            // running unit index over (expert in ring order, block)
  for (int i = 0; i < a.E; ++i) {
    const int e = (g.rank * g.El + i) % a.E;          // own experts first: their tiles feed the local GEMM soonest
    const int cnt = a.counts[e];
    const int blocks = (cnt + 127) >> 7;
    for (int j = 0; j < blocks; ++j, ++u) {
      if (u % stride != first) continue;
      const int dest = e / g.El, le = e - dest * g.El;
      const int valid = cnt - j * 128 < 128 ? cnt - j * 128 : 128;
      const int row0 = g.rank * g.C + j * 128;         // first destination row inside the expert's buffer
      uint8_t* dq = static_cast<uint8_t*>(a.xq.ptr[dest]) + ((long long)le * g.R + row0) * a.H;
      uint8_t* dsf = static_cast<uint8_t*>(a.xsf.ptr[dest]) + (long long)le * a.xsf_batch;
      const int* toks = a.token_of + (long long)e * g.C + j * 128;
      for (int s = 0; s < slabs; ++s, ++n) {
        uint8_t* stg = smem + (n & 1) * STG_BYTES;
        uint8_t* stsf = stg + 128 * SLAB;
        tma_store_wait_read<1>();                      // the bulk stores that read this buffer two uses ago are done
        asm volatile("bar.sync 2, 192;" ::: "memory");
        // quantise: warp w takes rows w, w + 6, ... four at a time; a lane holds 16 columns, two lanes one scale group
#pragma unroll 1
        for (int r0 = warp; r0 < 128; r0 += 24) {
          uint4 in[4][2];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int r = r0 + k * 6;
            const int tok = (r < valid) ? toks[r] : -1;
            if (r < 128 && tok >= 0) {
              const uint4* src = reinterpret_cast<const uint4*>(a.x + (long long)tok * a.H + s * SLAB + lane * 16);
              in[k][0] = ld_nc_v4(src);
              in[k][1] = ld_nc_v4(src + 1);
            } else {
              in[k][0] = in[k][1] = make_uint4(0, 0, 0, 0);
            }
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int r = r0 + k * 6;
            if (r >= 128) break;                       // warp-uniform
            float v[16];
            float amax = 0.f;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const uint32_t w[4] = {in[k][h].x, in[k][h].y, in[k][h].z, in[k][h].w};
#pragma unroll
              for (int c = 0; c < 4; ++c) {
                const float2 f = unpack_bf16x2(w[c]);
                v[h * 8 + c * 2] = f.x;
                v[h * 8 + c * 2 + 1] = f.y;
                amax = fmaxf(amax, fmaxf(fabsf(f.x), fabsf(f.y)));
              }
            }
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
            float inv;
            const uint32_t e8 = ue8m0_scale(amax, inv);
            uint4 o;
            o.x = cvt_e4m3x2(v[0] * inv, v[1] * inv) | (cvt_e4m3x2(v[2] * inv, v[3] * inv) << 16);
            o.y = cvt_e4m3x2(v[4] * inv, v[5] * inv) | (cvt_e4m3x2(v[6] * inv, v[7] * inv) << 16);
            o.z = cvt_e4m3x2(v[8] * inv, v[9] * inv) | (cvt_e4m3x2(v[10] * inv, v[11] * inv) << 16);
            o.w = cvt_e4m3x2(v[12] * inv, v[13] * inv) | (cvt_e4m3x2(v[14] * inv, v[15] * inv) << 16);
            *reinterpret_cast<uint4*>(stg + r * SLAB + lane * 16) = o;
            if ((lane & 1) == 0)   // k-block lane / 8 of this slab, scale group (lane % 8) / 2 inside it
              stsf[(lane >> 3) * 512 + (r & 31) * 16 + (r >> 5) * 4 + ((lane & 7) >> 1)] = (uint8_t)e8;
          }
        }
        fence_proxy_async();                           // staging writes -> bulk (async proxy) reads
        asm volatile("bar.sync 2, 192;" ::: "memory");
        if (tid < 128) {
          // rows beyond `valid` travel as zeros: reductions over the owner's rows (weight gradients) may include them
          bulk_store_1d(dq + (long long)tid * a.H + s * SLAB, stg + tid * SLAB, SLAB);
        } else if (tid < 132) {
          const int kb = s * (SLAB / 128) + (tid - 128);
          bulk_store_1d(dsf + ((long long)kb * a.atoms_x + (row0 >> 7)) * 512, stsf + (tid - 128) * 512, 512);
        }
        tma_store_commit();
      }
      // the block is complete on the owner once every thread's bulk stores have landed
      tma_store_wait<0>();
      asm volatile("fence.proxy.async;" ::: "memory");
      asm volatile("bar.sync 2, 192;" ::: "memory");
      if (tid == 0) {
        __threadfence_system();
        st_release_sys(static_cast<uint32_t*>(a.flags.ptr[dest]) + le * (g.R >> 7) + (row0 >> 7), g.epoch);
      }
    }
  }
  tma_store_wait<0>();
  asm volatile("bar.sync 2, 192;" ::: "memory");       // staging is about to become the operand ring
}

__global__ void __launch_bounds__(NUM_THREADS, 1)
b2f_moe_up_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                  const gemm::Params p, const MxScales sc, const UpArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  race_delay();
  if (a.x != nullptr) dispatch_phase(a, smem, blockIdx.x, gridDim.x);
  UpPolicy pol{a, (p.N + BN - 1) / BN};
  gemm_mx_body(&tma_a, &tma_b, p, sc, pol, blockIdx.x, gridDim.x, smem_raw);
}

// ---------------------------------------------------------------------------------------------- DOWN
constexpr int DN_STAGES = 4;                         // one ring stage traded for the 48 KB row staging block
constexpr int DN_PITCH = BN * 2;                     // 384 B of bf16 per staged row
constexpr int DN_SMEM = DN_STAGES * (STAGE_BYTES + SF_STAGE_BYTES) + 1024 + 256 + BM * DN_PITCH;
static_assert(DN_SMEM <= 227 * 1024, "shared memory budget of the down kernel");

struct DownArgs {
  const uint32_t* my_cnt;        // [world][El] counts received in the UP kernel of this call (already tagged)
  const __nv_bfloat16* bias;     // [El][N] or nullptr
  PeerPtrs rows;                 // sources' buffers: bf16 [E][C][N] (expert-major, slot order)
  PeerPtrs done;                 // sources' u32 counter (zeroed for this call's parity)
  MoeGeom g;
  int E;
};

struct DownPolicy {
  DownArgs a;
  uint8_t* staging;
  int N;
  mutable int pend_src = -1;
  __device__ __forceinline__ void map_tile(int t, int num_m, int nn, int& b, int& mb, int& nb) const {
    const int mi = t / nn;
    nb = t - mi * nn;
    const int per_src = a.g.El * a.g.tiles_per_src;
    const int si = mi / per_src;
    const int src = (a.g.rank + 1 + si) % a.g.world;   // peers first: their rows have the longest way home
    const int r = mi - si * per_src;
    b = r / a.g.tiles_per_src;
    mb = src * a.g.tiles_per_src + (r - b * a.g.tiles_per_src);
  }
  __device__ __forceinline__ int count_of(int b, int mb, int& src, int& j) const {
    src = mb / a.g.tiles_per_src;
    j = mb - src * a.g.tiles_per_src;
    return (int)(a.my_cnt[src * a.g.El + b] & 0xFFFFFu);
  }
  __device__ __forceinline__ bool tile_active(int b, int mb, int) const {
    int src, j;
    const int cnt = count_of(b, mb, src, j);
    return j * 128 < cnt;
  }
  __device__ __forceinline__ void wait_operands(int, int, int) const {}
  // the counter of a shipped tile is bumped one tile later, when its bulk stores have long landed
  __device__ __forceinline__ void flush() const {
    if (pend_src < 0) return;
    tma_store_wait<0>();
    asm volatile("fence.proxy.async;" ::: "memory");
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (threadIdx.x == 64) {
      __threadfence_system();
      red_add_release_sys(static_cast<uint32_t*>(a.done.ptr[pend_src]), 1u);
    }
    pend_src = -1;
  }
  __device__ __forceinline__ void before_epilogue(int, int, int) const {
    if ((threadIdx.x & 31) == 0) race_delay();
    __syncwarp();
    flush();
    tma_store_wait_read<0>();                  // this thread's staging row of the previous tile has been read out
  }
  __device__ __forceinline__ void store_chunk(const Params& p, int b, int mb, int nb, int m, int n0,
                                              const uint32_t (&acc)[32]) const {
    float v[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(acc[i]);
    if (a.bias) {
      const uint4* bp = reinterpret_cast<const uint4*>(a.bias + (long long)b * p.N + n0);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint4 u = __ldg(bp + q);
        const float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y), f2 = unpack_bf16x2(u.z), f3 = unpack_bf16x2(u.w);
        v[q * 8 + 0] += f0.x; v[q * 8 + 1] += f0.y; v[q * 8 + 2] += f1.x; v[q * 8 + 3] += f1.y;
        v[q * 8 + 4] += f2.x; v[q * 8 + 5] += f2.y; v[q * 8 + 6] += f3.x; v[q * 8 + 7] += f3.y;
      }
    }
    uint8_t* dst = staging + (size_t)(m - mb * BM) * DN_PITCH + (size_t)(n0 - nb * BN) * 2;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint4 u;
      u.x = pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]); u.y = pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]);
      u.z = pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]); u.w = pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7]);
      *reinterpret_cast<uint4*>(dst + q * 16) = u;
    }
  }
  __device__ __forceinline__ void tile_done(int b, int mb, int nb) const {
    // epilogue thread = accumulator row = slot: its finished row goes home with one bulk store
    int src, j;
    const int cnt = count_of(b, mb, src, j);
    const int r = (int)threadIdx.x - 64;
    const int row_in_tile = (((r >> 5) + 2) & 3) * 32 + (r & 31);     // TMEM lane quarter of this warp
    const int slot = j * 128 + row_in_tile;
    const int n0 = nb * BN;
    const int cols = N - n0 < BN ? N - n0 : BN;
    asm volatile("fence.proxy.async;" ::: "memory");
    if (slot < cnt) {
      __nv_bfloat16* home = static_cast<__nv_bfloat16*>(a.rows.ptr[src]) +
                            ((long long)(a.g.rank * a.g.El + b) * a.g.C + slot) * N + n0;
      bulk_store_1d(home, staging + (size_t)row_in_tile * DN_PITCH, (uint32_t)cols * 2);
    }
    tma_store_commit();
    pend_src = src;
  }
  __device__ __forceinline__ void finish() const {
    flush();
    tma_store_wait<0>();
  }
};

__global__ void __launch_bounds__(NUM_THREADS, 1)
b2f_moe_down_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                    const gemm::Params p, const MxScales sc, const DownArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  DownPolicy pol{a, smem + DN_STAGES * (STAGE_BYTES + SF_STAGE_BYTES) + 1024, p.N};
  gemm_mx_body<DownPolicy, DN_STAGES>(&tma_a, &tma_b, p, sc, pol, blockIdx.x, gridDim.x, smem_raw);
}

// One thread waits until every tile of this GPU's tokens has come home (expected = column tiles x kept 128-slot
// blocks over all experts); what follows on the stream then reads complete rows.
__global__ void moe_wait_rows_kernel(const uint32_t* done, const int* counts, int E, int num_n) {
  int expect = 0;
  for (int e = 0; e < E; ++e) expect += (counts[e] + 127) >> 7;
  expect *= num_n;
  wait_flag_ge(done, (uint32_t)expect);
  __threadfence_system();
}

// slot -> token map of one rank's routing decision: token_of[e][slot] = t (buffer pre-filled with -1).
__global__ void moe_token_of_kernel(const int* __restrict__ expert, const int* __restrict__ slot, int T, int C,
                                    int* __restrict__ token_of) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const int e = expert[t];
  if (e >= 0) token_of[(long long)e * C + slot[t]] = t;
}

// Zero the rows of inactive 128-slot blocks of a bf16 [El][R][N] buffer (the fused kernels never touch them) so that
// reductions over all rows (weight and bias gradients) see zeros there.
__global__ void moe_zero_inactive_kernel(__nv_bfloat16* buf, const uint32_t* cnt, int El, int world, int C, int N) {
  const int tps = C >> 7;
  const int blocks = El * world * tps;
  for (int blk = blockIdx.x; blk < blocks; blk += gridDim.x) {
    const int le = blk / (world * tps);
    const int r = blk - le * world * tps;
    const int src = r / tps, j = r - src * tps;
    const int c = (int)(cnt[src * El + le] & 0xFFFFFu);
    int first = c - j * 128;                 // rows [first, 128) of this block hold no token
    first = first < 0 ? 0 : first;
    if (first >= 128) continue;
    uint4* base = reinterpret_cast<uint4*>(buf + ((long long)le * world * C + (long long)src * C + j * 128 + first) * N);
    const long long n16 = (long long)(128 - first) * N / 8;
    for (long long i = threadIdx.x; i < n16; i += blockDim.x) base[i] = make_uint4(0, 0, 0, 0);
  }
}

int set_smem_attr(const void* fn, int bytes, bool* done) {
  int dev = 0;
  B2F_CHECK_CUDA(cudaGetDevice(&dev));
  if (!done[dev & 63]) {
    B2F_CHECK_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    done[dev & 63] = true;
  }
  return 0;
}

void fill_peers(PeerPtrs& p, const long long* src, int world) {
  for (int i = 0; i < MAX_WORLD; ++i) p.ptr[i] = i < world ? reinterpret_cast<void*>(src[i]) : nullptr;
}

}  // namespace

extern "C" int b2f_moe_token_of(const int* expert, const int* slot, int T, int C, int* token_of, void* stream) {
  moe_token_of_kernel<<<(T + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(expert, slot, T, C, token_of);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// UP: dispatch this GPU's tokens (x [T, H]; pass x = NULL to skip, e.g. when the rows were placed by other means) and
// run  keep, hq = epilogue( xq (x) Wq^T )  over this GPU's experts.  Peer arrays hold `world` device pointers.
//   Wq / wsf : [El][N, H] e4m3 + scale atoms (b2f_quantize_mx of the weight with batch = El)
//   xq / xsf : symmetric token buffers (this call's parity), flags u32 [El][R/128], cnt u32 [world][El]
extern "C" int b2f_moe_up(const void* x, const int* token_of, const int* counts, int T, int H, int E, int C,
                          const long long* xq_ptrs, const long long* xsf_ptrs, const long long* flag_ptrs,
                          const long long* cnt_ptrs, const void* Wq, const void* wsf, int N, const void* bias,
                          void* keep, const void* pre, void* hq, void* hsf, int backward, int rank, int world,
                          unsigned epoch, void* stream) {
  B2F_REQUIRE(world >= 1 && world <= MAX_WORLD && E % world == 0 && E <= MAX_E, "moe_up: bad expert-parallel geometry");
  B2F_REQUIRE(C % 128 == 0 && H % SLAB == 0 && N % 128 == 0, "moe_up: capacity % 128, hidden % 512, N % 128 must be 0");
  MoeGeom g;
  g.world = world; g.rank = rank; g.El = E / world; g.C = C; g.R = world * C; g.tiles_per_src = C / 128;
  g.epoch = epoch;
  UpArgs a{};
  a.x = static_cast<const __nv_bfloat16*>(x);
  a.token_of = token_of; a.counts = counts;
  fill_peers(a.xq, xq_ptrs, world); fill_peers(a.xsf, xsf_ptrs, world);
  fill_peers(a.flags, flag_ptrs, world); fill_peers(a.cnt, cnt_ptrs, world);
  a.H = H; a.E = E;
  a.atoms_x = b2f_mx_sf_atoms(g.R);
  a.xsf_batch = b2f_mx_sf_bytes(g.R, H);
  a.my_flags = reinterpret_cast<const uint32_t*>(flag_ptrs[rank]);
  a.my_cnt = reinterpret_cast<const uint32_t*>(cnt_ptrs[rank]);
  a.bias = static_cast<const __nv_bfloat16*>(bias);
  a.keep = static_cast<__nv_bfloat16*>(keep);
  a.pre = static_cast<const __nv_bfloat16*>(pre);
  a.hq = static_cast<uint8_t*>(hq); a.hsf = static_cast<uint8_t*>(hsf);
  a.atoms_h = b2f_mx_sf_atoms(g.R);
  a.hsf_batch = b2f_mx_sf_bytes(g.R, N);
  a.backward = backward;
  a.g = g;
  B2F_REQUIRE(!backward || pre != nullptr, "moe_up: the backward epilogue needs the saved pre-activation");
  gemm::Params p{};
  p.M = g.R; p.N = N; p.K = H; p.batch = g.El; p.alpha = 1.f;
  MxScales sc;
  sc.sfa = reinterpret_cast<const uint8_t*>(xsf_ptrs[rank]);
  sc.sfb = static_cast<const uint8_t*>(wsf);
  sc.a_atoms = a.atoms_x; sc.b_atoms = b2f_mx_sf_atoms(N);
  sc.sfa_batch = a.xsf_batch; sc.sfb_batch = b2f_mx_sf_bytes(N, H);
  CUtensorMap ma, mb;   // both maps cover this GPU's memory only
  if (int rc = make_mx_maps(&ma, &mb, reinterpret_cast<const void*>(xq_ptrs[rank]), Wq, g.R, N, H, g.El)) return rc;
  static bool attr[64] = {false};
  if (int rc = set_smem_attr(reinterpret_cast<const void*>(b2f_moe_up_kernel), UP_SMEM, attr)) return rc;
  b2f_moe_up_kernel<<<sm_count(), NUM_THREADS, UP_SMEM, static_cast<cudaStream_t>(stream)>>>(ma, mb, p, sc, a);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// DOWN: rows_home[src][(rank * El + le) * C + slot, :] = hq[le][src * C + slot, :] (x) Wq[le]^T (+bias) for every kept
// slot, written straight into the source GPU's buffer; every finished tile bumps the source's `done` counter.
extern "C" int b2f_moe_down(const void* hq, const void* hsf, const void* Wq, const void* wsf, int K, int N,
                            const void* bias, const long long* rows_ptrs, const long long* done_ptrs,
                            const void* my_cnt, int E, int C, int rank, int world, unsigned epoch, void* stream) {
  B2F_REQUIRE(world >= 1 && world <= MAX_WORLD && E % world == 0, "moe_down: bad expert-parallel geometry");
  B2F_REQUIRE(C % 128 == 0 && K % 128 == 0 && N % 8 == 0, "moe_down: capacity % 128, K % 128, N % 8 must be 0");
  MoeGeom g;
  g.world = world; g.rank = rank; g.El = E / world; g.C = C; g.R = world * C; g.tiles_per_src = C / 128;
  g.epoch = epoch;
  DownArgs a{};
  a.my_cnt = static_cast<const uint32_t*>(my_cnt);
  a.bias = static_cast<const __nv_bfloat16*>(bias);
  fill_peers(a.rows, rows_ptrs, world); fill_peers(a.done, done_ptrs, world);
  a.g = g; a.E = E;
  gemm::Params p{};
  p.M = g.R; p.N = N; p.K = K; p.batch = g.El; p.alpha = 1.f;
  MxScales sc;
  sc.sfa = static_cast<const uint8_t*>(hsf);
  sc.sfb = static_cast<const uint8_t*>(wsf);
  sc.a_atoms = b2f_mx_sf_atoms(g.R); sc.b_atoms = b2f_mx_sf_atoms(N);
  sc.sfa_batch = b2f_mx_sf_bytes(g.R, K); sc.sfb_batch = b2f_mx_sf_bytes(N, K);
  CUtensorMap ma, mb;
  if (int rc = make_mx_maps(&ma, &mb, hq, Wq, g.R, N, K, g.El)) return rc;
  static bool attr[64] = {false};
  if (int rc = set_smem_attr(reinterpret_cast<const void*>(b2f_moe_down_kernel), DN_SMEM, attr)) return rc;
  b2f_moe_down_kernel<<<sm_count(), NUM_THREADS, DN_SMEM, static_cast<cudaStream_t>(stream)>>>(ma, mb, p, sc, a);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_moe_wait_rows(const void* done, const int* counts, int E, int N, void* stream) {
  moe_wait_rows_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const uint32_t*>(done), counts, E,
                                                                        (N + BN - 1) / BN);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_moe_zero_inactive(void* buf, const void* my_cnt, int El, int world, int C, int N, void* stream) {
  B2F_REQUIRE(C % 128 == 0 && N % 8 == 0, "moe_zero_inactive: capacity % 128 and N % 8 must be 0");
  moe_zero_inactive_kernel<<<sm_count() * 2, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<__nv_bfloat16*>(buf), static_cast<const uint32_t*>(my_cnt), El, world, C, N);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
