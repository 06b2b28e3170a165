// b200flow K7 — flash attention forward + backward on tcgen05 / TMEM (sm_100a).
//
// Layout: q/k/v/o are [batch, seq, heads, D] views (token stride and batch stride
// free, heads packed, D in {64, 128}), so the kernels read straight out of a
// fused QKV projection buffer.  128x128 score tiles.
//
// Forward (one CTA per 256 query rows of one head, 384 threads): warpgroup 0 = TMA
// warp + MMA warp, warpgroups 1 and 2 = softmax of the two 128-row query tiles,
// which ping-pong on the tensor core.  One query row per thread with all 128
// scores in registers (no cross-thread reductions); the bf16 probabilities are
// written back over S in TMEM and the P V product consumes them as a TMEM
// A-operand (tcgen05.mma TS form), so P never touches shared memory.  O
// accumulates in TMEM; the running max is raised lazily (see the kernel).
//
// Backward is a statistics pre-pass plus two kernels that recompute P from the
// saved log-sum-exp, software-pipelined over half score tiles (128 x 64):
//   dQ  kernel (CTA per query tile):  S, dP -> dS in TMEM -> dQ += dS K
//   dKV kernel (CTA per key tile):    S^T, dP^T -> P^T, dS^T in TMEM ->
//                                     dV += P^T dO ; dK += dS^T Q
// Both keep their accumulators in TMEM for the whole loop: no atomics, and the
// result is deterministic.
//
// The MMA warp walks its loop as a whole and one elect.sync lane issues, with
// descriptors built as a constant high word + an incremented low word: a
// 128x128x16 MMA lasts 64 cycles, so the issue sequence must be shorter than that.
//
// Capability parity: flash-attn v2 / Megatron scaled_(upper_triang_)masked_softmax
// attention core.
#include <stdlib.h>

#include <type_traits>

#include "../common/ptx.cuh"
#include "../common/host.h"
#include "../../include/b200flow.h"

using namespace b2f;

namespace {

constexpr int TQ = 128;           // rows per tile (queries or keys)
constexpr int CHUNK_B = 128 * 128;  // one 128-row x 64-col bf16 chunk, 128 B rows = 16 KB
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

struct AttnParams {
  int B, H, HKV, S;
  float scale, scale_log2;
  int causal;
  const int* kv_lens;  // [B] valid key length per batch row, or null
  // forward output / backward inputs
  __nv_bfloat16* O;
  long long o_ts, o_bs;
  float* lse;    // [B, H, S] natural-log LSE (+inf for rows with no valid key)
  float* delta;  // [B, H, S] rowsum(dO * O)
  // backward outputs
  __nv_bfloat16 *dQ, *dK, *dV;
  long long dq_ts, dq_bs, dk_ts, dk_bs, dv_ts, dv_bs;
  int dbg_mode;  // B2F_ATTN_DEBUG builds only: 1 = softmax does no math, 2 = also no P V, 3 = also no Q K^T
  // attention dropout (Megatron --attention-dropout): probability element (b, h, q, k) is kept iff its byte of a
  // counter-based hash word is >= drop_thr (= round(256 p)); kept elements are scaled by drop_rinv = 1 / (1 - p)
  uint32_t drop_thr;
  float drop_rinv;
  uint32_t drop_seed;
};

// Dropout mask: one 32-bit word per (head, query, group of 4 consecutive keys), one byte per key.  A function of the
// indices only, so the forward kernel (query rows), the dQ kernel (query rows) and the dK/dV kernel (key rows)
// regenerate the same mask with no memory traffic, and a recomputed forward (activation checkpointing) sees it again.
__device__ __forceinline__ uint32_t drop_mix(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}
__device__ __forceinline__ uint32_t drop_seed_bh(uint32_t seed, int bh) { return drop_mix(seed + (uint32_t)bh * 0x632BE5ABu); }
__device__ __forceinline__ uint32_t drop_seed_q(uint32_t seed_bh, int q) { return seed_bh + (uint32_t)q * 0x9E3779B1u; }
__device__ __forceinline__ uint32_t drop_word(uint32_t seed_q, int k4) { return drop_mix(seed_q + (uint32_t)k4 * 0x85EBCA77u); }

template <int D>
struct Cfg {
  static constexpr int TILE_B = TQ * D * 2;
};

// UMMA shared-memory descriptors, split so the MMA-issuing thread only adds a constant to the
// low word per K step (a 128x128x16 MMA lasts 64 cycles: the issue loop has to be shorter).
//   high word: SBO = 1024 B, descriptor version 1, SWIZZLE_128B;  low word: address >> 4 | LBO << 16
constexpr uint32_t DESC_HI = (1024u >> 4) | (1u << 14) | (2u << 29);
__device__ __forceinline__ uint64_t mk_desc(uint32_t lo) {
  uint64_t d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(DESC_HI));
  return d;
}
// K-major operand (rows = M/N index, contraction along the 128 B rows): LBO unused (16 B).
__device__ __forceinline__ uint32_t lo_k(uint32_t tile) { return ((tile & 0x3FFFF) >> 4) | (1u << 16); }
__device__ __forceinline__ uint64_t desc_k(uint32_t lo, int k) {
  return mk_desc(lo + (k >> 2) * (CHUNK_B >> 4) + (k & 3) * 2);
}
// MN-major operand (rows = contraction index, 16 rows per UMMA_K step): LBO = chunk stride.
__device__ __forceinline__ uint32_t lo_mn(uint32_t tile) {
  return ((tile & 0x3FFFF) >> 4) | (static_cast<uint32_t>(CHUNK_B >> 4) << 16);
}
__device__ __forceinline__ uint64_t desc_mn(uint32_t lo, int k) { return mk_desc(lo + k * (2048 >> 4)); }

// Every kernel here allocates all 512 TMEM columns, so the allocation starts at column 0; the
// kernels check that once and then use literal TMEM addresses (keeps them in uniform registers).
__device__ __forceinline__ void require_tmem_base0(uint32_t base) {
  if (base != 0) {
    if (threadIdx.x == 0) printf("b200flow: attention expects TMEM base 0, got %u\n", base);
    __trap();
  }
}

template <int D>
__device__ __forceinline__ void load_tile(uint8_t* dst, const CUtensorMap* m, uint64_t* bar,
                                          int row0, int head, int batch) {
#pragma unroll
  for (int c = 0; c < D / 64; ++c) tma_load_4d(dst + c * CHUNK_B, m, bar, c * 64, head, row0, batch);
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ uint8_t* align1024(uint8_t* p) {
  return reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(p) + 1023) &
                                    ~static_cast<uintptr_t>(1023));
}

// ============================================================== forward
// Two query tiles per CTA, ping-ponged on the tensor core.
// warp 0 = TMA, warp 1 = MMA issuer, warps 4-7 = softmax of query tile A,
// warps 8-11 = softmax of tile B (registers moved to them with setmaxnreg).
// While one group exponentiates its scores the MMA pipe runs the other tile's Q K^T / P V.  Each softmax thread owns one query
// row and holds all 128 scores in registers (one TMEM read).  O accumulates in TMEM
// across key tiles; the running max is only raised when it grows by more than 2^8
// (then that warp rescales its O rows in TMEM), so most tiles skip the rescale.
#ifdef B2F_ATTN_DEBUG
// cycle counters summed over CTAs (read with b2f_attn_debug): softmax-A lane 0: [0] loop total,
// [1] waiting for scores, [2] TMEM read + max, [3] exp + P store, [4] tiles; MMA thread: [5] loop total,
// [6] waiting for P, [7] waiting for K/V.
__device__ unsigned long long g_attn_dbg[16];
#define DBG_T(var) const long long var = clock64()
#define DBG_ADD(i, v) do { if ((threadIdx.x & 31) == 0) atomicAdd(&g_attn_dbg[i], (unsigned long long)(v)); } while (0)
#else
#define DBG_T(var)
#define DBG_ADD(i, v)
#endif
constexpr int FWD2_THREADS = 384;  // warpgroup 0: TMA + MMA (+2 idle warps); 1, 2: softmax A, B
constexpr float RESCALE_THRESHOLD = 8.f;
constexpr int ATTN_POLY_DEFAULT = 0;   // pairs of four whose exp2 runs on the FMA pipe (B2F_ATTN_POLY overrides)

struct Fwd2Bars {
  uint64_t q_full[2], k_full[2], v_full[2], k_empty[2], v_empty[2], s_full[2], o_full[2];
  uint64_t p_full[2][4];   // per query tile, per 32-key quarter of P: the P V product starts on the first quarter
  uint32_t tmem_base;
};

// POLY: of every four pairs of probabilities of an unmasked tile, this many take their exp2 on the FMA pipe
// (ex2_poly2) and the rest on the special-function unit; masked tiles (diagonal, sequence end) use the unit only,
// so that masked elements stay exact zeros.
template <int D, bool DROP = false, int POLY = 0>
__global__ void __launch_bounds__(FWD2_THREADS, 1)
attn_fwd2_kernel(const __grid_constant__ CUtensorMap mq, const __grid_constant__ CUtensorMap mk,
                 const __grid_constant__ CUtensorMap mv, const AttnParams p) {
  DBG_T(t_cta0);
#ifdef B2F_ATTN_DEBUG
  unsigned long long ns0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns0));
#endif
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  constexpr int TILE_B = Cfg<D>::TILE_B;
  uint8_t* sQ = smem;                   // 2 query tiles
  uint8_t* sK = smem + 2 * TILE_B;      // 2 stages
  uint8_t* sV = smem + 4 * TILE_B;      // 2 stages
  Fwd2Bars* bars = reinterpret_cast<Fwd2Bars*>(smem + 6 * TILE_B);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nblk = (p.S + 2 * TQ - 1) / (2 * TQ);
  // grid: x = (head, batch) fastest, y = query block, longest (last) causal blocks first, so
  // CTAs are handed out longest-job-first across the whole launch
  const int q_blk = p.causal ? nblk - 1 - (int)blockIdx.y : (int)blockIdx.y;
  const int h = blockIdx.x % p.H, b = blockIdx.x / p.H;
  const int hkv = h / (p.H / p.HKV);
  const int q0 = q_blk * 2 * TQ;
  int kv_len = p.kv_lens ? p.kv_lens[b] : p.S;
  kv_len = kv_len < p.S ? kv_len : p.S;
  // key tiles needed by each query tile (tile B may lie wholly past the sequence end)
  int n_w[2];
#pragma unroll
  for (int w = 0; w < 2; ++w) {
    const int qs = q0 + w * TQ;
    int kv_end = p.causal ? (kv_len < qs + TQ ? kv_len : qs + TQ) : kv_len;
    int nn = (kv_end + TQ - 1) / TQ;
    if (nn < 1) nn = 1;
    n_w[w] = qs < p.S ? nn : 0;
  }
  const int nA = n_w[0], nB = n_w[1];
  const int n = nA > nB ? nA : nB;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&mq); tma_prefetch_desc(&mk); tma_prefetch_desc(&mv);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->q_full[i], 1);
      mbar_init(&bars->k_full[i], 1);
      mbar_init(&bars->v_full[i], 1);
      mbar_init(&bars->k_empty[i], 1);
      mbar_init(&bars->v_empty[i], 1);
      mbar_init(&bars->s_full[i], 1);
      for (int c = 0; c < 4; ++c) mbar_init(&bars->p_full[i][c], 128);
      mbar_init(&bars->o_full[i], 1);
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(&bars->tmem_base, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  require_tmem_base0(bars->tmem_base);
  constexpr uint32_t tmem = 0;
#ifdef B2F_ATTN_DEBUG
  if (threadIdx.x == 0) { DBG_T(t_pro); DBG_ADD(9, t_pro - t_cta0); }
#endif
  // columns: S_A [0,128) S_B [128,256) O_A [256,256+D) O_B [384,384+D)

  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 72;");
    if (warp == 0 && lane == 0) {
      for (int w = 0; w < 2; ++w)
        if (n_w[w] > 0) {
          mbar_arrive_expect_tx(&bars->q_full[w], TILE_B);
          load_tile<D>(sQ + w * TILE_B, &mq, &bars->q_full[w], q0 + w * TQ, h, b);
        }
      for (int j = 0; j < n; ++j) {
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        mbar_wait(&bars->k_empty[st], ph ^ 1);
        mbar_arrive_expect_tx(&bars->k_full[st], TILE_B);
        load_tile<D>(sK + st * TILE_B, &mk, &bars->k_full[st], j * TQ, hkv, b);
        mbar_wait(&bars->v_empty[st], ph ^ 1);
        mbar_arrive_expect_tx(&bars->v_full[st], TILE_B);
        load_tile<D>(sV + st * TILE_B, &mv, &bars->v_full[st], j * TQ, hkv, b);
      }
    } else if (warp == 1) {
      // all 32 lanes walk the loop (uniform control flow keeps the issue sequence in uniform
      // registers); one elected lane issues the MMAs and commits
      const uint32_t idesc_s = make_idesc(B2F_MMA_FMT, B2F_MMA_FMT, 128, 128, 0, 0);
      const uint32_t idesc_pv = make_idesc(B2F_MMA_FMT, B2F_MMA_FMT, 128, D, 0, 1);
      // S_w(j) = Q_w K_j^T ; the last reader of K_j releases its stage.
      auto issue_s = [&](int w, int j) {
        const int st = j & 1;
        const uint32_t aQ = lo_k(smem_u32(sQ + w * TILE_B)), aK = lo_k(smem_u32(sK + st * TILE_B));
        const bool last_reader = (w == 1) || j >= nB;
        if (elect_one()) {
#ifdef B2F_ATTN_DEBUG
          if (p.dbg_mode != 3)
#endif
#pragma unroll
          for (int k = 0; k < D / 16; ++k)
            umma_f16_ss(tmem + w * 128, desc_k(aQ, k), desc_k(aK, k), idesc_s, k > 0);
          umma_commit(&bars->s_full[w]);
          if (last_reader) umma_commit(&bars->k_empty[st]);
        }
        __syncwarp();
      };
      // O_w += P_w(j) V_j, one quarter of P (32 keys, two K=16 MMAs) at a time as the softmax warps publish it
      auto issue_pv = [&](int w, int j, int c) {
        const int st = j & 1;
        const uint32_t aV = lo_mn(smem_u32(sV + st * TILE_B));
        const bool last_reader = (w == 1) || j >= nB;
        if (elect_one()) {
#ifdef B2F_ATTN_DEBUG
          if (p.dbg_mode != 2)
#endif
#pragma unroll
          for (int k = 2 * c; k < 2 * c + 2; ++k)
            umma_f16_ts(tmem + 256 + w * 128, tmem + w * 128 + k * 8, desc_mn(aV, k), idesc_pv,
                        (j | k) != 0);
          if (c == 3) {
            if (last_reader) umma_commit(&bars->v_empty[st]);
            if (j + 1 == n_w[w]) umma_commit(&bars->o_full[w]);
          }
        }
        __syncwarp();
      };
      mbar_wait(&bars->k_full[0], 0);
      tc_fence_after();
      if (nA > 0) { mbar_wait(&bars->q_full[0], 0); tc_fence_after(); issue_s(0, 0); }
      if (nB > 0) { mbar_wait(&bars->q_full[1], 0); tc_fence_after(); issue_s(1, 0); }
      DBG_T(t_loop0);
      for (int j = 0; j < n; ++j) {
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
#pragma unroll
        for (int w = 0; w < 2; ++w) {
          if (j >= n_w[w]) continue;
          DBG_T(t0);
          mbar_wait(&bars->p_full[w][0], j & 1);
          DBG_T(t1);
          mbar_wait(&bars->v_full[st], ph);
          tc_fence_after();
          DBG_T(t2);
          issue_pv(w, j, 0);
#pragma unroll
          for (int c = 1; c < 4; ++c) {
            mbar_wait(&bars->p_full[w][c], j & 1);
            tc_fence_after();
            issue_pv(w, j, c);
          }
          DBG_T(t3);
          if (j + 1 < n_w[w]) {
            mbar_wait(&bars->k_full[st ^ 1], ((j + 1) >> 1) & 1);
            tc_fence_after();
            DBG_T(t4);
            issue_s(w, j + 1);
            DBG_ADD(7, t4 - t3);
          }
          DBG_ADD(6, t1 - t0);
          DBG_ADD(7, t2 - t1);
        }
      }
      DBG_T(t_loop1);
      DBG_ADD(5, t_loop1 - t_loop0);
    }
    __syncwarp();
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
    const int w = (warp - 4) >> 2;   // 0: tile A, 1: tile B
    const int qw = warp & 3;         // TMEM lane quadrant this warp may touch
    const int r = qw * 32 + lane;
    const int qs = q0 + w * TQ;
    const int qi = qs + r;
    const int nw = n_w[w];
    const uint32_t lane_base = static_cast<uint32_t>(qw * 32) << 16;
    const uint32_t T_S = tmem + lane_base + w * 128;
    const uint32_t T_O = tmem + lane_base + 256 + w * 128;
    float m = 0.f, l = 0.f;
    const uint32_t dseed = DROP ? drop_seed_q(drop_seed_bh(p.drop_seed, b * p.H + h), qi) : 0u;
    DBG_T(t_sm0);
    for (int j = 0; j < nw; ++j) {
      const int kv0 = j * TQ;
      DBG_T(ts0);
      mbar_wait(&bars->s_full[w], j & 1);
      tc_fence_after();
      DBG_T(ts1);
#ifdef B2F_ATTN_DEBUG
      if (p.dbg_mode >= 1) {
        tc_fence_before();
        for (int c = 0; c < 4; ++c) mbar_arrive(&bars->p_full[w][c]);
        continue;
      }
#endif
      uint32_t s[128];
#pragma unroll
      for (int c = 0; c < 4; ++c)
        tmem_ld_32x32b_x32(T_S + c * 32, *reinterpret_cast<uint32_t(*)[32]>(s + c * 32));
      tmem_ld_wait();
      const bool need_mask = (p.causal && kv0 + TQ - 1 > qs) || (kv0 + TQ > kv_len);
      if (need_mask) {
        int lim = kv_len;
        if (p.causal && qi + 1 < lim) lim = qi + 1;
        lim -= kv0;
#pragma unroll
        for (int i = 0; i < 128; ++i)
          if (i >= lim) s[i] = 0xff800000u;  // -inf
      }
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
      for (int i = 0; i < 128; i += 8) {      // FMNMX3: two new values per instruction
        mx0 = max3(mx0, __uint_as_float(s[i]), __uint_as_float(s[i + 4]));
        mx1 = max3(mx1, __uint_as_float(s[i + 1]), __uint_as_float(s[i + 5]));
        mx2 = max3(mx2, __uint_as_float(s[i + 2]), __uint_as_float(s[i + 6]));
        mx3 = max3(mx3, __uint_as_float(s[i + 3]), __uint_as_float(s[i + 7]));
      }
      const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * p.scale_log2;
      if (j == 0) {
        m = (mx == -INFINITY) ? 0.f : mx;
      } else {
        const bool grow = mx > m + RESCALE_THRESHOLD;
        if (__any_sync(0xffffffffu, grow)) {
          const float alpha = grow ? ex2(m - mx) : 1.f;
          if (grow) m = mx;
          l *= alpha;
#pragma unroll
          for (int c = 0; c < D / 32; ++c) {
            uint32_t acc[32];
            tmem_ld_32x32b_x32(T_O + c * 32, acc);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) acc[i] = __float_as_uint(__uint_as_float(acc[i]) * alpha);
            tmem_st_32x32b_x32(T_O + c * 32, acc);
          }
        }
      }
      DBG_T(ts2);
      const uint64_t sc2 = f2pack(p.scale_log2, p.scale_log2), nm2 = f2pack(-m, -m);
      uint64_t sum2 = f2pack(0.f, 0.f);
      auto probabilities = [&](auto mix_tag) {
        constexpr bool MIX = decltype(mix_tag)::value;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const uint64_t t = fma2(f2pack(__uint_as_float(s[c * 32 + 2 * i]), __uint_as_float(s[c * 32 + 2 * i + 1])), sc2, nm2);
            float p0, p1;
            if (MIX && (i & 3) < POLY) {
              ex2_poly2(t, p0, p1);
            } else {
              float t0, t1;
              f2unpack(t, t0, t1);
              p0 = ex2(t0);
              p1 = ex2(t1);
            }
            sum2 = add2(sum2, f2pack(p0, p1));       // the normaliser is the sum of the undropped probabilities
            if constexpr (DROP) {
              const uint32_t wd = drop_word(dseed, (kv0 + c * 32 + 2 * i) >> 2) >> ((i & 1) * 16);
              p0 = (wd & 0xFF) >= p.drop_thr ? p0 : 0.f;
              p1 = ((wd >> 8) & 0xFF) >= p.drop_thr ? p1 : 0.f;
            }
            pk[i] = pack_bf16x2(p0, p1);
          }
          tmem_st_32x32b_x16(T_S + c * 16, pk);
          tmem_st_wait();
          tc_fence_before();
          mbar_arrive(&bars->p_full[w][c]);
        }
      };
      if (POLY > 0 && !need_mask) probabilities(std::true_type{});
      else probabilities(std::false_type{});
      float sum0, sum1;
      f2unpack(sum2, sum0, sum1);
      l += sum0 + sum1;
#ifdef B2F_ATTN_DEBUG
      if (threadIdx.x == 128) {
        DBG_T(ts3);
        DBG_ADD(1, ts1 - ts0); DBG_ADD(2, ts2 - ts1); DBG_ADD(3, ts3 - ts2); DBG_ADD(4, 1);
      }
#endif
    }
#ifdef B2F_ATTN_DEBUG
    DBG_T(t_sm1);
    if (threadIdx.x == 128) { DBG_ADD(0, t_sm1 - t_sm0); }
#endif
    if (nw > 0) {
      mbar_wait(&bars->o_full[w], 0);
      tc_fence_after();
      const float inv = (l > 0.f ? 1.f / l : 0.f) * (DROP ? p.drop_rinv : 1.f);
      __nv_bfloat16* orow = p.O + (long long)b * p.o_bs + (long long)qi * p.o_ts + (long long)h * D;
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        uint32_t acc[32];
        tmem_ld_32x32b_x32(T_O + c * 32, acc);
        tmem_ld_wait();
        if (qi < p.S) {
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4) {
            uint4 u;
            float f[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(acc[q4 * 8 + i]) * inv;
            u.x = pack_bf16x2(f[0], f[1]); u.y = pack_bf16x2(f[2], f[3]);
            u.z = pack_bf16x2(f[4], f[5]); u.w = pack_bf16x2(f[6], f[7]);
            *reinterpret_cast<uint4*>(orow + c * 32 + q4 * 8) = u;
          }
        }
      }
      if (qi < p.S && p.lse)
        p.lse[((long long)b * p.H + h) * p.S + qi] = l > 0.f ? (m * LN2 + __logf(l)) : INFINITY;
    }
#ifdef B2F_ATTN_DEBUG
    if (threadIdx.x == 128) { DBG_T(t_ep); DBG_ADD(10, t_ep - t_sm1); }
#endif
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
#ifdef B2F_ATTN_DEBUG
  if (threadIdx.x == 0) {
    DBG_T(t_cta1);
    unsigned long long ns1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns1));
    DBG_ADD(8, t_cta1 - t_cta0);
    DBG_ADD(11, ns1 - ns0);
    DBG_ADD(12, 1);
  }
#endif
}

// ============================================================== backward
// Pre-pass over every (padded) query row: nlse2 = -lse * log2(e) and ndlt = -rowsum(O * dO) (negated
// so the main kernels feed them straight into packed FMA / ADD), written to the fp32 scratch
// [2][B*H][S_pad] (S_pad = S rounded up to 128).  Padding rows get nlse2 = -inf and ndlt = 0, so the
// main kernels can bulk-copy whole 128-row slices and see P = 0 there.
template <int D>
__global__ void __launch_bounds__(256)
attn_delta_kernel(const __nv_bfloat16* __restrict__ O, long long o_ts, long long o_bs,
                  const __nv_bfloat16* __restrict__ dO, long long do_ts, long long do_bs,
                  const float* __restrict__ lse, float* __restrict__ scr, int B, int H, int S,
                  int S_pad) {
  constexpr int TPR = D / 8;  // threads per row
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long row = gid / TPR;  // over B*S_pad*H (token-major, then head)
  const int part = (int)(gid % TPR);
  const long long total = (long long)B * S_pad * H;
  const int h = (int)(row % H);
  const long long tok = row / H;
  const int s = (int)(tok % S_pad);
  const int b = (int)(tok / S_pad);
  const bool live = row < total && s < S;
  float acc = 0.f;
  if (live) {
    const uint4 a = *reinterpret_cast<const uint4*>(O + b * o_bs + (long long)s * o_ts + (long long)h * D + part * 8);
    const uint4 g = *reinterpret_cast<const uint4*>(dO + b * do_bs + (long long)s * do_ts + (long long)h * D + part * 8);
    float2 a0 = unpack_bf16x2(a.x), a1 = unpack_bf16x2(a.y), a2 = unpack_bf16x2(a.z), a3 = unpack_bf16x2(a.w);
    float2 g0 = unpack_bf16x2(g.x), g1 = unpack_bf16x2(g.y), g2 = unpack_bf16x2(g.z), g3 = unpack_bf16x2(g.w);
    acc = a0.x * g0.x + a0.y * g0.y + a1.x * g1.x + a1.y * g1.y + a2.x * g2.x + a2.y * g2.y +
          a3.x * g3.x + a3.y * g3.y;
  }
#pragma unroll
  for (int o = TPR / 2; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (row < total && part == 0) {
    const long long idx = ((long long)b * H + h) * S_pad + s;
    scr[idx] = live ? -lse[((long long)b * H + h) * S + s] * LOG2E : -INFINITY;
    scr[(long long)B * H * S_pad + idx] = -acc;
  }
}

// Backward kernels: 384 threads.  Warpgroup 0 = TMA warp + MMA warp; warpgroups 1 and 2 each take
// half of the columns of a score (half) tile, one tile row per thread.  Both groups read their fp32
// columns of S / dP into registers, meet at a named barrier, and only then overwrite the front of
// those TMEM regions with the bf16 P / dS operands.  The softmax scale is folded into the epilogue.
constexpr int BWD_THREADS = 384;

struct BwdBars {
  uint64_t fixed_full, st_full[2], st_empty[2], s_full[2], dp_full, ds_full, acc_full;
  uint32_t tmem_base;
  alignas(128) float lse2[2][128];  // dkv kernel: per-stage statistics of the staged query tile
  alignas(128) float dlt[2][128];
};

// Accumulator rows -> bf16 global rows; the two compute warpgroups split the D columns.
template <int D>
__device__ __forceinline__ void store_acc_half(uint32_t taddr, __nv_bfloat16* row, int cg, bool ok, float mul) {
#pragma unroll
  for (int c = 0; c < D / 64; ++c) {
    uint32_t acc[32];
    const int col = cg * (D / 2) + c * 32;
    tmem_ld_32x32b_x32(taddr + col, acc);
    tmem_ld_wait();
    if (ok) {
#pragma unroll
      for (int q4 = 0; q4 < 4; ++q4) {
        uint4 u;
        u.x = pack_bf16x2(__uint_as_float(acc[q4 * 8 + 0]) * mul, __uint_as_float(acc[q4 * 8 + 1]) * mul);
        u.y = pack_bf16x2(__uint_as_float(acc[q4 * 8 + 2]) * mul, __uint_as_float(acc[q4 * 8 + 3]) * mul);
        u.z = pack_bf16x2(__uint_as_float(acc[q4 * 8 + 4]) * mul, __uint_as_float(acc[q4 * 8 + 5]) * mul);
        u.w = pack_bf16x2(__uint_as_float(acc[q4 * 8 + 6]) * mul, __uint_as_float(acc[q4 * 8 + 7]) * mul);
        *reinterpret_cast<uint4*>(row + col + q4 * 8) = u;
      }
    }
  }
}

// Both backward kernels work on half score tiles (128 rows x 64 columns) so that two of them fit
// next to the accumulators and the loop is software-pipelined: while the compute groups turn half h
// into its bf16 operands, the tensor core already produces S / dP of half h+1 in the other buffer,
// and the accumulate MMAs of half h follow as soon as its operands are stored.
//   TMEM: buffer b = columns [128 b, 128 b + 128): S half at +0, dP half at +64; the bf16 operands
//   overwrite the front of S / dP (32 columns each).

// ---- dQ: CTA per (q tile, head, batch); loops over key half-tiles.  dQ accumulates at [256, 256+D).
template <int D, bool DROP = false>
__global__ void __launch_bounds__(BWD_THREADS, 1)
attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap mq, const __grid_constant__ CUtensorMap mk,
                   const __grid_constant__ CUtensorMap mv, const __grid_constant__ CUtensorMap mdo,
                   const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  constexpr int TILE_B = Cfg<D>::TILE_B;
  uint8_t* sQ = smem;
  uint8_t* sdO = smem + TILE_B;
  uint8_t* sK = smem + 2 * TILE_B;  // 2 stages
  uint8_t* sV = smem + 4 * TILE_B;  // 2 stages
  BwdBars* bars = reinterpret_cast<BwdBars*>(smem + 6 * TILE_B);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (p.S + TQ - 1) / TQ;
  const int q_tile = p.causal ? nq - 1 - (int)blockIdx.y : (int)blockIdx.y;
  const int h = blockIdx.x % p.H, b = blockIdx.x / p.H;
  const int hkv = h / (p.H / p.HKV);
  const int q0 = q_tile * TQ;
  int kv_len = p.kv_lens ? p.kv_lens[b] : p.S;
  kv_len = kv_len < p.S ? kv_len : p.S;
  int kv_end = p.causal ? (kv_len < q0 + TQ ? kv_len : q0 + TQ) : kv_len;
  int n = (kv_end + TQ - 1) / TQ;
  if (n < 1) n = 1;
  const int H2 = 2 * n;  // half tiles

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&mq); tma_prefetch_desc(&mk); tma_prefetch_desc(&mv); tma_prefetch_desc(&mdo);
    mbar_init(&bars->fixed_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->st_full[i], 1);
      mbar_init(&bars->st_empty[i], 1);
      mbar_init(&bars->s_full[i], 1);
    }
    mbar_init(&bars->ds_full, 256);
    mbar_init(&bars->acc_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) { tmem_alloc(&bars->tmem_base, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  require_tmem_base0(bars->tmem_base);
  constexpr uint32_t T_DQ = 256;

  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 72;");
    if (warp == 0 && lane == 0) {
      mbar_arrive_expect_tx(&bars->fixed_full, 2 * TILE_B);
      load_tile<D>(sQ, &mq, &bars->fixed_full, q0, h, b);
      load_tile<D>(sdO, &mdo, &bars->fixed_full, q0, h, b);
      for (int j = 0; j < n; ++j) {
        const int st = j & 1;
        mbar_wait(&bars->st_empty[st], ((j >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(&bars->st_full[st], 2 * TILE_B);
        load_tile<D>(sK + st * TILE_B, &mk, &bars->st_full[st], j * TQ, hkv, b);
        load_tile<D>(sV + st * TILE_B, &mv, &bars->st_full[st], j * TQ, hkv, b);
      }
    } else if (warp == 1) {
      const uint32_t idesc_s = make_idesc(B2F_MMA_FMT, B2F_MMA_FMT, 128, 64, 0, 0);
      const uint32_t idesc_dq = make_idesc(B2F_MMA_FMT, B2F_MMA_FMT, 128, D, 0, 1);
      const uint32_t aQ = lo_k(smem_u32(sQ)), adO = lo_k(smem_u32(sdO));
      // S_h = Q K_h^T and dP_h = dO V_h^T for key half hh (64 keys = rows [64 half, +64) of the stage)
      auto issue_sdp = [&](int hh) {
        const int j = hh >> 1, half = hh & 1, st = j & 1;
        if (half == 0) {
          mbar_wait(&bars->st_full[st], (j >> 1) & 1);
          tc_fence_after();
        }
        const uint32_t aK = lo_k(smem_u32(sK + st * TILE_B + half * 8192));
        const uint32_t aV = lo_k(smem_u32(sV + st * TILE_B + half * 8192));
        const uint32_t tb = (hh & 1) * 128;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < D / 16; ++k) umma_f16_ss(tb, desc_k(aQ, k), desc_k(aK, k), idesc_s, k > 0);
#pragma unroll
          for (int k = 0; k < D / 16; ++k) umma_f16_ss(tb + 64, desc_k(adO, k), desc_k(aV, k), idesc_s, k > 0);
          umma_commit(&bars->s_full[hh & 1]);
        }
        __syncwarp();
      };
      mbar_wait(&bars->fixed_full, 0);
      issue_sdp(0);
      for (int hh = 0; hh < H2; ++hh) {
        const int j = hh >> 1, half = hh & 1, st = j & 1;
        if (hh + 1 < H2) issue_sdp(hh + 1);
        mbar_wait(&bars->ds_full, hh & 1);
        tc_fence_after();
        const uint32_t aK = lo_mn(smem_u32(sK + st * TILE_B));
        const uint32_t tds = (hh & 1) * 128 + 64;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16_ts(T_DQ, tds + k * 8, desc_mn(aK, half * 4 + k), idesc_dq, (hh | k) != 0);
          if (half == 1) umma_commit(&bars->st_empty[st]);
          if (hh + 1 == H2) umma_commit(&bars->acc_full);
        }
        __syncwarp();
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
    const int cg = (warp - 4) >> 2;  // column group: 32 of the 64 key columns of a half tile
    const int qw = warp & 3;
    const int r = qw * 32 + lane;
    const int qi = q0 + r;
    const uint32_t lane_base = static_cast<uint32_t>(qw * 32) << 16;
    const int S_pad = nq * TQ;
    const long long sidx = ((long long)b * p.H + h) * S_pad + qi;
    const float nlse2 = p.delta[sidx];
    const float ndlt = p.delta[(long long)p.B * p.H * S_pad + sidx];
    const uint64_t sc2 = f2pack(p.scale_log2, p.scale_log2), nl2 = f2pack(nlse2, nlse2), nd2 = f2pack(ndlt, ndlt);
    const uint32_t dseed = DROP ? drop_seed_q(drop_seed_bh(p.drop_seed, b * p.H + h), qi) : 0u;
    for (int hh = 0; hh < H2; ++hh) {
      const int kv0 = hh * 64 + cg * 32;  // first key column of this thread's 32
      // dropout: dS = P o (mask o dP / (1 - p) - delta): the kept dP entries are rescaled, the dropped ones zeroed
      auto drop_dp = [&](int i, float& a0, float& a1) {
        const uint32_t wd = drop_word(dseed, (kv0 + 2 * i) >> 2) >> ((i & 1) * 16);
        a0 = (wd & 0xFF) >= p.drop_thr ? a0 * p.drop_rinv : 0.f;
        a1 = ((wd >> 8) & 0xFF) >= p.drop_thr ? a1 * p.drop_rinv : 0.f;
      };
      const uint32_t tb = lane_base + (hh & 1) * 128;
      DBG_T(tq0);
      mbar_wait(&bars->s_full[hh & 1], (hh >> 1) & 1);
      tc_fence_after();
      DBG_T(tq1);
      uint32_t sa[32], da[32];
      tmem_ld_32x32b_x32(tb + cg * 32, sa);
      tmem_ld_32x32b_x32(tb + 64 + cg * 32, da);
      tmem_ld_wait();
      tc_fence_before();
      named_bar_sync(2, 256);  // every fp32 dP column is in registers: dS may now overwrite them
      tc_fence_after();
      DBG_T(tq2);
      const int kvh0 = hh * 64;
      const bool need_mask = (p.causal && kvh0 + 63 > q0) || (kvh0 + 64 > kv_len);
      uint32_t pk[16];
      if (!need_mask) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const uint64_t t = fma2(f2pack(__uint_as_float(sa[2 * i]), __uint_as_float(sa[2 * i + 1])), sc2, nl2);
          float t0, t1;
          f2unpack(t, t0, t1);
          float a0 = __uint_as_float(da[2 * i]), a1 = __uint_as_float(da[2 * i + 1]);
          if constexpr (DROP) drop_dp(i, a0, a1);
          const uint64_t d = add2(f2pack(a0, a1), nd2);
          float d0, d1;
          f2unpack(mul2(f2pack(ex2(t0), ex2(t1)), d), d0, d1);
          pk[i] = pack_bf16x2(d0, d1);
        }
      } else {
        int lim = kv_len;
        if (p.causal && qi + 1 < lim) lim = qi + 1;
        lim -= kv0;  // first masked column among this thread's 32
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const uint64_t t = fma2(f2pack(__uint_as_float(sa[2 * i]), __uint_as_float(sa[2 * i + 1])), sc2, nl2);
          float t0, t1;
          f2unpack(t, t0, t1);
          const float p0 = 2 * i < lim ? ex2(t0) : 0.f;
          const float p1 = 2 * i + 1 < lim ? ex2(t1) : 0.f;
          float a0 = __uint_as_float(da[2 * i]), a1 = __uint_as_float(da[2 * i + 1]);
          if constexpr (DROP) drop_dp(i, a0, a1);
          const uint64_t d = add2(f2pack(a0, a1), nd2);
          float d0, d1;
          f2unpack(mul2(f2pack(p0, p1), d), d0, d1);
          pk[i] = pack_bf16x2(d0, d1);
        }
      }
      tmem_st_32x32b_x16(tb + 64 + cg * 16, pk);
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(&bars->ds_full);
#ifdef B2F_ATTN_DEBUG
      if (threadIdx.x == 128) {
        DBG_T(tq3);
        DBG_ADD(5, tq1 - tq0); DBG_ADD(6, tq2 - tq1); DBG_ADD(7, tq3 - tq2); DBG_ADD(8, 1);
      }
#endif
    }
    mbar_wait(&bars->acc_full, 0);
    tc_fence_after();
    __nv_bfloat16* drow = p.dQ + (long long)b * p.dq_bs + (long long)qi * p.dq_ts + (long long)h * D;
    store_acc_half<D>(T_DQ + lane_base, drow, cg, qi < p.S, p.scale);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(0, 512); }
}

// ---- dK/dV: CTA per (key tile, kv head, batch); loops over the query heads of the group and over
// query half-tiles (S^T formulation: rows = keys, columns = 64 queries, so P^T / dS^T are TMEM
// A-operands).  dV accumulates at [256, 256+D), dK at [384, 384+D).  The TMA warp also stages each
// query tile's -lse2 / -delta rows (two 512 B bulk copies).
template <int D, bool DROP = false>
__global__ void __launch_bounds__(BWD_THREADS, 1)
attn_bwd_dkv_kernel(const __grid_constant__ CUtensorMap mq, const __grid_constant__ CUtensorMap mk,
                    const __grid_constant__ CUtensorMap mv, const __grid_constant__ CUtensorMap mdo,
                    const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  constexpr int TILE_B = Cfg<D>::TILE_B;
  uint8_t* sK = smem;
  uint8_t* sV = smem + TILE_B;
  uint8_t* sQ = smem + 2 * TILE_B;   // 2 stages
  uint8_t* sdO = smem + 4 * TILE_B;  // 2 stages
  BwdBars* bars = reinterpret_cast<BwdBars*>(smem + 6 * TILE_B);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (p.S + TQ - 1) / TQ;
  const int kv_tile = blockIdx.y;  // causal: key tile 0 sees every query tile, so ascending = longest first
  const int hkv = blockIdx.x % p.HKV, b = blockIdx.x / p.HKV;
  const int G = p.H / p.HKV;
  const int kv0 = kv_tile * TQ;
  int kv_len = p.kv_lens ? p.kv_lens[b] : p.S;
  kv_len = kv_len < p.S ? kv_len : p.S;
  const int i0 = p.causal ? kv_tile : 0;
  const int per_head = nq - i0;
  const int n = per_head * G;
  const int H2 = 2 * n;
  const int S_pad = nq * TQ;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&mq); tma_prefetch_desc(&mk); tma_prefetch_desc(&mv); tma_prefetch_desc(&mdo);
    mbar_init(&bars->fixed_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->st_full[i], 1);
      mbar_init(&bars->st_empty[i], 1);
      mbar_init(&bars->s_full[i], 1);
    }
    mbar_init(&bars->ds_full, 256);
    mbar_init(&bars->acc_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) { tmem_alloc(&bars->tmem_base, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  require_tmem_base0(bars->tmem_base);
  constexpr uint32_t T_DV = 256, T_DK = 384;

  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 72;");
    if (warp == 0 && lane == 0) {
      mbar_arrive_expect_tx(&bars->fixed_full, 2 * TILE_B);
      load_tile<D>(sK, &mk, &bars->fixed_full, kv0, hkv, b);
      load_tile<D>(sV, &mv, &bars->fixed_full, kv0, hkv, b);
      for (int it = 0; it < n; ++it) {
        const int st = it & 1;
        const int g = it / per_head, i = i0 + it % per_head;
        const int h = hkv * G + g;
        mbar_wait(&bars->st_empty[st], ((it >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(&bars->st_full[st], 2 * TILE_B + 2 * TQ * 4);
        load_tile<D>(sQ + st * TILE_B, &mq, &bars->st_full[st], i * TQ, h, b);
        load_tile<D>(sdO + st * TILE_B, &mdo, &bars->st_full[st], i * TQ, h, b);
        const float* stats = p.delta + ((long long)b * p.H + h) * S_pad + i * TQ;
        bulk_load_1d(bars->lse2[st], stats, TQ * 4, &bars->st_full[st]);
        bulk_load_1d(bars->dlt[st], stats + (long long)p.B * p.H * S_pad, TQ * 4, &bars->st_full[st]);
      }
    } else if (warp == 1) {
      const uint32_t idesc_s = make_idesc(B2F_MMA_FMT, B2F_MMA_FMT, 128, 64, 0, 0);
      const uint32_t idesc_acc = make_idesc(B2F_MMA_FMT, B2F_MMA_FMT, 128, D, 0, 1);
      const uint32_t aK = lo_k(smem_u32(sK)), aV = lo_k(smem_u32(sV));
      // S^T_h = K Q_h^T and dP^T_h = V dO_h^T for query half hh (rows [64 half, +64) of the stage)
      auto issue_sdp = [&](int hh) {
        const int it = hh >> 1, half = hh & 1, st = it & 1;
        if (half == 0) {
          mbar_wait(&bars->st_full[st], (it >> 1) & 1);
          tc_fence_after();
        }
        const uint32_t aQ = lo_k(smem_u32(sQ + st * TILE_B + half * 8192));
        const uint32_t adO = lo_k(smem_u32(sdO + st * TILE_B + half * 8192));
        const uint32_t tb = (hh & 1) * 128;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < D / 16; ++k) umma_f16_ss(tb, desc_k(aK, k), desc_k(aQ, k), idesc_s, k > 0);
#pragma unroll
          for (int k = 0; k < D / 16; ++k) umma_f16_ss(tb + 64, desc_k(aV, k), desc_k(adO, k), idesc_s, k > 0);
          umma_commit(&bars->s_full[hh & 1]);
        }
        __syncwarp();
      };
      mbar_wait(&bars->fixed_full, 0);
      if (H2 > 0) issue_sdp(0);
      for (int hh = 0; hh < H2; ++hh) {
        const int it = hh >> 1, half = hh & 1, st = it & 1;
        if (hh + 1 < H2) issue_sdp(hh + 1);
        mbar_wait(&bars->ds_full, hh & 1);
        tc_fence_after();
        const uint32_t aQ = lo_mn(smem_u32(sQ + st * TILE_B)), adO = lo_mn(smem_u32(sdO + st * TILE_B));
        const uint32_t tb = (hh & 1) * 128;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16_ts(T_DV, tb + k * 8, desc_mn(adO, half * 4 + k), idesc_acc, (hh | k) != 0);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16_ts(T_DK, tb + 64 + k * 8, desc_mn(aQ, half * 4 + k), idesc_acc, (hh | k) != 0);
          if (half == 1) umma_commit(&bars->st_empty[st]);
          if (hh + 1 == H2) umma_commit(&bars->acc_full);
        }
        __syncwarp();
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
    const int cg = (warp - 4) >> 2;  // column group: 32 of the 64 query columns of a half tile
    const int qw = warp & 3;
    const int r = qw * 32 + lane;
    const int ki = kv0 + r;
    const uint32_t lane_base = static_cast<uint32_t>(qw * 32) << 16;
    const bool row_dead = ki >= kv_len;
    const uint64_t sc2 = f2pack(p.scale_log2, p.scale_log2);
    for (int hh = 0; hh < H2; ++hh) {
      const int it = hh >> 1, half = hh & 1, st = it & 1;
      const int i = i0 + it % per_head;
      const int qh0 = i * TQ + half * 64;      // first query of this half tile
      const uint32_t tb = lane_base + (hh & 1) * 128;
      DBG_T(tk0);
      if (half == 0) mbar_wait(&bars->st_full[st], (it >> 1) & 1);  // statistics of this query tile are staged
      mbar_wait(&bars->s_full[hh & 1], (hh >> 1) & 1);
      tc_fence_after();
      DBG_T(tk1);
      uint32_t sa[32], da[32];
      tmem_ld_32x32b_x32(tb + cg * 32, sa);
      tmem_ld_32x32b_x32(tb + 64 + cg * 32, da);
      tmem_ld_wait();
      tc_fence_before();
      named_bar_sync(2, 256);
      tc_fence_after();
      DBG_T(tk2);
      const bool need_mask = (p.causal && qh0 < kv0 + TQ - 1) || (kv0 + TQ > kv_len);
      // causal: query column c is visible to key row ki iff qh0 + cg*32 + c >= ki
      const int first_ok = row_dead ? 64 : (p.causal ? ki - qh0 - cg * 32 : 0);
      const uint32_t l_addr = smem_u32(&bars->lse2[st][half * 64 + cg * 32]);
      const uint32_t d_addr = smem_u32(&bars->dlt[st][half * 64 + cg * 32]);
      uint32_t pp[16], pd[16];
      // dropout: this thread owns key ki and walks over queries, so every element has its own mask word; the kept
      // probabilities (dV operand) and the kept dP entries (inside dS) are scaled by 1 / (1 - p)
      const uint32_t dseed_bh = DROP ? drop_seed_bh(p.drop_seed, b * p.H + hkv * G + it / per_head) : 0u;
      auto keep_at = [&](int c) -> float {
        const uint32_t wd = drop_word(drop_seed_q(dseed_bh, qh0 + cg * 32 + c), ki >> 2) >> ((ki & 3) * 8);
        return (wd & 0xFF) >= p.drop_thr ? p.drop_rinv : 0.f;
      };
      if (!need_mask) {
#pragma unroll
        for (int i4 = 0; i4 < 8; ++i4) {
          const ulonglong2 nl = lds_v2_u64(l_addr + i4 * 16), nd = lds_v2_u64(d_addr + i4 * 16);
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int c = i4 * 4 + e * 2;
            const uint64_t t = fma2(f2pack(__uint_as_float(sa[c]), __uint_as_float(sa[c + 1])), sc2, e ? nl.y : nl.x);
            float t0, t1;
            f2unpack(t, t0, t1);
            const float p0 = ex2(t0), p1 = ex2(t1);
            float a0 = __uint_as_float(da[c]), a1 = __uint_as_float(da[c + 1]);
            float k0 = 1.f, k1 = 1.f;
            if constexpr (DROP) { k0 = keep_at(c); k1 = keep_at(c + 1); a0 *= k0; a1 *= k1; }
            const uint64_t d = add2(f2pack(a0, a1), e ? nd.y : nd.x);
            float d0, d1;
            f2unpack(mul2(f2pack(p0, p1), d), d0, d1);
            pp[i4 * 2 + e] = pack_bf16x2(p0 * k0, p1 * k1);
            pd[i4 * 2 + e] = pack_bf16x2(d0, d1);
          }
        }
      } else {
#pragma unroll
        for (int i4 = 0; i4 < 8; ++i4) {
          const ulonglong2 nl = lds_v2_u64(l_addr + i4 * 16), nd = lds_v2_u64(d_addr + i4 * 16);
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int c = i4 * 4 + e * 2;
            const uint64_t t = fma2(f2pack(__uint_as_float(sa[c]), __uint_as_float(sa[c + 1])), sc2, e ? nl.y : nl.x);
            float t0, t1;
            f2unpack(t, t0, t1);
            const float p0 = c >= first_ok ? ex2(t0) : 0.f;
            const float p1 = c + 1 >= first_ok ? ex2(t1) : 0.f;
            float a0 = __uint_as_float(da[c]), a1 = __uint_as_float(da[c + 1]);
            float k0 = 1.f, k1 = 1.f;
            if constexpr (DROP) { k0 = keep_at(c); k1 = keep_at(c + 1); a0 *= k0; a1 *= k1; }
            const uint64_t d = add2(f2pack(a0, a1), e ? nd.y : nd.x);
            float d0, d1;
            f2unpack(mul2(f2pack(p0, p1), d), d0, d1);
            pp[i4 * 2 + e] = pack_bf16x2(p0 * k0, p1 * k1);
            pd[i4 * 2 + e] = pack_bf16x2(d0, d1);
          }
        }
      }
      tmem_st_32x32b_x16(tb + cg * 16, pp);
      tmem_st_32x32b_x16(tb + 64 + cg * 16, pd);
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(&bars->ds_full);
#ifdef B2F_ATTN_DEBUG
      if (threadIdx.x == 128) {
        DBG_T(tk3);
        DBG_ADD(1, tk1 - tk0); DBG_ADD(2, tk2 - tk1); DBG_ADD(3, tk3 - tk2); DBG_ADD(4, 1);
      }
#endif
    }
    __nv_bfloat16* vrow = p.dV + (long long)b * p.dv_bs + (long long)ki * p.dv_ts + (long long)hkv * D;
    __nv_bfloat16* krow = p.dK + (long long)b * p.dk_bs + (long long)ki * p.dk_ts + (long long)hkv * D;
    if (n > 0) {
      mbar_wait(&bars->acc_full, 0);
      tc_fence_after();
      store_acc_half<D>(T_DV + lane_base, vrow, cg, ki < p.S, 1.f);
      store_acc_half<D>(T_DK + lane_base, krow, cg, ki < p.S, p.scale);
    } else if (ki < p.S) {
      for (int c = cg * (D / 2); c < (cg + 1) * (D / 2); c += 8) {
        *reinterpret_cast<uint4*>(vrow + c) = make_uint4(0, 0, 0, 0);
        *reinterpret_cast<uint4*>(krow + c) = make_uint4(0, 0, 0, 0);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(0, 512); }
}

int make_map(CUtensorMap* m, const void* ptr, int B, int S, int heads, int D, long long ts,
             long long bs) {
  return make_tma_4d(m, ptr, TmaDtype::BF16, B, heads, S, D, bs, D, ts, TQ, 64, 128);
}

template <typename K>
int set_smem(K kernel, int bytes) {
  B2F_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  return 0;
}

int check_common(int B, int S, int H, int HKV, int D) {
  B2F_REQUIRE(D == 64 || D == 128, "attention: head dim must be 64 or 128");
  B2F_REQUIRE(B > 0 && S > 0 && H > 0 && HKV > 0 && H % HKV == 0, "attention: bad shape");
  return 0;
}

}  // namespace

#ifdef B2F_ATTN_DEBUG
extern "C" int b2f_attn_debug(unsigned long long* out, int reset) {
  if (reset) {
    unsigned long long z[16] = {0};
    B2F_CHECK_CUDA(cudaMemcpyToSymbol(g_attn_dbg, z, sizeof(z)));
  } else {
    B2F_CHECK_CUDA(cudaMemcpyFromSymbol(out, g_attn_dbg, sizeof(unsigned long long) * 16));
  }
  return 0;
}
#endif

namespace {
int set_dropout(AttnParams& p, float p_drop, unsigned seed) {
  B2F_REQUIRE(p_drop >= 0.f && p_drop < 0.99f, "attention dropout probability must be in [0, 0.99)");
  p.drop_thr = (uint32_t)(p_drop * 256.f + 0.5f);            // resolution 1/256
  p.drop_rinv = 256.f / (256.f - (float)p.drop_thr);          // 1 / (1 - p) for the probability actually applied
  p.drop_seed = seed;
  return 0;
}
}  // namespace

static int attn_fwd_impl(const void* q, const void* k, const void* v, void* o, float* lse, int B,
                         int S, int H, int HKV, int D, long long q_ts, long long q_bs,
                         long long k_ts, long long k_bs, long long v_ts, long long v_bs,
                         long long o_ts, long long o_bs, float scale, int causal,
                         const int* kv_lens, float p_drop, unsigned seed, void* stream) {
  if (int rc = check_common(B, S, H, HKV, D)) return rc;
  B2F_REQUIRE((reinterpret_cast<uintptr_t>(o) & 15) == 0 && o_ts % 8 == 0 && o_bs % 8 == 0,
              "attention: output must be 16-byte aligned");
  CUtensorMap mq, mk, mv;
  if (int rc = make_map(&mq, q, B, S, H, D, q_ts, q_bs)) return rc;
  if (int rc = make_map(&mk, k, B, S, HKV, D, k_ts, k_bs)) return rc;
  if (int rc = make_map(&mv, v, B, S, HKV, D, v_ts, v_bs)) return rc;
  AttnParams p{};
  p.B = B; p.H = H; p.HKV = HKV; p.S = S;
  p.scale = scale; p.scale_log2 = scale * LOG2E; p.causal = causal; p.kv_lens = kv_lens;
  p.O = static_cast<__nv_bfloat16*>(o); p.o_ts = o_ts; p.o_bs = o_bs; p.lse = lse;
#ifdef B2F_ATTN_DEBUG
  if (const char* e = getenv("B2F_ATTN_DBG_MODE")) p.dbg_mode = atoi(e);
#endif
  if (int rc = set_dropout(p, p_drop, seed)) return rc;
  const bool drop = p.drop_thr > 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  dim3 grid2(H * B, (S + 2 * TQ - 1) / (2 * TQ), 1);
  static const int poly = [] { const char* e = getenv("B2F_ATTN_POLY"); return e ? atoi(e) : ATTN_POLY_DEFAULT; }();
  if (D == 128) {
    const int smem = 6 * Cfg<128>::TILE_B + 1024 + sizeof(Fwd2Bars);
    if (drop) {
      if (int rc = set_smem(attn_fwd2_kernel<128, true>, smem)) return rc;
      attn_fwd2_kernel<128, true><<<grid2, FWD2_THREADS, smem, st>>>(mq, mk, mv, p);
    } else if (poly == 0) {
      if (int rc = set_smem(attn_fwd2_kernel<128>, smem)) return rc;
      attn_fwd2_kernel<128><<<grid2, FWD2_THREADS, smem, st>>>(mq, mk, mv, p);
    } else if (poly == 1) {
      if (int rc = set_smem(attn_fwd2_kernel<128, false, 1>, smem)) return rc;
      attn_fwd2_kernel<128, false, 1><<<grid2, FWD2_THREADS, smem, st>>>(mq, mk, mv, p);
    } else {
      if (int rc = set_smem(attn_fwd2_kernel<128, false, 2>, smem)) return rc;
      attn_fwd2_kernel<128, false, 2><<<grid2, FWD2_THREADS, smem, st>>>(mq, mk, mv, p);
    }
  } else {
    const int smem = 6 * Cfg<64>::TILE_B + 1024 + sizeof(Fwd2Bars);
    if (drop) {
      if (int rc = set_smem(attn_fwd2_kernel<64, true>, smem)) return rc;
      attn_fwd2_kernel<64, true><<<grid2, FWD2_THREADS, smem, st>>>(mq, mk, mv, p);
    } else if (poly == 0) {
      if (int rc = set_smem(attn_fwd2_kernel<64>, smem)) return rc;
      attn_fwd2_kernel<64><<<grid2, FWD2_THREADS, smem, st>>>(mq, mk, mv, p);
    } else if (poly == 1) {
      if (int rc = set_smem(attn_fwd2_kernel<64, false, 1>, smem)) return rc;
      attn_fwd2_kernel<64, false, 1><<<grid2, FWD2_THREADS, smem, st>>>(mq, mk, mv, p);
    } else {
      if (int rc = set_smem(attn_fwd2_kernel<64, false, 2>, smem)) return rc;
      attn_fwd2_kernel<64, false, 2><<<grid2, FWD2_THREADS, smem, st>>>(mq, mk, mv, p);
    }
  }
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B,
                            int S, int H, int HKV, int D, long long q_ts, long long q_bs,
                            long long k_ts, long long k_bs, long long v_ts, long long v_bs,
                            long long o_ts, long long o_bs, float scale, int causal,
                            const int* kv_lens, void* stream) {
  return attn_fwd_impl(q, k, v, o, lse, B, S, H, HKV, D, q_ts, q_bs, k_ts, k_bs, v_ts, v_bs, o_ts, o_bs, scale,
                       causal, kv_lens, 0.f, 0u, stream);
}

// Forward with attention dropout of probability p_drop (applied to the softmax probabilities, Megatron's
// --attention-dropout); the mask is a function of (seed, batch, head, query, key) only.
extern "C" int b2f_attn_fwd_drop(const void* q, const void* k, const void* v, void* o, float* lse, int B,
                                 int S, int H, int HKV, int D, long long q_ts, long long q_bs,
                                 long long k_ts, long long k_bs, long long v_ts, long long v_bs,
                                 long long o_ts, long long o_bs, float scale, int causal,
                                 const int* kv_lens, float p_drop, unsigned seed, void* stream) {
  return attn_fwd_impl(q, k, v, o, lse, B, S, H, HKV, D, q_ts, q_bs, k_ts, k_bs, v_ts, v_bs, o_ts, o_bs, scale,
                       causal, kv_lens, p_drop, seed, stream);
}

static int attn_bwd_impl(const void* q, const void* k, const void* v, const void* o,
                         const void* d_o, const float* lse, float* delta, void* dq, void* dk,
                         void* dv, int B, int S, int H, int HKV, int D, long long q_ts,
                         long long q_bs, long long k_ts, long long k_bs, long long v_ts,
                         long long v_bs, long long o_ts, long long o_bs, long long do_ts,
                         long long do_bs, long long dq_ts, long long dq_bs, long long dk_ts,
                         long long dk_bs, long long dv_ts, long long dv_bs, float scale,
                         int causal, const int* kv_lens, float p_drop, unsigned seed, void* stream) {
  if (int rc = check_common(B, S, H, HKV, D)) return rc;
  B2F_REQUIRE(delta != nullptr && lse != nullptr, "attention bwd: lse and delta scratch required");
  B2F_REQUIRE((reinterpret_cast<uintptr_t>(delta) & 15) == 0, "attention bwd: scratch must be 16-byte aligned");
  CUtensorMap mq, mk, mv, mdo;
  if (int rc = make_map(&mq, q, B, S, H, D, q_ts, q_bs)) return rc;
  if (int rc = make_map(&mk, k, B, S, HKV, D, k_ts, k_bs)) return rc;
  if (int rc = make_map(&mv, v, B, S, HKV, D, v_ts, v_bs)) return rc;
  if (int rc = make_map(&mdo, d_o, B, S, H, D, do_ts, do_bs)) return rc;
  AttnParams p{};
  p.B = B; p.H = H; p.HKV = HKV; p.S = S;
  p.scale = scale; p.scale_log2 = scale * LOG2E; p.causal = causal; p.kv_lens = kv_lens;
  p.lse = const_cast<float*>(lse); p.delta = delta;
  p.dQ = static_cast<__nv_bfloat16*>(dq); p.dK = static_cast<__nv_bfloat16*>(dk);
  p.dV = static_cast<__nv_bfloat16*>(dv);
  p.dq_ts = dq_ts; p.dq_bs = dq_bs; p.dk_ts = dk_ts; p.dk_bs = dk_bs; p.dv_ts = dv_ts; p.dv_bs = dv_bs;
  if (int rc = set_dropout(p, p_drop, seed)) return rc;
  const bool drop = p.drop_thr > 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nq = (S + TQ - 1) / TQ;
  const int S_pad = nq * TQ;
  const long long delta_threads = (long long)B * S_pad * H * (D / 8);
  const unsigned dgrid = (unsigned)((delta_threads + 255) / 256);
  if (D == 128) {
    attn_delta_kernel<128><<<dgrid, 256, 0, st>>>((const __nv_bfloat16*)o, o_ts, o_bs,
                                                  (const __nv_bfloat16*)d_o, do_ts, do_bs, lse, delta,
                                                  B, H, S, S_pad);
    const int smem = 6 * Cfg<128>::TILE_B + 1024 + sizeof(BwdBars);
    if (drop) {
      if (int rc = set_smem(attn_bwd_dq_kernel<128, true>, smem)) return rc;
      if (int rc = set_smem(attn_bwd_dkv_kernel<128, true>, smem)) return rc;
      attn_bwd_dq_kernel<128, true><<<dim3(H * B, nq, 1), BWD_THREADS, smem, st>>>(mq, mk, mv, mdo, p);
      attn_bwd_dkv_kernel<128, true><<<dim3(HKV * B, nq, 1), BWD_THREADS, smem, st>>>(mq, mk, mv, mdo, p);
    } else {
      if (int rc = set_smem(attn_bwd_dq_kernel<128>, smem)) return rc;
      if (int rc = set_smem(attn_bwd_dkv_kernel<128>, smem)) return rc;
      attn_bwd_dq_kernel<128><<<dim3(H * B, nq, 1), BWD_THREADS, smem, st>>>(mq, mk, mv, mdo, p);
      attn_bwd_dkv_kernel<128><<<dim3(HKV * B, nq, 1), BWD_THREADS, smem, st>>>(mq, mk, mv, mdo, p);
    }
  } else {
    attn_delta_kernel<64><<<dgrid, 256, 0, st>>>((const __nv_bfloat16*)o, o_ts, o_bs,
                                                 (const __nv_bfloat16*)d_o, do_ts, do_bs, lse, delta,
                                                 B, H, S, S_pad);
    const int smem = 6 * Cfg<64>::TILE_B + 1024 + sizeof(BwdBars);
    if (drop) {
      if (int rc = set_smem(attn_bwd_dq_kernel<64, true>, smem)) return rc;
      if (int rc = set_smem(attn_bwd_dkv_kernel<64, true>, smem)) return rc;
      attn_bwd_dq_kernel<64, true><<<dim3(H * B, nq, 1), BWD_THREADS, smem, st>>>(mq, mk, mv, mdo, p);
      attn_bwd_dkv_kernel<64, true><<<dim3(HKV * B, nq, 1), BWD_THREADS, smem, st>>>(mq, mk, mv, mdo, p);
    } else {
      if (int rc = set_smem(attn_bwd_dq_kernel<64>, smem)) return rc;
      if (int rc = set_smem(attn_bwd_dkv_kernel<64>, smem)) return rc;
      attn_bwd_dq_kernel<64><<<dim3(H * B, nq, 1), BWD_THREADS, smem, st>>>(mq, mk, mv, mdo, p);
      attn_bwd_dkv_kernel<64><<<dim3(HKV * B, nq, 1), BWD_THREADS, smem, st>>>(mq, mk, mv, mdo, p);
    }
  }
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// delta: fp32 scratch of 2 * B * H * S_pad floats supplied by the caller (S_pad = S rounded up to
// 128); the pre-pass fills it with lse * log2(e) and rowsum(O * dO) for the two main kernels.
extern "C" int b2f_attn_bwd(const void* q, const void* k, const void* v, const void* o,
                            const void* d_o, const float* lse, float* delta, void* dq, void* dk,
                            void* dv, int B, int S, int H, int HKV, int D, long long q_ts,
                            long long q_bs, long long k_ts, long long k_bs, long long v_ts,
                            long long v_bs, long long o_ts, long long o_bs, long long do_ts,
                            long long do_bs, long long dq_ts, long long dq_bs, long long dk_ts,
                            long long dk_bs, long long dv_ts, long long dv_bs, float scale,
                            int causal, const int* kv_lens, void* stream) {
  return attn_bwd_impl(q, k, v, o, d_o, lse, delta, dq, dk, dv, B, S, H, HKV, D, q_ts, q_bs, k_ts, k_bs, v_ts, v_bs,
                       o_ts, o_bs, do_ts, do_bs, dq_ts, dq_bs, dk_ts, dk_bs, dv_ts, dv_bs, scale, causal, kv_lens,
                       0.f, 0u, stream);
}

// Backward of b2f_attn_fwd_drop: same p_drop and seed regenerate the forward's mask inside both kernels.
extern "C" int b2f_attn_bwd_drop(const void* q, const void* k, const void* v, const void* o,
                                 const void* d_o, const float* lse, float* delta, void* dq, void* dk,
                                 void* dv, int B, int S, int H, int HKV, int D, long long q_ts,
                                 long long q_bs, long long k_ts, long long k_bs, long long v_ts,
                                 long long v_bs, long long o_ts, long long o_bs, long long do_ts,
                                 long long do_bs, long long dq_ts, long long dq_bs, long long dk_ts,
                                 long long dk_bs, long long dv_ts, long long dv_bs, float scale,
                                 int causal, const int* kv_lens, float p_drop, unsigned seed, void* stream) {
  return attn_bwd_impl(q, k, v, o, d_o, lse, delta, dq, dk, dv, B, S, H, HKV, D, q_ts, q_bs, k_ts, k_bs, v_ts, v_bs,
                       o_ts, o_bs, do_ts, do_bs, dq_ts, dq_bs, dk_ts, dk_bs, dv_ts, dv_bs, scale, causal, kv_lens,
                       p_drop, seed, stream);
}
