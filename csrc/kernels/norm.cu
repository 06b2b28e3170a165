// b200flow K8 — LayerNorm / RMSNorm forward + backward for sm_100a.
//
// One CTA per row (persistent over rows in backward), 16-byte vector loads, the
// row lives in registers between the statistics pass and the normalise pass, so
// HBM traffic is the algorithmic minimum (read x [+residual], write y [+h]).
// Forward optionally fuses the residual add that precedes the norm
// (h = x + residual; y = norm(h)) — the bias-dropout-add + LN pairing of the
// Megatron layer.  Backward produces dx and per-CTA
// partial dgamma/dbeta rows in a scratch buffer, reduced by a second kernel
// (the two-stage reduction apex's fused_layer_norm_cuda uses; the reference
// pulls apex in at containers/megatron-deepspeed/Dockerfile:6-11).
#include "../common/ptx.cuh"
#include "../common/host.h"
#include "../../include/b200flow.h"

using namespace b2f;

namespace {

constexpr int NORM_THREADS = 256;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-wide sum of up to two values; every thread gets the result.
template <int NV>
__device__ __forceinline__ void block_sum(float (&v)[NV], float* red /*[NV*8]*/) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < NV; ++i) v[i] = warp_sum(v[i]);
  __syncthreads();  // protect `red` from the previous use
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < NV; ++i) red[i * 8 + warp] = v[i];
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < NORM_THREADS / 32; ++w) s += red[i * 8 + w];
    v[i] = s;
  }
}

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z),
         d = unpack_bf16x2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 u;
  u.x = pack_bf16x2(f[0], f[1]); u.y = pack_bf16x2(f[2], f[3]);
  u.z = pack_bf16x2(f[4], f[5]); u.w = pack_bf16x2(f[6], f[7]);
  return u;
}

// VPT = 16-byte vectors per thread; H <= VPT * 256 * 8.
template <int VPT, bool RMS>
__global__ void __launch_bounds__(NORM_THREADS)
norm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ residual,
                const __nv_bfloat16* __restrict__ gamma, const __nv_bfloat16* __restrict__ beta,
                __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ h_out,
                float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows, int H,
                float eps) {
  __shared__ float red[16];
  const int nvec = H >> 3;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)row * H);
    const uint4* rr = residual ? reinterpret_cast<const uint4*>(residual + (size_t)row * H) : nullptr;
    float v[VPT][8];
    float s[2] = {0.f, 0.f};
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int c = threadIdx.x + i * NORM_THREADS;
      if (c < nvec) {
        unpack8(ld_nc_v4(xr + c), v[i]);
        if (rr) {
          float r[8];
          unpack8(ld_nc_v4(rr + c), r);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[i][j] += r[j];
          if (h_out) {
            // round-trip through bf16 so that y is computed from exactly what is stored
            uint4 hp = pack8(v[i]);
            reinterpret_cast<uint4*>(h_out + (size_t)row * H)[c] = hp;
            unpack8(hp, v[i]);
          }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) { s[0] += v[i][j]; s[1] += v[i][j] * v[i][j]; }
      }
    }
    block_sum<2>(s, red);
    float mean = RMS ? 0.f : s[0] / H;
    float var = RMS ? s[1] / H : fmaxf(s[1] / H - mean * mean, 0.f);
    float rstd = rsqrtf(var + eps);
    if (threadIdx.x == 0) {
      if (mean_out) mean_out[row] = mean;
      rstd_out[row] = rstd;
    }
    uint4* yr = reinterpret_cast<uint4*>(y + (size_t)row * H);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int c = threadIdx.x + i * NORM_THREADS;
      if (c < nvec) {
        float g[8], o[8];
        unpack8(__ldg(reinterpret_cast<const uint4*>(gamma) + c), g);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * g[j];
        if (!RMS && beta) {
          float bb[8];
          unpack8(__ldg(reinterpret_cast<const uint4*>(beta) + c), bb);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += bb[j];
        }
        yr[c] = pack8(o);
      }
    }
  }
}

template <int VPT, bool RMS>
__global__ void __launch_bounds__(NORM_THREADS)
norm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ mean_in,
                const float* __restrict__ rstd_in, const __nv_bfloat16* __restrict__ dres_in,
                __nv_bfloat16* __restrict__ dx, float* __restrict__ partial /*[grid,2,H]*/,
                int rows, int H) {
  __shared__ float red[16];
  const int nvec = H >> 3;
  float dg[VPT][8], db[VPT][8];
  float g[VPT][8];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int c = threadIdx.x + i * NORM_THREADS;
#pragma unroll
    for (int j = 0; j < 8; ++j) { dg[i][j] = 0.f; db[i][j] = 0.f; g[i][j] = 0.f; }
    if (c < nvec) unpack8(__ldg(reinterpret_cast<const uint4*>(gamma) + c), g[i]);
  }
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const uint4* dyr = reinterpret_cast<const uint4*>(dy + (size_t)row * H);
    const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)row * H);
    const float mean = RMS ? 0.f : mean_in[row];
    const float rstd = rstd_in[row];
    float xh[VPT][8], gy[VPT][8];
    float s[2] = {0.f, 0.f};
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int c = threadIdx.x + i * NORM_THREADS;
      if (c < nvec) {
        float d[8], xv[8];
        unpack8(ld_nc_v4(dyr + c), d);
        unpack8(ld_nc_v4(xr + c), xv);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] = (xv[j] - mean) * rstd;
          gy[i][j] = d[j] * g[i][j];
          dg[i][j] += d[j] * xh[i][j];
          db[i][j] += d[j];
          s[0] += gy[i][j];
          s[1] += gy[i][j] * xh[i][j];
        }
      }
    }
    block_sum<2>(s, red);
    const float m1 = RMS ? 0.f : s[0] / H;
    const float m2 = s[1] / H;
    uint4* dxr = reinterpret_cast<uint4*>(dx + (size_t)row * H);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int c = threadIdx.x + i * NORM_THREADS;
      if (c < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * (gy[i][j] - m1 - xh[i][j] * m2);
        if (dres_in) {  // gradient arriving through the fused residual branch
          float r[8];
          unpack8(ld_nc_v4(reinterpret_cast<const uint4*>(dres_in + (size_t)row * H) + c), r);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += r[j];
        }
        dxr[c] = pack8(o);
      }
    }
  }
  float* pg = partial + (size_t)blockIdx.x * 2 * H;
  float* pb = pg + H;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int c = threadIdx.x + i * NORM_THREADS;
    if (c < nvec) {
      float4* o = reinterpret_cast<float4*>(pg + c * 8);
      o[0] = make_float4(dg[i][0], dg[i][1], dg[i][2], dg[i][3]);
      o[1] = make_float4(dg[i][4], dg[i][5], dg[i][6], dg[i][7]);
      if (!RMS) {
        float4* ob = reinterpret_cast<float4*>(pb + c * 8);
        ob[0] = make_float4(db[i][0], db[i][1], db[i][2], db[i][3]);
        ob[1] = make_float4(db[i][4], db[i][5], db[i][6], db[i][7]);
      }
    }
  }
}

// Reduce the per-CTA partial rows: out[h] (+)= sum_p partial[p][which][h].
// Block = 32 columns x 8 row slices (coalesced 128 B reads, 8 independent partial sums per column).
__global__ void __launch_bounds__(256)
norm_bwd_reduce_kernel(const float* __restrict__ partial, int nparts, int H,
                       float* __restrict__ dgamma, float* __restrict__ dbeta, int accumulate) {
  __shared__ float sg[8][33], sb[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int h = blockIdx.x * 32 + tx;
  float ag = 0.f, ab = 0.f;
  if (h < H) {
    for (int p = ty; p < nparts; p += 8) {
      ag += partial[(size_t)p * 2 * H + h];
      if (dbeta) ab += partial[(size_t)p * 2 * H + H + h];
    }
  }
  sg[ty][tx] = ag;
  sb[ty][tx] = ab;
  __syncthreads();
  if (ty == 0 && h < H) {
    float g = 0.f, b = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { g += sg[i][tx]; b += sb[i][tx]; }
    dgamma[h] = accumulate ? dgamma[h] + g : g;
    if (dbeta) dbeta[h] = accumulate ? dbeta[h] + b : b;
  }
}

// ---- group-per-row kernels (H <= 8192).  A row is owned by G threads (G = 32 for H <= 1024, so a
// warp per row and shuffle-only reductions; up to 256 for H = 8192), every thread holds at most
// four 16-byte vectors, and a 256-thread CTA works on 256/G rows at once.  Compared with one CTA
// per row this keeps 4-8x more rows in flight per SM, which is what HBM needs at small H.
constexpr int GVPT = 4;

template <int G>
__device__ __forceinline__ void group_sum2(float& a, float& b, float (*red)[8][2], int grp, int t) {
  a = warp_sum(a);
  b = warp_sum(b);
  if (G > 32) {
    const int w = t >> 5;
    __syncthreads();  // previous use of `red` is over
    if ((t & 31) == 0) { red[grp][w][0] = a; red[grp][w][1] = b; }
    __syncthreads();
    a = 0.f; b = 0.f;
#pragma unroll
    for (int i = 0; i < G / 32; ++i) { a += red[grp][i][0]; b += red[grp][i][1]; }
  }
}

template <int G, bool RMS>
__global__ void __launch_bounds__(NORM_THREADS)
norm_fwd_g_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ residual,
                  const __nv_bfloat16* __restrict__ gamma, const __nv_bfloat16* __restrict__ beta,
                  __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ h_out,
                  float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows, int H,
                  float eps) {
  constexpr int RPC = NORM_THREADS / G;
  __shared__ float red[RPC][8][2];
  const int grp = threadIdx.x / G, t = threadIdx.x % G;
  const int nvec = H >> 3;
  const float inv_h = 1.f / H;
  for (int base = blockIdx.x * RPC; base < rows; base += gridDim.x * RPC) {
    const int row = base + grp;
    const bool active = row < rows;
    const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)row * H);
    const uint4* rr = reinterpret_cast<const uint4*>(residual + (size_t)row * H);
    uint4 raw[GVPT];
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int i = 0; i < GVPT; ++i) {
      const int c = t + i * G;
      raw[i] = make_uint4(0, 0, 0, 0);
      if (active && c < nvec) {
        raw[i] = ld_nc_v4(xr + c);
        if (residual) {
          float v[8], r[8];
          unpack8(raw[i], v);
          unpack8(ld_nc_v4(rr + c), r);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] += r[j];
          raw[i] = pack8(v);  // y is computed from exactly the bf16 sum that is stored
          if (h_out) reinterpret_cast<uint4*>(h_out + (size_t)row * H)[c] = raw[i];
        }
        float v[8];
        unpack8(raw[i], v);
#pragma unroll
        for (int j = 0; j < 8; ++j) { s0 += v[j]; s1 += v[j] * v[j]; }
      }
    }
    group_sum2<G>(s0, s1, red, grp, t);
    const float mean = RMS ? 0.f : s0 * inv_h;
    const float var = RMS ? s1 * inv_h : fmaxf(s1 * inv_h - mean * mean, 0.f);
    const float rstd = rsqrtf(var + eps);
    if (active && t == 0) {
      if (mean_out) mean_out[row] = mean;
      rstd_out[row] = rstd;
    }
    uint4* yr = reinterpret_cast<uint4*>(y + (size_t)row * H);
#pragma unroll
    for (int i = 0; i < GVPT; ++i) {
      const int c = t + i * G;
      if (active && c < nvec) {
        float v[8], g[8], o[8];
        unpack8(raw[i], v);
        unpack8(__ldg(reinterpret_cast<const uint4*>(gamma) + c), g);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (v[j] - mean) * rstd * g[j];
        if (!RMS && beta) {
          float bb[8];
          unpack8(__ldg(reinterpret_cast<const uint4*>(beta) + c), bb);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += bb[j];
        }
        yr[c] = pack8(o);
      }
    }
  }
}

template <int G, bool RMS>
__global__ void __launch_bounds__(NORM_THREADS, 2)
norm_bwd_g_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                  const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ mean_in,
                  const float* __restrict__ rstd_in, const __nv_bfloat16* __restrict__ dres_in,
                  __nv_bfloat16* __restrict__ dx, float* __restrict__ partial /*[grid,2,H]*/,
                  int rows, int H) {
  constexpr int RPC = NORM_THREADS / G;
  __shared__ float red[RPC][8][2];
  __shared__ float part[8192];  // cross-group reduction of the dgamma / dbeta partials
  const int grp = threadIdx.x / G, t = threadIdx.x % G;
  const int nvec = H >> 3;
  const float inv_h = 1.f / H;
  float dg[GVPT][8], db[GVPT][8];
#pragma unroll
  for (int i = 0; i < GVPT; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) { dg[i][j] = 0.f; db[i][j] = 0.f; }
  }
  const uint4* gvec = reinterpret_cast<const uint4*>(gamma);  // re-read per use: stays in L1
  for (int base = blockIdx.x * RPC; base < rows; base += gridDim.x * RPC) {
    const int row = base + grp;
    const bool active = row < rows;
    const uint4* dyr = reinterpret_cast<const uint4*>(dy + (size_t)row * H);
    const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)row * H);
    const float mean = (RMS || !active) ? 0.f : mean_in[row];
    const float rstd = active ? rstd_in[row] : 0.f;
    uint4 draw[GVPT], xraw[GVPT];
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int i = 0; i < GVPT; ++i) {
      const int c = t + i * G;
      draw[i] = make_uint4(0, 0, 0, 0);
      xraw[i] = make_uint4(0, 0, 0, 0);
      if (active && c < nvec) {
        draw[i] = ld_nc_v4(dyr + c);
        xraw[i] = ld_nc_v4(xr + c);
        float d[8], xv[8], g[8];
        unpack8(draw[i], d);
        unpack8(xraw[i], xv);
        unpack8(__ldg(gvec + c), g);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xh = (xv[j] - mean) * rstd;
          const float gy = d[j] * g[j];
          dg[i][j] += d[j] * xh;
          db[i][j] += d[j];
          s0 += gy;
          s1 += gy * xh;
        }
      }
    }
    group_sum2<G>(s0, s1, red, grp, t);
    const float m1 = RMS ? 0.f : s0 * inv_h;
    const float m2 = s1 * inv_h;
    uint4* dxr = reinterpret_cast<uint4*>(dx + (size_t)row * H);
#pragma unroll
    for (int i = 0; i < GVPT; ++i) {
      const int c = t + i * G;
      if (active && c < nvec) {
        float d[8], xv[8], g[8], o[8];
        unpack8(draw[i], d);
        unpack8(xraw[i], xv);
        unpack8(__ldg(gvec + c), g);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xh = (xv[j] - mean) * rstd;
          o[j] = rstd * (d[j] * g[j] - m1 - xh * m2);
        }
        if (dres_in) {  // gradient arriving through the fused residual branch
          float r[8];
          unpack8(ld_nc_v4(reinterpret_cast<const uint4*>(dres_in + (size_t)row * H) + c), r);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += r[j];
        }
        dxr[c] = pack8(o);
      }
    }
  }
  // CTA partial = sum over its row groups, through shared memory (RPC * H <= 8192 floats)
#pragma unroll
  for (int which = 0; which < (RMS ? 1 : 2); ++which) {
    __syncthreads();
#pragma unroll
    for (int i = 0; i < GVPT; ++i) {
      const int c = t + i * G;
      if (c < nvec) {
        float4* o = reinterpret_cast<float4*>(part + grp * H + c * 8);
        const float(&src)[8] = which ? db[i] : dg[i];
        o[0] = make_float4(src[0], src[1], src[2], src[3]);
        o[1] = make_float4(src[4], src[5], src[6], src[7]);
      }
    }
    __syncthreads();
    float* dst = partial + (size_t)blockIdx.x * 2 * H + which * H;
    for (int h = threadIdx.x; h < H; h += NORM_THREADS) {
      float acc = 0.f;
#pragma unroll
      for (int g2 = 0; g2 < RPC; ++g2) acc += part[g2 * H + h];
      dst[h] = acc;
    }
  }
}

// Threads per row for the group kernels, 0 = use the CTA-per-row fallback.
static int group_size(int H) {
  const int nvec = H / 8;
  for (int g = 32; g <= 256; g *= 2)
    if (nvec <= GVPT * g) return g;
  return 0;
}

template <bool RMS>
int launch_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
               void* h_out, float* mean, float* rstd, int rows, int H, float eps,
               cudaStream_t st) {
  if (const int G = group_size(H)) {
    const int rpc = NORM_THREADS / G;
    const int want = (rows + rpc - 1) / rpc;
    const int cap = sm_count() * 8;
    const int ggrid = want < cap ? want : cap;
#define B2F_NORM_FWD_G(GG)                                                                      \
  norm_fwd_g_kernel<GG, RMS><<<ggrid, NORM_THREADS, 0, st>>>(                                    \
      (const __nv_bfloat16*)x, (const __nv_bfloat16*)residual, (const __nv_bfloat16*)gamma,     \
      (const __nv_bfloat16*)beta, (__nv_bfloat16*)y, (__nv_bfloat16*)h_out, mean, rstd, rows, H, \
      eps)
    switch (G) {
      case 32: B2F_NORM_FWD_G(32); break;
      case 64: B2F_NORM_FWD_G(64); break;
      case 128: B2F_NORM_FWD_G(128); break;
      default: B2F_NORM_FWD_G(256); break;
    }
#undef B2F_NORM_FWD_G
    B2F_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  const int vpt = (H / 8 + NORM_THREADS - 1) / NORM_THREADS;
  const int grid = rows < 148 * 8 ? rows : 148 * 8;
#define B2F_NORM_FWD(V)                                                                         \
  norm_fwd_kernel<V, RMS><<<grid, NORM_THREADS, 0, st>>>(                                       \
      (const __nv_bfloat16*)x, (const __nv_bfloat16*)residual, (const __nv_bfloat16*)gamma,     \
      (const __nv_bfloat16*)beta, (__nv_bfloat16*)y, (__nv_bfloat16*)h_out, mean, rstd, rows, H, \
      eps)
  switch (vpt) {
    case 1: B2F_NORM_FWD(1); break;
    case 2: B2F_NORM_FWD(2); break;
    case 3: case 4: B2F_NORM_FWD(4); break;
    default: B2F_NORM_FWD(8); break;
  }
#undef B2F_NORM_FWD
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

template <bool RMS>
int launch_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
               const float* rstd, const void* dres, void* dx, float* dgamma, float* dbeta,
               float* scratch, int nparts, int rows, int H, int accumulate, cudaStream_t st) {
  if (const int G = group_size(H)) {
#define B2F_NORM_BWD_G(GG)                                                                    \
  norm_bwd_g_kernel<GG, RMS><<<nparts, NORM_THREADS, 0, st>>>(                                 \
      (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)gamma, mean,   \
      rstd, (const __nv_bfloat16*)dres, (__nv_bfloat16*)dx, scratch, rows, H)
    switch (G) {
      case 32: B2F_NORM_BWD_G(32); break;
      case 64: B2F_NORM_BWD_G(64); break;
      case 128: B2F_NORM_BWD_G(128); break;
      default: B2F_NORM_BWD_G(256); break;
    }
#undef B2F_NORM_BWD_G
    B2F_CHECK_CUDA(cudaGetLastError());
    norm_bwd_reduce_kernel<<<(H + 31) / 32, 256, 0, st>>>(scratch, nparts, H, dgamma,
                                                            RMS ? nullptr : dbeta, accumulate);
    B2F_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  const int vpt = (H / 8 + NORM_THREADS - 1) / NORM_THREADS;
#define B2F_NORM_BWD(V)                                                                       \
  norm_bwd_kernel<V, RMS><<<nparts, NORM_THREADS, 0, st>>>(                                   \
      (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)gamma, mean,   \
      rstd, (const __nv_bfloat16*)dres, (__nv_bfloat16*)dx, scratch, rows, H)
  switch (vpt) {
    case 1: B2F_NORM_BWD(1); break;
    case 2: B2F_NORM_BWD(2); break;
    case 3: case 4: B2F_NORM_BWD(4); break;
    default: B2F_NORM_BWD(8); break;
  }
#undef B2F_NORM_BWD
  B2F_CHECK_CUDA(cudaGetLastError());
  norm_bwd_reduce_kernel<<<(H + 31) / 32, 256, 0, st>>>(scratch, nparts, H, dgamma,
                                                          RMS ? nullptr : dbeta, accumulate);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

// ---------------------------------------------------------------------------
// Allocation-free variant of the backward pass.
//
// Some hosts of this library may not allocate device memory at all once they are
// running: a serving or training process that carved the whole GPU into its own
// pool at start-up and hands libraries nothing but raw pointers, a process
// confined to a fixed memory partition where a late cudaMalloc can fail, or a
// latency-critical loop that cannot afford the device-wide synchronisation an
// allocation may imply.  For them the partial-sum scratch of the norm backward
// lives in this file-scope device array instead of a caller-supplied buffer.
// The price: every GPU that loads the library keeps these bytes until the
// process exits, and the size is a hard limit.
constexpr int NOALLOC_MAX_H = 8192;
constexpr long long NOALLOC_MAX_FLOATS = 296LL * 2 * 4096;  // 2,424,832 floats = 9.25 MiB
static __device__ __align__(256) float g_norm_bwd_scratch[NOALLOC_MAX_FLOATS];

// The array's name is only a host-side handle; the device address has to be asked
// for, once per device, with that device current.
static int noalloc_scratch_ptr(float** out) {
  static float* cache[64] = {nullptr};
  int dev = 0;
  B2F_CHECK_CUDA(cudaGetDevice(&dev));
  B2F_REQUIRE(dev >= 0 && dev < 64, "noalloc: device ordinal out of range");
  if (!cache[dev]) {
    float* p = nullptr;
    B2F_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_norm_bwd_scratch));
    B2F_REQUIRE(p != nullptr, "noalloc: cudaGetSymbolAddress returned a null scratch address");
    cache[dev] = p;
  }
  *out = cache[dev];
  return 0;
}

extern "C" int b2f_norm_bwd_noalloc_max_h() { return NOALLOC_MAX_H; }

// Same contract and same kernels as b2f_norm_bwd, minus the `scratch` argument.
//  * Limits: H <= 8192 (b2f_norm_bwd_noalloc_max_h()); larger H is refused with an
//    error, nothing is launched.  Any number of rows.
//  * The first call on each device looks up the scratch address (one
//    cudaGetSymbolAddress); make that first call outside any stream capture.
//  * Calls no cudaMalloc / cudaFree / allocating library routine.
//  * All calls on a device share the one array and it keeps whatever the last
//    call left in it.  Nothing in it needs to start at zero: every partial row
//    that the reduce step reads is first fully written by this same call.  Two
//    calls that may overlap on the GPU (different streams, no ordering between
//    them) race on the array — serialise them.
extern "C" int b2f_norm_bwd_noalloc(const void* dy, const void* x, const void* gamma,
                                    const float* mean, const float* rstd, const void* dres,
                                    void* dx, float* dgamma, float* dbeta, int rows, int H,
                                    int rms, int accumulate, void* stream) {
  B2F_REQUIRE(H % 8 == 0 && H > 0, "norm bwd (noalloc): H must be a positive multiple of 8");
  B2F_REQUIRE(H <= NOALLOC_MAX_H,
              "norm bwd (noalloc): H exceeds the built-in scratch limit of 8192 columns; "
              "use b2f_norm_bwd with a caller-provided scratch buffer");
  B2F_REQUIRE(rows > 0, "norm bwd (noalloc): rows must be positive");
  float* scratch = nullptr;
  if (int rc = noalloc_scratch_ptr(&scratch)) return rc;
  int nparts = b2f::sm_count() * 2;
  if (nparts > rows) nparts = rows;
  const long long fit = NOALLOC_MAX_FLOATS / (2LL * H);
  if (nparts > fit) nparts = static_cast<int>(fit);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return rms ? launch_bwd<true>(dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, scratch, nparts,
                                rows, H, accumulate, st)
             : launch_bwd<false>(dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, scratch,
                                 nparts, rows, H, accumulate, st);
}

extern "C" int b2f_norm_bwd_num_partials(int rows) {
  const int cap = b2f::sm_count() * 2;
  return rows < cap ? rows : cap;
}

extern "C" int b2f_norm_fwd(const void* x, const void* residual, const void* gamma,
                            const void* beta, void* y, void* h_out, float* mean, float* rstd,
                            int rows, int H, float eps, int rms, void* stream) {
  B2F_REQUIRE(H % 8 == 0 && H <= 8 * NORM_THREADS * 8, "norm: H must be a multiple of 8, <= 16384");
  B2F_REQUIRE(rows > 0, "norm: rows must be positive");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return rms ? launch_fwd<true>(x, residual, gamma, beta, y, h_out, mean, rstd, rows, H, eps, st)
             : launch_fwd<false>(x, residual, gamma, beta, y, h_out, mean, rstd, rows, H, eps, st);
}

// scratch: fp32 [b2f_norm_bwd_num_partials(rows), 2, H] owned by the caller.
extern "C" int b2f_norm_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
                            const float* rstd, const void* dres, void* dx, float* dgamma,
                            float* dbeta, float* scratch, int rows, int H, int rms,
                            int accumulate, void* stream) {
  B2F_REQUIRE(H % 8 == 0 && H <= 8 * NORM_THREADS * 8, "norm: H must be a multiple of 8, <= 16384");
  B2F_REQUIRE(rows > 0 && scratch != nullptr, "norm bwd: rows > 0 and scratch required");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nparts = b2f_norm_bwd_num_partials(rows);
  return rms ? launch_bwd<true>(dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, scratch, nparts,
                                rows, H, accumulate, st)
             : launch_bwd<false>(dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, scratch,
                                 nparts, rows, H, accumulate, st);
}
