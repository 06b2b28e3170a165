// b200flow K6 / K11 — fused AdamW and the multi-tensor l2norm / scale /
// overflow-check family for sm_100a.
//
// Parameters live in flat buffers (one per dtype group), so "multi-tensor apply"
// is a single grid-stride launch per buffer; a segment-table variant covers
// lists of separately allocated tensors (the torch DDP / Accelerate path).
// AdamW reads bf16 or fp32 gradients, updates the fp32 master weights and both
// moments in place and writes the bf16 working copy in the same pass; the
// gradient scale (1/loss_scale x clip coefficient) and the skip-on-overflow flag
// are read from device memory so the step needs no host synchronisation.
//
// Capability parity: apex amp_C multi_tensor_{adam,l2norm,scale} and DeepSpeed
// FusedAdam (containers/megatron-deepspeed/Dockerfile:6-12).
#include "../common/ptx.cuh"
#include "../common/host.h"
#include "../../include/b200flow.h"

using namespace b2f;

namespace {

struct AdamHyper {
  float lr, beta1, beta2, eps, weight_decay, bc1, bc2;  // bc = 1 - beta^t
};

__device__ __forceinline__ void adam_update(float& p, float& m, float& v, float g,
                                            const AdamHyper& h) {
  m = h.beta1 * m + (1.f - h.beta1) * g;
  v = h.beta2 * v + (1.f - h.beta2) * g * g;
  const float mhat = m / h.bc1;
  const float vhat = v / h.bc2;
  p = p - h.lr * (mhat / (sqrtf(vhat) + h.eps) + h.weight_decay * p);
}

// 4 elements per thread per iteration; n_vec4 = n / 4, tail handled separately.
template <bool GRAD_BF16>
__global__ void __launch_bounds__(256)
adamw_flat_kernel(float* __restrict__ master, float* __restrict__ m, float* __restrict__ v,
                  const void* __restrict__ grad, __nv_bfloat16* __restrict__ param_out,
                  long long n, AdamHyper h, const float* __restrict__ grad_scale,
                  const int* __restrict__ skip) {
  if (skip && *skip) return;
  const float gs = grad_scale ? *grad_scale : 1.f;
  const long long nv = n >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += stride) {
    float4 p4 = reinterpret_cast<float4*>(master)[i];
    float4 m4 = reinterpret_cast<float4*>(m)[i];
    float4 v4 = reinterpret_cast<float4*>(v)[i];
    float g[4];
    if (GRAD_BF16) {
      uint2 u = reinterpret_cast<const uint2*>(grad)[i];
      float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y);
      g[0] = a.x; g[1] = a.y; g[2] = b.x; g[3] = b.y;
    } else {
      float4 g4 = reinterpret_cast<const float4*>(grad)[i];
      g[0] = g4.x; g[1] = g4.y; g[2] = g4.z; g[3] = g4.w;
    }
    adam_update(p4.x, m4.x, v4.x, g[0] * gs, h);
    adam_update(p4.y, m4.y, v4.y, g[1] * gs, h);
    adam_update(p4.z, m4.z, v4.z, g[2] * gs, h);
    adam_update(p4.w, m4.w, v4.w, g[3] * gs, h);
    reinterpret_cast<float4*>(master)[i] = p4;
    reinterpret_cast<float4*>(m)[i] = m4;
    reinterpret_cast<float4*>(v)[i] = v4;
    if (param_out) {
      uint2 o;
      o.x = pack_bf16x2(p4.x, p4.y);
      o.y = pack_bf16x2(p4.z, p4.w);
      reinterpret_cast<uint2*>(param_out)[i] = o;
    }
  }
  // tail (n % 4 elements)
  const long long base = nv << 2;
  const long long t = base + (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) {
    float g = GRAD_BF16 ? h2f(reinterpret_cast<const __nv_bfloat16*>(grad)[t])
                        : reinterpret_cast<const float*>(grad)[t];
    float p = master[t], mm = m[t], vv = v[t];
    adam_update(p, mm, vv, g * gs, h);
    master[t] = p; m[t] = mm; v[t] = vv;
    if (param_out) param_out[t] = f2h(p);
  }
}

// Segment table for lists of tensors: one row per tensor.
struct Segment {
  void* master;   // fp32
  void* m;        // fp32
  void* v;        // fp32
  const void* grad;
  void* param_out;  // bf16 or null
  long long n;
  long long first_chunk;  // exclusive prefix sum of chunk counts
};
constexpr int MT_CHUNK = 2048 * 4;

template <bool GRAD_BF16>
__global__ void __launch_bounds__(256)
adamw_multi_kernel(const Segment* __restrict__ segs, int nsegs, AdamHyper h,
                   const float* __restrict__ grad_scale, const int* __restrict__ skip) {
  if (skip && *skip) return;
  const float gs = grad_scale ? *grad_scale : 1.f;
  // binary search: last segment with first_chunk <= blockIdx.x
  int lo = 0, hi = nsegs - 1;
  while (lo < hi) {
    int mid = (lo + hi + 1) >> 1;
    if (segs[mid].first_chunk <= (long long)blockIdx.x) lo = mid; else hi = mid - 1;
  }
  const Segment s = segs[lo];
  const long long start = ((long long)blockIdx.x - s.first_chunk) * MT_CHUNK;
  const long long end = start + MT_CHUNK < s.n ? start + MT_CHUNK : s.n;
  float* master = static_cast<float*>(s.master);
  float* m = static_cast<float*>(s.m);
  float* v = static_cast<float*>(s.v);
  __nv_bfloat16* po = static_cast<__nv_bfloat16*>(s.param_out);
  for (long long i = start + threadIdx.x; i < end; i += blockDim.x) {
    float g = GRAD_BF16 ? h2f(static_cast<const __nv_bfloat16*>(s.grad)[i])
                        : static_cast<const float*>(s.grad)[i];
    float p = master[i], mm = m[i], vv = v[i];
    adam_update(p, mm, vv, g * gs, h);
    master[i] = p; m[i] = mm; v[i] = vv;
    if (po) po[i] = f2h(p);
  }
}

// ---- sum of squares -> per-block partials -> deterministic final reduce ----
template <int DT>  // 0 = bf16, 1 = fp32
__global__ void __launch_bounds__(256)
sumsq_partial_kernel(const void* __restrict__ x, long long n, float* __restrict__ partial) {
  float acc = 0.f;
  const long long stride = (long long)gridDim.x * blockDim.x;
  if (DT == 0) {
    const long long nv = n >> 3;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += stride) {
      uint4 u = ld_nc_v4(reinterpret_cast<const uint4*>(x) + i);
      float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z),
             d = unpack_bf16x2(u.w);
      acc += a.x * a.x + a.y * a.y + b.x * b.x + b.y * b.y + c.x * c.x + c.y * c.y + d.x * d.x +
             d.y * d.y;
    }
    const long long t = (nv << 3) + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) {
      float f = h2f(reinterpret_cast<const __nv_bfloat16*>(x)[t]);
      acc += f * f;
    }
  } else {
    const long long nv = n >> 2;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += stride) {
      float4 f = reinterpret_cast<const float4*>(x)[i];
      acc += f.x * f.x + f.y * f.y + f.z * f.z + f.w * f.w;
    }
    const long long t = (nv << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) {
      float f = reinterpret_cast<const float*>(x)[t];
      acc += f * f;
    }
  }
  __shared__ float red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w];
    partial[blockIdx.x] = s;
  }
}

// out[0] (+)= sum(partials); overflow[0] |= !isfinite(sum).
__global__ void sumsq_final_kernel(const float* __restrict__ partial, int nparts,
                                   float* __restrict__ out, int* __restrict__ overflow,
                                   int accumulate) {
  __shared__ float red[32];
  float acc = 0.f;
  for (int i = threadIdx.x; i < nparts; i += blockDim.x) acc += partial[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w];
    if (overflow && !isfinite(s)) *overflow = 1;
    out[0] = accumulate ? out[0] + s : s;
  }
}

// coef = min(1, max_norm / (sqrt(sumsq) + 1e-6)) * extra ; also writes the norm.
__global__ void clip_coef_kernel(const float* __restrict__ sumsq, float max_norm, float extra,
                                 float* __restrict__ coef_out, float* __restrict__ norm_out) {
  const float nrm = sqrtf(sumsq[0]);
  if (norm_out) norm_out[0] = nrm;
  float c = 1.f;
  if (max_norm > 0.f) c = fminf(1.f, max_norm / (nrm + 1e-6f));
  coef_out[0] = c * extra;
}

template <int DT>
__global__ void __launch_bounds__(256)
scale_kernel(void* __restrict__ x, long long n, const float* __restrict__ scale_ptr, float scale) {
  const float s = scale_ptr ? scale_ptr[0] * scale : scale;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    if (DT == 0) {
      __nv_bfloat16* p = static_cast<__nv_bfloat16*>(x);
      p[i] = f2h(h2f(p[i]) * s);
    } else {
      float* p = static_cast<float*>(x);
      p[i] *= s;
    }
  }
}

// Momentum SGD on a flat buffer (tensorpack's MomentumOptimizer(0.9) of the Mask R-CNN
// workloads): buf = mu * buf + (g * scale + wd * p); p -= lr * buf; bf16 working copy out.
template <bool GRAD_BF16>
__global__ void __launch_bounds__(256)
sgd_momentum_flat_kernel(float* __restrict__ master, float* __restrict__ buf,
                         const void* __restrict__ grad, __nv_bfloat16* __restrict__ param_out,
                         long long n, float lr, float mu, float wd,
                         const float* __restrict__ grad_scale) {
  const float gs = grad_scale ? *grad_scale : 1.f;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float g = (GRAD_BF16 ? h2f(static_cast<const __nv_bfloat16*>(grad)[i])
                               : static_cast<const float*>(grad)[i]) * gs;
    float p = master[i];
    const float b = mu * buf[i] + g + wd * p;
    p -= lr * b;
    buf[i] = b;
    master[i] = p;
    if (param_out) param_out[i] = f2h(p);
  }
}

inline int grid_for(long long n, int per_thread) {
  long long blocks = (n / per_thread + 255) / 256;
  const long long cap = (long long)b2f::sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

}  // namespace

extern "C" int b2f_adamw_flat(float* master, float* m, float* v, const void* grad, int grad_bf16,
                              void* param_out, long long n, float lr, float beta1, float beta2,
                              float eps, float weight_decay, int step, const float* grad_scale,
                              const int* skip, void* stream) {
  B2F_REQUIRE(n >= 0 && step >= 1, "adamw: n >= 0 and step >= 1");
  if (n == 0) return 0;
  AdamHyper h{lr, beta1, beta2, eps, weight_decay, 1.f - powf(beta1, (float)step),
              1.f - powf(beta2, (float)step)};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = grid_for(n, 4);
  if (grad_bf16)
    adamw_flat_kernel<true><<<grid, 256, 0, st>>>(master, m, v, grad, (__nv_bfloat16*)param_out, n,
                                                  h, grad_scale, skip);
  else
    adamw_flat_kernel<false><<<grid, 256, 0, st>>>(master, m, v, grad, (__nv_bfloat16*)param_out,
                                                   n, h, grad_scale, skip);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// segs: device array of `nsegs` rows x 7 int64 {master, m, v, grad, param_out, n, first_chunk}.
extern "C" int b2f_adamw_multi(const void* segs, int nsegs, long long total_chunks, int grad_bf16,
                               float lr, float beta1, float beta2, float eps, float weight_decay,
                               int step, const float* grad_scale, const int* skip, void* stream) {
  B2F_REQUIRE(nsegs > 0 && total_chunks > 0 && step >= 1, "adamw_multi: empty segment table");
  static_assert(sizeof(Segment) == 7 * 8, "segment row must be 7 x int64");
  AdamHyper h{lr, beta1, beta2, eps, weight_decay, 1.f - powf(beta1, (float)step),
              1.f - powf(beta2, (float)step)};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (grad_bf16)
    adamw_multi_kernel<true><<<(unsigned)total_chunks, 256, 0, st>>>((const Segment*)segs, nsegs, h,
                                                                     grad_scale, skip);
  else
    adamw_multi_kernel<false><<<(unsigned)total_chunks, 256, 0, st>>>((const Segment*)segs, nsegs,
                                                                      h, grad_scale, skip);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_sgd_momentum_flat(float* master, float* buf, const void* grad, int grad_bf16,
                                     void* param_out, long long n, float lr, float momentum,
                                     float weight_decay, const float* grad_scale, void* stream) {
  if (n <= 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = grid_for(n, 1);
  if (grad_bf16)
    sgd_momentum_flat_kernel<true><<<grid, 256, 0, st>>>(master, buf, grad, (__nv_bfloat16*)param_out, n,
                                                         lr, momentum, weight_decay, grad_scale);
  else
    sgd_momentum_flat_kernel<false><<<grid, 256, 0, st>>>(master, buf, grad, (__nv_bfloat16*)param_out,
                                                          n, lr, momentum, weight_decay, grad_scale);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_multi_tensor_chunk() { return MT_CHUNK; }

extern "C" int b2f_sumsq_num_partials(long long n) { return grid_for(n, 8); }

// scratch: fp32 [b2f_sumsq_num_partials(n)] owned by the caller.
extern "C" int b2f_sumsq(const void* x, long long n, int dtype_fp32, float* scratch, float* out,
                         int* overflow, int accumulate, void* stream) {
  B2F_REQUIRE(n > 0 && scratch && out, "sumsq: n > 0, scratch and out required");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = grid_for(n, 8);
  if (dtype_fp32) sumsq_partial_kernel<1><<<grid, 256, 0, st>>>(x, n, scratch);
  else sumsq_partial_kernel<0><<<grid, 256, 0, st>>>(x, n, scratch);
  B2F_CHECK_CUDA(cudaGetLastError());
  sumsq_final_kernel<<<1, 1024, 0, st>>>(scratch, grid, out, overflow, accumulate);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_clip_coef(const float* sumsq, float max_norm, float extra_scale, float* coef_out,
                             float* norm_out, void* stream) {
  clip_coef_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(sumsq, max_norm, extra_scale,
                                                                   coef_out, norm_out);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_scale(void* x, long long n, int dtype_fp32, const float* scale_ptr, float scale,
                         void* stream) {
  if (n <= 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = grid_for(n, 1);
  if (dtype_fp32) scale_kernel<1><<<grid, 256, 0, st>>>(x, n, scale_ptr, scale);
  else scale_kernel<0><<<grid, 256, 0, st>>>(x, n, scale_ptr, scale);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
