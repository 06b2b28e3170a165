// b200flow K10 (+ RoPE, cross-entropy) — bandwidth-bound fused elementwise ops
// for sm_100a: SwiGLU / GELU forward+backward, bias-dropout-add, rotary
// embedding, and a fused softmax-cross-entropy that produces the loss and the
// logit gradient in one launch.  All kernels move 16 bytes per thread per access
// and touch each tensor exactly once per direction.
//
// Capability parity: Megatron fused_bias_gelu / bias_dropout_add, apex
// fused_rotary_positional_embedding and the cross-entropy of pretrain_gpt
// (SURVEY.md §2H C10, C20, C21).
#include "../common/ptx.cuh"
#include "../common/host.h"
#include "../../include/b200flow.h"

using namespace b2f;

namespace {

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
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }

inline int ew_grid(long long nvec) {
  long long b = (nvec + 255) / 256;
  const long long cap = (long long)b2f::sm_count() * 16;
  if (b > cap) b = cap;
  return b < 1 ? 1 : (int)b;
}

// gu: [rows, 2F] with gate in [:, :F] and up in [:, F:]; out: [rows, F].
__global__ void __launch_bounds__(256)
swiglu_fwd_kernel(const __nv_bfloat16* __restrict__ gu, __nv_bfloat16* __restrict__ out,
                  long long rows, int F) {
  const int fv = F >> 3;
  const long long nvec = rows * fv;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / fv;
    const int c = (int)(i - r * fv);
    const uint4* row = reinterpret_cast<const uint4*>(gu + r * 2 * (long long)F);
    float g[8], u[8], o[8];
    unpack8(ld_nc_v4(row + c), g);
    unpack8(ld_nc_v4(row + fv + c), u);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = g[j] * sigmoidf_(g[j]) * u[j];
    reinterpret_cast<uint4*>(out + r * (long long)F)[c] = pack8(o);
  }
}

__global__ void __launch_bounds__(256)
swiglu_bwd_kernel(const __nv_bfloat16* __restrict__ gu, const __nv_bfloat16* __restrict__ dout,
                  __nv_bfloat16* __restrict__ dgu, long long rows, int F) {
  const int fv = F >> 3;
  const long long nvec = rows * fv;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / fv;
    const int c = (int)(i - r * fv);
    const uint4* row = reinterpret_cast<const uint4*>(gu + r * 2 * (long long)F);
    float g[8], u[8], d[8], dg[8], du[8];
    unpack8(ld_nc_v4(row + c), g);
    unpack8(ld_nc_v4(row + fv + c), u);
    unpack8(ld_nc_v4(reinterpret_cast<const uint4*>(dout + r * (long long)F) + c), d);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float s = sigmoidf_(g[j]);
      const float silu = g[j] * s;
      du[j] = d[j] * silu;
      dg[j] = d[j] * u[j] * (s + silu * (1.f - s));
    }
    uint4* orow = reinterpret_cast<uint4*>(dgu + r * 2 * (long long)F);
    orow[c] = pack8(dg);
    orow[fv + c] = pack8(du);
  }
}

__device__ __forceinline__ float gelu_tanh_f(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  return 0.5f * x * (1.f + tanhf(u));
}
__device__ __forceinline__ float gelu_tanh_grad(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  const float t = tanhf(u);
  const float du = 0.7978845608028654f * (1.f + 3.f * 0.044715f * x * x);
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * du;
}

// y = gelu(x + bias) ; dx = dy * gelu'(x + bias)   (x is the saved pre-activation)
template <bool BWD>
__global__ void __launch_bounds__(256)
gelu_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ bias,
            const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ out, long long rows,
            int H) {
  const int hv = H >> 3;
  const long long nvec = rows * hv;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % hv);
    float xv[8], o[8];
    unpack8(ld_nc_v4(reinterpret_cast<const uint4*>(x) + i), xv);
    if (bias) {
      float b[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(bias) + c), b);
#pragma unroll
      for (int j = 0; j < 8; ++j) xv[j] += b[j];
    }
    if (BWD) {
      float d[8];
      unpack8(ld_nc_v4(reinterpret_cast<const uint4*>(dy) + i), d);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = d[j] * gelu_tanh_grad(xv[j]);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = gelu_tanh_f(xv[j]);
    }
    reinterpret_cast<uint4*>(out)[i] = pack8(o);
  }
}

// Counter-based RNG (splitmix-style hash of (seed, element index)); one 64-bit
// hash yields the keep decisions of 8 consecutive elements.
__device__ __forceinline__ uint64_t hash64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

// out = residual + dropout(x + bias) ; mask (1 byte / element) saved when p > 0.
__global__ void __launch_bounds__(256)
bias_dropout_add_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ bias,
                        const __nv_bfloat16* __restrict__ residual, __nv_bfloat16* __restrict__ out,
                        uint8_t* __restrict__ mask, long long rows, int H, float p,
                        unsigned long long seed) {
  const int hv = H >> 3;
  const long long nvec = rows * hv;
  const float keep_scale = p > 0.f ? 1.f / (1.f - p) : 1.f;
  const uint32_t thresh = (uint32_t)(p * 256.f);  // 8-bit resolution per element
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % hv);
    float xv[8], r[8], o[8];
    unpack8(ld_nc_v4(reinterpret_cast<const uint4*>(x) + i), xv);
    if (bias) {
      float b[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(bias) + c), b);
#pragma unroll
      for (int j = 0; j < 8; ++j) xv[j] += b[j];
    }
    if (residual) unpack8(ld_nc_v4(reinterpret_cast<const uint4*>(residual) + i), r);
    else {
#pragma unroll
      for (int j = 0; j < 8; ++j) r[j] = 0.f;
    }
    uint64_t bits = p > 0.f ? hash64(seed ^ (uint64_t)i * 0xD1B54A32D192ED03ull) : 0ull;
    uint64_t mk = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const bool keep = p <= 0.f || ((bits >> (8 * j)) & 0xFF) >= thresh;
      o[j] = r[j] + (keep ? xv[j] * keep_scale : 0.f);
      mk |= (uint64_t)(keep ? 1 : 0) << (8 * j);
    }
    reinterpret_cast<uint4*>(out)[i] = pack8(o);
    if (mask) reinterpret_cast<uint64_t*>(mask)[i] = mk;
  }
}

// dx = dy * mask * 1/(1-p)
__global__ void __launch_bounds__(256)
dropout_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const uint8_t* __restrict__ mask,
                   __nv_bfloat16* __restrict__ dx, long long nvec, float p) {
  const float keep_scale = 1.f / (1.f - p);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
       i += (long long)gridDim.x * blockDim.x) {
    float d[8];
    unpack8(ld_nc_v4(reinterpret_cast<const uint4*>(dy) + i), d);
    const uint64_t mk = reinterpret_cast<const uint64_t*>(mask)[i];
#pragma unroll
    for (int j = 0; j < 8; ++j) d[j] = ((mk >> (8 * j)) & 1) ? d[j] * keep_scale : 0.f;
    reinterpret_cast<uint4*>(dx)[i] = pack8(d);
  }
}

// Rotate-half RoPE applied in place to x[b, s, nh, d] (token-major), using
// cos/sin tables [s_max, d/2] fp32; `sign` = +1 forward, -1 backward.
__global__ void __launch_bounds__(256)
rope_kernel(__nv_bfloat16* __restrict__ x, const float* __restrict__ cos_t,
            const float* __restrict__ sin_t, long long tokens, int seq, int nh, int d,
            long long token_stride, int pos_offset, int pos_div, float sign) {
  const int half = d >> 1;
  const int hv = half >> 3;  // vectors per half head
  const long long nvec = tokens * nh * hv;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % hv);
    const long long th = i / hv;
    const int head = (int)(th % nh);
    const long long tok = th / nh;
    const int pos = (int)((tok / pos_div) % seq) + pos_offset;
    __nv_bfloat16* base = x + tok * token_stride + (long long)head * d;
    uint4* p1 = reinterpret_cast<uint4*>(base) + c;
    uint4* p2 = reinterpret_cast<uint4*>(base + half) + c;
    float a[8], b[8], oa[8], ob[8];
    unpack8(*p1, a);
    unpack8(*p2, b);
    const float* cp = cos_t + (long long)pos * half + c * 8;
    const float* sp = sin_t + (long long)pos * half + c * 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float cs = cp[j], sn = sp[j] * sign;
      oa[j] = a[j] * cs - b[j] * sn;
      ob[j] = b[j] * cs + a[j] * sn;
    }
    *p1 = pack8(oa);
    *p2 = pack8(ob);
  }
}

// Fused softmax cross-entropy: one CTA per token row.
//   loss[t] = logsumexp(logits[t]) - logits[t, target]   (0 where target == ignore_index)
//   dlogits[t, :] = (softmax - onehot) * grad_scale       (written in place when dlogits == logits)
__global__ void __launch_bounds__(512)
cross_entropy_kernel(const __nv_bfloat16* __restrict__ logits, const long long* __restrict__ targets,
                     float* __restrict__ loss, __nv_bfloat16* __restrict__ dlogits, int V,
                     long long ld, float grad_scale, long long ignore_index) {
  __shared__ float red_m[16], red_s[16];
  const long long row = blockIdx.x;
  const __nv_bfloat16* lr = logits + row * ld;
  const long long tgt = targets[row];
  const int vv = V >> 3;
  float mx = -INFINITY, sm = 0.f;
  for (int c = threadIdx.x; c < vv; c += blockDim.x) {
    float f[8];
    unpack8(ld_v4(reinterpret_cast<const uint4*>(lr) + c), f);
    float cm = f[0];
#pragma unroll
    for (int j = 1; j < 8; ++j) cm = fmaxf(cm, f[j]);
    const float nm = fmaxf(mx, cm);
    float add = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) add += __expf(f[j] - nm);
    sm = sm * __expf(mx - nm) + add;
    mx = nm;
  }
  for (int c = (vv << 3) + threadIdx.x; c < V; c += blockDim.x) {  // tail when V % 8 != 0
    const float f = h2f(lr[c]);
    const float nm = fmaxf(mx, f);
    sm = sm * __expf(mx - nm) + __expf(f - nm);
    mx = nm;
  }
  // combine (max, sum) pairs across the block
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, mx, o);
    const float os = __shfl_xor_sync(0xffffffffu, sm, o);
    const float nm = fmaxf(mx, om);
    sm = (mx == -INFINITY ? 0.f : sm * __expf(mx - nm)) + (om == -INFINITY ? 0.f : os * __expf(om - nm));
    mx = nm;
  }
  if ((threadIdx.x & 31) == 0) { red_m[threadIdx.x >> 5] = mx; red_s[threadIdx.x >> 5] = sm; }
  __syncthreads();
  float gm = -INFINITY, gs = 0.f;
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) {
    const float om = red_m[w], os = red_s[w];
    const float nm = fmaxf(gm, om);
    gs = (gm == -INFINITY ? 0.f : gs * __expf(gm - nm)) + (om == -INFINITY ? 0.f : os * __expf(om - nm));
    gm = nm;
  }
  const bool ignored = (tgt == ignore_index) || tgt < 0 || tgt >= V;
  const float lse = gm + __logf(gs);
  if (threadIdx.x == 0)
    loss[row] = ignored ? 0.f : lse - h2f(lr[tgt]);
  if (!dlogits) return;
  __nv_bfloat16* dr = dlogits + row * ld;
  const float inv = 1.f / gs;
  const float scale = ignored ? 0.f : grad_scale;
  for (int c = threadIdx.x; c < vv; c += blockDim.x) {
    float f[8];
    unpack8(ld_v4(reinterpret_cast<const uint4*>(lr) + c), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float pj = __expf(f[j] - gm) * inv;
      if ((long long)(c * 8 + j) == tgt) pj -= 1.f;
      f[j] = pj * scale;
    }
    reinterpret_cast<uint4*>(dr)[c] = pack8(f);
  }
  for (int c = (vv << 3) + threadIdx.x; c < V; c += blockDim.x) {
    float pj = __expf(h2f(lr[c]) - gm) * inv;
    if ((long long)c == tgt) pj -= 1.f;
    dr[c] = f2h(pj * scale);
  }
}

// ---- K9: cross-entropy over vocab-sharded logits (Megatron vocab_parallel_cross_entropy, SURVEY.md
// §2H C10) as two kernels around ONE exchange of per-row statistics.
//   stats  : per row {local max, sum exp(z - local max), z[target] if the target is in this shard}
//            written to slot `rank` of the statistics buffer [world][T][3] — of this GPU, or of
//            every peer (mapped symmetric memory), which makes the exchange a store pattern.
//   finish : combines the `world` slots (global max / sum / target logit), writes the loss and the
//            gradient (softmax - onehot) * scale over the logits in place.
struct CEPeers {
  float* ptr[8];
};

__global__ void __launch_bounds__(512)
vocab_ce_stats_kernel(const __nv_bfloat16* __restrict__ logits, const long long* __restrict__ targets,
                      int Vl, long long ld, long long lo, long long T, int rank, int world,
                      CEPeers out) {
  __shared__ float red_m[16], red_s[16];
  const long long row = blockIdx.x;
  const __nv_bfloat16* lr = logits + row * ld;
  const int vv = Vl >> 3;
  float mx = -INFINITY, sm = 0.f;
  for (int c = threadIdx.x; c < vv; c += blockDim.x) {
    float f[8];
    unpack8(ld_v4(reinterpret_cast<const uint4*>(lr) + c), f);
    float cm = f[0];
#pragma unroll
    for (int j = 1; j < 8; ++j) cm = fmaxf(cm, f[j]);
    const float nm = fmaxf(mx, cm);
    float add = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) add += __expf(f[j] - nm);
    sm = sm * __expf(mx - nm) + add;
    mx = nm;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, mx, o);
    const float os = __shfl_xor_sync(0xffffffffu, sm, o);
    const float nm = fmaxf(mx, om);
    sm = (mx == -INFINITY ? 0.f : sm * __expf(mx - nm)) + (om == -INFINITY ? 0.f : os * __expf(om - nm));
    mx = nm;
  }
  if ((threadIdx.x & 31) == 0) { red_m[threadIdx.x >> 5] = mx; red_s[threadIdx.x >> 5] = sm; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float gm = -INFINITY, gs = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) {
      const float om = red_m[w], os = red_s[w];
      const float nm = fmaxf(gm, om);
      gs = (gm == -INFINITY ? 0.f : gs * __expf(gm - nm)) + (om == -INFINITY ? 0.f : os * __expf(om - nm));
      gm = nm;
    }
    const long long tgt = targets[row] - lo;
    const float tl = (tgt >= 0 && tgt < Vl) ? h2f(lr[tgt]) : 0.f;
    for (int r = 0; r < world; ++r) {
      if (!out.ptr[r]) continue;
      float* dst = out.ptr[r] + ((long long)rank * T + row) * 3;
      dst[0] = gm; dst[1] = gs; dst[2] = tl;
    }
  }
}

__global__ void __launch_bounds__(512)
vocab_ce_finish_kernel(__nv_bfloat16* __restrict__ logits, const long long* __restrict__ targets,
                       int Vl, long long ld, long long lo, long long T, int world,
                       const float* __restrict__ stats /*[world][T][3]*/, long long ignore_index,
                       float grad_scale, float* __restrict__ loss) {
  const long long row = blockIdx.x;
  float gm = -INFINITY;
  for (int r = 0; r < world; ++r) gm = fmaxf(gm, stats[((long long)r * T + row) * 3]);
  float gs = 0.f, tl = 0.f;
  for (int r = 0; r < world; ++r) {
    const float* st = stats + ((long long)r * T + row) * 3;
    gs += st[1] * __expf(st[0] - gm);
    tl += st[2];
  }
  const long long tgt_g = targets[row];
  const bool ignored = tgt_g == ignore_index;
  if (threadIdx.x == 0) loss[row] = ignored ? 0.f : gm + __logf(gs) - tl;
  const long long tgt = tgt_g - lo;
  const float inv = 1.f / gs;
  const float scale = ignored ? 0.f : grad_scale;
  __nv_bfloat16* lr = logits + row * ld;
  const int vv = Vl >> 3;
  for (int c = threadIdx.x; c < vv; c += blockDim.x) {
    float f[8];
    unpack8(ld_v4(reinterpret_cast<const uint4*>(lr) + c), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float pj = __expf(f[j] - gm) * inv;
      if ((long long)(c * 8 + j) == tgt) pj -= 1.f;
      f[j] = pj * scale;
    }
    reinterpret_cast<uint4*>(lr)[c] = pack8(f);
  }
}

// Row softmax with optional causal / additive mask — the standalone
// scaled_(upper_triang_)masked_softmax of Megatron; the
// training path uses the fused attention kernel, this is kept for parity.
__global__ void __launch_bounds__(256)
scaled_masked_softmax_kernel(const __nv_bfloat16* __restrict__ x, const uint8_t* __restrict__ mask,
                             __nv_bfloat16* __restrict__ y, int sq, int sk, float scale,
                             int causal, long long mask_batch_stride, int heads) {
  __shared__ float red[8];
  const long long row = blockIdx.x;  // over [b*heads*sq]
  const int q = (int)(row % sq);
  const long long bh = row / sq;
  const long long b = bh / heads;
  const __nv_bfloat16* xr = x + row * sk;
  const uint8_t* mr = mask ? mask + b * mask_batch_stride + (long long)q * sk : nullptr;
  const int limit = causal ? (q + 1 + (sk - sq)) : sk;
  float mx = -INFINITY;
  for (int c = threadIdx.x; c < sk; c += blockDim.x) {
    const bool dead = c >= limit || (mr && mr[c]);
    if (!dead) mx = fmaxf(mx, h2f(xr[c]) * scale);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
  __syncthreads();
  mx = red[0];
  for (int w = 1; w < 8; ++w) mx = fmaxf(mx, red[w]);
  __syncthreads();
  float sm = 0.f;
  for (int c = threadIdx.x; c < sk; c += blockDim.x) {
    const bool dead = c >= limit || (mr && mr[c]);
    if (!dead) sm += __expf(h2f(xr[c]) * scale - mx);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sm += __shfl_xor_sync(0xffffffffu, sm, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sm;
  __syncthreads();
  sm = 0.f;
  for (int w = 0; w < 8; ++w) sm += red[w];
  const float inv = sm > 0.f ? 1.f / sm : 0.f;
  for (int c = threadIdx.x; c < sk; c += blockDim.x) {
    const bool dead = c >= limit || (mr && mr[c]);
    y[row * sk + c] =
        f2h(dead ? 0.f : __expf(h2f(xr[c]) * scale - mx) * inv);
  }
}

// fp32 -> bf16 cast with optional scale (bucket cast of the DDP path).
__global__ void __launch_bounds__(256)
cast_f32_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, long long n,
                     float scale) {
  const long long nv = n >> 3;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
       i += (long long)gridDim.x * blockDim.x) {
    float4 a = reinterpret_cast<const float4*>(x)[2 * i], b = reinterpret_cast<const float4*>(x)[2 * i + 1];
    float f[8] = {a.x * scale, a.y * scale, a.z * scale, a.w * scale,
                  b.x * scale, b.y * scale, b.z * scale, b.w * scale};
    reinterpret_cast<uint4*>(y)[i] = pack8(f);
  }
  const long long t = (nv << 3) + (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) y[t] = f2h(x[t] * scale);
}

}  // namespace

extern "C" int b2f_swiglu_fwd(const void* gu, void* out, long long rows, int F, void* stream) {
  B2F_REQUIRE(F % 8 == 0 && rows > 0, "swiglu: F must be a multiple of 8");
  swiglu_fwd_kernel<<<ew_grid(rows * (F / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)gu, (__nv_bfloat16*)out, rows, F);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b2f_swiglu_bwd(const void* gu, const void* dout, void* dgu, long long rows, int F,
                              void* stream) {
  B2F_REQUIRE(F % 8 == 0 && rows > 0, "swiglu: F must be a multiple of 8");
  swiglu_bwd_kernel<<<ew_grid(rows * (F / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)gu, (const __nv_bfloat16*)dout, (__nv_bfloat16*)dgu, rows, F);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b2f_gelu_fwd(const void* x, const void* bias, void* y, long long rows, int H,
                            void* stream) {
  B2F_REQUIRE(H % 8 == 0 && rows > 0, "gelu: H must be a multiple of 8");
  gelu_kernel<false><<<ew_grid(rows * (H / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, (const __nv_bfloat16*)bias, nullptr, (__nv_bfloat16*)y, rows, H);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b2f_gelu_bwd(const void* x, const void* bias, const void* dy, void* dx,
                            long long rows, int H, void* stream) {
  B2F_REQUIRE(H % 8 == 0 && rows > 0, "gelu: H must be a multiple of 8");
  gelu_kernel<true><<<ew_grid(rows * (H / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, (const __nv_bfloat16*)bias, (const __nv_bfloat16*)dy,
      (__nv_bfloat16*)dx, rows, H);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b2f_bias_dropout_add(const void* x, const void* bias, const void* residual,
                                    void* out, void* mask, long long rows, int H, float p,
                                    unsigned long long seed, void* stream) {
  B2F_REQUIRE(H % 8 == 0 && rows > 0 && p >= 0.f && p < 1.f, "bias_dropout_add: bad arguments");
  bias_dropout_add_kernel<<<ew_grid(rows * (H / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, (const __nv_bfloat16*)bias, (const __nv_bfloat16*)residual,
      (__nv_bfloat16*)out, (uint8_t*)mask, rows, H, p, seed);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b2f_dropout_bwd(const void* dy, const void* mask, void* dx, long long n, float p,
                               void* stream) {
  B2F_REQUIRE(n % 8 == 0 && n > 0 && p >= 0.f && p < 1.f, "dropout_bwd: n must be a multiple of 8");
  dropout_bwd_kernel<<<ew_grid(n / 8), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dy, (const uint8_t*)mask, (__nv_bfloat16*)dx, n / 8, p);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b2f_rope(void* x, const float* cos_t, const float* sin_t, long long tokens, int seq,
                        int nh, int d, long long token_stride, int pos_offset, int pos_div,
                        int backward, void* stream) {
  B2F_REQUIRE(d % 16 == 0 && tokens > 0 && token_stride % 8 == 0, "rope: head dim must be a multiple of 16");
  rope_kernel<<<ew_grid(tokens * nh * (d / 16)), 256, 0, (cudaStream_t)stream>>>(
      (__nv_bfloat16*)x, cos_t, sin_t, tokens, seq, nh, d, token_stride, pos_offset,
      pos_div > 0 ? pos_div : 1, backward ? -1.f : 1.f);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b2f_cross_entropy(const void* logits, const long long* targets, float* loss,
                                 void* dlogits, long long rows, int V, long long ld,
                                 float grad_scale, long long ignore_index, void* stream) {
  B2F_REQUIRE(rows > 0 && V > 0 && ld % 8 == 0, "cross_entropy: ld must be a multiple of 8");
  cross_entropy_kernel<<<(unsigned)rows, 512, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)logits, targets, loss, (__nv_bfloat16*)dlogits, V, ld, grad_scale,
      ignore_index);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
// stats_ptrs: `world` device addresses of the [world][T][3] fp32 statistics buffer on every rank
// (symmetric memory), or only entry `rank` non-zero when the exchange is done by a collective.
extern "C" int b2f_vocab_ce_stats(const void* logits, const long long* targets, long long T, int Vl,
                                  long long ld, long long lo, int rank, int world,
                                  const long long* stats_ptrs, void* stream) {
  B2F_REQUIRE(T > 0 && Vl > 0 && Vl % 8 == 0 && ld % 8 == 0, "vocab CE: shard width and ld must be multiples of 8");
  B2F_REQUIRE(world >= 1 && world <= 8 && rank >= 0 && rank < world, "vocab CE: world must be 1..8");
  CEPeers out;
  for (int r = 0; r < 8; ++r) out.ptr[r] = r < world ? reinterpret_cast<float*>(stats_ptrs[r]) : nullptr;
  vocab_ce_stats_kernel<<<(unsigned)T, 512, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)logits, targets, Vl, ld, lo, T, rank, world, out);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b2f_vocab_ce_finish(void* logits, const long long* targets, long long T, int Vl, long long ld,
                                   long long lo, int world, const float* stats, long long ignore_index,
                                   float grad_scale, float* loss, void* stream) {
  B2F_REQUIRE(T > 0 && Vl > 0 && Vl % 8 == 0 && ld % 8 == 0, "vocab CE: shard width and ld must be multiples of 8");
  vocab_ce_finish_kernel<<<(unsigned)T, 512, 0, (cudaStream_t)stream>>>(
      (__nv_bfloat16*)logits, targets, Vl, ld, lo, T, world, stats, ignore_index, grad_scale, loss);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b2f_scaled_masked_softmax(const void* x, const void* mask, void* y, long long batch,
                                         int heads, int sq, int sk, float scale, int causal,
                                         void* stream) {
  B2F_REQUIRE(batch > 0 && heads > 0 && sq > 0 && sk > 0, "softmax: bad shape");
  scaled_masked_softmax_kernel<<<(unsigned)(batch * heads * sq), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, (const uint8_t*)mask, (__nv_bfloat16*)y, sq, sk, scale, causal,
      (long long)sq * sk, heads);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b2f_cast_f32_bf16(const float* x, void* y, long long n, float scale, void* stream) {
  if (n <= 0) return 0;
  cast_f32_bf16_kernel<<<ew_grid((n + 7) / 8), 256, 0, (cudaStream_t)stream>>>(
      x, (__nv_bfloat16*)y, n, scale);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
