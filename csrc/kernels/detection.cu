// b200flow K12-K15 — Mask R-CNN detection kernels for sm_100a.
//
//  K12 RoIAlign forward/backward over an FPN pyramid: one launch covers all ROIs, the
//      pyramid level k = floor(4 + log2(sqrt(area)/224)) is chosen in the kernel, features
//      are NHWC bf16 so every bilinear tap is a 16-byte load of 8 channels.
//  K13 box decode + clip, and batched greedy NMS: IoU bit-matrix kernel plus an on-device
//      sweep (no host round trip), one segment per (image, level).
//  K14 proposal <-> ground-truth matcher: max / argmax IoU per proposal without ever
//      materialising the N x M matrix.
//  K15 mask-target crop (GT mask -> 28x28 in its ROI) and mask paste (28x28 -> box in the
//      image, threshold 0.5).  Raster-style arithmetic, plain CUDA, no graphics API.
//
// Capability parity: tf.image.crop_and_resize / non_max_suppression / tensorpack's
// model_box.py, model_frcnn.py, model_mrcnn.py used by the reference's Mask R-CNN
// workloads (examples/maskrcnn/*.yaml) and the notebook's mask
// paste (containers/aws-samples-maskrcnn/notebooks/mask-rcnn-tensorflow-viz.ipynb:190-193).
#include "../common/ptx.cuh"
#include "../common/host.h"
#include "../../include/b200flow.h"

using namespace b2f;

namespace {

struct Pyramid {
  const __nv_bfloat16* feat[4];   // NHWC, levels P2..P5
  float* grad[4];                 // NHWC fp32 (backward)
  int H[4], W[4];
  float scale[4];                 // 1/4, 1/8, 1/16, 1/32
  int C;
};

__device__ __forceinline__ int fpn_level(float x1, float y1, float x2, float y2) {
  const float area = fmaxf(x2 - x1, 0.f) * fmaxf(y2 - y1, 0.f);
  int k = (int)floorf(4.f + log2f(sqrtf(area) / 224.f + 1e-6f));
  k = k < 2 ? 2 : (k > 5 ? 5 : k);
  return k - 2;
}

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}

struct Taps {
  int y0, y1, x0, x1;
  float w00, w01, w10, w11;
  bool valid;
};
__device__ __forceinline__ Taps bilinear_taps(float y, float x, int H, int W) {
  Taps t;
  t.valid = !(y < -1.f || y > H || x < -1.f || x > W);
  y = fminf(fmaxf(y, 0.f), (float)(H - 1));
  x = fminf(fmaxf(x, 0.f), (float)(W - 1));
  t.y0 = (int)y; t.x0 = (int)x;
  t.y1 = t.y0 + 1 < H ? t.y0 + 1 : H - 1;
  t.x1 = t.x0 + 1 < W ? t.x0 + 1 : W - 1;
  const float ly = y - t.y0, lx = x - t.x0;
  t.w00 = (1 - ly) * (1 - lx); t.w01 = (1 - ly) * lx; t.w10 = ly * (1 - lx); t.w11 = ly * lx;
  return t;
}

// p[0..8) += g[0..8) * w as two 16-byte reductions (REDG.ADD.F32x4): a quarter of the atomic operations of the
// scalar form; p is 32-byte aligned (8-channel groups of an NHWC fp32 buffer).
__device__ __forceinline__ void red_add8(float* p, const float (&g)[8], float w) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(g[0] * w), "f"(g[1] * w),
               "f"(g[2] * w), "f"(g[3] * w) : "memory");
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p + 4), "f"(g[4] * w), "f"(g[5] * w),
               "f"(g[6] * w), "f"(g[7] * w) : "memory");
}

// rois: [R, 5] = (batch, x1, y1, x2, y2) in image coordinates. out: [R, P, P, C] bf16 (NHWC).
// One thread = one (roi, ph, pw, 8-channel group).  `aligned`: half-pixel shift.
template <bool BWD>
__global__ void __launch_bounds__(256)
roi_align_kernel(Pyramid pyr, const float* __restrict__ rois, int R, int P, int sampling,
                 __nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ dout,
                 int aligned, int fixed_level) {
  const int cg = pyr.C >> 3;
  const long long total = (long long)R * P * P * cg;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int c8 = (int)(idx % cg);
    long long t = idx / cg;
    const int pw = (int)(t % P); t /= P;
    const int ph = (int)(t % P);
    const int r = (int)(t / P);
    const float* roi = rois + (long long)r * 5;
    const int b = (int)roi[0];
    const int lv = fixed_level >= 0 ? fixed_level : fpn_level(roi[1], roi[2], roi[3], roi[4]);
    const float sc = pyr.scale[lv];
    const float off = aligned ? 0.5f : 0.f;
    const float x1 = roi[1] * sc - off, y1 = roi[2] * sc - off;
    float rw = roi[3] * sc - off - x1, rh = roi[4] * sc - off - y1;
    if (!aligned) { rw = fmaxf(rw, 1.f); rh = fmaxf(rh, 1.f); }
    const float bw = rw / P, bh = rh / P;
    const int gh = sampling > 0 ? sampling : (int)ceilf(rh / P);
    const int gw = sampling > 0 ? sampling : (int)ceilf(rw / P);
    const float inv = 1.f / fmaxf((float)(gh * gw), 1.f);
    const int H = pyr.H[lv], W = pyr.W[lv], C = pyr.C;
    const __nv_bfloat16* f = pyr.feat[lv] + (long long)b * H * W * C + c8 * 8;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    float g[8];
    if (BWD) {
      unpack8(*reinterpret_cast<const uint4*>(dout + (((long long)r * P + ph) * P + pw) * C + c8 * 8), g);
#pragma unroll
      for (int j = 0; j < 8; ++j) g[j] *= inv;
    }
    for (int iy = 0; iy < gh; ++iy) {
      const float y = y1 + ph * bh + (iy + 0.5f) * bh / gh;
      for (int ix = 0; ix < gw; ++ix) {
        const float x = x1 + pw * bw + (ix + 0.5f) * bw / gw;
        const Taps tp = bilinear_taps(y, x, H, W);
        if (!tp.valid) continue;
        if (!BWD) {
          float v00[8], v01[8], v10[8], v11[8];
          unpack8(__ldg(reinterpret_cast<const uint4*>(f + ((long long)tp.y0 * W + tp.x0) * C)), v00);
          unpack8(__ldg(reinterpret_cast<const uint4*>(f + ((long long)tp.y0 * W + tp.x1) * C)), v01);
          unpack8(__ldg(reinterpret_cast<const uint4*>(f + ((long long)tp.y1 * W + tp.x0) * C)), v10);
          unpack8(__ldg(reinterpret_cast<const uint4*>(f + ((long long)tp.y1 * W + tp.x1) * C)), v11);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            acc[j] += tp.w00 * v00[j] + tp.w01 * v01[j] + tp.w10 * v10[j] + tp.w11 * v11[j];
        } else {
          float* gb = pyr.grad[lv] + (long long)b * H * W * C + c8 * 8;
          red_add8(gb + ((long long)tp.y0 * W + tp.x0) * C, g, tp.w00);
          red_add8(gb + ((long long)tp.y0 * W + tp.x1) * C, g, tp.w01);
          red_add8(gb + ((long long)tp.y1 * W + tp.x0) * C, g, tp.w10);
          red_add8(gb + ((long long)tp.y1 * W + tp.x1) * C, g, tp.w11);
        }
      }
    }
    if (!BWD) {
      uint4 o;
      o.x = pack_bf16x2(acc[0] * inv, acc[1] * inv); o.y = pack_bf16x2(acc[2] * inv, acc[3] * inv);
      o.z = pack_bf16x2(acc[4] * inv, acc[5] * inv); o.w = pack_bf16x2(acc[6] * inv, acc[7] * inv);
      *reinterpret_cast<uint4*>(out + (((long long)r * P + ph) * P + pw) * C + c8 * 8) = o;
    }
  }
}

// ---------------------------------------------------------------- conv epilogue for the folded-BN backbone
// y = act(y + bias[c] (+ residual)) in place over an NHWC half-precision tensor, eight channels per thread.  cuDNN
// leaves a convolution's bias to a separate broadcast add (an unvectorised kernel on channels-last data) and the
// ReLU and the residual add to two more passes; this is all three in one pass over the activation.
__global__ void __launch_bounds__(256)
bias_act_nhwc_kernel(__nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ bias,
                     const __nv_bfloat16* __restrict__ residual, long long n8, int C8, int relu) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8;
       i += (long long)gridDim.x * blockDim.x) {
    float v[8], b[8];
    uint4* py = reinterpret_cast<uint4*>(y) + i;
    unpack8(*py, v);
    unpack8(__ldg(reinterpret_cast<const uint4*>(bias) + (int)(i % C8)), b);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] += b[j];
    if (residual) {
      unpack8(__ldg(reinterpret_cast<const uint4*>(residual) + i), b);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] += b[j];
    }
    if (relu) {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], 0.f);
    }
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
    o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
    *py = o;
  }
}

// Column sums of a half-precision [rows, C] matrix (bias gradients: of a linear layer, or of a convolution over an NHWC
// gradient) as per-block partial sums: partial[blockIdx.x][c] = sum over this block's rows.  A block covers up to 256
// groups of eight channels (blockIdx.y selects which) with 256 / groups row lanes; fp32 accumulation in registers, one
// shared-memory reduction over the lanes, no atomics - the caller adds the few hundred partial rows, so the result
// does not depend on scheduling.
__global__ void __launch_bounds__(256)
colsum_partial_kernel(const __nv_bfloat16* __restrict__ g, float* __restrict__ partial, long long rows, int C8) {
  __shared__ float part[256 * 8];
  const int g0 = blockIdx.y * 256;
  const int G = C8 - g0 < 256 ? C8 - g0 : 256;
  const int lanes = 256 / G;
  const int c8 = threadIdx.x % G, lane = threadIdx.x / G;
  const long long per = (rows + gridDim.x - 1) / gridDim.x;
  const long long r0 = (long long)blockIdx.x * per, r1 = r0 + per < rows ? r0 + per : rows;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (lane < lanes) {
#pragma unroll 4
    for (long long r = r0 + lane; r < r1; r += lanes) {
      float v[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(g) + r * C8 + g0 + c8), v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) part[threadIdx.x * 8 + j] = acc[j];
  __syncthreads();
  if (lane == 0) {
    float* dst = partial + (long long)blockIdx.x * C8 * 8 + (long long)(g0 + c8) * 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float t = 0.f;
      for (int l = 0; l < lanes; ++l) t += part[(l * G + c8) * 8 + j];
      dst[j] = t;
    }
  }
}

// ---------------------------------------------------------------- K13 decode
// boxes = decode(anchors, deltas) with weights (wx, wy, ww, wh), clipped to the image.
__global__ void __launch_bounds__(256)
box_decode_kernel(const float* __restrict__ anchors, const float* __restrict__ deltas,
                  float* __restrict__ out, long long n, float wx, float wy, float ww, float wh,
                  float img_w, float img_h, float max_ratio) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const float4 a = reinterpret_cast<const float4*>(anchors)[i];
    const float4 d = reinterpret_cast<const float4*>(deltas)[i];
    const float w = a.z - a.x, h = a.w - a.y;
    const float cx = a.x + 0.5f * w, cy = a.y + 0.5f * h;
    const float dw = fminf(d.z / ww, max_ratio), dh = fminf(d.w / wh, max_ratio);
    const float pcx = d.x / wx * w + cx, pcy = d.y / wy * h + cy;
    const float pw = __expf(dw) * w, ph = __expf(dh) * h;
    float4 o;
    o.x = fminf(fmaxf(pcx - 0.5f * pw, 0.f), img_w);
    o.y = fminf(fmaxf(pcy - 0.5f * ph, 0.f), img_h);
    o.z = fminf(fmaxf(pcx + 0.5f * pw, 0.f), img_w);
    o.w = fminf(fmaxf(pcy + 0.5f * ph, 0.f), img_h);
    reinterpret_cast<float4*>(out)[i] = o;
  }
}

__device__ __forceinline__ float iou(const float4& a, const float4& b) {
  const float iw = fminf(a.z, b.z) - fmaxf(a.x, b.x), ih = fminf(a.w, b.w) - fmaxf(a.y, b.y);
  const float inter = fmaxf(iw, 0.f) * fmaxf(ih, 0.f);
  const float ua = (a.z - a.x) * (a.w - a.y) + (b.z - b.x) * (b.w - b.y) - inter;
  return ua > 0.f ? inter / ua : 0.f;
}

// ------------------------------------------------------------------- K13 NMS
// Segments: boxes [total, 4] sorted by descending score inside each segment;
// seg_off[s] .. seg_off[s+1].  mask: u64 [total, words] with words = ceil(max_seg / 64).
constexpr int NMS_T = 64;
__global__ void __launch_bounds__(NMS_T)
nms_mask_kernel(const float* __restrict__ boxes, const int* __restrict__ seg_off, float thr,
                unsigned long long* __restrict__ mask, int words) {
  const int s = blockIdx.z;
  const int base = seg_off[s], n = seg_off[s + 1] - base;
  const int row0 = blockIdx.y * NMS_T, col0 = blockIdx.x * NMS_T;
  // blocks strictly below the diagonal are never read by the sweep (a box can only be
  // suppressed by a higher-scoring one), so they are skipped entirely
  if (row0 >= n || col0 >= n || col0 + NMS_T <= row0) return;
  __shared__ float4 cb[NMS_T];
  const int nc = n - col0 < NMS_T ? n - col0 : NMS_T;
  if ((int)threadIdx.x < nc) cb[threadIdx.x] = reinterpret_cast<const float4*>(boxes)[base + col0 + threadIdx.x];
  __syncthreads();
  const int i = row0 + threadIdx.x;
  if (i < n) {
    const float4 a = reinterpret_cast<const float4*>(boxes)[base + i];
    unsigned long long bits = 0ull;
    const int start = row0 == col0 ? threadIdx.x + 1 : 0;
    for (int j = start; j < nc; ++j)
      if (iou(a, cb[j]) > thr) bits |= 1ull << j;
    mask[(long long)(base + i) * words + blockIdx.x] = bits;
  }
}

// One block per segment walks the rows in score order; `removed` lives in shared memory.
__global__ void __launch_bounds__(256)
nms_sweep_kernel(const unsigned long long* __restrict__ mask, const int* __restrict__ seg_off,
                 int words, uint8_t* __restrict__ keep) {
  extern __shared__ unsigned long long removed[];
  const int s = blockIdx.x;
  const int base = seg_off[s], n = seg_off[s + 1] - base;
  const int nw = (n + 63) / 64;
  for (int w = threadIdx.x; w < nw; w += blockDim.x) removed[w] = 0ull;
  __syncthreads();
  for (int i = 0; i < n; ++i) {
    const bool dead = (removed[i >> 6] >> (i & 63)) & 1ull;   // uniform read
    if (!dead) {
      const unsigned long long* row = mask + (long long)(base + i) * words;
      for (int w = (i >> 6) + threadIdx.x; w < nw; w += blockDim.x) removed[w] |= row[w];
    }
    if (threadIdx.x == 0) keep[base + i] = dead ? 0 : 1;
    __syncthreads();
  }
}

// ------------------------------------------------------------------- K14
// For every proposal: best IoU over the image's ground-truth boxes and its index.
// gt_off[b] .. gt_off[b+1] index gt boxes of image b; prop_batch[i] = image of proposal i.
__global__ void __launch_bounds__(256)
match_kernel(const float* __restrict__ props, const int* __restrict__ prop_batch, long long n,
             const float* __restrict__ gts, const int* __restrict__ gt_off,
             float* __restrict__ best_iou, int* __restrict__ best_idx) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const float4 p = reinterpret_cast<const float4*>(props)[i];
    const int b = prop_batch ? prop_batch[i] : 0;
    float bi = -1.f;
    int bj = -1;
    for (int j = gt_off[b]; j < gt_off[b + 1]; ++j) {
      const float v = iou(p, reinterpret_cast<const float4*>(gts)[j]);
      if (v > bi) { bi = v; bj = j; }
    }
    best_iou[i] = bj >= 0 ? bi : 0.f;
    best_idx[i] = bj;
  }
}

// Dense IoU matrix (anchor labelling needs per-gt maxima as well).
__global__ void __launch_bounds__(256)
pairwise_iou_kernel(const float* __restrict__ a, long long na, const float* __restrict__ b, int nb,
                    float* __restrict__ out) {
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < na * nb;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long i = idx / nb;
    const int j = (int)(idx - i * nb);
    out[idx] = iou(reinterpret_cast<const float4*>(a)[i], reinterpret_cast<const float4*>(b)[j]);
  }
}

// ------------------------------------------------------------------- K15
// Mask targets: sample GT instance masks (u8 [G, H, W]) inside ROIs -> [R, M, M] float (0/1 after >= 0.5).
__global__ void __launch_bounds__(256)
mask_crop_kernel(const uint8_t* __restrict__ masks, int H, int W, const float* __restrict__ rois,
                 const int* __restrict__ gt_index, int R, int M, float* __restrict__ out) {
  const long long total = (long long)R * M * M;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int px = (int)(idx % M), py = (int)((idx / M) % M), r = (int)(idx / ((long long)M * M));
    const float* roi = rois + (long long)r * 4;
    const uint8_t* m = masks + (long long)gt_index[r] * H * W;
    const float x = roi[0] + (px + 0.5f) * (roi[2] - roi[0]) / M - 0.5f;
    const float y = roi[1] + (py + 0.5f) * (roi[3] - roi[1]) / M - 0.5f;
    const Taps t = bilinear_taps(y, x, H, W);
    float v = 0.f;
    if (t.valid)
      v = t.w00 * m[(long long)t.y0 * W + t.x0] + t.w01 * m[(long long)t.y0 * W + t.x1] +
          t.w10 * m[(long long)t.y1 * W + t.x0] + t.w11 * m[(long long)t.y1 * W + t.x1];
    out[idx] = v >= 0.5f ? 1.f : 0.f;
  }
}

// Mask paste: each detection's M x M probability map is bilinearly resized into its box and
// thresholded; out: u8 [R, H, W].
__global__ void __launch_bounds__(256)
mask_paste_kernel(const float* __restrict__ probs, int M, const float* __restrict__ boxes, int R,
                  int H, int W, float thr, uint8_t* __restrict__ out) {
  const long long total = (long long)R * H * W;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(idx % W), y = (int)((idx / W) % H), r = (int)(idx / ((long long)H * W));
    const float* b = boxes + (long long)r * 4;
    const float bw = fmaxf(b[2] - b[0], 1.f), bh = fmaxf(b[3] - b[1], 1.f);
    uint8_t o = 0;
    if (x + 0.5f >= b[0] && x + 0.5f <= b[2] && y + 0.5f >= b[1] && y + 0.5f <= b[3]) {
      const float mx = (x + 0.5f - b[0]) / bw * M - 0.5f, my = (y + 0.5f - b[1]) / bh * M - 0.5f;
      const Taps t = bilinear_taps(my, mx, M, M);
      const float* p = probs + (long long)r * M * M;
      const float v = t.w00 * p[t.y0 * M + t.x0] + t.w01 * p[t.y0 * M + t.x1] +
                      t.w10 * p[t.y1 * M + t.x0] + t.w11 * p[t.y1 * M + t.x1];
      o = v > thr ? 1 : 0;
    }
    out[idx] = o;
  }
}

inline int grid1d(long long n) {
  long long b = (n + 255) / 256;
  const long long cap = (long long)b2f::sm_count() * 16;
  return (int)(b > cap ? cap : (b < 1 ? 1 : b));
}

int fill_pyramid(Pyramid* p, const void* const* feats, float* const* grads, const int* H, const int* W,
                 const float* scales, int levels, int C) {
  B2F_REQUIRE(levels >= 1 && levels <= 4 && C % 8 == 0, "roi_align: 1..4 levels, C multiple of 8");
  for (int i = 0; i < 4; ++i) {
    const int j = i < levels ? i : levels - 1;
    p->feat[i] = static_cast<const __nv_bfloat16*>(feats[j]);
    p->grad[i] = grads ? grads[j] : nullptr;
    p->H[i] = H[j]; p->W[i] = W[j]; p->scale[i] = scales[j];
  }
  p->C = C;
  return 0;
}

// ---- K16: polygon -> binary mask (pycocotools maskApi frPoly): one thread per pixel of
// one object; even-odd crossing count at the pixel centre over all edges of the object's polygons
// (a COCO object may have several parts; a pixel inside any part is set).
//   xy          float [sum(points), 2] vertices of all polygons, object after object
//   poly_off    int [P + 1]  first vertex of each polygon
//   obj_poly    int [N + 1]  first polygon of each object
__global__ void __launch_bounds__(256)
poly_to_mask_kernel(const float* __restrict__ xy, const int* __restrict__ poly_off,
                    const int* __restrict__ obj_poly, int H, int W, unsigned char* __restrict__ out) {
  const int obj = blockIdx.y;
  const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= (long long)H * W) return;
  const float px = (float)(pix % W) + 0.5f, py = (float)(pix / W) + 0.5f;
  bool inside = false;
  for (int pl = obj_poly[obj]; pl < obj_poly[obj + 1] && !inside; ++pl) {
    const int v0 = poly_off[pl], n = poly_off[pl + 1] - v0;
    bool odd = false;
    for (int i = 0, j = n - 1; i < n; j = i++) {
      const float xi = xy[2 * (v0 + i)], yi = xy[2 * (v0 + i) + 1];
      const float xj = xy[2 * (v0 + j)], yj = xy[2 * (v0 + j) + 1];
      if ((yi > py) != (yj > py)) {
        const float xc = (xj - xi) * (py - yi) / (yj - yi) + xi;
        if (px < xc) odd = !odd;
      }
    }
    inside = odd;
  }
  out[(long long)obj * H * W + pix] = inside ? 1 : 0;
}

}  // namespace

extern "C" int b2f_roi_align_fwd(const void* const* feats, const int* H, const int* W,
                                 const float* scales, int levels, int C, const float* rois, int R,
                                 int P, int sampling, int aligned, void* out, void* stream) {
  if (R == 0) return 0;
  Pyramid p;
  if (int rc = fill_pyramid(&p, feats, nullptr, H, W, scales, levels, C)) return rc;
  roi_align_kernel<false><<<grid1d((long long)R * P * P * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      p, rois, R, P, sampling, (__nv_bfloat16*)out, nullptr, aligned, levels == 1 ? 0 : -1);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// grads: fp32 NHWC buffers (zeroed by the caller) the same shape as the features.
extern "C" int b2f_roi_align_bwd(const void* const* feats, float* const* grads, const int* H,
                                 const int* W, const float* scales, int levels, int C,
                                 const float* rois, int R, int P, int sampling, int aligned,
                                 const void* dout, void* stream) {
  if (R == 0) return 0;
  Pyramid p;
  if (int rc = fill_pyramid(&p, feats, grads, H, W, scales, levels, C)) return rc;
  roi_align_kernel<true><<<grid1d((long long)R * P * P * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      p, rois, R, P, sampling, nullptr, (const __nv_bfloat16*)dout, aligned, levels == 1 ? 0 : -1);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_bias_act_nhwc(void* y, const void* bias, const void* residual, long long numel, int C,
                                 int relu, void* stream) {
  if (numel == 0) return 0;
  B2F_REQUIRE(C > 0 && C % 8 == 0 && numel % C == 0, "bias_act_nhwc: channels must be a multiple of 8");
  bias_act_nhwc_kernel<<<grid1d(numel / 8), 256, 0, (cudaStream_t)stream>>>(
      (__nv_bfloat16*)y, (const __nv_bfloat16*)bias, (const __nv_bfloat16*)residual, numel / 8, C / 8, relu);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Number of partial rows b2f_colsum_partial writes for a matrix of `rows` rows and C columns.
extern "C" int b2f_colsum_num_partials(long long rows, int C) {
  const int ychunks = (C / 8 + 255) / 256;
  long long blocks = (rows + 127) / 128;
  long long cap = (long long)sm_count() * 4 / ychunks;
  if (cap < 1) cap = 1;
  if (blocks > cap) blocks = cap;
  return blocks < 1 ? 1 : (int)blocks;
}

// partial: fp32 [b2f_colsum_num_partials(rows, C), C]; every element is written.  C % 8 == 0.
extern "C" int b2f_colsum_partial(const void* g, float* partial, long long rows, int C, void* stream) {
  B2F_REQUIRE(C > 0 && C % 8 == 0 && rows >= 0, "colsum: columns must be a multiple of 8");
  dim3 grid(b2f_colsum_num_partials(rows, C), (C / 8 + 255) / 256);
  colsum_partial_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)g, partial, rows, C / 8);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_box_decode(const float* anchors, const float* deltas, float* out, long long n,
                              float wx, float wy, float ww, float wh, float img_w, float img_h,
                              void* stream) {
  if (n == 0) return 0;
  box_decode_kernel<<<grid1d(n), 256, 0, (cudaStream_t)stream>>>(anchors, deltas, out, n, wx, wy, ww, wh,
                                                                img_w, img_h, 4.135166556742356f);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// mask scratch: u64 [total * ceil(max_seg/64)] supplied by the caller; keep: u8 [total].
extern "C" int b2f_nms_batched(const float* boxes, const int* seg_off, int num_segs, int max_seg,
                               float iou_thr, unsigned long long* mask_scratch, unsigned char* keep,
                               void* stream) {
  if (num_segs == 0 || max_seg == 0) return 0;
  const int words = (max_seg + 63) / 64;
  B2F_REQUIRE(words * 8 <= 200 * 1024, "nms: segment too long for the on-device sweep");
  dim3 grid(words, words, num_segs);
  cudaStream_t st = (cudaStream_t)stream;
  nms_mask_kernel<<<grid, NMS_T, 0, st>>>(boxes, seg_off, iou_thr, mask_scratch, words);
  B2F_CHECK_CUDA(cudaGetLastError());
  nms_sweep_kernel<<<num_segs, 256, words * 8, st>>>(mask_scratch, seg_off, words, keep);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_match_proposals(const float* props, const int* prop_batch, long long n,
                                   const float* gts, const int* gt_off, float* best_iou,
                                   int* best_idx, void* stream) {
  if (n == 0) return 0;
  match_kernel<<<grid1d(n), 256, 0, (cudaStream_t)stream>>>(props, prop_batch, n, gts, gt_off, best_iou,
                                                            best_idx);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_pairwise_iou(const float* a, long long na, const float* b, int nb, float* out,
                                void* stream) {
  if (na == 0 || nb == 0) return 0;
  pairwise_iou_kernel<<<grid1d(na * nb), 256, 0, (cudaStream_t)stream>>>(a, na, b, nb, out);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_mask_crop(const unsigned char* masks, int H, int W, const float* rois,
                             const int* gt_index, int R, int M, float* out, void* stream) {
  if (R == 0) return 0;
  mask_crop_kernel<<<grid1d((long long)R * M * M), 256, 0, (cudaStream_t)stream>>>(masks, H, W, rois,
                                                                                 gt_index, R, M, out);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b2f_mask_paste(const float* probs, int M, const float* boxes, int R, int H, int W,
                              float thr, unsigned char* out, void* stream) {
  if (R == 0) return 0;
  mask_paste_kernel<<<grid1d((long long)R * H * W), 256, 0, (cudaStream_t)stream>>>(probs, M, boxes, R, H,
                                                                                  W, thr, out);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// masks: uint8 [N, H, W].  Vertices in pixel coordinates (x right, y down), as in COCO annotations.
extern "C" int b2f_poly_to_mask(const float* xy, const int* poly_off, const int* obj_poly, int N, int H,
                                int W, unsigned char* masks, void* stream) {
  B2F_REQUIRE(N > 0 && H > 0 && W > 0, "poly_to_mask: bad shape");
  const unsigned gx = (unsigned)(((long long)H * W + 255) / 256);
  poly_to_mask_kernel<<<dim3(gx, (unsigned)N), 256, 0, (cudaStream_t)stream>>>(xy, poly_off, obj_poly, H, W, masks);
  B2F_CHECK_CUDA(cudaGetLastError());
  return 0;
}
