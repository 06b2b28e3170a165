// b200flow — host-side helpers shared by every launcher in the CUDA library:
// error reporting through the C ABI, and TMA tensor-map construction through
// the driver entry point (no link-time dependency on libcuda, so the library
// also loads on GPU-less build machines).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

namespace b2f {

void set_error(const std::string& msg);  // defined in common/host.cpp

#define B2F_CHECK_CUDA(expr)                                                              \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      b2f::set_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e) + " at " + \
                     __FILE__ + ":" + std::to_string(__LINE__));                          \
      return -1;                                                                          \
    }                                                                                     \
  } while (0)

#define B2F_REQUIRE(cond, msg)                                                      \
  do {                                                                              \
    if (!(cond)) {                                                                  \
      b2f::set_error(std::string("b200flow: ") + (msg) + " [" #cond "] at " +       \
                     __FILE__ + ":" + std::to_string(__LINE__));                    \
      return -2;                                                                    \
    }                                                                               \
  } while (0)

enum class TmaDtype { BF16, F16, U8, F32 };

// Encode a 2-D row-major tensor map: `rows` x `cols` elements, `ld` elements
// between rows; box = box_rows x box_cols; swizzle: 0 none / 128 = 128 B.
// Returns 0 on success. Out-of-bounds box elements are zero-filled.
int make_tma_2d(CUtensorMap* out, const void* base, TmaDtype dt, uint64_t rows, uint64_t cols,
                uint64_t ld, uint32_t box_rows, uint32_t box_cols, int swizzle_bytes);

// 3-D variant: [batch, rows, cols] with element strides (batch_stride, ld, 1).
int make_tma_3d(CUtensorMap* out, const void* base, TmaDtype dt, uint64_t batch, uint64_t rows,
                uint64_t cols, uint64_t batch_stride, uint64_t ld, uint32_t box_rows,
                uint32_t box_cols, int swizzle_bytes);

// 4-D variant for attention operands: element strides (s3, s2, s1, 1) over dims
// [d3, d2, rows, cols]; box = box_rows x box_cols in the two innermost dims.
int make_tma_4d(CUtensorMap* out, const void* base, TmaDtype dt, uint64_t d3, uint64_t d2,
                uint64_t rows, uint64_t cols, uint64_t s3, uint64_t s2, uint64_t s1,
                uint32_t box_rows, uint32_t box_cols, int swizzle_bytes);

int sm_count();  // SMs of the current device (cached per device)

}  // namespace b2f
