// ORACLE — TEST INFRASTRUCTURE ONLY.  See common.h.
#pragma once
#include <string>

#include "gemm.h"

namespace oracle {

// src/ops/mod.rs:665-705 OpError variants (codes shared with include/rten_gpu.h).
enum OpErrorCode {
  OP_OK = 0,
  OP_INCORRECT_INPUT_TYPE = 1,
  OP_INCORRECT_OUTPUT_TYPE = 2,
  OP_INCOMPATIBLE_INPUT_SHAPES = 3,
  OP_MISSING_INPUTS = 4,
  OP_INVALID_VALUE = 5,
  OP_UNSUPPORTED_VALUE = 6,
  OP_PANIC = 100,  // where the reference would panic (assert!/expect)
};

struct OpError {
  int code = OP_OK;
  const char* msg = "";
  bool ok() const { return code == OP_OK; }
};

// src/ops/mod.rs:127-166 Padding::{Same, Fixed}
struct Padding {
  bool same = false;
  size_t pads[4] = {0, 0, 0, 0};  // top, left, bottom, right
};

// src/ops/pooling.rs:27-89
OpError calc_output_size_and_padding(size_t in_h, size_t in_w, size_t k_h, size_t k_w, size_t stride_h,
                                     size_t stride_w, const Padding& padding, size_t dilation_y, size_t dilation_x,
                                     size_t* out_h, size_t* out_w, size_t fixed[4]);

// src/ops/conv/im2col.rs:75-181 (tables) + 190-260 (masked gather pack_b)
struct VirtualIm2Col : VirtualMatrix {
  const float* image;
  size_t chans, h, w;
  ptrdiff_t stride_c, stride_h, stride_w;
  std::vector<int32_t> row_chan, row_y, row_x;  // RowOffsets
  std::vector<int32_t> col_y, col_x;            // ColOffsets (padded to panel width)
  size_t n_cols;
  int32_t max_y_offset, max_x_offset;

  VirtualIm2Col(const float* image, size_t chans, size_t h, size_t w, ptrdiff_t sc, ptrdiff_t sh, ptrdiff_t sw,
                size_t k_h, size_t k_w, const size_t padding[4], size_t stride_h, size_t stride_w, size_t dil_y,
                size_t dil_x, size_t panel_width);
  size_t rows() const override { return row_chan.size(); }
  size_t cols() const override { return n_cols; }
  void pack_b(float* out, size_t panel_width, size_t r0, size_t r1, size_t c0, size_t c1) const override;
};

// src/ops/conv/depthwise.rs:24-38
void min_max_out_x_coords(size_t k_x, size_t in_w, size_t pad_left, size_t stride, size_t dilation, size_t out_w,
                          size_t* min_out_x, size_t* max_out_x);

// src/ops/conv.rs:86-280.  input NCHW contiguous [n,c,h,w], kernel [out_c, k_in_c, k_h, k_w],
// bias [out_c] or null.  `out` holds n*out_c*out_h*out_w floats; pass null to only validate and get out_shape.
OpError conv2d(const Kernel* gemm_kernel, const float* input, const size_t in_shape[4], const float* kernel,
               const size_t k_shape[4], const float* bias, size_t bias_len, const Padding& padding, size_t groups,
               const size_t strides[2], const size_t dilations[2], float* out, size_t out_shape[4]);

}  // namespace oracle
