// ORACLE — TEST INFRASTRUCTURE ONLY.  See common.h.
// Restates src/ops/conv.rs, src/ops/conv/im2col.rs, src/ops/conv/depthwise.rs and
// src/ops/pooling.rs:27-89 of the reference.
#include "conv.h"

#include <algorithm>
#include <cstring>

namespace oracle {

static inline size_t sat_sub(size_t a, size_t b) { return a > b ? a - b : 0; }

// src/ops/pooling.rs:27-89
OpError calc_output_size_and_padding(size_t in_h, size_t in_w, size_t k_h, size_t k_w, size_t stride_h,
                                     size_t stride_w, const Padding& padding, size_t dilation_y, size_t dilation_x,
                                     size_t* out_h, size_t* out_w, size_t fixed[4]) {
  if (dilation_y == 0 || dilation_x == 0) return {OP_INVALID_VALUE, "Dilations must be > 0"};
  if (stride_h == 0 || stride_w == 0) return {OP_INVALID_VALUE, "Strides must be > 0"};
  if (padding.same) {
    size_t oh = div_ceil(in_h, stride_h), ow = div_ceil(in_w, stride_w);
    size_t pad_total_h = sat_sub((oh - 1) * stride_h + (k_h - 1) * dilation_y + 1, in_h);
    size_t pad_total_w = sat_sub((ow - 1) * stride_w + (k_w - 1) * dilation_x + 1, in_w);
    fixed[0] = pad_total_h / 2;
    fixed[1] = pad_total_w / 2;
    fixed[2] = div_ceil(pad_total_h, 2);  // SAME_UPPER: odd unit goes to the end
    fixed[3] = div_ceil(pad_total_w, 2);
    *out_h = oh;
    *out_w = ow;
  } else {
    size_t padded_in_h = in_h + padding.pads[0] + padding.pads[2];
    size_t padded_in_w = in_w + padding.pads[1] + padding.pads[3];
    size_t dilated_k_h = k_h + (k_h - 1) * (dilation_y - 1);
    size_t dilated_k_w = k_w + (k_w - 1) * (dilation_x - 1);
    if (padded_in_h < dilated_k_h || padded_in_w < dilated_k_w)
      return {OP_INVALID_VALUE, "Input too small for kernel size"};
    *out_h = (padded_in_h - dilation_y * (k_h - 1) - 1) / stride_h + 1;
    *out_w = (padded_in_w - dilation_x * (k_w - 1) - 1) / stride_w + 1;
    for (int i = 0; i < 4; i++) fixed[i] = padding.pads[i];
  }
  return {};
}

// src/ops/conv/im2col.rs:75-181
VirtualIm2Col::VirtualIm2Col(const float* image_, size_t chans_, size_t h_, size_t w_, ptrdiff_t sc, ptrdiff_t sh,
                             ptrdiff_t sw, size_t k_h, size_t k_w, const size_t padding[4], size_t stride_h_,
                             size_t stride_w_, size_t dil_y, size_t dil_x, size_t panel_width)
    : image(image_), chans(chans_), h(h_), w(w_), stride_c(sc), stride_h(sh), stride_w(sw) {
  Padding p;
  for (int i = 0; i < 4; i++) p.pads[i] = padding[i];
  size_t y_patches = 0, x_patches = 0, fixed[4];
  calc_output_size_and_padding(h, w, k_h, k_w, stride_h_, stride_w_, p, dil_y, dil_x, &y_patches, &x_patches, fixed);
  const int32_t isc = (int32_t)sc, ish = (int32_t)sh, isw = (int32_t)sw;
  const int32_t pad_top = (int32_t)padding[0], pad_left = (int32_t)padding[1];

  for (size_t chan = 0; chan < chans; chan++) {
    for (size_t i = 0; i < k_h * k_w; i++) row_chan.push_back((int32_t)chan * isc);
    for (size_t k_y = 0; k_y < k_h; k_y++) {
      for (size_t k_x = 0; k_x < k_w; k_x++) {
        row_y.push_back(ish * (int32_t)k_y * (int32_t)dil_y);
        row_x.push_back(isw * (int32_t)k_x * (int32_t)dil_x);
      }
    }
  }

  n_cols = x_patches * y_patches;
  size_t n_cols_padded = round_up(n_cols, panel_width);
  for (size_t patch_y = 0; patch_y < y_patches; patch_y++) {
    int32_t img_y = (int32_t)patch_y * (int32_t)stride_h_ - pad_top;
    for (size_t patch_x = 0; patch_x < x_patches; patch_x++) {
      int32_t img_x = (int32_t)patch_x * (int32_t)stride_w_ - pad_left;
      col_y.push_back(img_y * ish);
      col_x.push_back(img_x * isw);
    }
  }
  for (size_t col = n_cols; col < n_cols_padded; col++) {
    int32_t patch_y = (int32_t)col / (int32_t)x_patches, patch_x = (int32_t)col % (int32_t)x_patches;
    col_y.push_back((patch_y * (int32_t)stride_h_ - pad_top) * ish);
    col_x.push_back((patch_x * (int32_t)stride_w_ - pad_left) * isw);
  }
  max_y_offset = (int32_t)((h - 1) * (size_t)sh);
  max_x_offset = (int32_t)((w - 1) * (size_t)sw);
}

// src/ops/conv/im2col.rs:190-260.  `gather_mask` semantics: masked-off lanes are 0
// (rten-simd/src/arch/scalar.rs:190-196).  Columns past n_cols inside the last
// panel are gathered like the reference does (they never reach the output), but
// only when in range of the image so the oracle itself stays memory-safe.
void VirtualIm2Col::pack_b(float* out, size_t panel_width, size_t r0, size_t r1, size_t c0, size_t c1) const {
  size_t c1p = round_up(c1, panel_width);
  size_t out_offset = 0;
  for (size_t start_col = c0; start_col < c1p; start_col += panel_width) {
    for (size_t r = r0; r < r1; r++) {
      const int32_t rc = row_chan[r], ry = row_y[r], rx = row_x[r];
      for (size_t i = 0; i < panel_width; i++) {
        int32_t y_off = col_y[start_col + i] + ry;
        int32_t x_off = col_x[start_col + i] + rx;
        bool in_range = y_off >= 0 && y_off <= max_y_offset && x_off >= 0 && x_off <= max_x_offset;
        out[out_offset++] = in_range ? image[rc + y_off + x_off] : 0.f;
      }
    }
  }
}

// src/ops/conv/depthwise.rs:24-38
void min_max_out_x_coords(size_t k_x, size_t in_w, size_t pad_left, size_t stride, size_t dilation, size_t out_w,
                          size_t* min_out_x, size_t* max_out_x) {
  *min_out_x = sat_sub(pad_left, k_x * dilation);
  *max_out_x = std::min(div_ceil(sat_sub(in_w + pad_left, k_x * dilation), stride), out_w);
}


// src/ops/conv/depthwise.rs:49-120.  Bias-initialised rows, then for k_y, k_x an AXPY
// over the valid column range: `dest[i] += src[i*stride_w] * scale` (mul then add,
// not fused — compile with -ffp-contract=off).
static void conv_2d_depthwise_block(float* out /* [chunk, out_h, out_w] */, size_t chan0, size_t chan1,
                                    const float* input /* [C, in_h, in_w] */, size_t in_h, size_t in_w,
                                    const float* kernel /* [C, 1, k_h, k_w] */, size_t k_h, size_t k_w,
                                    const float* bias, const size_t padding[4], size_t stride_h, size_t stride_w,
                                    size_t dilation_y, size_t out_h, size_t out_w,
                                    const std::vector<size_t>& in_lo, const std::vector<size_t>& in_hi,
                                    const std::vector<size_t>& out_lo, const std::vector<size_t>& out_hi) {
  const size_t pad_top = padding[0];
  for (size_t c = chan0; c < chan1; c++) {
    const float* kview = kernel + c * k_h * k_w;
    float* out_chan = out + (c - chan0) * out_h * out_w;
    const float* in_chan = input + c * in_h * in_w;
    const float init_value = bias ? bias[c] : 0.f;
    for (size_t out_y = 0; out_y < out_h; out_y++) {
      float* out_row = out_chan + out_y * out_w;
      for (size_t x = 0; x < out_w; x++) out_row[x] = init_value;
      for (size_t k_y = 0; k_y < k_h; k_y++) {
        size_t in_y = out_y * stride_h + k_y * dilation_y;
        if (in_y < pad_top || in_y >= in_h + pad_top) continue;
        const float* in_row = in_chan + (in_y - pad_top) * in_w;
        for (size_t k_x = 0; k_x < k_w; k_x++) {
          float* dest = out_row + out_lo[k_x];
          const float* src = in_row + in_lo[k_x];
          const float scale = kview[k_y * k_w + k_x];
          size_t src_len = in_hi[k_x] > in_lo[k_x] ? in_hi[k_x] - in_lo[k_x] : 0;
          size_t src_els = div_ceil(src_len, stride_w);
          for (size_t i = 0; i < src_els; i++) dest[i] += src[i * stride_w] * scale;
        }
      }
    }
  }
}

// src/ops/conv/depthwise.rs:127-203
static void conv_2d_depthwise(const float* input, size_t batch, size_t in_c, size_t in_h, size_t in_w,
                              const float* kernel, size_t out_c, size_t k_h, size_t k_w, const float* bias,
                              const size_t padding[4], const size_t strides[2], const size_t dilations[2],
                              size_t out_h, size_t out_w, float* output) {
  const size_t pad_left = padding[1], stride_w = strides[1], dilation_x = dilations[1];
  std::vector<size_t> in_lo(k_w), in_hi(k_w), out_lo(k_w), out_hi(k_w);
  for (size_t k_x = 0; k_x < k_w; k_x++) {
    size_t mn, mx;
    min_max_out_x_coords(k_x, in_w, pad_left, stride_w, dilation_x, out_w, &mn, &mx);
    out_lo[k_x] = mn;
    out_hi[k_x] = mx;
    size_t min_in_x = mn * stride_w + k_x * dilation_x - pad_left;
    size_t max_in_x = (mn >= mx) ? min_in_x : (mx - 1) * stride_w + k_x * dilation_x - pad_left + 1;
    in_lo[k_x] = min_in_x;
    in_hi[k_x] = max_in_x;
  }
  const size_t target_chunk_size = 32 * 1024;
  size_t channel_chunk_size = std::max<size_t>(1, std::min(target_chunk_size / (out_h * out_w), out_c));
  size_t n_chunks = div_ceil(out_c, channel_chunk_size);
  for (size_t n = 0; n < batch; n++) {
    float* out_n = output + n * out_c * out_h * out_w;
    const float* in_n = input + n * in_c * in_h * in_w;
    ThreadPool::instance().parallel_for(n_chunks, [&](size_t chunk) {
      size_t c0 = chunk * channel_chunk_size, c1 = std::min(c0 + channel_chunk_size, out_c);
      conv_2d_depthwise_block(out_n + c0 * out_h * out_w, c0, c1, in_n, in_h, in_w, kernel, k_h, k_w, bias, padding,
                              strides[0], stride_w, dilations[0], out_h, out_w, in_lo, in_hi, out_lo, out_hi);
    });
  }
}

// src/ops/conv.rs:24-68.  Images serial; per image out[n] [O,HW] = W[O,C] x in[n] [C,HW] + bias.
static int conv_2d_pointwise(const Kernel* gk, const float* input, size_t batch, size_t in_c_actual, size_t hw,
                             const float* kernel, size_t out_c, size_t in_c, const float* bias, float* output) {
  Matrix kernel_mat{kernel, out_c, in_c, (ptrdiff_t)in_c, 1};
  for (size_t n = 0; n < batch; n++) {
    Matrix in_mat{input + n * in_c_actual * hw, in_c_actual, hw, (ptrdiff_t)hw, 1};
    GemmInputA a;
    a.unpacked = &kernel_mat;
    GemmInputB b;
    b.unpacked = &in_mat;
    int rc = gemm_impl(gk, output + n * out_c * hw, out_c * hw, hw, a, b, 1.f, 0.f, bias, out_c);
    if (rc != 0) return rc;
  }
  return 0;
}


// src/ops/conv.rs:141-280 (4-D path; the 1-D lift at :98-139 is done by the caller
// via reshape, see capi.cpp).  Check ORDER follows the reference: output size,
// pointwise fast path, channel/group checks, depthwise, grouped im2col.
OpError conv2d(const Kernel* gk, const float* input, const size_t in_shape[4], const float* kernel,
               const size_t k_shape[4], const float* bias, size_t bias_len, const Padding& padding, size_t groups,
               const size_t strides[2], const size_t dilations[2], float* out, size_t out_shape[4]) {
  const size_t batch = in_shape[0], in_c = in_shape[1], in_h = in_shape[2], in_w = in_shape[3];
  const size_t out_c = k_shape[0], k_in_c = k_shape[1], k_h = k_shape[2], k_w = k_shape[3];
  size_t out_h, out_w, fixed[4];
  OpError err = calc_output_size_and_padding(in_h, in_w, k_h, k_w, strides[0], strides[1], padding, dilations[0],
                                             dilations[1], &out_h, &out_w, fixed);
  if (!err.ok()) return err;
  const bool has_padding = fixed[0] > 0 || fixed[1] > 0 || fixed[2] > 0 || fixed[3] > 0;
  out_shape[0] = batch;
  out_shape[1] = out_c;
  out_shape[2] = out_h;
  out_shape[3] = out_w;

  if (k_h == 1 && k_w == 1 && !has_padding && groups == 1 && strides[0] == 1 && strides[1] == 1 &&
      dilations[0] == 1 && dilations[1] == 1) {
    // The reference reshapes in[n] to [in_c(kernel), HW] and panics inside gemm on mismatch.
    if (in_c != k_in_c) return {OP_PANIC, "Columns of matrix `a` must match rows of matrix `b`"};
    if (bias && bias_len != out_c) return {OP_PANIC, "Bias vector length must match rows of matrix `a`"};
    if (batch * out_c * in_h * in_w == 0 || !out) return {};
    int rc = conv_2d_pointwise(gk, input, batch, in_c, in_h * in_w, kernel, out_c, k_in_c, bias, out);
    if (rc != 0) return {OP_PANIC, "gemm panic"};
    return {};
  }

  if (groups == 0) return {OP_PANIC, "attempt to divide by zero"};
  const size_t out_channels_per_group = out_c / groups, in_channels_per_group = in_c / groups;
  if (in_channels_per_group != k_in_c)
    return {OP_INCOMPATIBLE_INPUT_SHAPES, "Input channels (per group) does not match kernel input channels"};
  if (in_c % groups != 0 || out_c % groups != 0)
    return {OP_INCOMPATIBLE_INPUT_SHAPES, "Input channels and output channels must be divisible by group count"};

  if (batch * out_c * out_h * out_w == 0 || !out) return {};

  if (in_c == out_c && groups == in_c) {
    conv_2d_depthwise(input, batch, in_c, in_h, in_w, kernel, out_c, k_h, k_w, bias, fixed, strides, dilations, out_h,
                      out_w, out);
    return {};
  }

  const size_t n_patches = out_h * out_w;
  const size_t kdim = in_channels_per_group * k_h * k_w;
  for (size_t group = 0; group < groups; group++) {
    const size_t in_chan_start = group * in_channels_per_group, out_chan_start = group * out_channels_per_group;
    Matrix kernel_mat{kernel + out_chan_start * kdim, out_channels_per_group, kdim, (ptrdiff_t)kdim, 1};
    PackedAMatrix prepacked;
    const bool use_prepacked = batch > 1;
    if (use_prepacked) prepacked = prepack_a(gk, kernel_mat);
    ThreadPool::instance().parallel_for(batch, [&](size_t n) {
      const float* in_item = input + (n * in_c + in_chan_start) * in_h * in_w;
      float* out_mat = out + (n * out_c + out_chan_start) * n_patches;
      VirtualIm2Col im2col(in_item, in_channels_per_group, in_h, in_w, (ptrdiff_t)(in_h * in_w), (ptrdiff_t)in_w, 1,
                           k_h, k_w, fixed, strides[0], strides[1], dilations[0], dilations[1], gk->nr);
      GemmInputA a;
      if (use_prepacked) a.packed = &prepacked;
      else a.unpacked = &kernel_mat;
      GemmInputB b;
      b.virt = &im2col;
      gemm_impl(gk, out_mat, out_channels_per_group * n_patches, n_patches, a, b, 1.f, 0.f,
                bias ? bias + out_chan_start : nullptr, out_channels_per_group);
    });
  }
  return {};
}

}  // namespace oracle
