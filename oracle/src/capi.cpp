// ORACLE — TEST INFRASTRUCTURE ONLY.  See common.h.
// Flat C entry points so tests / bench.py can drive the oracle through ctypes.
#include <cstring>

#include "ops.h"

using namespace oracle;

static thread_local const char* g_last_msg = "";
static int fail(const OpError& e) {
  g_last_msg = e.msg;
  return e.code;
}

extern "C" {

const char* orc_last_error() { return g_last_msg; }
void orc_set_num_threads(int n) { ThreadPool::instance().set_num_threads(n); }
int orc_num_threads() { return ThreadPool::instance().num_threads(); }

// rten-tensor/src/rng.rs:11-32 XorShiftRng.  `state` is updated in place.
void orc_rng_fill(uint64_t* state, float* out, size_t n) {
  uint64_t s = *state;
  const float scale = 1.0f / (float)(1ull << 40);
  for (size_t i = 0; i < n; i++) {
    s ^= s << 13;
    s ^= s >> 7;
    s ^= s << 17;
    out[i] = (float)(s >> 24) * scale;
  }
  *state = s;
}

int orc_kernel_info(int kernel_type, size_t* mr, size_t* nr) {
  const Kernel* k = select_kernel((KernelType)kernel_type);
  if (!k) return -1;
  *mr = k->mr;
  *nr = k->nr;
  return (int)k->type;
}

void orc_pack_a_block(size_t mr, float* out, const float* a, size_t rows, size_t cols, ptrdiff_t rs, ptrdiff_t cs,
                      size_t r0, size_t r1, size_t c0, size_t c1) {
  pack_a_block(mr, out, Matrix{a, rows, cols, rs, cs}, r0, r1, c0, c1);
}
void orc_pack_b_block(size_t nr, float* out, const float* b, size_t rows, size_t cols, ptrdiff_t rs, ptrdiff_t cs,
                      size_t r0, size_t r1, size_t c0, size_t c1) {
  pack_b_block(nr, out, Matrix{b, rows, cols, rs, cs}, r0, r1, c0, c1);
}
void orc_block_sizes(size_t m, size_t n, size_t k, size_t mr, size_t nr, size_t* mc, size_t* nc, size_t* kc) {
  *mc = row_block_size(m, mr);
  *nc = col_block_size(n, nr);
  *kc = depth_block_size(k);
}

// GemmExecutor::{gemm, gemm_bias, gemm_uninit_bias} (src/gemm.rs:448-542).
// prepack: bit0 = prepack A, bit1 = prepack B.  Returns 0, or <0 where the reference panics.
int orc_gemm(int kernel_type, float* out, size_t out_len, size_t out_row_stride, const float* a, size_t m, size_t k,
             ptrdiff_t a_rs, ptrdiff_t a_cs, const float* b, size_t k2, size_t n, ptrdiff_t b_rs, ptrdiff_t b_cs,
             float alpha, float beta, const float* bias, size_t bias_len, int prepack) {
  const Kernel* kern = select_kernel((KernelType)kernel_type);
  if (!kern) return -10;
  Matrix am{a, m, k, a_rs, a_cs}, bm{b, k2, n, b_rs, b_cs};
  PackedAMatrix pa;
  PackedBMatrix pb;
  GemmInputA ia;
  GemmInputB ib;
  if (prepack & 1) {
    pa = prepack_a(kern, am);
    ia.packed = &pa;
  } else {
    ia.unpacked = &am;
  }
  if (prepack & 2) {
    pb = prepack_b(kern, bm);
    ib.packed = &pb;
  } else {
    ib.unpacked = &bm;
  }
  return gemm_impl(kern, out, out_len, out_row_stride, ia, ib, alpha, beta, bias, bias_len);
}

static Padding make_padding(int same, const size_t pads[4]) {
  Padding p;
  p.same = same != 0;
  for (int i = 0; i < 4; i++) p.pads[i] = pads[i];
  return p;
}

int orc_conv_out_size(size_t in_h, size_t in_w, size_t k_h, size_t k_w, size_t stride_h, size_t stride_w, int same,
                      const size_t pads[4], size_t dil_y, size_t dil_x, size_t* out_h, size_t* out_w,
                      size_t fixed[4]) {
  return fail(calc_output_size_and_padding(in_h, in_w, k_h, k_w, stride_h, stride_w, make_padding(same, pads), dil_y,
                                           dil_x, out_h, out_w, fixed));
}

// conv() for 4-D NCHW input (src/ops/conv.rs:141-280).  out may be null (shape / validation only).
int orc_conv2d(int kernel_type, const float* input, const size_t in_shape[4], const float* kernel,
               const size_t k_shape[4], const float* bias, size_t bias_len, int same, const size_t pads[4],
               size_t groups, const size_t strides[2], const size_t dilations[2], float* out, size_t out_shape[4]) {
  const Kernel* kern = select_kernel((KernelType)kernel_type);
  if (!kern) return -10;
  return fail(conv2d(kern, input, in_shape, kernel, k_shape, bias, bias_len, make_padding(same, pads), groups, strides,
                     dilations, out, out_shape));
}

// VirtualIm2Col tables + pack_b (src/ops/conv/im2col.rs).  Table outputs may be null.
// out receives rows x round_up(cols, panel_width) packed floats when non-null.
int orc_im2col(const float* image, size_t chans, size_t h, size_t w, size_t k_h, size_t k_w, const size_t pads[4],
               size_t stride_h, size_t stride_w, size_t dil_y, size_t dil_x, size_t panel_width, int32_t* row_chan,
               int32_t* row_y, int32_t* row_x, int32_t* col_y, int32_t* col_x, size_t* n_rows, size_t* n_cols,
               float* out) {
  VirtualIm2Col v(image, chans, h, w, (ptrdiff_t)(h * w), (ptrdiff_t)w, 1, k_h, k_w, pads, stride_h, stride_w, dil_y,
                  dil_x, panel_width);
  *n_rows = v.rows();
  *n_cols = v.cols();
  if (row_chan) std::memcpy(row_chan, v.row_chan.data(), v.row_chan.size() * 4);
  if (row_y) std::memcpy(row_y, v.row_y.data(), v.row_y.size() * 4);
  if (row_x) std::memcpy(row_x, v.row_x.data(), v.row_x.size() * 4);
  if (col_y) std::memcpy(col_y, v.col_y.data(), v.col_y.size() * 4);
  if (col_x) std::memcpy(col_x, v.col_x.data(), v.col_x.size() * 4);
  if (out) v.pack_b(out, panel_width, 0, v.rows(), 0, v.cols());
  return 0;
}

void orc_unary(int op, const float* in, float* out, size_t n, float lo, float hi) {
  unary((UnaryOp)op, in, out, n, lo, hi);
}
float orc_vm_scalar(int op, float x) {
  switch (op) {
    case 0: return vm_exp(x);
    case 1: return vm_sigmoid(x);
    case 2: return vm_erf(x);
    default: return vm_gelu(x);
  }
}
int orc_binary(int op, const float* a, const size_t* a_shape, size_t a_nd, const float* b, const size_t* b_shape,
               size_t b_nd, float* out, size_t* out_shape, size_t* out_nd) {
  return binary((BinaryOp)op, a, a_shape, a_nd, b, b_shape, b_nd, out, out_shape, out_nd);
}
int orc_pool2d(int is_max, const float* input, const size_t in_shape[4], const size_t ksize[2],
               const size_t strides[2], int same, const size_t pads[4], int count_include_pad, float* out,
               size_t out_shape[4]) {
  return fail(pool2d(is_max != 0, input, in_shape, ksize, strides, make_padding(same, pads), count_include_pad != 0,
                     out, out_shape));
}
void orc_global_average_pool(const float* input, size_t batch, size_t chans, size_t h, size_t w, float* out) {
  global_average_pool(input, batch, chans, h, w, out);
}

static TensorView make_view(const float* data, size_t nd, const size_t* shape, const ptrdiff_t* strides) {
  TensorView t{};
  t.data = data;
  t.ndim = nd;
  for (size_t i = 0; i < nd && i < 8; i++) {
    t.shape[i] = shape[i];
    t.strides[i] = strides[i];
  }
  return t;
}

// matmul() (src/ops/matmul.rs:119-239).  out may be null to query the shape only.
int orc_matmul(int kernel_type, const float* a, size_t a_nd, const size_t* a_shape, const ptrdiff_t* a_strides,
               const float* b, size_t b_nd, const size_t* b_shape, const ptrdiff_t* b_strides, float* out,
               size_t out_cap, size_t* out_shape, size_t* out_nd) {
  const Kernel* kern = select_kernel((KernelType)kernel_type);
  if (!kern) return -10;
  std::vector<float> o;
  std::vector<size_t> os;
  OpError e = matmul(kern, make_view(a, a_nd, a_shape, a_strides), make_view(b, b_nd, b_shape, b_strides), &o, &os);
  if (!e.ok()) return fail(e);
  *out_nd = os.size();
  for (size_t i = 0; i < os.size(); i++) out_shape[i] = os[i];
  if (out) {
    if (out_cap < o.size()) return -11;
    std::memcpy(out, o.data(), o.size() * sizeof(float));
  }
  return 0;
}

// gemm_op() (src/ops/matmul.rs:27-81).
int orc_gemm_op(int kernel_type, const float* a, const size_t* a_shape, const ptrdiff_t* a_strides, const float* b,
                const size_t* b_shape, const ptrdiff_t* b_strides, const float* c, size_t c_nd, const size_t* c_shape,
                const ptrdiff_t* c_strides, float alpha, float beta, int transpose_a, int transpose_b, float* out,
                size_t out_cap, size_t out_shape[2]) {
  const Kernel* kern = select_kernel((KernelType)kernel_type);
  if (!kern) return -10;
  TensorView cv{};
  if (c) cv = make_view(c, c_nd, c_shape, c_strides);
  std::vector<float> o;
  OpError e = gemm_op(kern, make_view(a, 2, a_shape, a_strides), make_view(b, 2, b_shape, b_strides), c ? &cv : nullptr,
                      alpha, beta, transpose_a != 0, transpose_b != 0, &o, out_shape);
  if (!e.ok()) return fail(e);
  if (out) {
    if (out_cap < o.size()) return -11;
    std::memcpy(out, o.data(), o.size() * sizeof(float));
  }
  return 0;
}

}  // extern "C"
