// ORACLE — TEST INFRASTRUCTURE ONLY.  See common.h.
#pragma once
#include "common.h"

namespace oracle {

// src/gemm.rs:240-259 KernelType (x86-64 subset + Base).
enum KernelType { KERNEL_BASE = 0, KERNEL_FMA = 1, KERNEL_AVX512 = 2, KERNEL_AUTO = -1 };

struct Kernel {
  KernelType type;
  size_t mr, nr;
  const char* name;
  // src/gemm/kernels.rs:374-383 Kernel::kernel
  void (*kernel)(float* tile, size_t tile_row_stride, const float* a, const float* b, size_t depth, float alpha,
                 float beta);
  // src/gemm/kernels.rs:401-406 Kernel::gemv_kernel
  void (*gemv_kernel)(float* out, size_t out_len, const float* a, size_t a_len, Matrix b, float alpha, float beta);
};

// src/gemm.rs:263-282 GemmExecutor::new (kernel preference order) / with_kernel.
const Kernel* select_kernel(KernelType hint);

// src/gemm/packing.rs:20-90 / 106-186
void pack_a_block(size_t MR, float* out, Matrix a, size_t row_start, size_t row_end, size_t col_start, size_t col_end);
void pack_b_block(size_t NR, float* out, Matrix b, size_t row_start, size_t row_end, size_t col_start, size_t col_end);

// src/gemm.rs:546-571
size_t depth_block_size(size_t a_cols);
size_t col_block_size(size_t b_cols, size_t nr);
size_t row_block_size(size_t a_rows, size_t mr);

// src/gemm.rs:133-161 VirtualMatrix
struct VirtualMatrix {
  virtual ~VirtualMatrix() {}
  virtual size_t rows() const = 0;
  virtual size_t cols() const = 0;
  virtual void pack_b(float* out, size_t panel_width, size_t row_start, size_t row_end, size_t col_start,
                      size_t col_end) const = 0;
};

// src/gemm.rs:26-92 PackedAMatrix / PackedBMatrix
struct PackedAMatrix {
  std::vector<float> data;
  size_t panel_len, row_blocks, rows, cols;
  const float* block(size_t row_block, size_t depth_block) const {
    return data.data() + (depth_block * row_blocks + row_block) * panel_len;
  }
};
struct PackedBMatrix {
  std::vector<float> data;
  size_t panel_len, depth_blocks, rows, cols;
  const float* block(size_t col_block, size_t depth_block) const {
    return data.data() + (col_block * depth_blocks + depth_block) * panel_len;
  }
};

// src/gemm.rs:96-193 GemmInputA / GemmInputB
struct GemmInputA {
  const Matrix* unpacked = nullptr;
  const PackedAMatrix* packed = nullptr;
  size_t rows() const { return unpacked ? unpacked->rows : packed->rows; }
  size_t cols() const { return unpacked ? unpacked->cols : packed->cols; }
};
struct GemmInputB {
  const Matrix* unpacked = nullptr;
  const PackedBMatrix* packed = nullptr;
  const VirtualMatrix* virt = nullptr;
  size_t rows() const { return unpacked ? unpacked->rows : packed ? packed->rows : virt->rows(); }
  size_t cols() const { return unpacked ? unpacked->cols : packed ? packed->cols : virt->cols(); }
};

PackedAMatrix prepack_a(const Kernel* k, Matrix a);  // src/gemm.rs:337-379
PackedBMatrix prepack_b(const Kernel* k, Matrix b);  // src/gemm.rs:396-438

// src/gemm.rs:733-930 gemm_impl.  Returns 0 on success or a negative code where
// the reference panics (-1 K mismatch, -2 bias length, -3 output too short).
int gemm_impl(const Kernel* kernel, float* out_data, size_t out_len, size_t out_row_stride, GemmInputA a, GemmInputB b,
              float alpha, float beta, const float* bias, size_t bias_len);

}  // namespace oracle
