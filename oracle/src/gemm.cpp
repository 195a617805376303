// ORACLE — TEST INFRASTRUCTURE ONLY.  See common.h.
// Restates src/gemm.rs (GemmExecutor, gemm_impl, gemm_block, gemv) and
// src/gemm/packing.rs of the reference.
#include "gemm.h"

#include <algorithm>
#include <cstring>

namespace oracle {

extern const Kernel kBaseKernel;
extern const Kernel kFmaKernel;
extern const Kernel kAvx512Kernel;

// src/gemm.rs:263-282, 298-318.  KERNEL_AUTO follows the reference's DEFAULT build
// (no `avx512` feature): Fma if AVX2+FMA are present, else Base.
const Kernel* select_kernel(KernelType hint) {
  bool has_fma = __builtin_cpu_supports("avx2") && __builtin_cpu_supports("fma");
  bool has_512 = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512vl");
  switch (hint) {
    case KERNEL_BASE: return &kBaseKernel;
    case KERNEL_FMA: return has_fma ? &kFmaKernel : nullptr;
    case KERNEL_AVX512: return has_512 ? &kAvx512Kernel : nullptr;
    default: return has_fma ? &kFmaKernel : &kBaseKernel;
  }
}

// src/gemm/packing.rs:20-90.  Row panels of MR rows, column-major inside a panel,
// zero padded past the last row.
void pack_a_block(size_t MR, float* out, Matrix a, size_t r0, size_t r1, size_t c0, size_t c1) {
  size_t a_rows = r1 - r0, a_cols = c1 - c0;
  size_t n_panels = round_up(a_rows, MR) / MR;
  for (size_t panel = 0; panel < n_panels; panel++) {
    size_t panel_offset = panel * a_cols * MR;
    size_t panel_start_row = panel * MR;
    for (size_t col = 0; col < a_cols; col++) {
      for (size_t row = 0; row < MR; row++) {
        size_t a_row = r0 + panel_start_row + row;
        out[panel_offset + col * MR + row] = a_row < r1 ? a.at(a_row, c0 + col) : 0.f;
      }
    }
  }
}

// src/gemm/packing.rs:106-186.  Column panels of NR columns, row-major inside a
// panel, zero padded past the last column.
void pack_b_block(size_t NR, float* out, Matrix b, size_t r0, size_t r1, size_t c0, size_t c1) {
  size_t b_cols = c1 - c0, b_rows = r1 - r0;
  size_t n_panels = round_up(b_cols, NR) / NR;
  for (size_t panel = 0; panel < n_panels; panel++) {
    size_t panel_offset = panel * b_rows * NR;
    size_t panel_start_col = panel * NR;
    if (b_cols - panel_start_col >= NR && b.col_stride == 1) {
      for (size_t row = 0; row < b_rows; row++) {
        const float* src = b.data + (ptrdiff_t)(r0 + row) * b.row_stride + (c0 + panel_start_col);
        std::memcpy(out + panel_offset + row * NR, src, NR * sizeof(float));
      }
      continue;
    }
    for (size_t row = 0; row < b_rows; row++) {
      for (size_t col = 0; col < NR; col++) {
        size_t out_col = panel_start_col + col;
        out[panel_offset + row * NR + col] = out_col < b_cols ? b.at(r0 + row, c0 + out_col) : 0.f;
      }
    }
  }
}

// src/gemm.rs:546-571
size_t depth_block_size(size_t a_cols) { return std::min<size_t>(256, a_cols); }
size_t col_block_size(size_t b_cols, size_t nr) {
  size_t parallelism = (size_t)ThreadPool::instance().num_threads();
  size_t lower_bound = std::min<size_t>(128, b_cols);
  size_t unrounded = std::min<size_t>(std::max(b_cols / parallelism, lower_bound), 1024);
  return round_up(unrounded, nr);
}
size_t row_block_size(size_t a_rows, size_t mr) { return round_up(std::min<size_t>(64, a_rows), mr); }

// src/gemm.rs:337-379
PackedAMatrix prepack_a(const Kernel* k, Matrix a) {
  size_t kc = depth_block_size(a.cols), mr = k->mr, mc = row_block_size(a.rows, mr);
  PackedAMatrix p;
  p.panel_len = kc * mc;
  p.row_blocks = div_ceil(a.rows, mc);
  p.rows = a.rows;
  p.cols = a.cols;
  size_t depth_blocks = div_ceil(a.cols, kc);
  p.data.assign(depth_blocks * p.row_blocks * p.panel_len, 0.f);
  size_t idx = 0;
  for (size_t d0 = 0; d0 < a.cols; d0 += kc) {
    size_t d1 = std::min(d0 + kc, a.cols);
    for (size_t r0 = 0; r0 < a.rows; r0 += mc) {
      size_t r1 = std::min(r0 + mc, a.rows);
      pack_a_block(mr, p.data.data() + idx * p.panel_len, a, r0, r1, d0, d1);
      idx++;
    }
  }
  return p;
}

// src/gemm.rs:396-438
PackedBMatrix prepack_b(const Kernel* k, Matrix b) {
  size_t nr = k->nr, nc = col_block_size(b.cols, nr), kc = depth_block_size(b.rows);
  PackedBMatrix p;
  p.panel_len = nc * kc;
  p.depth_blocks = div_ceil(b.rows, kc);
  p.rows = b.rows;
  p.cols = b.cols;
  size_t col_blocks = div_ceil(b.cols, nc);
  p.data.assign(col_blocks * p.depth_blocks * p.panel_len, 0.f);
  size_t idx = 0;
  for (size_t c0 = 0; c0 < b.cols; c0 += nc) {
    size_t c1 = std::min(c0 + nc, b.cols);
    for (size_t d0 = 0; d0 < b.rows; d0 += kc) {
      size_t d1 = std::min(d0 + kc, b.rows);
      pack_b_block(nr, p.data.data() + idx * p.panel_len, b, d0, d1, c0, c1);
      idx++;
    }
  }
  return p;
}

// src/gemm.rs:651-704
static void gemv(const Kernel* kernel, const float* a_data, size_t a_cols, Matrix b, float* out_data, float alpha,
                 float beta, const float* bias0) {
  size_t b_cols = b.cols;
  size_t threads = (size_t)ThreadPool::instance().num_threads();
  size_t b_block_size = std::max<size_t>(div_ceil(b_cols, threads), 128);
  size_t k_block_size = b.row_stride == 1 ? 512 : 8;
  size_t n_blocks = div_ceil(b_cols, b_block_size);
  ThreadPool::instance().parallel_for(n_blocks, [&](size_t blk) {
    size_t c0 = blk * b_block_size, c1 = std::min((blk + 1) * b_block_size, b_cols);
    float* out_chunk = out_data + c0;
    float effective_beta = beta;
    for (size_t k0 = 0; k0 < a_cols; k0 += k_block_size) {
      size_t k1 = std::min(k0 + k_block_size, a_cols);
      Matrix b_block = b;
      b_block.data = b.data + (ptrdiff_t)k0 * b.row_stride + (ptrdiff_t)c0 * b.col_stride;
      b_block.rows = k1 - k0;
      b_block.cols = c1 - c0;
      kernel->gemv_kernel(out_chunk, c1 - c0, a_data + k0, k1 - k0, b_block, alpha, effective_beta);
      effective_beta = 1.f;
    }
    if (bias0)
      for (size_t i = 0; i < c1 - c0; i++) out_chunk[i] += *bias0;
  });
}

// src/gemm.rs:941-1050
static void gemm_block(const Kernel* kernel, float* out, size_t out_rows, size_t out_cols, size_t out_row_stride,
                       size_t col_tile0, size_t col_tile1, size_t row_tile0, size_t row_tile1, bool first_update,
                       const float* packed_a, const float* packed_b, size_t panel_length, float alpha, float beta,
                       const float* bias) {
  constexpr size_t MAX_MR = 8, MAX_NR = 32;
  const size_t mr = kernel->mr, nr = kernel->nr;
  const size_t b_panel_size = panel_length * nr, a_panel_size = mr * panel_length;
  for (size_t ct = col_tile0; ct < col_tile1; ct++) {
    const float* b_panel = packed_b + (ct - col_tile0) * b_panel_size;
    for (size_t rt = row_tile0; rt < row_tile1; rt++) {
      const float* a_panel = packed_a + (rt - row_tile0) * a_panel_size;
      size_t start_row = rt * mr, start_col = ct * nr;
      float* tile = out + start_row * out_row_stride + start_col;
      size_t used_rows = std::min(out_rows - start_row, mr), used_cols = std::min(out_cols - start_col, nr);
      if (used_rows == mr && used_cols == nr) {
        kernel->kernel(tile, out_row_stride, a_panel, b_panel, panel_length, alpha, beta);
      } else {
        float tmp_tile[MAX_MR * MAX_NR];
        kernel->kernel(tmp_tile, nr, a_panel, b_panel, panel_length, alpha, 0.f);
        for (size_t i = 0; i < used_rows; i++)
          for (size_t j = 0; j < used_cols; j++) {
            float* el = tile + out_row_stride * i + j;
            float t = beta == 0.f ? 0.f : *el;
            *el = beta * t + tmp_tile[i * nr + j];
          }
      }
      if (bias && first_update) {
        for (size_t row = 0; row < used_rows; row++)
          for (size_t col = 0; col < used_cols; col++) tile[row * out_row_stride + col] += bias[rt * mr + row];
      }
    }
  }
}

// src/gemm.rs:733-930
int gemm_impl(const Kernel* kernel, float* out_data, size_t out_len, size_t out_row_stride, GemmInputA a, GemmInputB b,
              float alpha, float beta, const float* bias, size_t bias_len) {
  if (a.cols() != b.rows()) return -1;  // "Columns of matrix `a` must match rows of matrix `b`"
  if (bias && bias_len != a.rows()) return -2;  // "Bias vector length must match rows of matrix `a`"
  const size_t M = a.rows(), N = b.cols(), K = a.cols();
  if (M == 0 || N == 0) return 0;
  if (K == 0) {
    for (size_t i = 0; i < out_len; i++) {
      float t = beta == 0.f ? 0.f : out_data[i];
      out_data[i] = beta * t;
    }
    return 0;
  }
  // MatrixMut::from_data_with_strides validation: last element must be in range.
  if (out_row_stride < N || (M - 1) * out_row_stride + N > out_len) return -3;

  if (M == 1 && a.unpacked && b.unpacked) {
    // a.slice::<1,_>(0).to_contiguous()
    std::vector<float> a_row(K);
    for (size_t k = 0; k < K; k++) a_row[k] = a.unpacked->at(0, k);
    gemv(kernel, a_row.data(), K, *b.unpacked, out_data, alpha, beta, bias ? &bias[0] : nullptr);
    return 0;
  }

  const size_t mr = kernel->mr, nr = kernel->nr;
  const size_t nc = col_block_size(N, nr), mc = row_block_size(M, mr), kc = depth_block_size(K);
  const size_t n_col_blocks = div_ceil(N, nc), n_row_blocks = div_ceil(M, mc);
  ThreadPool& pool = ThreadPool::instance();

  pool.parallel_for(n_col_blocks, [&](size_t col_idx) {
    size_t col_start = col_idx * nc, col_end = std::min(col_start + nc, N);
    std::vector<float> packed_b_buf;
    size_t depth_idx = 0;
    for (size_t d0 = 0; d0 < K; d0 += kc, depth_idx++) {
      size_t d1 = std::min(d0 + kc, K), panel_length = d1 - d0;
      const float* packed_b;
      if (b.packed) {
        packed_b = b.packed->block(col_idx, depth_idx);
      } else {
        packed_b_buf.resize(round_up(col_end - col_start, nr) * panel_length);
        if (b.unpacked) pack_b_block(nr, packed_b_buf.data(), *b.unpacked, d0, d1, col_start, col_end);
        else b.virt->pack_b(packed_b_buf.data(), nr, d0, d1, col_start, col_end);
        packed_b = packed_b_buf.data();
      }
      float effective_beta = d0 == 0 ? beta : 1.f;
      pool.parallel_for(n_row_blocks, [&](size_t row_idx) {
        size_t row_start = row_idx * mc, row_end = std::min(row_start + mc, M);
        static thread_local std::vector<float> packed_a_buf;
        const float* packed_a;
        if (a.packed) {
          packed_a = a.packed->block(row_idx, depth_idx);
        } else {
          packed_a_buf.resize(round_up(row_end - row_start, mr) * panel_length);
          pack_a_block(mr, packed_a_buf.data(), *a.unpacked, row_start, row_end, d0, d1);
          packed_a = packed_a_buf.data();
        }
        gemm_block(kernel, out_data, M, N, out_row_stride, col_start / nr, div_ceil(col_end, nr), row_start / mr,
                   div_ceil(row_end, mr), d0 == 0, packed_a, packed_b, panel_length, alpha, effective_beta, bias);
      });
    }
  });
  return 0;
}

}  // namespace oracle
