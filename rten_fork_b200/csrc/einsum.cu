// Einsum (reference: src/ops/einsum.rs).  Each pairwise step is lowered the way the reference's einsum_matmul does it
// (:330-420): the operands are permuted to [batch.., M.., K..] and [batch.., K.., N..], multiplied by the contraction
// engine as one batched MatMul, and the [batch.., M.., N..] product is permuted into the requested label order.
// Terms are folded left to right (einsum_path :440-520).  Not built on this path: "..." broadcasting, repeated labels
// inside a term (diagonals) and single-operand reductions; they fail with UnsupportedValue rather than fall back.
#include <string>

#include "operators.h"
#include "ops.h"

namespace rtg {
namespace {

struct Labeled {        // a tensor view whose dimension i carries label term[i]
  std::string term;
  rtg_tensor t;
};

int find_label(const std::string& s, char c) {
  const size_t p = s.find(c);
  return p == std::string::npos ? -1 : static_cast<int>(p);
}

// View of `x` with one dimension per label of `order`; labels `x` does not have are broadcast (stride 0, `sizes`).
rtg_tensor view_in_order(const Labeled& x, const std::string& order, const int64_t* sizes) {
  rtg_tensor v = x.t;
  v.ndim = static_cast<int32_t>(order.size());
  for (size_t i = 0; i < order.size(); ++i) {
    const int k = find_label(x.term, order[i]);
    v.shape[i] = sizes[order[i] - 'a'];
    v.strides[i] = k >= 0 ? x.t.strides[k] : 0;
  }
  return v;
}

// Dense copy of `view` reshaped to [d0, d1, d2] in pool scratch (or the view itself when it already is that).
rtg_status dense_3d(Context* ctx, const rtg_tensor& view, int64_t d0, int64_t d1, int64_t d2, Scratch* tmp, rtg_tensor* out) {
  rtg_tensor r = view;
  r.ndim = 3;
  r.shape[0] = d0, r.shape[1] = d1, r.shape[2] = d2;
  set_contiguous_strides(r);
  if (!is_contiguous(view) && d0 * d1 * d2 > 0) {
    RTG_TRY(tmp->alloc(ctx, static_cast<size_t>(d0 * d1 * d2) * 4));
    rtg_tensor dst = view;
    dst.data = tmp->p;
    set_contiguous_strides(dst);
    RTG_TRY(copy_view(ctx, dst, view));
    r.data = tmp->p;
  }
  *out = r;
  return RTG_OK;
}

// out[labels `out_term`] = sum over the labels absent from `out_term` of a * b
rtg_status contract(Context* ctx, const Labeled& a, const Labeled& b, const std::string& out_term, const int64_t* sizes,
                    Value* result) {
  std::string batch, ms, ns, ks;
  for (char c : out_term) {
    const bool in_a = find_label(a.term, c) >= 0, in_b = find_label(b.term, c) >= 0;
    if (in_a && in_b) batch += c;
    else if (in_a) ms += c;
    else ns += c;
  }
  for (char c : a.term + b.term)
    if (find_label(out_term, c) < 0 && find_label(ks, c) < 0) ks += c;
  auto prod = [&](const std::string& s) {
    int64_t p = 1;
    for (char c : s) p *= sizes[c - 'a'];
    return p;
  };
  const int64_t nb = prod(batch), m = prod(ms), n = prod(ns), k = prod(ks);
  if (batch.size() + ms.size() + ks.size() > RTG_MAX_DIMS || batch.size() + ks.size() + ns.size() > RTG_MAX_DIMS)
    return set_error(RTG_ERR_UNSUPPORTED_VALUE, "Einsum input or term has too many dimensions");
  Scratch ta, tb, tp;
  rtg_tensor a3, b3;
  RTG_TRY(dense_3d(ctx, view_in_order(a, batch + ms + ks, sizes), nb, m, k, &ta, &a3));
  RTG_TRY(dense_3d(ctx, view_in_order(b, batch + ks + ns, sizes), nb, k, n, &tb, &b3));
  int64_t oshape[RTG_MAX_DIMS];
  for (size_t i = 0; i < out_term.size(); ++i) oshape[i] = sizes[out_term[i] - 'a'];
  Value y;
  RTG_TRY(alloc_value(ctx, RTG_F32, static_cast<int>(out_term.size()), oshape, false, &y));
  const std::string prod_term = batch + ms + ns;
  rtg_tensor p3 = a3;  // [nb, m, n] product, written straight into the result when the label order already matches
  p3.shape[0] = nb, p3.shape[1] = m, p3.shape[2] = n;
  set_contiguous_strides(p3);
  p3.data = y.t.data;
  rtg_status st = RTG_OK;
  if (prod_term != out_term && nb * m * n > 0) {
    st = tp.alloc(ctx, static_cast<size_t>(nb * m * n) * 4);
    p3.data = tp.p;
  }
  if (st == RTG_OK) st = matmul_forward(ctx, a3, b3, p3, nullptr);
  if (st == RTG_OK && prod_term != out_term && nb * m * n > 0) {
    Labeled p{prod_term, y.t};
    p.t.data = tp.p;
    p.t.ndim = static_cast<int32_t>(prod_term.size());
    for (size_t i = 0; i < prod_term.size(); ++i) p.t.shape[i] = sizes[prod_term[i] - 'a'];
    set_contiguous_strides(p.t);
    st = copy_view(ctx, y.t, view_in_order(p, out_term, sizes));
  }
  if (st != RTG_OK) {
    release_value(ctx, &y);
    return st;
  }
  *result = y;
  return RTG_OK;
}

}  // namespace

rtg_status EinsumOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  // ---- EinsumExpr::parse (:27-93)
  std::string eq;
  for (char c : equation)
    if (c != ' ' && c != '\t' && c != '\n' && c != '\r') eq += c;
  const size_t arrow = eq.find("->");
  const std::string lhs = arrow == std::string::npos ? eq : eq.substr(0, arrow);
  if (lhs.empty()) return set_error(RTG_ERR_INVALID_VALUE, "Einsum equation must have at least one term");
  std::vector<std::string> terms(1);
  for (char c : lhs) {
    if (c == ',') terms.emplace_back();
    else terms.back() += c;
  }
  std::string output = arrow == std::string::npos ? std::string() : eq.substr(arrow + 2);
  for (const std::string& t : terms) {
    if (t.find("...") != std::string::npos)
      return set_error(RTG_ERR_UNSUPPORTED_VALUE, "Einsum broadcasting (\"...\") is not supported on this path");
    for (char c : t)
      if (c < 'a' || c > 'z') return set_error(RTG_ERR_INVALID_VALUE, "Input term is invalid");
  }
  if (output.find("...") != std::string::npos)
    return set_error(RTG_ERR_UNSUPPORTED_VALUE, "Einsum broadcasting (\"...\") is not supported on this path");
  int count[26] = {};
  for (const std::string& t : terms)
    for (char c : t) count[c - 'a']++;
  if (arrow == std::string::npos) {  // implicit output: the labels that occur once, in alphabetical order
    for (int i = 0; i < 26; ++i)
      if (count[i] == 1) output += static_cast<char>('a' + i);
  }
  for (size_t i = 0; i < output.size(); ++i) {
    const char c = output[i];
    if (c < 'a' || c > 'z') return set_error(RTG_ERR_INVALID_VALUE, "Output term is invalid");
    if (count[c - 'a'] == 0) return set_error(RTG_ERR_INVALID_VALUE, "Output term contains a label no input has");
    if (output.find(c) != i) return set_error(RTG_ERR_INVALID_VALUE, "Output term has repeated labels");
  }
  if (output.size() > RTG_MAX_DIMS) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "Einsum input or term has too many dimensions");
  // ---- einsum() argument checks (:151-203)
  if (terms.size() != in.v.size())
    return set_error(RTG_ERR_INVALID_VALUE, "Number of terms in Einsum equation does not match input tensor count");
  std::vector<Labeled> xs;
  int64_t sizes[26];
  for (int i = 0; i < 26; ++i) sizes[i] = -1;
  for (size_t i = 0; i < terms.size(); ++i) {
    const Value* v;
    RTG_TRY(in.require(static_cast<int>(i), &v));
    if (v->t.dtype != RTG_F32) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
    if (static_cast<int>(terms[i].size()) != v->t.ndim)
      return set_error(RTG_ERR_INVALID_VALUE, "Einsum term dimension count does not match input tensor");
    for (size_t d = 0; d < terms[i].size(); ++d) {
      const char c = terms[i][d];
      if (terms[i].find(c) != d)
        return set_error(RTG_ERR_UNSUPPORTED_VALUE, "Einsum diagonals (repeated labels in a term) are not supported on this path");
      int64_t& s = sizes[c - 'a'];
      if (s >= 0 && s != v->t.shape[d])
        return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES, "Dimension sizes for the same Einsum label do not match");
      s = v->t.shape[d];
    }
    xs.push_back(Labeled{terms[i], v->t});
  }
  if (xs.size() == 1) {
    for (char c : xs[0].term)
      if (find_label(output, c) < 0)
        return set_error(RTG_ERR_UNSUPPORTED_VALUE, "single-operand Einsum reductions are not supported on this path");
    int64_t oshape[RTG_MAX_DIMS];
    for (size_t i = 0; i < output.size(); ++i) oshape[i] = sizes[output[i] - 'a'];
    Value y;
    RTG_TRY(alloc_value(ctx, RTG_F32, static_cast<int>(output.size()), oshape, false, &y));
    rtg_status st = copy_view(ctx, y.t, view_in_order(xs[0], output, sizes));
    if (st != RTG_OK) {
      release_value(ctx, &y);
      return st;
    }
    out->push_back(y);
    return RTG_OK;
  }
  // ---- fold the terms left to right; an intermediate keeps the labels a later term or the output still needs
  Labeled cur = xs[0];
  Value held;
  bool have = false;
  for (size_t i = 1; i < xs.size(); ++i) {
    std::string step_out = output;
    if (i + 1 < xs.size()) {
      step_out.clear();
      for (char c : cur.term + xs[i].term) {
        bool needed = find_label(output, c) >= 0;
        for (size_t j = i + 1; j < xs.size() && !needed; ++j) needed = find_label(xs[j].term, c) >= 0;
        if (needed && find_label(step_out, c) < 0) step_out += c;
      }
    }
    Value y;
    rtg_status st = contract(ctx, cur, xs[i], step_out, sizes, &y);
    if (have) release_value(ctx, &held);
    if (st != RTG_OK) return st;
    held = y, have = true;
    cur = Labeled{step_out, y.t};
  }
  out->push_back(held);
  return RTG_OK;
}

}  // namespace rtg
