// Minimal, bounds-checked FlatBuffers reader (wire format only; no generated code, no flatc).
// A FlatBuffer is: u32 root offset, then tables.  A table starts with an i32 offset back to its vtable;
// the vtable is u16 vtable_size, u16 table_size, then one u16 slot per field (0 = field absent => default).
// Vectors are u32 length + elements; strings are vectors of bytes with a trailing NUL; unions are a u8 type field
// followed by an offset field.  Reference: the .rten schema, src/schema.fbs.
#pragma once
#include <stdint.h>
#include <string.h>

#include <stdexcept>
#include <string>

namespace rtg {
namespace fb {

struct Error : std::runtime_error {
  using std::runtime_error::runtime_error;
};

struct Buf {
  const uint8_t* p = nullptr;
  size_t n = 0;
  void need(size_t off, size_t len) const {
    if (off > n || len > n - off) throw Error("flatbuffer: offset out of bounds");
  }
  template <class T>
  T rd(size_t off) const {
    need(off, sizeof(T));
    T v;
    memcpy(&v, p + off, sizeof(T));
    return v;
  }
};

template <class T>
struct Vec {
  Buf b;
  size_t pos = 0;  // first element
  uint32_t len = 0;
  bool present = false;
  uint32_t size() const { return len; }
  T operator[](uint32_t i) const {
    if (i >= len) throw Error("flatbuffer: vector index out of range");
    return b.rd<T>(pos + static_cast<size_t>(i) * sizeof(T));
  }
  const uint8_t* data() const { return b.p + pos; }
};

struct Table {
  Buf b;
  size_t pos = 0;
  size_t vt = 0;
  uint16_t vt_size = 0;
  bool valid = false;

  static Table at(const Buf& b, size_t pos) {
    Table t;
    t.b = b;
    t.pos = pos;
    const int32_t so = b.rd<int32_t>(pos);
    const int64_t vt = static_cast<int64_t>(pos) - so;
    if (vt < 0) throw Error("flatbuffer: bad vtable offset");
    t.vt = static_cast<size_t>(vt);
    t.vt_size = b.rd<uint16_t>(t.vt);
    if (t.vt_size < 4 || (t.vt_size & 1)) throw Error("flatbuffer: bad vtable size");
    b.need(t.vt, t.vt_size);
    t.valid = true;
    return t;
  }
  // Byte offset of field `id` inside the table, or 0 when absent.
  size_t field(int id) const {
    const size_t slot = 4 + 2 * static_cast<size_t>(id);
    if (!valid || slot + 2 > vt_size) return 0;
    const uint16_t off = b.rd<uint16_t>(vt + slot);
    return off ? pos + off : 0;
  }
  bool has(int id) const { return field(id) != 0; }
  template <class T>
  T scalar(int id, T def) const {
    const size_t f = field(id);
    return f ? b.rd<T>(f) : def;
  }
  size_t indirect(int id) const {
    const size_t f = field(id);
    if (!f) return 0;
    return f + b.rd<uint32_t>(f);
  }
  Table table(int id) const {
    const size_t t = indirect(id);
    return t ? Table::at(b, t) : Table();
  }
  std::string str(int id) const {
    const size_t s = indirect(id);
    if (!s) return std::string();
    const uint32_t len = b.rd<uint32_t>(s);
    b.need(s + 4, len);
    return std::string(reinterpret_cast<const char*>(b.p + s + 4), len);
  }
  template <class T>
  Vec<T> vec(int id) const {
    Vec<T> v;
    const size_t s = indirect(id);
    if (!s) return v;
    v.b = b;
    v.len = b.rd<uint32_t>(s);
    v.pos = s + 4;
    b.need(v.pos, static_cast<size_t>(v.len) * sizeof(T));
    v.present = true;
    return v;
  }
  // element i of a vector of tables
  Table vec_table(const Vec<uint32_t>& v, uint32_t i) const {
    if (i >= v.len) throw Error("flatbuffer: vector index out of range");
    const size_t slot = v.pos + static_cast<size_t>(i) * 4;
    return Table::at(b, slot + b.rd<uint32_t>(slot));
  }
};

inline Table root(const uint8_t* p, size_t n) {
  Buf b{p, n};
  const uint32_t off = b.rd<uint32_t>(0);
  return Table::at(b, off);
}

}  // namespace fb
}  // namespace rtg
