// Protobuf wire-format cursor shared by the checkpoint reader (tf_bundle.cc) and the TFRecord
// reader (tfrecord.cc).  Header-only, host code.
#pragma once
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>

namespace gnnqc {
namespace wire {

[[noreturn]] inline void wire_fail(const char* what) { throw std::runtime_error(std::string("protobuf wire: ") + what); }

// Bounded cursor over a byte range; all protobuf / table decoding goes through it.
struct Cursor {
  const uint8_t* p;
  const uint8_t* end;
  bool done() const { return p >= end; }
  uint64_t varint() {
    uint64_t v = 0;
    for (int shift = 0; shift < 64; shift += 7) {
      if (p >= end) wire_fail("truncated varint");
      uint8_t c = *p++;
      v |= static_cast<uint64_t>(c & 0x7f) << shift;
      if (!(c & 0x80)) return v;
    }
    wire_fail("varint too long");
  }
  Cursor sub(size_t n) {
    if (static_cast<size_t>(end - p) < n) wire_fail("truncated length-delimited field");
    Cursor c{p, p + n};
    p += n;
    return c;
  }
  uint32_t fixed32() {
    if (end - p < 4) wire_fail("truncated fixed32");
    uint32_t v;
    std::memcpy(&v, p, 4);
    p += 4;
    return v;
  }
  void skip(int wire) {
    switch (wire) {
      case 0: varint(); break;
      case 1: sub(8); break;
      case 2: sub(varint()); break;
      case 5: sub(4); break;
      default: wire_fail("unsupported protobuf wire type");
    }
  }
};


}  // namespace wire
}  // namespace gnnqc
