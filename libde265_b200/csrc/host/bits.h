// bits.h — RBSP bit reader and NAL-unit de-escaping for the host-side HEVC syntax parser.
// Replaces the reference's bitstream.cc (get_bits/get_uvlc/get_svlc) and the emulation-prevention
// removal of nal-parser.cc:151-260 (which also records the removed byte positions, needed to map
// slice entry_point offsets, slice.cc:815-850).
#pragma once
#include <stdint.h>
#include <stddef.h>
#include <vector>

namespace b265h {

struct BitReader {
  const uint8_t* d = nullptr;
  size_t n = 0;      // bytes
  size_t pos = 0;    // bit position
  bool err = false;

  BitReader() {}
  BitReader(const uint8_t* data, size_t bytes) : d(data), n(bytes) {}

  size_t bits_left() const { return n * 8 > pos ? n * 8 - pos : 0; }
  inline uint32_t u1() {
    if (pos >= n * 8) { err = true; return 0; }
    uint32_t v = (d[pos >> 3] >> (7 - (pos & 7))) & 1;
    pos++;
    return v;
  }
  inline uint32_t u(int nb) {
    uint32_t v = 0;
    while (nb > 0) {
      if (pos >= n * 8) { err = true; return v << nb; }
      int avail = 8 - (int)(pos & 7);
      int take = nb < avail ? nb : avail;
      uint32_t byte = d[pos >> 3];
      v = (v << take) | ((byte >> (avail - take)) & ((1u << take) - 1));
      pos += take;
      nb -= take;
    }
    return v;
  }
  inline void skip(int nb) { pos += nb; if (pos > n * 8) err = true; }
  inline uint32_t ue() {
    int z = 0;
    while (u1() == 0) {
      if (err || ++z > 32) { err = true; return 0; }
    }
    if (z == 0) return 0;
    return ((1u << z) - 1) + u(z);
  }
  inline int32_t se() {
    uint32_t v = ue();
    return (v & 1) ? (int32_t)((v + 1) >> 1) : -(int32_t)(v >> 1);
  }
  inline bool aligned() const { return (pos & 7) == 0; }
  inline void align() { pos = (pos + 7) & ~(size_t)7; }
  // more_rbsp_data(): anything before the trailing stop bit?
  bool more_rbsp_data() const {
    if (pos >= n * 8) return false;
    size_t last = n;
    while (last > 0 && d[last - 1] == 0) last--;
    if (last == 0) return false;
    int b = d[last - 1], tz = 0;
    while (!((b >> tz) & 1)) tz++;
    size_t stop_pos = (last - 1) * 8 + (7 - tz);
    return pos < stop_pos;
  }
};

// Remove emulation prevention bytes (00 00 03 -> 00 00); records, for every removed byte, its index
// in the OUTPUT (number of output bytes produced before it).
inline void unescape_nal(const uint8_t* in, size_t n, std::vector<uint8_t>& out, std::vector<uint32_t>& removed) {
  out.clear();
  removed.clear();
  out.reserve(n);
  int zeros = 0;
  for (size_t i = 0; i < n; i++) {
    uint8_t b = in[i];
    if (zeros >= 2 && b == 3) {
      removed.push_back((uint32_t)out.size());
      zeros = 0;
      continue;
    }
    out.push_back(b);
    zeros = (b == 0) ? zeros + 1 : 0;
  }
}

}  // namespace b265h
