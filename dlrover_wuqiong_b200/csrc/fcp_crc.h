// fcp_crc.h — host-side GF(2) arithmetic for CRC-32C used by the product
// library (table generation, finishing of raw device values).  This is NOT
// the oracle: oracle/ holds an independent bytewise implementation that the
// tests check this against.
//
// Reflected CRC-32C state convention: processing one zero byte from state c
// is the linear map shift1(c); shiftN is its N-th power.  A "raw" CRC is the
// CRC with init 0 and no final xor; it is linear in the message and
//   raw(A || B) = shift_{|B|}(raw(A)) ^ raw(B)
//   std(M)      = raw(M) ^ shift_{|M|}(0xFFFFFFFF) ^ 0xFFFFFFFF
#pragma once
#include <stdint.h>

namespace fcpcrc {

constexpr uint32_t kPoly = 0x82F63B78u;

struct Mat {
  uint32_t col[32];  // col[i] = image of the state with only bit i set
};

inline uint32_t apply(const Mat& m, uint32_t v) {
  uint32_t r = 0;
  for (int i = 0; i < 32; ++i)
    if ((v >> i) & 1u) r ^= m.col[i];
  return r;
}

inline Mat mul(const Mat& a, const Mat& b) {  // a after b
  Mat r;
  for (int i = 0; i < 32; ++i) r.col[i] = apply(a, b.col[i]);
  return r;
}

inline Mat identity() {
  Mat r;
  for (int i = 0; i < 32; ++i) r.col[i] = 1u << i;
  return r;
}

inline uint32_t shift_bit(uint32_t c) {
  return (c >> 1) ^ (kPoly & (0u - (c & 1u)));
}

// inverse of shift_bit: kPoly has bit 31 set and (c >> 1) never has, so bit 31
// of the result tells whether the polynomial was folded in.
inline uint32_t unshift_bit(uint32_t c) {
  return (c & 0x80000000u) ? (((c ^ kPoly) << 1) | 1u) : (c << 1);
}

inline Mat pow_bytes(bool inverse, uint64_t nbytes) {
  Mat base;
  for (int i = 0; i < 32; ++i) {
    uint32_t c = 1u << i;
    for (int k = 0; k < 8; ++k) c = inverse ? unshift_bit(c) : shift_bit(c);
    base.col[i] = c;
  }
  Mat r = identity();
  while (nbytes) {
    if (nbytes & 1) r = mul(base, r);
    base = mul(base, base);
    nbytes >>= 1;
  }
  return r;
}

inline Mat shift_bytes(uint64_t n) { return pow_bytes(false, n); }
inline Mat unshift_bytes(uint64_t n) { return pow_bytes(true, n); }

// 4 x 256 lookup table of a linear map: tab[j*256 + b] = m(b << 8j)
inline void make_table(const Mat& m, uint32_t* tab) {
  for (int j = 0; j < 4; ++j)
    for (uint32_t b = 0; b < 256; ++b) tab[j * 256 + b] = apply(m, b << (8 * j));
}

}  // namespace fcpcrc
