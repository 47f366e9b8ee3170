// fcp_kernels.cuh — sm_100a kernels of the flash-checkpoint data path.
//
// Work decomposition: the packed image is cut into 16 KiB
// regions = 32 rows of 512 B.  One warp owns one region at a time; in a row,
// lane l owns the 16-byte word at row*512 + l*16, so every global access to
// the packed image is a fully coalesced 512 B warp transaction.  There is no
// CTA-level cooperation besides the shared CRC tables, so the grid is a
// persistent set of warps striding over regions.
//
// Fused CRC-32C: lane l keeps a Horner accumulator over its column of words
//     acc <- shift_496B(acc) ; acc <- crc32c_update(acc, word)       per row
// (the 496 B shift brings the state from the end of the lane's previous word
// to the start of its next one).  After 32 rows a 5-level warp tree combines
// the 32 column values into the raw CRC of the region; a tiny second kernel
// Horner-combines the regions of each chunk.  All of it is exact GF(2) linear
// algebra, so the result is the standard CRC-32C of the chunk.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/fcp.h"

namespace fcpk {

constexpr int kRowBytes = 512;
constexpr int kRows = 32;
constexpr int kRegion = kRowBytes * kRows;  // == FCP_REGION_BYTES
static_assert(kRegion == FCP_REGION_BYTES, "region size mismatch");

// word offsets inside the device table blob
constexpr int kTabT = 0;            // update by 4 data bytes  (shift 4 B)
constexpr int kTabK = 1024;         // shift 496 B (row stride - word)
constexpr int kTabC = 2048;         // 5 tables: shift 16,32,64,128,256 B
constexpr int kTabF = 2048 + 5120;  // shift one region (16 KiB)
constexpr int kTabWords = kTabF + 1024;

// per-device constant tables, filled once by fcp_init()
static __device__ uint32_t g_fcp_tab[kTabWords];

// ---------------------------------------------------------------- memory ops
__device__ __forceinline__ uint4 ld16_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ld16(const void* p) {
  uint4 r;
  asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st16(void* p, const uint4& v) {
  asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// bytes [mis, mis+16) of the 32-byte concatenation a||b, 0 < mis < 16
__device__ __forceinline__ uint4 funnel16(const uint4& a, const uint4& b,
                                          uint32_t mis) {
  uint32_t x0, x1, x2, x3, x4;
  switch (mis >> 2) {
    case 0: x0 = a.x; x1 = a.y; x2 = a.z; x3 = a.w; x4 = b.x; break;
    case 1: x0 = a.y; x1 = a.z; x2 = a.w; x3 = b.x; x4 = b.y; break;
    case 2: x0 = a.z; x1 = a.w; x2 = b.x; x3 = b.y; x4 = b.z; break;
    default: x0 = a.w; x1 = b.x; x2 = b.y; x3 = b.z; x4 = b.w; break;
  }
  const uint32_t sh = (mis & 3u) * 8u;
  uint4 r;
  r.x = __funnelshift_r(x0, x1, sh);
  r.y = __funnelshift_r(x1, x2, sh);
  r.z = __funnelshift_r(x2, x3, sh);
  r.w = __funnelshift_r(x3, x4, sh);
  return r;
}

// store 16 bytes at an arbitrarily aligned address with the widest stores the
// alignment allows
__device__ __forceinline__ void st16_unaligned(uintptr_t t, const uint4& v) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  if ((t & 15u) == 0) {
    st16(reinterpret_cast<void*>(t), v);
  } else if ((t & 3u) == 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) reinterpret_cast<uint32_t*>(t)[i] = w[i];
  } else if ((t & 1u) == 0) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      reinterpret_cast<uint16_t*>(t)[i] =
          static_cast<uint16_t>(w[i >> 1] >> ((i & 1) * 16));
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i)
      reinterpret_cast<uint8_t*>(t)[i] =
          static_cast<uint8_t>(w[i >> 2] >> ((i & 3) * 8));
  }
}

// ------------------------------------------------------------------- CRC ops
// unreplicated 4x256 table lookup of a linear map
__device__ __forceinline__ uint32_t lin4(const uint32_t* tab, uint32_t x) {
  return tab[x & 0xffu] ^ tab[256 + ((x >> 8) & 0xffu)] ^
         tab[512 + ((x >> 16) & 0xffu)] ^ tab[768 + (x >> 24)];
}
// REPL-way replicated copy (entry e of lane-copy r at e*REPL + r): with
// REPL == 32 lane l only ever touches bank l, so lookups never conflict.
template <int REPL>
__device__ __forceinline__ uint32_t lin4r(const uint32_t* tab_lane,
                                          uint32_t x) {
  return tab_lane[(x & 0xffu) * REPL] ^
         tab_lane[(256 + ((x >> 8) & 0xffu)) * REPL] ^
         tab_lane[(512 + ((x >> 16) & 0xffu)) * REPL] ^
         tab_lane[(768 + (x >> 24)) * REPL];
}

template <int REPL>
__device__ __forceinline__ uint32_t crc_word(const uint32_t* sT_lane,
                                             const uint32_t* sK, uint32_t acc,
                                             const uint4& w) {
  uint32_t s = lin4(sK, acc);
  s = lin4r<REPL>(sT_lane, s ^ w.x);
  s = lin4r<REPL>(sT_lane, s ^ w.y);
  s = lin4r<REPL>(sT_lane, s ^ w.z);
  s = lin4r<REPL>(sT_lane, s ^ w.w);
  return s;
}

// combine the 32 column accumulators of a region; result valid in lane 31
__device__ __forceinline__ uint32_t warp_combine(const uint32_t* sC,
                                                 uint32_t acc, int lane) {
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const uint32_t recv = __shfl_up_sync(0xffffffffu, acc, 1u << k);
    const uint32_t sh = lin4(sC + k * 1024, recv);
    const uint32_t m = (2u << k) - 1u;
    if ((static_cast<uint32_t>(lane) & m) == m) acc ^= sh;
  }
  return acc;
}

// -------------------------------------------------------- descriptor cursor
struct Cur {
  uint64_t off, end;
  uintptr_t ptr;
  uint32_t idx;
};

__device__ __forceinline__ void cur_load(Cur& c, const fcp_desc* __restrict__ d,
                                         uint32_t n) {
  if (c.idx < n) {
    const unsigned long long* q =
        reinterpret_cast<const unsigned long long*>(d + c.idx);
    c.ptr = static_cast<uintptr_t>(__ldg(q));
    c.off = __ldg(q + 1);
    c.end = c.off + __ldg(q + 2);
  } else {
    c.ptr = 0;
    c.off = ~0ull;
    c.end = ~0ull;
  }
}
// advance to the first descriptor whose end is beyond byte p
__device__ __forceinline__ void cur_seek(Cur& c, const fcp_desc* __restrict__ d,
                                         uint32_t n, uint64_t p) {
  while (c.end <= p) {
    ++c.idx;
    cur_load(c, d, n);
  }
}

// first descriptor index whose end is > p; warp-cooperative 32-ary search
__device__ __forceinline__ uint32_t warp_first_desc(
    const fcp_desc* __restrict__ d, uint32_t n, uint64_t p, int lane) {
  uint32_t lo = 0, hi = n;
  while (hi > lo) {
    const uint32_t span = hi - lo;
    const uint32_t step = (span + 31u) / 32u;
    const uint32_t idx = lo + static_cast<uint32_t>(lane) * step;
    bool pred = false;
    if (idx < hi) {
      const unsigned long long* q =
          reinterpret_cast<const unsigned long long*>(d + idx);
      pred = (__ldg(q + 1) + __ldg(q + 2)) <= p;
    }
    const uint32_t cnt = __popc(__ballot_sync(0xffffffffu, pred));
    if (cnt == 0) {
      hi = lo;
    } else {
      const uint32_t nlo = lo + (cnt - 1u) * step + 1u;
      const uint32_t cand = lo + cnt * step;
      hi = (cnt < 32u && cand < hi) ? cand : hi;
      lo = nlo;
    }
  }
  return lo;
}

struct Args {
  const fcp_desc* descs;
  uint32_t n_desc;
  uint64_t total;
  uint64_t region_begin, region_end;
  uint8_t* packed;
  uint32_t* region_sums;
};

}  // namespace fcpk
