// fcp_kernels.cuh — sm_100a kernels of the flash-checkpoint data path.
//
// Work decomposition: the packed image is cut into 16 KiB
// regions = 16 rows of 1 KiB.  One warp owns one region at a time; in a row,
// lane l owns the 32-byte word at row*1024 + l*32 and moves it with one
// 256-bit access (LDG.E.256 / STG.E.256, new on sm_100), so every global
// access to the packed image is a fully coalesced 1 KiB warp transaction.
// There is no CTA-level cooperation besides the shared CRC tables, so the
// grid is a persistent set of warps striding over regions.
//
// Fused CRC-32C: lane l keeps a Horner accumulator over its column of words
//     acc <- shift_992B(acc) ; acc <- crc32c_update(acc, 8 x u32)     per row
// (the 992 B shift brings the state from the end of the lane's previous word
// to the start of its next one).  After 16 rows a 5-level warp tree combines
// the 32 column values into the raw CRC of the region; a tiny second kernel
// Horner-combines the regions of each chunk.  All of it is exact GF(2) linear
// algebra, so the result is the standard CRC-32C of the chunk.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/fcp.h"

namespace fcpk {

constexpr int kWordBytes = 32;
constexpr int kRowBytes = 32 * kWordBytes;  // 1 KiB
constexpr int kRows = 16;
constexpr int kRegion = kRowBytes * kRows;  // == FCP_REGION_BYTES
static_assert(kRegion == FCP_REGION_BYTES, "region size mismatch");

// word offsets inside the device table blob
constexpr int kTabT = 0;            // update by 4 data bytes  (shift 4 B)
constexpr int kTabK = 1024;         // shift 992 B (row stride - word)
constexpr int kTabC = 2048;         // 5 tables: shift 32,64,128,256,512 B
constexpr int kTabF = 2048 + 5120;  // shift one region (16 KiB)
constexpr int kTabWords = kTabF + 1024;
// shift by 2^j work units of 4 KiB, j = 0 .. kRegionShiftTabs-1 (chunk_bytes is
// a uint32: at most 2^20 such units per chunk); used by the chunk finaliser.
// A 16 KiB region is 2^2 units (kCoarseUnitLog); images too small to give every
// warp of a full grid a region are worked in 4 KiB units instead (kFineRows
// rows per warp visit: four times the warps, see profiles/r02_small_images.md).
constexpr int kRegionShiftTabs = 21;
constexpr int kUnitBytes = 4096;
constexpr int kFineRows = 4;
constexpr int kCoarseUnitLog = 2;

// per-device constant tables, filled once by fcp_init()
static __device__ __align__(16) uint32_t g_fcp_tab[kTabWords];
static __device__ __align__(16) uint32_t g_fcp_tab_regions[kRegionShiftTabs * 1024];

struct W8 {
  uint32_t v[8];
};

// ---------------------------------------------------------------- memory ops
__device__ __forceinline__ W8 ld32_stream(const void* p) {
  W8 r;
  asm volatile(
      "ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3]), "=r"(r.v[4]),
        "=r"(r.v[5]), "=r"(r.v[6]), "=r"(r.v[7])
      : "l"(p));
  return r;
}
__device__ __forceinline__ W8 ld32(const void* p) {
  W8 r;
  asm volatile("ld.global.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3]),
                 "=r"(r.v[4]), "=r"(r.v[5]), "=r"(r.v[6]), "=r"(r.v[7])
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st32(void* p, const W8& w) {
  asm volatile("st.global.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p),
               "r"(w.v[0]), "r"(w.v[1]), "r"(w.v[2]), "r"(w.v[3]), "r"(w.v[4]),
               "r"(w.v[5]), "r"(w.v[6]), "r"(w.v[7])
               : "memory");
}
__device__ __forceinline__ void st16(void* p, uint32_t a, uint32_t b,
                                     uint32_t c, uint32_t d) {
  asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(a), "r"(b),
               "r"(c), "r"(d)
               : "memory");
}

// ask the TMA unit to pull [p, p+bytes) into L2 (fire and forget); p and
// bytes must be multiples of 16
__device__ __forceinline__ void prefetch_l2(uintptr_t p, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes)
               : "memory");
}

// the same, cut off at `lim`: nothing at or beyond that address is requested
// (the read side ends there — end of the tensor or of the packed image — and
// what follows may not be mapped)
__device__ __forceinline__ void prefetch_l2_upto(uintptr_t p, uint32_t bytes,
                                                 uintptr_t lim) {
  if (p >= lim) return;
  const uint64_t room = (lim - p) & ~static_cast<uint64_t>(15);
  const uint32_t n = room < bytes ? static_cast<uint32_t>(room) : bytes;
  if (n) prefetch_l2(p, n);
}

// bytes [m, m+32) of the 64-byte concatenation lo||hi, 0 < m < 32
__device__ __forceinline__ W8 extract32(const W8& lo, const W8& hi,
                                        uint32_t m) {
  uint32_t x[9];
#define FCP_PICK(S)                                                        \
  {                                                                        \
    _Pragma("unroll") for (int i = 0; i < 9; ++i) x[i] =                   \
        (S + i < 8) ? lo.v[(S + i) & 7] : hi.v[(S + i) & 7];               \
  }
  switch (m >> 2) {
    case 0: FCP_PICK(0) break;
    case 1: FCP_PICK(1) break;
    case 2: FCP_PICK(2) break;
    case 3: FCP_PICK(3) break;
    case 4: FCP_PICK(4) break;
    case 5: FCP_PICK(5) break;
    case 6: FCP_PICK(6) break;
    default: FCP_PICK(7) break;
  }
#undef FCP_PICK
  const uint32_t sh = (m & 3u) * 8u;
  W8 r;
#pragma unroll
  for (int i = 0; i < 8; ++i) r.v[i] = __funnelshift_r(x[i], x[i + 1], sh);
  return r;
}

// store 32 bytes at an arbitrarily aligned address with the widest stores the
// alignment allows
__device__ __forceinline__ void st32_unaligned(uintptr_t t, const W8& w) {
  if ((t & 31u) == 0) {
    st32(reinterpret_cast<void*>(t), w);
  } else if ((t & 15u) == 0) {
    st16(reinterpret_cast<void*>(t), w.v[0], w.v[1], w.v[2], w.v[3]);
    st16(reinterpret_cast<void*>(t + 16), w.v[4], w.v[5], w.v[6], w.v[7]);
  } else if ((t & 3u) == 0) {
#pragma unroll
    for (int i = 0; i < 8; ++i) reinterpret_cast<uint32_t*>(t)[i] = w.v[i];
  } else if ((t & 1u) == 0) {
#pragma unroll
    for (int i = 0; i < 16; ++i)
      reinterpret_cast<uint16_t*>(t)[i] =
          static_cast<uint16_t>(w.v[i >> 1] >> ((i & 1) * 16));
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i)
      reinterpret_cast<uint8_t*>(t)[i] =
          static_cast<uint8_t>(w.v[i >> 2] >> ((i & 3) * 8));
  }
}

// ------------------------------------------------------------------- CRC ops
// The lookup tables live in shared memory and are addressed in the shared
// window directly (32-bit addresses, no generic-pointer arithmetic).  A table
// of a linear map has 4 parts (one per input byte) of 256 entries; it is
// replicated R ways with entry e of copy r at (e*R + r)*4 bytes.  With R == 32
// lane l only ever touches bank l: no conflicts at all; with R == 16 two
// lanes share a copy that spans two banks.
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}

template <int R>
struct TabGeom {
  static constexpr uint32_t kEntry = R * 4u;         // bytes between entries
  static constexpr uint32_t kMask = 0xffu * kEntry;  // scaled byte index
  static constexpr uint32_t kPart = 256u * kEntry;   // bytes per part
  static constexpr int kLog = R == 32 ? 7 : R == 16 ? 6 : R == 8 ? 5 : 2;
  static_assert((1u << kLog) == kEntry, "unsupported replication");
};

// `base` = shared address of the table, `lane_off` = (lane % R) * 4.  Each
// lookup is shift + (and|or) + LDS [reg + uniform base + imm].
template <int R>
__device__ __forceinline__ uint32_t lin4s(uint32_t base, uint32_t lane_off,
                                          uint32_t x) {
  using G = TabGeom<R>;
  const uint32_t o0 = ((x << G::kLog) & G::kMask) | lane_off;
  const uint32_t o1 = ((x >> (8 - G::kLog)) & G::kMask) | lane_off;
  const uint32_t o2 = ((x >> (16 - G::kLog)) & G::kMask) | lane_off;
  const uint32_t o3 = ((x >> (24 - G::kLog)) & G::kMask) | lane_off;
  return lds32(base + o0) ^ lds32(base + G::kPart + o1) ^
         lds32(base + 2 * G::kPart + o2) ^ lds32(base + 3 * G::kPart + o3);
}

struct Tabs {
  uint32_t T, K, C;     // shared addresses of the update, row-shift and
                        // combine tables
  uint32_t offT, offK;  // this lane's copy offsets
};

template <int RT, int RK>
__device__ __forceinline__ uint32_t crc_word(const Tabs& t, uint32_t acc,
                                             const W8& w) {
  uint32_t s = lin4s<RK>(t.K, t.offK, acc);
#pragma unroll
  for (int i = 0; i < 8; ++i) s = lin4s<RT>(t.T, t.offT, s ^ w.v[i]);
  return s;
}

// combine the 32 column accumulators of a region; result valid in lane 31
__device__ __forceinline__ uint32_t warp_combine(const Tabs& t, uint32_t acc,
                                                 int lane) {
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const uint32_t recv = __shfl_up_sync(0xffffffffu, acc, 1u << k);
    const uint32_t sh = lin4s<1>(t.C + k * 4096u, 0u, recv);
    const uint32_t m = (2u << k) - 1u;
    if ((static_cast<uint32_t>(lane) & m) == m) acc ^= sh;
  }
  return acc;
}

// global-memory flavour used by the finalize kernel (unreplicated table)
__device__ __forceinline__ uint32_t lin4(const uint32_t* tab, uint32_t x) {
  return tab[x & 0xffu] ^ tab[256 + ((x >> 8) & 0xffu)] ^
         tab[512 + ((x >> 16) & 0xffu)] ^ tab[768 + (x >> 24)];
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
  uint32_t run;    // regions per warp visit, set by the launcher
  uint32_t flags;  // kNoPrefetch: the read side is peer memory (never in local L2)
  uint32_t* work;  // two zeroed counters for dynamic run distribution, or nullptr
  // Scatter from an image whose pieces live in several buffers (peers' HBM):
  // region r is read through slice_base[r / slice_regions] instead of
  // `packed`.  Every entry is the address image byte 0 WOULD have in that
  // buffer (so absolute offsets apply); the buffer only has to hold the bytes
  // of the slices that point at it (a region never reads outside itself).
  const uint8_t* const* slice_base = nullptr;  // device array, or nullptr
  uint32_t slice_regions = 0;
  uint64_t rotate = 0;  // sliced scatter: run index this launch starts from
  // kFusedFinalize: the last warp to leave also combines the region values
  // of chunks [chunk_begin, chunk_end) into chunk_sums (small images: saves
  // the separate finalize launch)
  uint32_t* chunk_sums = nullptr;
  uint32_t regions_per_chunk = 0;
  uint64_t chunk_begin = 0, chunk_end = 0;
};
constexpr uint32_t kNoPrefetch = 1u;
constexpr uint32_t kFusedFinalize = 2u;
constexpr uint32_t kFineUnits = 4u;  // region_begin/end and region_sums count 4 KiB units
constexpr uint64_t kFuseMaxChunks = 2;  // the last warp combines them one after the other: 32 cost 40 us

}  // namespace fcpk
