// fcp_kernels2.cuh — the pack / scatter kernel body and the chunk finaliser.
#pragma once
#include "fcp_kernels.cuh"

namespace fcpk {

// ---- word-level slow paths (tensor boundaries, gaps, tail of the image) ----

// PACK: assemble the 16 packed bytes [o, o+16) from whatever descriptors
// cover them; uncovered / out-of-range bytes are zero.
__device__ __noinline__ uint4 gather_word(Cur& c, const fcp_desc* __restrict__ d,
                                          uint32_t n, uint64_t o,
                                          uint64_t total) {
  uint32_t w[4] = {0u, 0u, 0u, 0u};
#pragma unroll
  for (int b = 0; b < 16; ++b) {
    const uint64_t p = o + b;
    if (p < total) {
      cur_seek(c, d, n, p);
      if (c.off <= p) {
        const uint32_t v = *reinterpret_cast<const uint8_t*>(c.ptr + (p - c.off));
        w[b >> 2] |= v << ((b & 3) * 8);
      }
    }
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}

// SCATTER: hand the 16 packed bytes [o, o+16) to whatever descriptors cover
// them, byte by byte.
__device__ __noinline__ void scatter_word(Cur& c, const fcp_desc* __restrict__ d,
                                          uint32_t n, uint64_t o, uint64_t total,
                                          const uint4& v) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int b = 0; b < 16; ++b) {
    const uint64_t p = o + b;
    if (p < total) {
      cur_seek(c, d, n, p);
      if (c.off <= p)
        *reinterpret_cast<uint8_t*>(c.ptr + (p - c.off)) =
            static_cast<uint8_t>(w[b >> 2] >> ((b & 3) * 8));
    }
  }
}

// zero the bytes of a packed word that lie at or beyond `total`
__device__ __forceinline__ uint4 mask_tail(const uint4& v, uint64_t o,
                                           uint64_t total) {
  if (o + 16 <= total) return v;
  uint32_t w[4] = {v.x, v.y, v.z, v.w};
  const int valid = (o >= total) ? 0 : static_cast<int>(total - o);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int vb = valid - i * 4;  // valid bytes of this 32-bit word
    if (vb <= 0) w[i] = 0u;
    else if (vb < 4) w[i] &= (1u << (vb * 8)) - 1u;
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}

// store a packed word whose tail may stick out of the image (PACK side)
__device__ __forceinline__ void store_packed(uint8_t* packed, uint64_t o,
                                             uint64_t total, const uint4& v) {
  if (o + 16 <= total) {
    st16(packed + o, v);
  } else {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    for (int b = 0; b < 16 && o + b < total; ++b)
      packed[o + b] = static_cast<uint8_t>(w[b >> 2] >> ((b & 3) * 8));
  }
}

// ---------------------------------------------------------------- the kernel
// PACK = true : tensors -> packed image   (desc.ptr is the source)
// PACK = false: packed image -> tensors   (desc.ptr is the target)
// REPL: replication of the 4-byte update table in shared memory (1 or 32).
template <bool PACK, bool CRC, int REPL, int THREADS>
__global__ void __launch_bounds__(THREADS, 1) fcp_kernel(const Args a) {
  extern __shared__ uint32_t smem[];
  uint32_t* sT = smem;                    // 1024 * REPL words
  uint32_t* sK = smem + 1024 * REPL;      // 1024 words
  uint32_t* sC = sK + 1024;               // 5120 words
  if (CRC) {
    for (int i = threadIdx.x; i < 1024 * REPL; i += THREADS)
      sT[i] = g_fcp_tab[kTabT + i / REPL];
    for (int i = threadIdx.x; i < 1024 + 5120; i += THREADS)
      sK[i] = g_fcp_tab[kTabK + i];
    __syncthreads();
  }
  const int lane = threadIdx.x & 31;
  const uint32_t* sT_lane = sT + (REPL == 32 ? lane : 0);
  const uint64_t warps_total =
      static_cast<uint64_t>(gridDim.x) * (THREADS / 32);
  const uint64_t gwarp =
      static_cast<uint64_t>(blockIdx.x) * (THREADS / 32) + (threadIdx.x >> 5);
  const fcp_desc* __restrict__ descs = a.descs;
  const uint32_t n = a.n_desc;
  const uint64_t total = a.total;

  for (uint64_t r = a.region_begin + gwarp; r < a.region_end; r += warps_total) {
    const uint64_t rbase = r * kRegion;
    Cur c;
    c.idx = warp_first_desc(descs, n, rbase, lane);
    cur_load(c, descs, n);
    uint32_t acc = 0;
    const uint64_t lane_o = rbase + static_cast<uint64_t>(lane) * 16;

    // Whole region inside one tensor (uniform across the warp): no per-word
    // descriptor logic, loads batched for memory-level parallelism.
    const bool whole = (c.off <= rbase) && (rbase + kRegion <= c.end);
    if (whole) {
      const uintptr_t t0 = c.ptr + (rbase - c.off) + lane * 16;  // tensor side
      uint8_t* p0 = a.packed + lane_o;                           // packed side
      const uint32_t mis = static_cast<uint32_t>(t0 & 15u);
      if (PACK) {
        if (mis == 0) {
#pragma unroll 1
          for (int row = 0; row < kRows; row += 8) {
            uint4 v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k)
              v[k] = ld16_stream(reinterpret_cast<const void*>(
                  t0 + (row + k) * kRowBytes));
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              st16(p0 + (row + k) * kRowBytes, v[k]);
              if (CRC) acc = crc_word<REPL>(sT_lane, sK, acc, v[k]);
            }
          }
        } else {
          const uintptr_t ta = t0 - mis;
#pragma unroll 1
          for (int row = 0; row < kRows; row += 4) {
            uint4 lo[4], hi[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              lo[k] = ld16(reinterpret_cast<const void*>(
                  ta + (row + k) * kRowBytes));
              hi[k] = ld16(reinterpret_cast<const void*>(
                  ta + (row + k) * kRowBytes + 16));
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint4 v = funnel16(lo[k], hi[k], mis);
              st16(p0 + (row + k) * kRowBytes, v);
              if (CRC) acc = crc_word<REPL>(sT_lane, sK, acc, v);
            }
          }
        }
      } else {
#pragma unroll 1
        for (int row = 0; row < kRows; row += 8) {
          uint4 v[8];
#pragma unroll
          for (int k = 0; k < 8; ++k)
            v[k] = ld16_stream(p0 + (row + k) * kRowBytes);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            st16_unaligned(t0 + (row + k) * kRowBytes, v[k]);
            if (CRC) acc = crc_word<REPL>(sT_lane, sK, acc, v[k]);
          }
        }
      }
    } else {
      // General path: every word decides for itself.
#pragma unroll 1
      for (int row = 0; row < kRows; ++row) {
        const uint64_t o = lane_o + static_cast<uint64_t>(row) * kRowBytes;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (o < total) {
          cur_seek(c, descs, n, o);
          const bool interior = (c.off <= o) && (o + 16 <= c.end);
          if (PACK) {
            if (interior) {
              const uintptr_t t = c.ptr + (o - c.off);
              const uint32_t mis = static_cast<uint32_t>(t & 15u);
              if (mis == 0) {
                v = ld16_stream(reinterpret_cast<const void*>(t));
              } else {
                const uint4 lo = ld16(reinterpret_cast<const void*>(t - mis));
                const uint4 hi =
                    ld16(reinterpret_cast<const void*>(t - mis + 16));
                v = funnel16(lo, hi, mis);
              }
              st16(a.packed + o, v);
            } else {
              v = gather_word(c, descs, n, o, total);
              store_packed(a.packed, o, total, v);
            }
          } else {
            v = mask_tail(ld16_stream(a.packed + o), o, total);
            if (interior)
              st16_unaligned(c.ptr + (o - c.off), v);
            else
              scatter_word(c, descs, n, o, total, v);
          }
        }
        if (CRC) acc = crc_word<REPL>(sT_lane, sK, acc, v);
      }
      __syncwarp();
    }
    if (CRC) {
      acc = warp_combine(sC, acc, lane);
      if (lane == 31) a.region_sums[r] = acc;
    }
  }
}

// One thread per chunk: Horner over the chunk's region values.  The value
// written is the raw CRC of the chunk zero-padded to a whole number of
// regions; the host strips the padding and applies init/xorout
// (fcp_finish_checksums).
__global__ void fcp_finalize_kernel(const uint32_t* __restrict__ region_sums,
                                    uint32_t* __restrict__ chunk_sums,
                                    uint64_t n_regions,
                                    uint32_t regions_per_chunk,
                                    uint64_t chunk_begin, uint64_t chunk_end) {
  const uint64_t c = chunk_begin + blockIdx.x * static_cast<uint64_t>(blockDim.x) +
                     threadIdx.x;
  if (c >= chunk_end) return;
  const uint64_t r0 = c * regions_per_chunk;
  uint64_t r1 = r0 + regions_per_chunk;
  if (r1 > n_regions) r1 = n_regions;
  const uint32_t* F = g_fcp_tab + kTabF;
  uint32_t acc = 0;
  // region values are independent of acc: fetch them 8 at a time so the
  // serial Horner chain only waits on (L1-resident) table lookups
  uint64_t r = r0;
  for (; r + 8 <= r1; r += 8) {
    uint32_t v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = __ldg(region_sums + r + k);
#pragma unroll
    for (int k = 0; k < 8; ++k) acc = lin4(F, acc) ^ v[k];
  }
  for (; r < r1; ++r) acc = lin4(F, acc) ^ __ldg(region_sums + r);
  chunk_sums[c] = acc;
}

}  // namespace fcpk
