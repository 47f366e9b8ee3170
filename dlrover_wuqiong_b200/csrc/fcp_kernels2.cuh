// fcp_kernels2.cuh — the pack / scatter kernel body and the chunk finaliser.
#pragma once
#include "fcp_kernels.cuh"

#ifndef FCP_ALIGNED_BATCH
#define FCP_ALIGNED_BATCH 2
#endif

namespace fcpk {
constexpr uint32_t kMaxRun = 8;  // consecutive regions per warp visit (<= 128 KiB)
}

namespace fcpk {

// ---- word-level slow paths (tensor boundaries, gaps, tail of the image) ----

// PACK: assemble the 32 packed bytes [o, o+32) from whatever descriptors
// cover them; uncovered / out-of-range bytes are zero.  One pass per
// descriptor that overlaps the word (typically two at a tensor boundary):
// its bytes are fetched with predicated byte loads at compile-time positions,
// so the word stays in registers and the loads go out back to back.
__device__ __noinline__ W8 gather_word(Cur& c, const fcp_desc* __restrict__ d,
                                       uint32_t n, uint64_t o, uint64_t total) {
  W8 w;
#pragma unroll
  for (int i = 0; i < 8; ++i) w.v[i] = 0u;
  const uint64_t end = o + kWordBytes < total ? o + kWordBytes : total;
  uint64_t p = o;
  while (p < end) {
    cur_seek(c, d, n, p);          // first descriptor that ends beyond p
    if (c.off >= end) break;       // nothing but a gap up to the end of the word
    const uint32_t a = static_cast<uint32_t>((c.off > p ? c.off : p) - o);
    const uint32_t b = static_cast<uint32_t>((c.end < end ? c.end : end) - o);
    const uint8_t* src = reinterpret_cast<const uint8_t*>(c.ptr + (o - c.off));  // byte i at src[i]
#pragma unroll
    for (int i = 0; i < kWordBytes; ++i)
      if (static_cast<uint32_t>(i) >= a && static_cast<uint32_t>(i) < b)
        w.v[i >> 2] |= static_cast<uint32_t>(src[i]) << ((i & 3) * 8);
    p = o + b;
  }
  return w;
}

// SCATTER: hand the 32 packed bytes [o, o+32) to whatever descriptors cover
// them; same structure as gather_word.  The word comes as scalars (a
// reference parameter would pin the caller's W8 in local memory).
__device__ __noinline__ void scatter_word_regs(Cur& c, const fcp_desc* __restrict__ d,
                                               uint32_t n, uint64_t o, uint64_t total,
                                               uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3,
                                               uint32_t w4, uint32_t w5, uint32_t w6, uint32_t w7) {
  const uint64_t end = o + kWordBytes < total ? o + kWordBytes : total;
  uint64_t p = o;
  while (p < end) {
    cur_seek(c, d, n, p);
    if (c.off >= end) break;
    const uint32_t a = static_cast<uint32_t>((c.off > p ? c.off : p) - o);
    const uint32_t b = static_cast<uint32_t>((c.end < end ? c.end : end) - o);
    uint8_t* dst = reinterpret_cast<uint8_t*>(c.ptr + (o - c.off));
#pragma unroll
    for (int i = 0; i < kWordBytes; ++i) {
      const uint32_t w = (i >> 2) == 0 ? w0 : (i >> 2) == 1 ? w1 : (i >> 2) == 2 ? w2
                       : (i >> 2) == 3 ? w3 : (i >> 2) == 4 ? w4 : (i >> 2) == 5 ? w5
                       : (i >> 2) == 6 ? w6 : w7;
      if (static_cast<uint32_t>(i) >= a && static_cast<uint32_t>(i) < b)
        dst[i] = static_cast<uint8_t>(w >> ((i & 3) * 8));
    }
    p = o + b;
  }
}

__device__ __forceinline__ void scatter_word(Cur& c, const fcp_desc* __restrict__ d,
                                             uint32_t n, uint64_t o, uint64_t total,
                                             const W8& w) {
  scatter_word_regs(c, d, n, o, total, w.v[0], w.v[1], w.v[2], w.v[3], w.v[4], w.v[5], w.v[6],
                    w.v[7]);
}

// store a packed word whose tail may stick out of the image (PACK side)
__device__ __forceinline__ void store_packed(uint8_t* packed, uint64_t o,
                                             uint64_t total, const W8& w) {
  if (o + kWordBytes <= total) {
    st32(packed + o, w);
  } else {
#pragma unroll
    for (int b = 0; b < kWordBytes; ++b)
      if (o + b < total)
        packed[o + b] = static_cast<uint8_t>(w.v[b >> 2] >> ((b & 3) * 8));
  }
}

// load the packed word at offset o (SCATTER side); a word that straddles the
// end of the image is read byte by byte so no byte beyond `total` is touched
// (bytes beyond it count as zero for the checksum)
__device__ __forceinline__ W8 load_packed(const uint8_t* packed, uint64_t o,
                                          uint64_t total) {
  if (o + kWordBytes <= total) return ld32_stream(packed + o);
  W8 w;
#pragma unroll
  for (int i = 0; i < 8; ++i) w.v[i] = 0u;
#pragma unroll
  for (int b = 0; b < kWordBytes; ++b)
    if (o + b < total)
      w.v[b >> 2] |= static_cast<uint32_t>(packed[o + b]) << ((b & 3) * 8);
  return w;
}

// bytes [from, from+n) of w to consecutive addresses starting at t
// Bytes [from, from+n) of a word, stored one by one at t.  Cold (twice per
// misaligned region).  The word comes as eight scalars on purpose: a
// `const W8&` parameter made the caller keep its W8s in local memory for the
// whole loop (two 32-byte local stores per payload word, ~15 % of the kernel).
__device__ __noinline__ void store_some_bytes_regs(uintptr_t t, uint32_t w0, uint32_t w1,
                                                   uint32_t w2, uint32_t w3, uint32_t w4,
                                                   uint32_t w5, uint32_t w6, uint32_t w7,
                                                   int from, int n) {
  uint8_t* out = reinterpret_cast<uint8_t*>(t);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const uint32_t w = i == 0 ? w0 : i == 1 ? w1 : i == 2 ? w2 : i == 3 ? w3
                     : i == 4 ? w4 : i == 5 ? w5 : i == 6 ? w6 : w7;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int b = i * 4 + j;
      if (b >= from && b < from + n) out[b - from] = static_cast<uint8_t>(w >> (j * 8));
    }
  }
}

__device__ __forceinline__ void store_some_bytes(uintptr_t t, const W8& w, int from, int n) {
  store_some_bytes_regs(t, w.v[0], w.v[1], w.v[2], w.v[3], w.v[4], w.v[5], w.v[6], w.v[7],
                        from, n);
}

// shift a raw CRC state by n regions of zeros (binary decomposition of n over
// the 2^j-region tables)
__device__ __forceinline__ uint32_t shift_regions(uint32_t x, uint32_t n, int unit_log) {
  for (int j = unit_log; n != 0u; ++j, n >>= 1)
    if (n & 1u) x = lin4(g_fcp_tab_regions + j * 1024, x);
  return x;
}

// One WARP combines the region values of chunk c into the raw CRC of the
// chunk (its last region zero-padded to 16 KiB; the host strips that padding
// and applies init/xorout, fcp_finish_checksums).  Lane l owns up to
// q = ceil(regions_per_chunk / 32) consecutive regions, right-aligned so that
// lane 31 owns the chunk's last ones (empty leading lanes contribute the raw
// CRC of nothing, 0, which no shift changes); q Horner steps, then a 5-level
// tree.  Region values are read from L2 (ld.cg): in the fused form they were
// written by other SMs of the same launch.
// `unit_log`: the values are those of 4 KiB << unit_log work units; `n_real`
// of them exist, and the chunk is taken as zero-padded up to `n_padded` units
// (the end of the image's last 16 KiB region, whatever the unit).
__device__ __forceinline__ void combine_chunk(const uint32_t* __restrict__ region_sums,
                                              uint32_t* __restrict__ chunk_sums,
                                              uint64_t n_real, uint64_t n_padded,
                                              uint32_t regions_per_chunk, uint64_t c, int lane,
                                              int unit_log) {
  const uint64_t r0 = c * regions_per_chunk;
  uint64_t r1 = r0 + regions_per_chunk;
  if (r1 > n_padded) r1 = n_padded;
  const int64_t n = static_cast<int64_t>(r1 - r0);
  const int64_t q = (static_cast<int64_t>(regions_per_chunk) + 31) / 32;
  int64_t hi = n - static_cast<int64_t>(31 - lane) * q;
  int64_t lo = hi - q;
  if (hi < 0) hi = 0;
  if (lo < 0) lo = 0;
  uint32_t acc = 0;
  const uint32_t* one = g_fcp_tab_regions + unit_log * 1024;  // shift by one unit
  int64_t j = lo;
  for (; j + 4 <= hi; j += 4) {  // values first: they do not depend on acc
    uint32_t v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
      v[k] = r0 + j + k < n_real ? __ldcg(region_sums + r0 + j + k) : 0u;
#pragma unroll
    for (int k = 0; k < 4; ++k) acc = lin4(one, acc) ^ v[k];
  }
  for (; j < hi; ++j)
    acc = lin4(one, acc) ^ (r0 + j < n_real ? __ldcg(region_sums + r0 + j) : 0u);
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const uint32_t recv = __shfl_up_sync(0xffffffffu, acc, 1u << k);
    const uint32_t sh = shift_regions(recv, static_cast<uint32_t>(q) << k, unit_log);
    const uint32_t m = (2u << k) - 1u;
    if ((static_cast<uint32_t>(lane) & m) == m) acc ^= sh;
  }
  if (lane == 31) chunk_sums[c] = acc;
}

// ---------------------------------------------------------------- the kernel
// PACK = true : tensors -> packed image   (desc.ptr is the source)
// PACK = false: packed image -> tensors   (desc.ptr is the target)
// REPL: replication of the 4-byte update table in shared memory (1 or 32);
// the row-shift table is replicated REPL/2 ways.
// SLICED: the image is read through Args.slice_base (scatter from several
// buffers); a separate instantiation so the ordinary kernels keep their
// register budget.
// ROWS: rows (KiB) of the image one warp visit covers: 16 = a region; 4 = the
// fine unit of small images (Args.region_begin/end and region_sums then count
// 4 KiB units; same code, a quarter of the rows).
template <bool PACK, bool CRC, int REPL, int THREADS, bool SLICED = false, int ROWS = 16>
__global__ void __launch_bounds__(THREADS, 1) fcp_kernel(const Args a) {
  constexpr int kRows = ROWS;                 // (shadow the 16-row constants
  constexpr int kRegion = kRowBytes * ROWS;   //  of the namespace)
  constexpr int RK = REPL > 1 ? REPL / 2 : 1;
  constexpr int kBatch = FCP_ALIGNED_BATCH;
  extern __shared__ uint32_t smem[];
  uint32_t* sT = smem;              // 1024 * REPL words
  uint32_t* sK = sT + 1024 * REPL;  // 1024 * RK words
  uint32_t* sC = sK + 1024 * RK;    // 5120 words
  if (CRC) {
    if (REPL == 1) {
      // unreplicated tables sit in shared memory in the order of the device
      // blob: one straight copy, all of a thread's 16-byte loads in flight at
      // once (a dependent word-by-word fill cost ~22 us per CTA on a cold L2,
      // most of the time of a small image)
      constexpr int kVec = (1024 + 1024 + 5120) / 4;
      constexpr int kPer = (kVec + THREADS - 1) / THREADS;
      const uint4* src = reinterpret_cast<const uint4*>(g_fcp_tab);
      uint4 tmp[kPer];
#pragma unroll
      for (int k = 0; k < kPer; ++k) {
        const int i = threadIdx.x + k * THREADS;
        if (i < kVec) tmp[k] = __ldg(src + i);
      }
#pragma unroll
      for (int k = 0; k < kPer; ++k) {
        const int i = threadIdx.x + k * THREADS;
        if (i < kVec) reinterpret_cast<uint4*>(smem)[i] = tmp[k];
      }
    } else {
      for (int i = threadIdx.x; i < 1024 * REPL; i += THREADS)
        sT[i] = g_fcp_tab[kTabT + i / REPL];
      for (int i = threadIdx.x; i < 1024 * RK; i += THREADS)
        sK[i] = g_fcp_tab[kTabK + i / RK];
      for (int i = threadIdx.x; i < 5120; i += THREADS) sC[i] = g_fcp_tab[kTabC + i];
    }
    __syncthreads();
  }
  const int lane = threadIdx.x & 31;
  Tabs tb;
  tb.T = static_cast<uint32_t>(__cvta_generic_to_shared(sT));
  tb.K = static_cast<uint32_t>(__cvta_generic_to_shared(sK));
  tb.C = static_cast<uint32_t>(__cvta_generic_to_shared(sC));
  tb.offT = static_cast<uint32_t>(lane % REPL) * 4u;
  tb.offK = static_cast<uint32_t>(lane % RK) * 4u;
  const uint64_t warps_total =
      static_cast<uint64_t>(gridDim.x) * (THREADS / 32);
  const uint64_t gwarp =
      static_cast<uint64_t>(blockIdx.x) * (THREADS / 32) + (threadIdx.x >> 5);
  const fcp_desc* __restrict__ descs = a.descs;
  const uint32_t n = a.n_desc;
  const uint64_t total = a.total;
  const bool pf = (a.flags & kNoPrefetch) == 0;

  // A warp takes runs of a.run consecutive regions: the descriptor search is
  // paid once per run, inside a run the cursor just walks forward.
  const uint64_t kRun = a.run;
  const uint64_t n_runs = (a.region_end - a.region_begin + kRun - 1) / kRun;
  // Work distribution.  The first run of a warp is its own index; after that
  // a.work == nullptr strides statically, otherwise the warp draws the next
  // run from a device-wide counter (a.work[0]): SMs do not stream at the same
  // rate (ncu: per-SM active cycles differ by up to 25 % under an equal
  // static split), and the drawn order keeps all of them busy to the end.
  // The draw for the next visit is issued before the current run is
  // processed, so its latency is hidden.
  uint32_t* const work = a.work;
  uint64_t run = gwarp;
  while (run < n_runs) {
   uint32_t drawn = 0;
   if (work != nullptr && lane == 0) drawn = atomicAdd(&work[0], 1u);
   // a sliced scatter starts at a rank-specific slice (a.rotate runs in):
   // the ranks of a group then read different peers at any moment instead of
   // all draining the same peer's NVLink egress together
   const uint64_t run_at = !SLICED ? run : (run + a.rotate) % n_runs;
   const uint64_t run_begin = a.region_begin + run_at * kRun;
   const uint64_t run_end =
       run_begin + kRun < a.region_end ? run_begin + kRun : a.region_end;
   Cur c;
   c.idx = warp_first_desc(descs, n, run_begin * kRegion, lane);
   cur_load(c, descs, n);
   for (uint64_t r = run_begin; r < run_end; ++r) {
    const uint64_t rbase = r * kRegion;
    // where this region of the image is read from / written to
    uint8_t* const pk = !SLICED
        ? a.packed
        : reinterpret_cast<uint8_t*>(static_cast<uintptr_t>(__ldg(
              reinterpret_cast<const unsigned long long*>(a.slice_base) +
              r / a.slice_regions)));
    cur_seek(c, descs, n, rbase);  // same answer in every lane
    uint32_t acc = 0;
    const uint64_t lane_o = rbase + static_cast<uint64_t>(lane) * kWordBytes;

    // Whole region inside one tensor (uniform across the warp): no per-word
    // descriptor logic, loads batched for memory-level parallelism.
    const bool whole = (c.off <= rbase) && (rbase + kRegion <= c.end);
    if (whole) {
      const uintptr_t t0 = c.ptr + (rbase - c.off) + lane * kWordBytes;  // tensor side
      uint8_t* p0 = pk + lane_o;                                   // packed side
      const uint32_t mis = static_cast<uint32_t>(t0 & 31u);
      // read side of this region, seen by lane 0 (16-byte aligned down): the
      // batch after the one being loaded is prefetched into L2 by one bulk
      // request, so the 256-bit loads mostly hit L2 instead of waiting on HBM
      const uintptr_t rd0 = (PACK ? c.ptr + (rbase - c.off)
                                  : reinterpret_cast<uintptr_t>(pk + rbase)) & ~uintptr_t(15);
      // where the read side ends: the prefetch of the last batch reaches a few
      // bytes past the region and must not leave the tensor / the image
      const uintptr_t rd_lim = PACK ? c.ptr + (c.end - c.off)
                                    : reinterpret_cast<uintptr_t>(pk) + total;
      if (PACK && mis != 0) {
        // source not 32-byte aligned relative to the image: fetch the two
        // aligned words around each word (the second is the next lane's
        // first: an L1 hit) and realign in registers
        const uintptr_t ta = t0 - mis;
#pragma unroll 1
        for (int row = 0; row < kRows; row += 2) {
          if (lane == 0 && pf && row + 2 < kRows)
            prefetch_l2_upto(rd0 + (row + 2) * kRowBytes, 2 * kRowBytes + 64, rd_lim);
          W8 lo[2], hi[2];
#pragma unroll
          for (int k = 0; k < 2; ++k) {
            lo[k] = ld32(reinterpret_cast<const void*>(ta + (row + k) * kRowBytes));
            hi[k] = ld32(reinterpret_cast<const void*>(
                ta + (row + k) * kRowBytes + kWordBytes));
          }
#pragma unroll
          for (int k = 0; k < 2; ++k) {
            const W8 v = extract32(lo[k], hi[k], mis);
            st32(p0 + (row + k) * kRowBytes, v);
            if (CRC) acc = crc_word<REPL, RK>(tb, acc, v);
          }
        }
      } else if (!PACK && mis != 0) {
        // target not 32-byte aligned relative to the image: every lane writes
        // the ALIGNED target word that starts inside its packed word, built
        // from the tail of its own word and the head of the next packed word
        // (a second aligned load; it is the right neighbour's own word, so an
        // L1 hit — cheaper than passing eight registers through shuffles).
        // Only the few bytes at the very start and end of the region go out
        // as byte stores.
        const uint32_t keep = 32u - mis;        // bytes of my word before the aligned word
        const uintptr_t tw = t0 + keep;         // my aligned target word, row 0
        const uintptr_t src = reinterpret_cast<uintptr_t>(p0);
#pragma unroll 1
        for (int row = 0; row < kRows; row += 2) {
          if (lane == 0 && pf && row + 2 < kRows)
            prefetch_l2_upto(rd0 + (row + 2) * kRowBytes, 2 * kRowBytes + 48, rd_lim);
          W8 v[2], hi[2];
#pragma unroll
          for (int k = 0; k < 2; ++k) {
            v[k] = ld32(reinterpret_cast<const void*>(src + (row + k) * kRowBytes));
            // the word after the region's last one may lie outside the image
            const bool tail = (row + k == kRows - 1) && lane == 31;
            if (!tail)
              hi[k] = ld32(reinterpret_cast<const void*>(
                  src + (row + k) * kRowBytes + kWordBytes));
            else
              hi[k] = v[k];  // unused: that lane writes its tail bytewise
          }
#pragma unroll
          for (int k = 0; k < 2; ++k) {
            if (row + k == 0 && lane == 0)
              store_some_bytes(t0, v[k], 0, static_cast<int>(keep));
            if ((row + k == kRows - 1) && lane == 31) {
              store_some_bytes(tw + (row + k) * kRowBytes, v[k], static_cast<int>(keep),
                               static_cast<int>(mis));
            } else {
              st32(reinterpret_cast<void*>(tw + (row + k) * kRowBytes),
                   extract32(v[k], hi[k], keep));
            }
            if (CRC) acc = crc_word<REPL, RK>(tb, acc, v[k]);
          }
        }
      } else {
        const uintptr_t src = PACK ? t0 : reinterpret_cast<uintptr_t>(p0);
#pragma unroll 1
        for (int row = 0; row < kRows; row += kBatch) {
          if (lane == 0 && pf && row + kBatch < kRows)
            prefetch_l2_upto(rd0 + (row + kBatch) * kRowBytes, kBatch * kRowBytes + 16, rd_lim);
          W8 v[kBatch];
#pragma unroll
          for (int k = 0; k < kBatch; ++k)
            v[k] = ld32(reinterpret_cast<const void*>(src + (row + k) * kRowBytes));
#pragma unroll
          for (int k = 0; k < kBatch; ++k) {
            if (PACK) st32(p0 + (row + k) * kRowBytes, v[k]);
            else st32_unaligned(t0 + (row + k) * kRowBytes, v[k]);
            if (CRC) acc = crc_word<REPL, RK>(tb, acc, v[k]);
          }
        }
      }
    } else {
      // General path: every word decides for itself.  (Keeping several rows
      // in flight here was tried for images of many small tensors: no gain
      // with a warm L2 — such images are bound by the two dependent launches,
      // ~24 us on the GPU, and by the host's ~30 us to enqueue a call.)
#pragma unroll 1
      for (int row = 0; row < kRows; ++row) {
        const uint64_t o = lane_o + static_cast<uint64_t>(row) * kRowBytes;
        W8 v;
#pragma unroll
        for (int i = 0; i < 8; ++i) v.v[i] = 0u;
        if (o < total) {
          cur_seek(c, descs, n, o);
          const bool interior = (c.off <= o) && (o + kWordBytes <= c.end);
          if (PACK) {
            if (interior) {
              const uintptr_t t = c.ptr + (o - c.off);
              const uint32_t mis = static_cast<uint32_t>(t & 31u);
              if (mis == 0) {
                v = ld32_stream(reinterpret_cast<const void*>(t));
              } else {
                const W8 lo = ld32(reinterpret_cast<const void*>(t - mis));
                const W8 hi = ld32(reinterpret_cast<const void*>(t - mis + kWordBytes));
                v = extract32(lo, hi, mis);
              }
              st32(pk + o, v);
            } else {
              v = gather_word(c, descs, n, o, total);
              store_packed(pk, o, total, v);
            }
          } else {
            v = load_packed(pk, o, total);
            if (interior)
              st32_unaligned(c.ptr + (o - c.off), v);
            else
              scatter_word(c, descs, n, o, total, v);
          }
        }
        if (CRC) acc = crc_word<REPL, RK>(tb, acc, v);
      }
      __syncwarp();
    }
    if (CRC) {
      acc = warp_combine(tb, acc, lane);
      if (lane == 31) a.region_sums[r] = acc;
    }
   }
   run = work != nullptr ? warps_total + __shfl_sync(0xffffffffu, drawn, 0)
                         : run + warps_total;
  }
  // The last warp to leave resets the counters, so launches that follow on
  // the same stream find them at zero (a.work[1] counts departed warps; a
  // warp increments it only after its final draw).
  if (work != nullptr) {
    bool last = false;
    if (THREADS < 1024 && CRC && (a.flags & kFusedFinalize)) {
      __threadfence();  // this warp's region values are visible before it signs off
      __syncwarp();
    }
    if (lane == 0) {
      const uint32_t gone = atomicAdd(&work[1], 1u);
      last = gone == static_cast<uint32_t>(warps_total) - 1u;
    }
    if (THREADS < 1024 && CRC && (a.flags & kFusedFinalize)) {
      // small image: the last warp out combines the chunks itself instead of
      // a second launch (every other warp has fenced its values and left)
      last = __shfl_sync(0xffffffffu, last, 0);
      if (last) {
        __threadfence();
        const uint64_t n_real = (total + kRegion - 1) / kRegion;
        const uint64_t n_padded =
            (total + FCP_REGION_BYTES - 1) / FCP_REGION_BYTES * (FCP_REGION_BYTES / kRegion);
        for (uint64_t c = a.chunk_begin; c < a.chunk_end; ++c)
          combine_chunk(a.region_sums, a.chunk_sums, n_real, n_padded, a.regions_per_chunk, c,
                        lane, ROWS == 16 ? kCoarseUnitLog : 0);
      }
    }
    if (lane == 0 && last) {
      work[0] = 0u;
      work[1] = 0u;
    }
  }
}

// Plain streaming copy used by FCP_MODE_STAGED_KERNEL to move the staged
// image into host-mapped memory with SM stores instead of the copy engine
// (a few CTAs are enough to fill the host link, so it can run beside the next
// training step).  Both pointers 32-byte aligned; n_words = bytes / 32.
__global__ void __launch_bounds__(256) fcp_copy_kernel(const uint8_t* __restrict__ src,
                                                       uint8_t* __restrict__ dst,
                                                       uint64_t n_words, uint64_t tail_bytes) {
  const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
  uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n_words; i += 4 * stride) {
    W8 v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = ld32_stream(src + (i + k * stride) * kWordBytes);
#pragma unroll
    for (int k = 0; k < 4; ++k) st32(dst + (i + k * stride) * kWordBytes, v[k]);
  }
  for (; i < n_words; i += stride) st32(dst + i * kWordBytes, ld32_stream(src + i * kWordBytes));
  if (blockIdx.x == 0 && threadIdx.x < tail_bytes)
    dst[n_words * kWordBytes + threadIdx.x] = src[n_words * kWordBytes + threadIdx.x];
}

// ------------------------------------------------------------------ XXH64
// Standard XXH64 (seed 0) of every chunk of a device-resident packed image.
// The algorithm is four serial lanes per stream, so a chunk is owned by four
// adjacent threads (one XXH lane each, 8 bytes of every 32-byte stripe); all
// chunks run concurrently.  Not fused into the pack kernel: XXH64 is not
// linear, partial results of different threads cannot be combined.
constexpr unsigned long long kX1 = 0x9E3779B185EBCA87ULL, kX2 = 0xC2B2AE3D27D4EB4FULL,
                             kX3 = 0x165667B19E3779F9ULL, kX4 = 0x85EBCA77C2B2AE63ULL,
                             kX5 = 0x27D4EB2F165667C5ULL;
__device__ __forceinline__ unsigned long long xrotl(unsigned long long x, int r) {
  return (x << r) | (x >> (64 - r));
}
__device__ __forceinline__ unsigned long long xround(unsigned long long acc,
                                                     unsigned long long in) {
  return xrotl(acc + in * kX2, 31) * kX1;
}
__device__ __forceinline__ unsigned long long xmerge(unsigned long long h,
                                                     unsigned long long v) {
  return (h ^ xround(0, v)) * kX1 + kX4;
}

__global__ void __launch_bounds__(128) fcp_xxh64_kernel(const uint8_t* __restrict__ packed,
                                                       uint64_t total, uint32_t chunk_bytes,
                                                       uint64_t chunk_begin, uint64_t chunk_end,
                                                       unsigned long long* __restrict__ out) {
  const uint64_t t = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const uint64_t chunk = chunk_begin + (t >> 2);
  const int xl = static_cast<int>(t & 3);
  const bool live = chunk < chunk_end;
  const uint64_t cs = live ? chunk * chunk_bytes : 0;
  uint64_t len = 0;
  if (live) len = (cs + chunk_bytes <= total) ? chunk_bytes : total - cs;
  const uint64_t n_stripes = len / 32;
  const uint8_t* p = packed + cs + xl * 8;
  unsigned long long v = xl == 0 ? kX1 + kX2 : xl == 1 ? kX2 : xl == 2 ? 0ULL : 0ULL - kX1;
  uint64_t s = 0;
  for (; s + 8 <= n_stripes; s += 8) {
    unsigned long long w[8];
#pragma unroll
    for (int k = 0; k < 8; ++k)
      w[k] = __ldg(reinterpret_cast<const unsigned long long*>(p + (s + k) * 32));
#pragma unroll
    for (int k = 0; k < 8; ++k) v = xround(v, w[k]);
  }
  for (; s < n_stripes; ++s)
    v = xround(v, __ldg(reinterpret_cast<const unsigned long long*>(p + s * 32)));
  // the four lanes of a chunk sit in lanes 4g..4g+3 of one warp
  const int base = (threadIdx.x & 31) & ~3;
  const unsigned long long v1 = __shfl_sync(0xffffffffu, v, base);
  const unsigned long long v2 = __shfl_sync(0xffffffffu, v, base + 1);
  const unsigned long long v3 = __shfl_sync(0xffffffffu, v, base + 2);
  const unsigned long long v4 = __shfl_sync(0xffffffffu, v, base + 3);
  if (!live || xl != 0) return;
  unsigned long long h;
  if (len >= 32) {
    h = xrotl(v1, 1) + xrotl(v2, 7) + xrotl(v3, 12) + xrotl(v4, 18);
    h = xmerge(xmerge(xmerge(xmerge(h, v1), v2), v3), v4);
  } else {
    h = kX5;
  }
  h += len;
  const uint8_t* q = packed + cs + n_stripes * 32;
  const uint8_t* end = packed + cs + len;
  for (; q + 8 <= end; q += 8)
    h = xrotl(h ^ xround(0, *reinterpret_cast<const unsigned long long*>(q)), 27) * kX1 + kX4;
  if (q + 4 <= end) {
    h = xrotl(h ^ (static_cast<unsigned long long>(*reinterpret_cast<const uint32_t*>(q)) * kX1),
              23) * kX2 + kX3;
    q += 4;
  }
  for (; q < end; ++q) h = xrotl(h ^ (static_cast<unsigned long long>(*q) * kX5), 11) * kX1;
  h ^= h >> 33;
  h *= kX2;
  h ^= h >> 29;
  h *= kX3;
  h ^= h >> 32;
  out[chunk] = h;
}

// One warp per chunk (it was one thread walking 128 dependent table lookups:
// 17 us, the floor of every small image).
__global__ void __launch_bounds__(128) fcp_finalize_kernel(
    const uint32_t* __restrict__ region_sums, uint32_t* __restrict__ chunk_sums,
    uint64_t n_real, uint64_t n_padded, uint32_t regions_per_chunk, uint64_t chunk_begin,
    uint64_t chunk_end, int unit_log) {
  const uint64_t c = chunk_begin + blockIdx.x * static_cast<uint64_t>(blockDim.x >> 5) +
                     (threadIdx.x >> 5);
  if (c >= chunk_end) return;  // uniform across the warp
  combine_chunk(region_sums, chunk_sums, n_real, n_padded, regions_per_chunk, c,
                threadIdx.x & 31, unit_log);
}

}  // namespace fcpk
