// fcp_host.cpp — host-only entry points of libfcp.so (no CUDA call in here):
// standard CRC-32C of the fixed windows of a packed image that lives in host
// memory.  The agent-side saver uses it to re-check the per-chunk sums the
// pack kernel recorded before it persists the shm image; raw image files get
// their per-file sums from it.
//
// x86-64 with SSE4.2: the crc32 instruction, three independent streams per
// window (latency 3, throughput 1) recombined with the GF(2) shift operators
// of fcp_crc.h.  Anything else: slicing-by-8 tables.
#include <errno.h>
#include <stdint.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <thread>
#include <vector>

#include "../../include/fcp.h"
#include "fcp_crc.h"

#if defined(__x86_64__)
#include <nmmintrin.h>
#define FCP_HAVE_HW_CRC 1
#endif

namespace {

// ---- portable: slicing-by-8 ------------------------------------------------
uint32_t g_slice[8][256];
std::once_flag g_slice_once;

void slice_init() {
  for (uint32_t b = 0; b < 256; ++b) {
    uint32_t c = b;
    for (int k = 0; k < 8; ++k) c = (c >> 1) ^ (fcpcrc::kPoly & (0u - (c & 1u)));
    g_slice[0][b] = c;
  }
  for (uint32_t b = 0; b < 256; ++b)
    for (int j = 1; j < 8; ++j)
      g_slice[j][b] = (g_slice[j - 1][b] >> 8) ^ g_slice[0][g_slice[j - 1][b] & 0xffu];
}

// raw update (no init / final xor) of state c over n bytes
uint32_t update_soft(uint32_t c, const uint8_t* p, uint64_t n) {
  std::call_once(g_slice_once, slice_init);
  while (n && (reinterpret_cast<uintptr_t>(p) & 7u)) {
    c = g_slice[0][(c ^ *p++) & 0xffu] ^ (c >> 8);
    --n;
  }
  for (; n >= 8; n -= 8, p += 8) {
    uint64_t w;
    memcpy(&w, p, 8);
    w ^= c;
    c = g_slice[7][w & 0xff] ^ g_slice[6][(w >> 8) & 0xff] ^ g_slice[5][(w >> 16) & 0xff] ^
        g_slice[4][(w >> 24) & 0xff] ^ g_slice[3][(w >> 32) & 0xff] ^
        g_slice[2][(w >> 40) & 0xff] ^ g_slice[1][(w >> 48) & 0xff] ^ g_slice[0][w >> 56];
  }
  while (n--) c = g_slice[0][(c ^ *p++) & 0xffu] ^ (c >> 8);
  return c;
}

#ifdef FCP_HAVE_HW_CRC
constexpr uint64_t kLane = 4096;  // bytes per stream of one 3-stream block
fcpcrc::Mat g_shift1, g_shift2;   // shift by kLane, 2*kLane bytes
std::once_flag g_hw_once;

__attribute__((target("sse4.2"))) uint32_t update_hw(uint32_t c, const uint8_t* p,
                                                     uint64_t n) {
  std::call_once(g_hw_once, [] {
    g_shift1 = fcpcrc::shift_bytes(kLane);
    g_shift2 = fcpcrc::shift_bytes(2 * kLane);
  });
  uint64_t s = c;
  while (n && (reinterpret_cast<uintptr_t>(p) & 7u)) {
    s = _mm_crc32_u8(static_cast<uint32_t>(s), *p++);
    --n;
  }
  while (n >= 3 * kLane) {
    uint64_t a = s, b = 0, d = 0;
    const uint64_t* q = reinterpret_cast<const uint64_t*>(p);
    for (uint64_t i = 0; i < kLane / 8; ++i) {
      a = _mm_crc32_u64(a, q[i]);
      b = _mm_crc32_u64(b, q[i + kLane / 8]);
      d = _mm_crc32_u64(d, q[i + 2 * (kLane / 8)]);
    }
    s = fcpcrc::apply(g_shift2, static_cast<uint32_t>(a)) ^
        fcpcrc::apply(g_shift1, static_cast<uint32_t>(b)) ^ static_cast<uint32_t>(d);
    p += 3 * kLane;
    n -= 3 * kLane;
  }
  for (; n >= 8; n -= 8, p += 8) {
    uint64_t w;
    memcpy(&w, p, 8);
    s = _mm_crc32_u64(s, w);
  }
  while (n--) s = _mm_crc32_u8(static_cast<uint32_t>(s), *p++);
  return static_cast<uint32_t>(s);
}
#endif

uint32_t crc32c_std(const uint8_t* p, uint64_t n) {
#ifdef FCP_HAVE_HW_CRC
  static const bool hw = __builtin_cpu_supports("sse4.2");
  if (hw) return update_hw(0xFFFFFFFFu, p, n) ^ 0xFFFFFFFFu;
#endif
  return update_soft(0xFFFFFFFFu, p, n) ^ 0xFFFFFFFFu;
}

}  // namespace

extern "C" int fcp_host_chunk_crc32c(const void* image, uint64_t total_bytes,
                                     uint32_t chunk_bytes, uint32_t* out,
                                     int n_threads) {
  if (chunk_bytes == 0) return -EINVAL;
  if (total_bytes == 0) return 0;
  if (!image || !out) return -EINVAL;
  const uint64_t n_chunks = (total_bytes + chunk_bytes - 1) / chunk_bytes;
  if (n_threads <= 0) {
    const unsigned hc = std::thread::hardware_concurrency();
    n_threads = hc >= 8 ? 4 : hc >= 2 ? 2 : 1;
  }
  if (static_cast<uint64_t>(n_threads) > n_chunks) n_threads = static_cast<int>(n_chunks);
  const uint8_t* base = static_cast<const uint8_t*>(image);
  std::atomic<uint64_t> next{0};
  auto work = [&]() {
    for (;;) {
      const uint64_t c = next.fetch_add(1);
      if (c >= n_chunks) return;
      const uint64_t b = c * chunk_bytes;
      const uint64_t n = total_bytes - b < chunk_bytes ? total_bytes - b : chunk_bytes;
      out[c] = crc32c_std(base + b, n);
    }
  };
  if (n_threads <= 1) {
    work();
    return 0;
  }
  std::vector<std::thread> pool;
  for (int t = 0; t < n_threads; ++t) pool.emplace_back(work);
  for (auto& t : pool) t.join();
  return 0;
}
