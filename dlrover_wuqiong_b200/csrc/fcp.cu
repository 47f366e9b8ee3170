// fcp.cu — host side of libfcp.so: C ABI declared in include/fcp.h.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -shared
//        -Xcompiler -fPIC  (see build.py next to this file)
#include <errno.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <mutex>
#include <thread>
#include <utility>
#include <vector>

#include "fcp_crc.h"
#include "fcp_kernels2.cuh"

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define FCP_CUDA(expr)                                                    \
  do {                                                                    \
    cudaError_t e_ = (expr);                                              \
    if (e_ != cudaSuccess)                                                \
      return fail(-EIO, "%s failed: %s (%s:%d)", #expr,                   \
                  cudaGetErrorString(e_), __FILE__, __LINE__);            \
  } while (0)

constexpr int kMaxDevices = 64;
std::mutex g_init_mu;
bool g_tab_ready[kMaxDevices] = {false};
int g_sm_count[kMaxDevices] = {0};
bool g_attr_done[kMaxDevices] = {false};

inline uint64_t ceil_div(uint64_t a, uint64_t b) { return (a + b - 1) / b; }

// RAII device switch
struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) ok = false;
    if (ok && prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

int current_device(int* dev) {
  FCP_CUDA(cudaGetDevice(dev));
  if (*dev < 0 || *dev >= kMaxDevices)
    return fail(-EINVAL, "device index %d out of range", *dev);
  return 0;
}

int init_device(int dev) {
  std::lock_guard<std::mutex> lk(g_init_mu);
  if (dev < 0 || dev >= kMaxDevices)
    return fail(-EINVAL, "device index %d out of range", dev);
  if (g_tab_ready[dev]) return 0;
  DeviceGuard g(dev);
  if (!g.ok) return fail(-EIO, "cannot select device %d", dev);
  std::vector<uint32_t> tab(fcpk::kTabWords);
  fcpcrc::make_table(fcpcrc::shift_bytes(4), &tab[fcpk::kTabT]);
  fcpcrc::make_table(fcpcrc::shift_bytes(fcpk::kRowBytes - fcpk::kWordBytes),
                     &tab[fcpk::kTabK]);
  for (int k = 0; k < 5; ++k)
    fcpcrc::make_table(fcpcrc::shift_bytes(static_cast<uint64_t>(fcpk::kWordBytes) << k),
                       &tab[fcpk::kTabC + k * 1024]);
  fcpcrc::make_table(fcpcrc::shift_bytes(fcpk::kRegion), &tab[fcpk::kTabF]);
  FCP_CUDA(cudaMemcpyToSymbol(fcpk::g_fcp_tab, tab.data(),
                              sizeof(uint32_t) * fcpk::kTabWords));
  std::vector<uint32_t> reg(fcpk::kRegionShiftTabs * 1024);
  for (int j = 0; j < fcpk::kRegionShiftTabs; ++j)
    fcpcrc::make_table(fcpcrc::shift_bytes(static_cast<uint64_t>(fcpk::kUnitBytes) << j),
                       &reg[j * 1024]);
  FCP_CUDA(cudaMemcpyToSymbol(fcpk::g_fcp_tab_regions, reg.data(),
                              sizeof(uint32_t) * reg.size()));
  int sms = 0;
  FCP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  g_sm_count[dev] = sms;
  g_tab_ready[dev] = true;
  return 0;
}

// ---- kernel launch -------------------------------------------------------
constexpr int kBigThreads = 1024;
constexpr int kSmallThreads = 256;
constexpr size_t kSmemBig = sizeof(uint32_t) * (1024 * 32 + 1024 * 16 + 5120);
constexpr size_t kSmemSmall = sizeof(uint32_t) * (1024 + 1024 + 5120);

template <bool PACK, bool CRC, bool SLICED = false>
int launch_dir(const fcpk::Args& a_in, int dev, cudaStream_t st, uint64_t first_region = 0,
               bool* fused = nullptr) {
  fcpk::Args a = a_in;
  const uint64_t regions = a.region_end - a.region_begin;
  if (regions == 0) return 0;
  const int sms = g_sm_count[dev] > 0 ? g_sm_count[dev] : 148;
  // longest run that still leaves every warp of a full grid >= 4 visits
  uint64_t run = regions / (static_cast<uint64_t>(sms) * 32 * 4);
  a.run = static_cast<uint32_t>(run < 1 ? 1 : run > fcpk::kMaxRun ? fcpk::kMaxRun : run);
  a.rotate = (first_region - a.region_begin) / a.run;  // in runs (sliced scatter only)
  static const bool no_pf = getenv("FCP_DEBUG_NO_PREFETCH") != nullptr;  // tuning aid
  if (no_pf) a.flags |= fcpk::kNoPrefetch;
  // Big shape: one 1024-thread CTA per SM with the conflict-free (32x
  // replicated) update table and a 16x replicated row-shift table (212 KiB of
  // shared memory).  Small shape for images too small to give every warp of
  // the big grid a region: 256-thread CTAs, unreplicated tables.
  static const bool force_big = getenv("FCP_DEBUG_FORCE_BIG") != nullptr;  // tuning aid
  const uint64_t runs = ceil_div(regions, a.run);
  const bool fine = !SLICED && (a.flags & fcpk::kFineUnits) != 0;  // 4 KiB units: small shape
  const bool big = !fine && (CRC || force_big) && runs >= static_cast<uint64_t>(sms) * 32;
  // the last-warp-out finalize exists in the 256-thread shape only, needs the
  // plan's counters and a checksum
  if (big || !CRC || a.work == nullptr) a.flags &= ~fcpk::kFusedFinalize;
  if (fused) *fused = (a.flags & fcpk::kFusedFinalize) != 0;
  if (big) {
    auto kern = fcpk::fcp_kernel<PACK, CRC, 32, kBigThreads, SLICED>;
    if (!g_attr_done[dev]) {
      FCP_CUDA(cudaFuncSetAttribute(fcpk::fcp_kernel<false, true, 32, kBigThreads, true>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)kSmemBig));
      FCP_CUDA(cudaFuncSetAttribute(fcpk::fcp_kernel<false, false, 32, kBigThreads, true>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)kSmemBig));
      FCP_CUDA(cudaFuncSetAttribute(fcpk::fcp_kernel<true, true, 32, kBigThreads>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)kSmemBig));
      FCP_CUDA(cudaFuncSetAttribute(fcpk::fcp_kernel<false, true, 32, kBigThreads>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)kSmemBig));
      FCP_CUDA(cudaFuncSetAttribute(fcpk::fcp_kernel<true, false, 32, kBigThreads>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)kSmemBig));
      FCP_CUDA(cudaFuncSetAttribute(fcpk::fcp_kernel<false, false, 32, kBigThreads>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)kSmemBig));
      g_attr_done[dev] = true;
    }
    kern<<<sms, kBigThreads, kSmemBig, st>>>(a);
  } else {
    const uint64_t want = ceil_div(runs, kSmallThreads / 32);
    static const int per_sm = getenv("FCP_DEBUG_SMALL_CTAS_PER_SM")
                                  ? atoi(getenv("FCP_DEBUG_SMALL_CTAS_PER_SM")) : 8;  // tuning aid
    const uint64_t cap = static_cast<uint64_t>(sms) * (per_sm > 0 ? per_sm : 8);
    const unsigned grid = static_cast<unsigned>(want < cap ? want : cap);
    if (fine) {
      auto kern = fcpk::fcp_kernel<PACK, CRC, 1, kSmallThreads, false, fcpk::kFineRows>;
      kern<<<grid, kSmallThreads, CRC ? kSmemSmall : 0, st>>>(a);
    } else {
      auto kern = fcpk::fcp_kernel<PACK, CRC, 1, kSmallThreads, SLICED>;
      kern<<<grid, kSmallThreads, CRC ? kSmemSmall : 0, st>>>(a);
    }
  }
  FCP_CUDA(cudaGetLastError());
  return 0;
}

int launch_any(const fcpk::Args& a, int direction, int checksum, int dev,
               cudaStream_t st, uint64_t first_region = 0, bool* fused = nullptr) {
  if (fused) *fused = false;
  if (a.slice_base != nullptr) {  // scatter only
    if (direction != 1) return fail(-EINVAL, "a sliced image can only be scattered");
    return checksum ? launch_dir<false, true, true>(a, dev, st, first_region, fused)
                    : launch_dir<false, false, true>(a, dev, st, first_region, fused);
  }
  if (direction == 0)
    return checksum ? launch_dir<true, true>(a, dev, st, 0, fused)
                    : launch_dir<true, false>(a, dev, st, 0, fused);
  return checksum ? launch_dir<false, true>(a, dev, st, 0, fused)
                  : launch_dir<false, false>(a, dev, st, 0, fused);
}

// `n_regions` / `regions_per_chunk` count 16 KiB regions; with fine = true the
// values in region_sums are those of 4 KiB units of an image of `total` bytes
int launch_finalize(const uint32_t* region_sums, uint32_t* chunk_sums,
                    uint64_t n_regions, uint32_t regions_per_chunk,
                    uint64_t chunk_begin, uint64_t chunk_end, cudaStream_t st,
                    bool fine = false, uint64_t total = 0) {
  const uint64_t nc = chunk_end - chunk_begin;
  if (nc == 0) return 0;
  const unsigned threads = 128;  // one warp per chunk
  const unsigned grid = static_cast<unsigned>(ceil_div(nc, threads / 32));
  if (fine) {
    const uint64_t per = FCP_REGION_BYTES / fcpk::kUnitBytes;
    fcpk::fcp_finalize_kernel<<<grid, threads, 0, st>>>(
        region_sums, chunk_sums, ceil_div(total, fcpk::kUnitBytes), n_regions * per,
        static_cast<uint32_t>(regions_per_chunk * per), chunk_begin, chunk_end, 0);
  } else {
    fcpk::fcp_finalize_kernel<<<grid, threads, 0, st>>>(
        region_sums, chunk_sums, n_regions, n_regions, regions_per_chunk, chunk_begin,
        chunk_end, fcpk::kCoarseUnitLog);
  }
  FCP_CUDA(cudaGetLastError());
  return 0;
}

int check_geometry(uint64_t total, uint32_t chunk_bytes) {
  if (chunk_bytes == 0 || chunk_bytes % FCP_REGION_BYTES != 0)
    return fail(-EINVAL, "chunk_bytes %u is not a multiple of %u", chunk_bytes,
                FCP_REGION_BYTES);
  (void)total;
  return 0;
}

}  // namespace

// ===================================================================== plan
struct fcp_plan {
  int device = 0;
  uint64_t total = 0;
  uint32_t chunk_bytes = 0;
  uint64_t n_desc = 0;
  uint64_t n_regions = 0;
  uint64_t n_chunks = 0;
  uint64_t slice_bytes = 1ull << 30;  // smallest slice; they grow x2 up to x4
  uint64_t last_launches = 0;
  uint64_t copy_piece = 0;  // split host copies at multiples of this (0 = never)
  uint32_t kernel_flags = 0;
  std::vector<fcp_desc> h_descs;
  fcp_desc* h_pinned = nullptr;  // pinned mirror of h_descs: source of the uploads
  bool descs_dirty = false;      // h_pinned is newer than d_descs
  bool upload_pending = false;   // ev_upload has been recorded
  cudaEvent_t ev_upload = nullptr;
  fcp_desc* d_descs = nullptr;
  uint32_t* d_region_sums = nullptr;
  uint32_t* d_work = nullptr;  // run-distribution counters of the kernels (Args.work)
  uint32_t* d_chunk_sums = nullptr;
  uint32_t* h_chunk_sums = nullptr;  // pinned
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_begin = nullptr, ev_packed = nullptr, ev_done = nullptr;
  cudaEvent_t ev_kern_b = nullptr, ev_kern_e = nullptr;
  std::vector<cudaEvent_t> ev_slice;  // one per slice, created on demand
  float t_kernel_ms = 0.f, t_copy_ms = 0.f, t_total_ms = 0.f;
  bool last_has_sums = false;
  bool save_in_flight = false;
  int save_mode = 0;
  bool fine = false;  // small image: kernels work in 4 KiB units (4x the warps)
  uint64_t win_begin = 0, win_end = ~0ull;  // chunk window of fcp_save_async
  // staged save whose device->host copies are enqueued apart from the pack
  bool defer_copies = false, copies_pending = false;
  void* sv_staging = nullptr;
  void* sv_host = nullptr;
  uint8_t* sv_alias = nullptr;
  int sv_checksum = 0;
  std::vector<std::pair<uint64_t, uint64_t>> sv_slices;
  const uint8_t** d_slice_bases = nullptr;  // table of fcp_restore_from_slices
  uint64_t slice_bases_cap = 0;
  // pinned bounce ring of fcp_restore_bounced: 2 slots per worker
  uint8_t* ring = nullptr;
  uint64_t ring_slot_bytes = 0;
  int ring_workers = 0;
};

namespace {

// Bring the device copy of the descriptor table up to date ON THE STREAM the
// kernels are about to be launched on (from pinned memory, so the copy is an
// ordinary stream operation): ordered before those kernels whatever kind of
// stream it is.  A synchronous pageable cudaMemcpy would only be ordered
// against the legacy default stream and blocking streams.
int plan_sync_descs(fcp_plan* p, cudaStream_t st) {
  if (!p->descs_dirty || p->n_desc == 0) return 0;
  FCP_CUDA(cudaMemcpyAsync(p->d_descs, p->h_pinned, sizeof(fcp_desc) * p->n_desc,
                           cudaMemcpyHostToDevice, st));
  FCP_CUDA(cudaEventRecord(p->ev_upload, st));
  p->upload_pending = true;
  p->descs_dirty = false;
  return 0;
}

int plan_slice_events(fcp_plan* p, size_t n) {
  while (p->ev_slice.size() < n) {
    cudaEvent_t e;
    FCP_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    p->ev_slice.push_back(e);
  }
  return 0;
}

fcpk::Args plan_args(const fcp_plan* p, void* packed, uint64_t rb, uint64_t re) {
  fcpk::Args a;
  a.descs = p->d_descs;
  a.n_desc = static_cast<uint32_t>(p->n_desc);
  a.total = p->total;
  a.region_begin = rb;
  a.region_end = re;
  a.packed = static_cast<uint8_t*>(packed);
  a.region_sums = p->d_region_sums;
  a.run = 1;
  a.flags = p->kernel_flags;
  a.work = std::getenv("FCP_DEBUG_STATIC_RUNS") ? nullptr : p->d_work;
  return a;
}

// chunk values of chunks [cb, ce) from the values the plan's launches left
int plan_finalize(fcp_plan* p, uint64_t cb, uint64_t ce, cudaStream_t st) {
  return launch_finalize(p->d_region_sums, p->d_chunk_sums, p->n_regions,
                         p->chunk_bytes / FCP_REGION_BYTES, cb, ce, st, p->fine, p->total);
}

// kernels for chunks [cb, ce) of the plan, on stream st; `finalize` = also
// combine the region values of those chunks into chunk values
int plan_launch(fcp_plan* p, void* packed, int direction, int checksum,
                uint64_t cb, uint64_t ce, cudaStream_t st, bool finalize = true) {
  const uint32_t rpc = p->chunk_bytes / FCP_REGION_BYTES;
  uint64_t rb = cb * rpc, re = ce * rpc;
  if (re > p->n_regions) re = p->n_regions;
  bool fused = false;
  if (re > rb) {
    fcpk::Args a = plan_args(p, packed, rb, re);
    uint32_t per_chunk = rpc;
    if (p->fine) {  // the same range in 4 KiB units
      const uint64_t per = FCP_REGION_BYTES / fcpk::kUnitBytes;
      const uint64_t n_units = ceil_div(p->total, fcpk::kUnitBytes);
      a.flags |= fcpk::kFineUnits;
      a.region_begin = rb * per;
      a.region_end = re * per < n_units ? re * per : n_units;
      per_chunk = static_cast<uint32_t>(rpc * per);
    }
    if (checksum && finalize && ce - cb <= fcpk::kFuseMaxChunks) {
      a.flags |= fcpk::kFusedFinalize;  // (kept only for the small shape, see launch_dir)
      a.chunk_sums = p->d_chunk_sums;
      a.regions_per_chunk = per_chunk;
      a.chunk_begin = cb;
      a.chunk_end = ce < p->n_chunks ? ce : p->n_chunks;
    }
    int rc = launch_any(a, direction, checksum, p->device, st, 0, &fused);
    if (rc) return rc;
    p->last_launches++;
  }
  if (checksum && finalize && ce > cb && !fused) {
    int rc = plan_finalize(p, cb, ce, st);
    if (rc) return rc;
    p->last_launches++;
  }
  return 0;
}

// Host<->device copy of image bytes [b0, b1).  When the host range was pinned
// piece by piece (fcp_plan_set_copy_piece_bytes) no copy may straddle a piece
// boundary: CUDA rejects a transfer that spans two registrations.
int copy_image_range(fcp_plan* p, uint8_t* dev, uint8_t* host, uint64_t b0,
                     uint64_t b1, bool to_host, cudaStream_t st) {
  while (b0 < b1) {
    uint64_t e = b1;
    if (p->copy_piece) {
      const uint64_t next = (b0 / p->copy_piece + 1) * p->copy_piece;
      if (next < e) e = next;
    }
    if (to_host)
      FCP_CUDA(cudaMemcpyAsync(host + b0, dev + b0, e - b0, cudaMemcpyDeviceToHost, st));
    else
      FCP_CUDA(cudaMemcpyAsync(dev + b0, host + b0, e - b0, cudaMemcpyHostToDevice, st));
    b0 = e;
  }
  return 0;
}

// Chunk ranges of the staged pipelines.  Sizes grow 4x per slice from
// slice_bytes up to 16x (1, 4, 16, 16 ... GiB by default): a save starts its
// first DMA after a small pack, and since the link is ~100x slower than the
// kernels the rest can go in few large, more efficient launches (measured on
// the 19 GB shard: pack roofline 0.88 with 1-2-4-4.. GiB slices, 0.93 with
// 8 GiB ones); a restore runs the schedule backwards so only a small scatter
// trails the last DMA.
std::vector<std::pair<uint64_t, uint64_t>> make_slices(const fcp_plan* p,
                                                       bool growing,
                                                       uint64_t first = 0,
                                                       uint64_t last = ~0ull) {
  std::vector<uint64_t> sizes;
  const uint64_t cps0 = p->slice_bytes / p->chunk_bytes;
  if (last > p->n_chunks) last = p->n_chunks;
  if (first > last) first = last;
  uint64_t left = last - first, cur = cps0 ? cps0 : 1;
  while (left) {
    const uint64_t take = cur < left ? cur : left;
    sizes.push_back(take);
    left -= take;
    if (cur < 16 * cps0) cur *= 4;
  }
  if (!growing) std::reverse(sizes.begin(), sizes.end());
  std::vector<std::pair<uint64_t, uint64_t>> out;
  uint64_t at = first;
  for (uint64_t sz : sizes) {
    out.emplace_back(at, at + sz);
    at += sz;
  }
  return out;
}

int finish_sums(const uint32_t* raw, uint64_t n_chunks, uint64_t total,
                uint32_t chunk_bytes, uint32_t* out) {
  if (n_chunks == 0) return 0;
  // full chunks: std = raw ^ shift_len(~0) ^ ~0
  const uint32_t full_c =
      fcpcrc::apply(fcpcrc::shift_bytes(chunk_bytes), 0xFFFFFFFFu) ^ 0xFFFFFFFFu;
  for (uint64_t c = 0; c + 1 < n_chunks; ++c) out[c] = raw[c] ^ full_c;
  // last chunk: the device value is the raw CRC of the data followed by zero
  // padding up to a whole region; strip the padding first.
  const uint64_t last_len = total - (n_chunks - 1) * chunk_bytes;
  const uint64_t padded = ceil_div(last_len, FCP_REGION_BYTES) * FCP_REGION_BYTES;
  uint32_t v = raw[n_chunks - 1];
  if (padded != last_len)
    v = fcpcrc::apply(fcpcrc::unshift_bytes(padded - last_len), v);
  out[n_chunks - 1] =
      v ^ fcpcrc::apply(fcpcrc::shift_bytes(last_len), 0xFFFFFFFFu) ^ 0xFFFFFFFFu;
  return 0;
}

}  // namespace

extern "C" {

int fcp_version(void) { return FCP_VERSION; }
const char* fcp_last_error(void) { return g_err; }
int fcp_init(int device) { return init_device(device); }

int fcp_plan_create(const fcp_desc* descs, uint64_t n, uint64_t total_bytes,
                    uint32_t chunk_bytes, int device, fcp_plan** out) {
  if (!out) return fail(-EINVAL, "out is NULL");
  *out = nullptr;
  int rc = check_geometry(total_bytes, chunk_bytes);
  if (rc) return rc;
  if (n > 0xFFFFFFF0ull) return fail(-EINVAL, "too many descriptors");
  uint64_t prev_end = 0;
  for (uint64_t i = 0; i < n; ++i) {
    if (descs[i].nbytes == 0)
      return fail(-EINVAL, "descriptor %llu is empty", (unsigned long long)i);
    if (descs[i].dst_off < prev_end)
      return fail(-EINVAL, "descriptor %llu overlaps or is out of order",
                  (unsigned long long)i);
    prev_end = descs[i].dst_off + descs[i].nbytes;
    if (prev_end > total_bytes)
      return fail(-EINVAL, "descriptor %llu ends beyond total_bytes",
                  (unsigned long long)i);
  }
  rc = init_device(device);
  if (rc) return rc;
  DeviceGuard g(device);
  if (!g.ok) return fail(-EIO, "cannot select device %d", device);
  fcp_plan* p = new fcp_plan();
  p->device = device;
  p->total = total_bytes;
  p->chunk_bytes = chunk_bytes;
  p->n_desc = n;
  p->n_regions = ceil_div(total_bytes, FCP_REGION_BYTES);
  p->n_chunks = ceil_div(total_bytes, chunk_bytes);
  p->h_descs.assign(descs, descs + n);
  auto bail = [&](int code) {
    fcp_plan_destroy(p);
    return code;
  };
#define FCP_TRY(expr)                                                       \
  do {                                                                      \
    cudaError_t e_ = (expr);                                                \
    if (e_ != cudaSuccess)                                                  \
      return bail(fail(-ENOMEM, "%s failed: %s", #expr, cudaGetErrorString(e_))); \
  } while (0)
  FCP_TRY(cudaMalloc(&p->d_descs, sizeof(fcp_desc) * (n ? n : 1)));
  FCP_TRY(cudaMallocHost(&p->h_pinned, sizeof(fcp_desc) * (n ? n : 1)));
  FCP_TRY(cudaEventCreateWithFlags(&p->ev_upload, cudaEventDisableTiming));
  // one value per 4 KiB unit: room for the fine work unit of small images
  FCP_TRY(cudaMalloc(&p->d_region_sums,
                     sizeof(uint32_t) * (FCP_REGION_BYTES / fcpk::kUnitBytes) *
                         (p->n_regions ? p->n_regions : 1)));
  // images too small to give every warp of a full 1024-thread grid a region
  // are worked in 4 KiB units by the 256-thread shape (4x the warps)
  p->fine = p->n_regions < static_cast<uint64_t>(g_sm_count[device] > 0 ? g_sm_count[device] : 148) * 32 &&
            getenv("FCP_DEBUG_COARSE") == nullptr;
  FCP_TRY(cudaMalloc(&p->d_chunk_sums,
                     sizeof(uint32_t) * (p->n_chunks ? p->n_chunks : 1)));
  FCP_TRY(cudaMalloc(&p->d_work, 256));
  FCP_TRY(cudaMemset(p->d_work, 0, 256));
  FCP_TRY(cudaMallocHost(&p->h_chunk_sums,
                         sizeof(uint32_t) * (p->n_chunks ? p->n_chunks : 1)));
  FCP_TRY(cudaStreamCreateWithFlags(&p->copy_stream, cudaStreamNonBlocking));
  FCP_TRY(cudaEventCreate(&p->ev_begin));
  FCP_TRY(cudaEventCreate(&p->ev_packed));
  FCP_TRY(cudaEventCreate(&p->ev_done));
  FCP_TRY(cudaEventCreate(&p->ev_kern_b));
  FCP_TRY(cudaEventCreate(&p->ev_kern_e));
#undef FCP_TRY
  if (n) {
    memcpy(p->h_pinned, descs, sizeof(fcp_desc) * n);
    p->descs_dirty = true;
    rc = plan_sync_descs(p, p->copy_stream);
    if (rc) return bail(rc);
    if (cudaStreamSynchronize(p->copy_stream) != cudaSuccess)
      return bail(fail(-EIO, "descriptor upload failed"));
  }
  *out = p;
  return 0;
}

int fcp_plan_update_ptrs(fcp_plan* p, const void* const* ptrs, uint64_t n) {
  if (!p) return fail(-EINVAL, "plan is NULL");
  if (n != p->n_desc)
    return fail(-EINVAL, "expected %llu pointers, got %llu",
                (unsigned long long)p->n_desc, (unsigned long long)n);
  bool changed = false;
  for (uint64_t i = 0; i < n; ++i)
    if (p->h_descs[i].ptr != ptrs[i]) {
      p->h_descs[i].ptr = ptrs[i];
      changed = true;
    }
  if (!changed) return 0;
  // the pinned mirror may still be the source of an upload in flight
  if (p->upload_pending) {
    FCP_CUDA(cudaEventSynchronize(p->ev_upload));
    p->upload_pending = false;
  }
  for (uint64_t i = 0; i < n; ++i) p->h_pinned[i].ptr = ptrs[i];
  p->descs_dirty = true;  // uploaded by the next launch, on its stream
  return 0;
}

void fcp_plan_destroy(fcp_plan* p) {
  if (!p) return;
  DeviceGuard g(p->device);
  if (p->copy_stream) cudaStreamSynchronize(p->copy_stream);
  if (p->d_descs) cudaFree(p->d_descs);
  if (p->d_region_sums) cudaFree(p->d_region_sums);
  if (p->d_work) cudaFree(p->d_work);
  if (p->d_chunk_sums) cudaFree(p->d_chunk_sums);
  if (p->h_chunk_sums) cudaFreeHost(p->h_chunk_sums);
  if (p->h_pinned) cudaFreeHost(p->h_pinned);
  if (p->ev_upload) cudaEventDestroy(p->ev_upload);
  if (p->ring) cudaFreeHost(p->ring);
  if (p->d_slice_bases) cudaFree(p->d_slice_bases);
  if (p->copy_stream) cudaStreamDestroy(p->copy_stream);
  cudaEvent_t evs[] = {p->ev_begin, p->ev_packed, p->ev_done, p->ev_kern_b,
                       p->ev_kern_e};
  for (cudaEvent_t e : evs)
    if (e) cudaEventDestroy(e);
  for (cudaEvent_t e : p->ev_slice) cudaEventDestroy(e);
  delete p;
}

uint64_t fcp_plan_total_bytes(const fcp_plan* p) { return p ? p->total : 0; }
uint64_t fcp_plan_num_chunks(const fcp_plan* p) { return p ? p->n_chunks : 0; }
uint64_t fcp_plan_num_regions(const fcp_plan* p) { return p ? p->n_regions : 0; }

int fcp_plan_set_slice_bytes(fcp_plan* p, uint64_t slice_bytes) {
  if (!p) return fail(-EINVAL, "plan is NULL");
  if (slice_bytes == 0 || slice_bytes % p->chunk_bytes != 0)
    return fail(-EINVAL, "slice_bytes must be a positive multiple of chunk_bytes");
  p->slice_bytes = slice_bytes;
  return 0;
}

int fcp_plan_set_read_prefetch(fcp_plan* p, int enable) {
  if (!p) return fail(-EINVAL, "plan is NULL");
  if (enable) p->kernel_flags &= ~fcpk::kNoPrefetch;
  else p->kernel_flags |= fcpk::kNoPrefetch;
  return 0;
}

int fcp_plan_set_copy_piece_bytes(fcp_plan* p, uint64_t piece_bytes) {
  if (!p) return fail(-EINVAL, "plan is NULL");
  p->copy_piece = piece_bytes;
  return 0;
}

int fcp_host_register(void* base, uint64_t bytes, void** dev_alias) {
  if (!base || bytes == 0) return fail(-EINVAL, "empty host range");
  cudaError_t e = cudaHostRegister(
      base, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(-ENOMEM, "cudaHostRegister(%p, %llu) failed: %s", base,
                (unsigned long long)bytes, cudaGetErrorString(e));
  }
  if (dev_alias) {
    void* d = nullptr;
    e = cudaHostGetDevicePointer(&d, base, 0);
    if (e != cudaSuccess) {
      cudaGetLastError();
      cudaHostUnregister(base);
      return fail(-EIO, "cudaHostGetDevicePointer failed: %s",
                  cudaGetErrorString(e));
    }
    *dev_alias = d;
  }
  return 0;
}

int fcp_enable_peer(int device, int peer) {
  if (device == peer) return 0;
  DeviceGuard g(device);
  if (!g.ok) return fail(-EIO, "cannot select device %d", device);
  int can = 0;
  FCP_CUDA(cudaDeviceCanAccessPeer(&can, device, peer));
  if (!can) return fail(-ENOTSUP, "device %d cannot access device %d", device, peer);
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return fail(-EIO, "cudaDeviceEnablePeerAccess(%d) failed: %s", peer,
                cudaGetErrorString(e));
  }
  cudaGetLastError();
  return 0;
}

// ---- CUDA IPC of a device buffer (fused peer recovery) -------------------
// cuMemGetAddressRange through the runtime's driver entry point lookup, so
// the library needs no link-time dependency on libcuda.
typedef int (*fcp_cuMemGetAddressRange_t)(unsigned long long*, size_t*, unsigned long long);

static int allocation_base(const void* p, uintptr_t* base) {
  static fcp_cuMemGetAddressRange_t fn = nullptr;
  if (!fn) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuMemGetAddressRange", &sym, cudaEnableDefault, &q);
    if (e != cudaSuccess || !sym) {
      cudaGetLastError();
      return fail(-ENOTSUP, "cuMemGetAddressRange is not available: %s", cudaGetErrorString(e));
    }
    fn = reinterpret_cast<fcp_cuMemGetAddressRange_t>(sym);
  }
  unsigned long long b = 0;
  size_t n = 0;
  const int rc = fn(&b, &n, reinterpret_cast<unsigned long long>(p));
  if (rc != 0) return fail(-EIO, "cuMemGetAddressRange failed with %d", rc);
  *base = static_cast<uintptr_t>(b);
  return 0;
}

int fcp_ipc_export(const void* dev_ptr, uint8_t* handle64, uint64_t* offset) {
  if (!dev_ptr || !handle64 || !offset) return fail(-EINVAL, "NULL argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "unexpected IPC handle size");
  uintptr_t base = 0;
  int rc = allocation_base(dev_ptr, &base);
  if (rc) return rc;
  cudaIpcMemHandle_t h;
  FCP_CUDA(cudaIpcGetMemHandle(&h, reinterpret_cast<void*>(base)));
  memcpy(handle64, &h, 64);
  *offset = reinterpret_cast<uintptr_t>(dev_ptr) - base;
  return 0;
}

int fcp_ipc_open(const uint8_t* handle64, void** base) {
  if (!handle64 || !base) return fail(-EINVAL, "NULL argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  FCP_CUDA(cudaIpcOpenMemHandle(base, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int fcp_ipc_close(void* base) {
  FCP_CUDA(cudaIpcCloseMemHandle(base));
  return 0;
}

int fcp_host_unregister(void* base) {
  cudaError_t e = cudaHostUnregister(base);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(-EIO, "cudaHostUnregister failed: %s", cudaGetErrorString(e));
  }
  return 0;
}

}  // extern "C"

// ============================================= kernels-only entry points
extern "C" {

int fcp_launch(const fcp_desc* descs_dev, uint64_t n, uint64_t total_bytes,
               uint32_t chunk_bytes, uint64_t region_begin, uint64_t region_end,
               void* packed, uint32_t* region_sums, uint32_t* chunk_sums,
               int direction, int checksum, void* stream) {
  int rc = check_geometry(total_bytes, chunk_bytes);
  if (rc) return rc;
  const uint32_t rpc = chunk_bytes / FCP_REGION_BYTES;
  const uint64_t n_regions = ceil_div(total_bytes, FCP_REGION_BYTES);
  if (region_begin > region_end || region_end > n_regions)
    return fail(-EINVAL, "bad region range");
  if (region_begin % rpc != 0 || (region_end % rpc != 0 && region_end != n_regions))
    return fail(-EINVAL, "region range is not a whole number of chunks");
  if (checksum && (!region_sums || !chunk_sums))
    return fail(-EINVAL, "checksum requested without sum buffers");
  if (direction != 0 && direction != 1) return fail(-EINVAL, "bad direction");
  int dev = 0;
  rc = current_device(&dev);
  if (rc) return rc;
  rc = init_device(dev);
  if (rc) return rc;
  fcpk::Args a;
  a.descs = descs_dev;
  a.n_desc = static_cast<uint32_t>(n);
  a.total = total_bytes;
  a.region_begin = region_begin;
  a.region_end = region_end;
  a.packed = static_cast<uint8_t*>(packed);
  a.region_sums = region_sums;
  a.run = 1;
  a.flags = 0;
  a.work = nullptr;  // static striding: no counters to own here
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  rc = launch_any(a, direction, checksum, dev, st);
  if (rc) return rc;
  if (checksum)
    rc = launch_finalize(region_sums, chunk_sums, n_regions, rpc,
                         region_begin / rpc, ceil_div(region_end, rpc), st);
  return rc;
}

int fcp_chunk_xxh64(const void* packed, uint64_t total_bytes, uint32_t chunk_bytes,
                    uint64_t* out, void* stream) {
  if (chunk_bytes == 0 || chunk_bytes % 32 != 0)
    return fail(-EINVAL, "chunk_bytes %u is not a multiple of 32", chunk_bytes);
  if (total_bytes == 0) return 0;
  if (!packed || !out) return fail(-EINVAL, "NULL argument");
  const uint64_t n_chunks = ceil_div(total_bytes, chunk_bytes);
  const unsigned threads = 128;
  const uint64_t blocks = ceil_div(n_chunks * 4, threads);
  if (blocks > 0x7fffffffull) return fail(-E2BIG, "too many chunks");
  fcpk::fcp_xxh64_kernel<<<static_cast<unsigned>(blocks), threads, 0,
                           static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint8_t*>(packed), total_bytes, chunk_bytes, 0, n_chunks,
      reinterpret_cast<unsigned long long*>(out));
  FCP_CUDA(cudaGetLastError());
  return 0;
}

int fcp_finish_checksums(const uint32_t* raw, uint64_t n_chunks,
                         uint64_t total_bytes, uint32_t chunk_bytes,
                         uint32_t* out) {
  int rc = check_geometry(total_bytes, chunk_bytes);
  if (rc) return rc;
  if (n_chunks != ceil_div(total_bytes, chunk_bytes))
    return fail(-EINVAL, "n_chunks does not match total_bytes/chunk_bytes");
  return finish_sums(raw, n_chunks, total_bytes, chunk_bytes, out);
}

static int plan_kernels_only(fcp_plan* p, void* packed, int direction,
                             int checksum, void* stream) {
  if (!p) return fail(-EINVAL, "plan is NULL");
  if (!packed && p->total) return fail(-EINVAL, "packed is NULL");
  DeviceGuard g(p->device);
  if (!g.ok) return fail(-EIO, "cannot select device %d", p->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = plan_sync_descs(p, st);
  if (rc) return rc;
  FCP_CUDA(cudaEventRecord(p->ev_kern_b, st));
  p->last_launches = 0;
  rc = plan_launch(p, packed, direction, checksum, 0, p->n_chunks, st);
  if (rc) return rc;
  FCP_CUDA(cudaEventRecord(p->ev_kern_e, st));
  p->last_has_sums = checksum != 0;
  return 0;
}

int fcp_pack(fcp_plan* p, void* packed, int checksum, void* stream) {
  return plan_kernels_only(p, packed, 0, checksum, stream);
}

int fcp_scatter(fcp_plan* p, const void* packed, int checksum, void* stream) {
  return plan_kernels_only(p, const_cast<void*>(packed), 1, checksum, stream);
}

int fcp_read_checksums(fcp_plan* p, uint32_t* out, void* stream) {
  if (!p || !out) return fail(-EINVAL, "NULL argument");
  if (!p->last_has_sums) return fail(-ENODATA, "last launch had no checksum");
  DeviceGuard g(p->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  FCP_CUDA(cudaMemcpyAsync(p->h_chunk_sums, p->d_chunk_sums,
                           sizeof(uint32_t) * p->n_chunks,
                           cudaMemcpyDeviceToHost, st));
  FCP_CUDA(cudaStreamSynchronize(st));
  return finish_sums(p->h_chunk_sums, p->n_chunks, p->total, p->chunk_bytes, out);
}

// ======================================================== transfer engine
// Device->host leg of a staged save: slice k leaves as soon as its pack
// kernel is done (ev_slice[k]); the chunk sums follow the finalize kernel.
static int enqueue_save_copies(fcp_plan* p) {
  for (size_t s = 0; s < p->sv_slices.size(); ++s) {
    const uint64_t b0 = p->sv_slices[s].first * p->chunk_bytes;
    uint64_t b1 = p->sv_slices[s].second * static_cast<uint64_t>(p->chunk_bytes);
    if (b1 > p->total) b1 = p->total;
    FCP_CUDA(cudaStreamWaitEvent(p->copy_stream, p->ev_slice[s], 0));
    if (p->save_mode == FCP_MODE_STAGED_KERNEL) {
      static const int ctas = getenv("FCP_COPY_CTAS") ? atoi(getenv("FCP_COPY_CTAS")) : 32;
      const uint64_t n = b1 - b0;  // b0 is a multiple of the chunk size
      fcpk::fcp_copy_kernel<<<ctas > 0 ? ctas : 32, 256, 0, p->copy_stream>>>(
          static_cast<const uint8_t*>(p->sv_staging) + b0, p->sv_alias + b0, n / 32, n % 32);
      FCP_CUDA(cudaGetLastError());
      p->last_launches++;
    } else {
      int rc = copy_image_range(p, static_cast<uint8_t*>(p->sv_staging),
                                static_cast<uint8_t*>(p->sv_host), b0, b1, true,
                                p->copy_stream);
      if (rc) return rc;
    }
  }
  FCP_CUDA(cudaStreamWaitEvent(p->copy_stream, p->ev_kern_e, 0));
  if (p->sv_checksum && p->n_chunks)
    FCP_CUDA(cudaMemcpyAsync(p->h_chunk_sums, p->d_chunk_sums,
                             sizeof(uint32_t) * p->n_chunks,
                             cudaMemcpyDeviceToHost, p->copy_stream));
  FCP_CUDA(cudaEventRecord(p->ev_done, p->copy_stream));
  p->copies_pending = false;
  return 0;
}

int fcp_save_async(fcp_plan* p, void* staging, void* host_dst, int mode,
                   int checksum, void* stream) {
  if (!p) return fail(-EINVAL, "plan is NULL");
  if (p->save_in_flight) return fail(-EBUSY, "a save is already in flight");
  if (!host_dst && p->total) return fail(-EINVAL, "host_dst is NULL");
  if (mode != FCP_MODE_DIRECT && !staging && p->total)
    return fail(-EINVAL, "staged modes need a staging buffer");
  if (mode != FCP_MODE_STAGED && mode != FCP_MODE_DIRECT && mode != FCP_MODE_STAGED_KERNEL)
    return fail(-EINVAL, "bad mode %d", mode);
  DeviceGuard g(p->device);
  if (!g.ok) return fail(-EIO, "cannot select device %d", p->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = plan_sync_descs(p, st);
  if (rc) return rc;
  FCP_CUDA(cudaEventRecord(p->ev_begin, st));
  p->last_launches = 0;
  if (mode == FCP_MODE_DIRECT) {
    if (p->win_begin != 0 || p->win_end < p->n_chunks)
      return fail(-EINVAL, "a chunk window needs a staged mode");
    void* alias = nullptr;
    FCP_CUDA(cudaHostGetDevicePointer(&alias, host_dst, 0));
    rc = plan_launch(p, alias, 0, checksum, 0, p->n_chunks, st);
    if (rc) return rc;
    FCP_CUDA(cudaEventRecord(p->ev_packed, st));
    FCP_CUDA(cudaStreamWaitEvent(p->copy_stream, p->ev_packed, 0));
  } else {
    uint8_t* host_alias = nullptr;
    if (mode == FCP_MODE_STAGED_KERNEL && p->total) {
      // every piece of the segment is registered as mapped; with unified
      // addressing its device alias is the host address itself
      void* a0 = nullptr;
      FCP_CUDA(cudaHostGetDevicePointer(&a0, host_dst, 0));
      if (a0 != host_dst)
        return fail(-ENOTSUP, "host pointer is not usable from the device as is");
      host_alias = static_cast<uint8_t*>(a0);
    }
    p->sv_slices = make_slices(p, true, p->win_begin, p->win_end);
    rc = plan_slice_events(p, p->sv_slices.size());
    if (rc) return rc;
    for (size_t s = 0; s < p->sv_slices.size(); ++s) {
      rc = plan_launch(p, staging, 0, checksum, p->sv_slices[s].first,
                       p->sv_slices[s].second, st, false);
      if (rc) return rc;
      FCP_CUDA(cudaEventRecord(p->ev_slice[s], st));
    }
    // the sources are free once the last pack kernel is done; combining the
    // region values into chunk values only touches the plan's own buffers
    FCP_CUDA(cudaEventRecord(p->ev_packed, st));
    const uint64_t wb = p->win_begin < p->n_chunks ? p->win_begin : p->n_chunks;
    const uint64_t we = p->win_end < p->n_chunks ? p->win_end : p->n_chunks;
    if (checksum && we > wb) {
      rc = plan_finalize(p, wb, we, st);
      if (rc) return rc;
      p->last_launches++;
    }
    FCP_CUDA(cudaEventRecord(p->ev_kern_e, st));
    p->sv_staging = staging;
    p->sv_host = host_dst;
    p->sv_alias = host_alias;
    p->sv_checksum = checksum;
    p->save_mode = mode;
    p->last_has_sums = checksum != 0;
    p->save_in_flight = true;
    p->copies_pending = true;
    return p->defer_copies ? 0 : enqueue_save_copies(p);
  }
  if (checksum && p->n_chunks)
    FCP_CUDA(cudaMemcpyAsync(p->h_chunk_sums, p->d_chunk_sums,
                             sizeof(uint32_t) * p->n_chunks,
                             cudaMemcpyDeviceToHost, p->copy_stream));
  FCP_CUDA(cudaEventRecord(p->ev_done, p->copy_stream));
  p->last_has_sums = checksum != 0;
  p->save_in_flight = true;
  p->save_mode = mode;
  return 0;
}

int fcp_plan_set_defer_copies(fcp_plan* p, int defer) {
  if (!p) return fail(-EINVAL, "plan is NULL");
  if (p->save_in_flight) return fail(-EBUSY, "a save is in flight");
  p->defer_copies = defer != 0;
  return 0;
}

int fcp_save_release_copies(fcp_plan* p) {
  if (!p) return fail(-EINVAL, "plan is NULL");
  if (!p->save_in_flight) return fail(-EINVAL, "no save in flight");
  if (!p->copies_pending) return 0;
  DeviceGuard g(p->device);
  if (!g.ok) return fail(-EIO, "cannot select device %d", p->device);
  return enqueue_save_copies(p);
}

int fcp_save_packed_wait(fcp_plan* p) {
  if (!p) return fail(-EINVAL, "plan is NULL");
  if (!p->save_in_flight) return fail(-EINVAL, "no save in flight");
  FCP_CUDA(cudaEventSynchronize(p->ev_packed));
  return 0;
}

int fcp_save_wait(fcp_plan* p, uint32_t* chunk_sums) {
  if (!p) return fail(-EINVAL, "plan is NULL");
  if (!p->save_in_flight) return fail(-EINVAL, "no save in flight");
  if (p->copies_pending)
    return fail(-EINVAL, "the copies of this save were deferred: call fcp_save_release_copies");
  p->save_in_flight = false;
  FCP_CUDA(cudaEventSynchronize(p->ev_done));
  FCP_CUDA(cudaEventElapsedTime(&p->t_kernel_ms, p->ev_begin, p->ev_packed));
  FCP_CUDA(cudaEventElapsedTime(&p->t_total_ms, p->ev_begin, p->ev_done));
  p->t_copy_ms = p->t_total_ms - p->t_kernel_ms;
  if (chunk_sums) {
    if (!p->last_has_sums) return fail(-ENODATA, "save ran without checksum");
    return finish_sums(p->h_chunk_sums, p->n_chunks, p->total, p->chunk_bytes,
                       chunk_sums);
  }
  return 0;
}

static int verify_sums(fcp_plan* p, const uint32_t* expected) {
  std::vector<uint32_t> got(p->n_chunks);
  int rc = finish_sums(p->h_chunk_sums, p->n_chunks, p->total, p->chunk_bytes,
                       got.data());
  if (rc) return rc;
  for (uint64_t c = 0; c < p->n_chunks; ++c)
    if (got[c] != expected[c])
      return fail(-EBADMSG,
                  "checksum mismatch in chunk %llu: expected %08x, got %08x",
                  (unsigned long long)c, expected[c], got[c]);
  return 0;
}

int fcp_restore(fcp_plan* p, const void* host_src, void* staging, int mode,
                int checksum, const uint32_t* expected, void* stream) {
  if (!p) return fail(-EINVAL, "plan is NULL");
  if (p->save_in_flight) return fail(-EBUSY, "a save is in flight");
  if (!host_src && p->total) return fail(-EINVAL, "host_src is NULL");
  if (expected && !checksum)
    return fail(-EINVAL, "expected sums given but checksum disabled");
  if (mode == FCP_MODE_STAGED_KERNEL) mode = FCP_MODE_STAGED;  // reads: copy engine
  if (mode == FCP_MODE_STAGED && !staging && p->total)
    return fail(-EINVAL, "staged mode needs a staging buffer");
  if (mode != FCP_MODE_STAGED && mode != FCP_MODE_DIRECT)
    return fail(-EINVAL, "bad mode %d", mode);
  DeviceGuard g(p->device);
  if (!g.ok) return fail(-EIO, "cannot select device %d", p->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = plan_sync_descs(p, st);
  if (rc) return rc;
  p->last_launches = 0;
  FCP_CUDA(cudaEventRecord(p->ev_begin, st));
  if (mode == FCP_MODE_DIRECT) {
    void* alias = nullptr;
    FCP_CUDA(cudaHostGetDevicePointer(&alias, const_cast<void*>(host_src), 0));
    rc = plan_launch(p, alias, 1, checksum, 0, p->n_chunks, st);
    if (rc) return rc;
  } else {
    // copies start once the caller's stream reaches this point
    FCP_CUDA(cudaStreamWaitEvent(p->copy_stream, p->ev_begin, 0));
    const auto slices = make_slices(p, false);
    rc = plan_slice_events(p, slices.size());
    if (rc) return rc;
    for (size_t s = 0; s < slices.size(); ++s) {
      const uint64_t cb = slices[s].first, ce = slices[s].second;
      const uint64_t b0 = cb * p->chunk_bytes;
      uint64_t b1 = ce * static_cast<uint64_t>(p->chunk_bytes);
      if (b1 > p->total) b1 = p->total;
      rc = copy_image_range(p, static_cast<uint8_t*>(staging),
                            const_cast<uint8_t*>(static_cast<const uint8_t*>(host_src)),
                            b0, b1, false, p->copy_stream);
      if (rc) return rc;
      FCP_CUDA(cudaEventRecord(p->ev_slice[s], p->copy_stream));
      FCP_CUDA(cudaStreamWaitEvent(st, p->ev_slice[s], 0));
      rc = plan_launch(p, staging, 1, checksum, cb, ce, st, false);
      if (rc) return rc;
    }
    if (checksum && p->n_chunks) {
      rc = plan_finalize(p, 0, p->n_chunks, st);
      if (rc) return rc;
      p->last_launches++;
    }
  }
  if (checksum && p->n_chunks)
    FCP_CUDA(cudaMemcpyAsync(p->h_chunk_sums, p->d_chunk_sums,
                             sizeof(uint32_t) * p->n_chunks,
                             cudaMemcpyDeviceToHost, st));
  FCP_CUDA(cudaEventRecord(p->ev_done, st));
  FCP_CUDA(cudaEventSynchronize(p->ev_done));
  FCP_CUDA(cudaEventElapsedTime(&p->t_total_ms, p->ev_begin, p->ev_done));
  p->t_kernel_ms = 0.f;
  p->t_copy_ms = 0.f;
  p->last_has_sums = checksum != 0;
  if (expected) return verify_sums(p, expected);
  return 0;
}

int fcp_restore_from_device(fcp_plan* p, const void* packed, int checksum,
                            const uint32_t* expected, void* stream) {
  if (!p) return fail(-EINVAL, "plan is NULL");
  if (expected && !checksum)
    return fail(-EINVAL, "expected sums given but checksum disabled");
  DeviceGuard g(p->device);
  if (!g.ok) return fail(-EIO, "cannot select device %d", p->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = plan_sync_descs(p, st);
  if (rc) return rc;
  FCP_CUDA(cudaEventRecord(p->ev_begin, st));
  p->last_launches = 0;
  rc = plan_launch(p, const_cast<void*>(packed), 1, checksum, 0,
                   p->n_chunks, st);
  if (rc) return rc;
  if (checksum && p->n_chunks)
    FCP_CUDA(cudaMemcpyAsync(p->h_chunk_sums, p->d_chunk_sums,
                             sizeof(uint32_t) * p->n_chunks,
                             cudaMemcpyDeviceToHost, st));
  FCP_CUDA(cudaEventRecord(p->ev_done, st));
  FCP_CUDA(cudaEventSynchronize(p->ev_done));
  FCP_CUDA(cudaEventElapsedTime(&p->t_total_ms, p->ev_begin, p->ev_done));
  p->t_kernel_ms = p->t_total_ms;
  p->t_copy_ms = 0.f;
  p->last_has_sums = checksum != 0;
  if (expected) return verify_sums(p, expected);
  return 0;
}

// Copy image bytes [b0, b1) from an UNPINNED host image into `staging`
// through the plan's pinned ring, n_workers host threads; synchronous.
static int bounce_range(fcp_plan* p, const void* host_src, void* staging,
                        uint64_t b0, uint64_t b1, int n_workers) {
  if (n_workers <= 0) {
    unsigned hc = std::thread::hardware_concurrency();
    n_workers = hc >= 16 ? 8 : hc >= 4 ? static_cast<int>(hc / 2) : 1;
  }
  if (n_workers > 32) n_workers = 32;
  DeviceGuard g(p->device);
  if (!g.ok) return fail(-EIO, "cannot select device %d", p->device);
  // small ring: pinned allocation costs ~0.5 s/GiB and is on the critical
  // path of the first (cold) restore.  Measured on 16 GB, 8 workers: 1 MiB
  // slots 32 GB/s, 2 MiB 43, 4 MiB 51, 8 MiB 51, 16 MiB 47 (tools/cold_probe.py)
  static const uint64_t slot_mb =
      getenv("FCP_BOUNCE_SLOT_MB") ? strtoull(getenv("FCP_BOUNCE_SLOT_MB"), nullptr, 10) : 4;
  const uint64_t slot = (slot_mb ? slot_mb : 4) << 20;
  if (!p->ring || p->ring_workers < n_workers || p->ring_slot_bytes != slot) {
    if (p->ring) cudaFreeHost(p->ring);
    p->ring = nullptr;
    FCP_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&p->ring),
                           slot * 2 * static_cast<uint64_t>(n_workers),
                           cudaHostAllocPortable));
    p->ring_workers = n_workers;
    p->ring_slot_bytes = slot;
  }
  if (b1 <= b0) return 0;
  const uint64_t n_slices = ceil_div(b1 - b0, slot);
  std::atomic<uint64_t> next{0};
  std::atomic<int> failed{0};
  const uint8_t* src = static_cast<const uint8_t*>(host_src) + b0;
  uint8_t* dst = static_cast<uint8_t*>(staging) + b0;
  const int device = p->device;
  uint8_t* ring = p->ring;
  const uint64_t total = b1 - b0;
  auto worker = [&, device, ring, total](int w) {
    if (cudaSetDevice(device) != cudaSuccess) { failed = 1; return; }
    cudaStream_t ws = nullptr;
    cudaEvent_t ev[2] = {nullptr, nullptr};
    if (cudaStreamCreateWithFlags(&ws, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming) != cudaSuccess) {
      failed = 1;
    }
    uint64_t k = 0;
    while (!failed) {
      const uint64_t i = next.fetch_add(1);
      if (i >= n_slices) break;
      const int sl = static_cast<int>(k & 1);
      uint8_t* buf = ring + (static_cast<uint64_t>(w) * 2 + sl) * slot;
      if (k >= 2 && cudaEventSynchronize(ev[sl]) != cudaSuccess) { failed = 1; break; }
      const uint64_t off = i * slot;
      const uint64_t len = off + slot <= total ? slot : total - off;
      memcpy(buf, src + off, len);  // page cache -> pinned
      if (cudaMemcpyAsync(dst + off, buf, len, cudaMemcpyHostToDevice, ws) != cudaSuccess ||
          cudaEventRecord(ev[sl], ws) != cudaSuccess) {
        failed = 1;
        break;
      }
      ++k;
    }
    if (ws) cudaStreamSynchronize(ws);
    if (ev[0]) cudaEventDestroy(ev[0]);
    if (ev[1]) cudaEventDestroy(ev[1]);
    if (ws) cudaStreamDestroy(ws);
  };
  {
    std::vector<std::thread> pool;
    for (int w = 0; w < n_workers; ++w) pool.emplace_back(worker, w);
    for (auto& t : pool) t.join();
  }
  if (failed) {
    cudaGetLastError();
    return fail(-EIO, "bounced host->device copy failed");
  }
  return 0;
}

int fcp_restore_bounced(fcp_plan* p, const void* host_src, void* staging,
                        int checksum, const uint32_t* expected, int n_workers,
                        void* stream) {
  if (!p) return fail(-EINVAL, "plan is NULL");
  if (p->save_in_flight) return fail(-EBUSY, "a save is in flight");
  if ((!host_src || !staging) && p->total)
    return fail(-EINVAL, "host_src and staging are required");
  if (expected && !checksum)
    return fail(-EINVAL, "expected sums given but checksum disabled");
  DeviceGuard g(p->device);
  if (!g.ok) return fail(-EIO, "cannot select device %d", p->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = plan_sync_descs(p, st);
  if (rc) return rc;
  FCP_CUDA(cudaEventRecord(p->ev_begin, st));
  FCP_CUDA(cudaStreamSynchronize(st));  // staging is free, caller's work done
  rc = bounce_range(p, host_src, staging, 0, p->total, n_workers);
  if (rc) return rc;
  p->last_launches = 0;
  rc = plan_launch(p, staging, 1, checksum, 0, p->n_chunks, st);
  if (rc) return rc;
  if (checksum && p->n_chunks)
    FCP_CUDA(cudaMemcpyAsync(p->h_chunk_sums, p->d_chunk_sums,
                             sizeof(uint32_t) * p->n_chunks,
                             cudaMemcpyDeviceToHost, st));
  FCP_CUDA(cudaEventRecord(p->ev_done, st));
  FCP_CUDA(cudaEventSynchronize(p->ev_done));
  FCP_CUDA(cudaEventElapsedTime(&p->t_total_ms, p->ev_begin, p->ev_done));
  p->t_kernel_ms = 0.f;
  p->t_copy_ms = 0.f;
  p->last_has_sums = checksum != 0;
  if (expected) return verify_sums(p, expected);
  return 0;
}

int fcp_plan_set_window(fcp_plan* p, uint64_t chunk_begin, uint64_t chunk_end) {
  if (!p) return fail(-EINVAL, "plan is NULL");
  if (p->save_in_flight) return fail(-EBUSY, "a save is in flight");
  if (chunk_begin > chunk_end) return fail(-EINVAL, "bad chunk window");
  p->win_begin = chunk_begin;
  p->win_end = chunk_end;
  return 0;
}

int fcp_stage_range(fcp_plan* p, const void* host_src, void* staging,
                    uint64_t byte_begin, uint64_t byte_end, int registered,
                    int n_workers, void* stream) {
  if (!p) return fail(-EINVAL, "plan is NULL");
  if (p->save_in_flight) return fail(-EBUSY, "a save is in flight");
  if (byte_end > p->total) byte_end = p->total;
  if (byte_begin >= byte_end) return 0;
  if (!host_src || !staging) return fail(-EINVAL, "host_src and staging are required");
  DeviceGuard g(p->device);
  if (!g.ok) return fail(-EIO, "cannot select device %d", p->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  FCP_CUDA(cudaEventRecord(p->ev_begin, st));
  FCP_CUDA(cudaStreamSynchronize(st));  // the staging range is free
  int rc = 0;
  if (registered) {
    rc = copy_image_range(p, static_cast<uint8_t*>(staging),
                          const_cast<uint8_t*>(static_cast<const uint8_t*>(host_src)),
                          byte_begin, byte_end, false, p->copy_stream);
    if (rc) return rc;
    FCP_CUDA(cudaStreamSynchronize(p->copy_stream));
  } else {
    rc = bounce_range(p, host_src, staging, byte_begin, byte_end, n_workers);
    if (rc) return rc;
  }
  FCP_CUDA(cudaEventRecord(p->ev_done, st));
  FCP_CUDA(cudaEventSynchronize(p->ev_done));
  FCP_CUDA(cudaEventElapsedTime(&p->t_total_ms, p->ev_begin, p->ev_done));
  p->t_copy_ms = p->t_total_ms;
  p->t_kernel_ms = 0.f;
  p->last_launches = 0;
  return 0;
}

int fcp_restore_from_slices(fcp_plan* p, const void* const* slice_bases,
                            uint64_t n_slices, uint64_t slice_bytes, int checksum,
                            const uint32_t* expected, int read_prefetch,
                            uint64_t first_slice, void* stream) {
  if (!p) return fail(-EINVAL, "plan is NULL");
  if (expected && !checksum)
    return fail(-EINVAL, "expected sums given but checksum disabled");
  if (slice_bytes == 0 || slice_bytes % FCP_REGION_BYTES != 0)
    return fail(-EINVAL, "slice_bytes must be a positive multiple of %u", FCP_REGION_BYTES);
  if (!slice_bases || n_slices < ceil_div(p->total, slice_bytes))
    return fail(-EINVAL, "%llu slices of %llu bytes do not cover the image",
                (unsigned long long)n_slices, (unsigned long long)slice_bytes);
  if (slice_bytes / FCP_REGION_BYTES > 0xFFFFFFFFull)
    return fail(-EINVAL, "slice_bytes too large");
  DeviceGuard g(p->device);
  if (!g.ok) return fail(-EIO, "cannot select device %d", p->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = plan_sync_descs(p, st);
  if (rc) return rc;
  if (p->slice_bases_cap < n_slices) {
    if (p->d_slice_bases) cudaFree(p->d_slice_bases);
    p->d_slice_bases = nullptr;
    p->slice_bases_cap = 0;
    if (cudaMalloc(&p->d_slice_bases, sizeof(void*) * n_slices) != cudaSuccess) {
      cudaGetLastError();
      return fail(-ENOMEM, "cannot allocate the slice table");
    }
    p->slice_bases_cap = n_slices;
  }
  // pageable source: staged by the runtime before the call returns, then
  // ordered on `st` before the kernels
  FCP_CUDA(cudaMemcpyAsync(p->d_slice_bases, slice_bases, sizeof(void*) * n_slices,
                           cudaMemcpyHostToDevice, st));
  FCP_CUDA(cudaEventRecord(p->ev_begin, st));
  p->last_launches = 0;
  const uint32_t rpc = p->chunk_bytes / FCP_REGION_BYTES;
  if (p->n_regions) {
    fcpk::Args a = plan_args(p, nullptr, 0, p->n_regions);
    a.slice_base = p->d_slice_bases;
    a.slice_regions = static_cast<uint32_t>(slice_bytes / FCP_REGION_BYTES);
    if (!read_prefetch) a.flags |= fcpk::kNoPrefetch;
    rc = launch_any(a, 1, checksum, p->device, st, first_slice * a.slice_regions);
    if (rc) return rc;
    p->last_launches++;
    if (checksum) {
      rc = launch_finalize(p->d_region_sums, p->d_chunk_sums, p->n_regions, rpc, 0,
                           p->n_chunks, st);
      if (rc) return rc;
      p->last_launches++;
    }
  }
  if (checksum && p->n_chunks)
    FCP_CUDA(cudaMemcpyAsync(p->h_chunk_sums, p->d_chunk_sums,
                             sizeof(uint32_t) * p->n_chunks,
                             cudaMemcpyDeviceToHost, st));
  FCP_CUDA(cudaEventRecord(p->ev_done, st));
  FCP_CUDA(cudaEventSynchronize(p->ev_done));
  FCP_CUDA(cudaEventElapsedTime(&p->t_total_ms, p->ev_begin, p->ev_done));
  p->t_kernel_ms = p->t_total_ms;
  p->t_copy_ms = 0.f;
  p->last_has_sums = checksum != 0;
  if (expected) return verify_sums(p, expected);
  return 0;
}

uint64_t fcp_plan_last_launches(const fcp_plan* p) {
  return p ? p->last_launches : 0;
}

int fcp_plan_timings(fcp_plan* p, float* out3) {
  if (!p || !out3) return fail(-EINVAL, "NULL argument");
  out3[0] = p->t_kernel_ms;
  out3[1] = p->t_copy_ms;
  out3[2] = p->t_total_ms;
  return 0;
}

}  // extern "C"

// ================================================= no-allocation entry point
// Scratch for fcp_launch_noalloc: one uint32 per 16 KiB region of the image.
// In host code this name is only a handle for cudaGetSymbolAddress; the
// kernels receive the device address looked up below.
static __device__ __align__(256) uint32_t
    g_fcp_noalloc_region_sums[FCP_NOALLOC_MAX_REGIONS];

namespace {
uint32_t* g_noalloc_ptr[kMaxDevices] = {nullptr};
std::mutex g_noalloc_mu;

// once per device, with that device current
int noalloc_scratch(int dev, uint32_t** out) {
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_fcp_noalloc_region_sums);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return fail(-EIO, "cudaGetSymbolAddress(scratch) failed on device %d: %s",
                  dev, cudaGetErrorString(e));
    }
    g_noalloc_ptr[dev] = static_cast<uint32_t*>(p);
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}
}  // namespace

extern "C" int fcp_launch_noalloc(const fcp_desc* descs_dev, uint64_t n,
                                  uint64_t total_bytes, uint32_t chunk_bytes,
                                  void* packed, uint32_t* chunk_sums,
                                  int direction, int checksum, void* stream) {
  int rc = check_geometry(total_bytes, chunk_bytes);
  if (rc) return rc;
  if (total_bytes > FCP_NOALLOC_MAX_BYTES)
    return fail(-E2BIG,
                "image of %llu bytes exceeds the fixed scratch of "
                "fcp_launch_noalloc (max %llu bytes); use fcp_launch with "
                "caller-provided scratch",
                (unsigned long long)total_bytes,
                (unsigned long long)FCP_NOALLOC_MAX_BYTES);
  if (direction != 0 && direction != 1) return fail(-EINVAL, "bad direction");
  if (checksum && !chunk_sums)
    return fail(-EINVAL, "checksum requested without chunk_sums");
  if (!packed && total_bytes) return fail(-EINVAL, "packed is NULL");
  int dev = 0;
  rc = current_device(&dev);
  if (rc) return rc;
  rc = init_device(dev);
  if (rc) return rc;
  uint32_t* scratch = nullptr;
  rc = noalloc_scratch(dev, &scratch);
  if (rc) return rc;
  const uint32_t rpc = chunk_bytes / FCP_REGION_BYTES;
  const uint64_t n_regions = ceil_div(total_bytes, FCP_REGION_BYTES);
  fcpk::Args a;
  a.descs = descs_dev;
  a.n_desc = static_cast<uint32_t>(n);
  a.total = total_bytes;
  a.region_begin = 0;
  a.region_end = n_regions;
  a.packed = static_cast<uint8_t*>(packed);
  a.region_sums = scratch;
  a.run = 1;
  a.flags = 0;
  a.work = nullptr;  // static striding: no counters to own here
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  rc = launch_any(a, direction, checksum, dev, st);
  if (rc) return rc;
  if (checksum)
    rc = launch_finalize(scratch, chunk_sums, n_regions, rpc, 0,
                         ceil_div(n_regions, rpc), st);
  return rc;
}
