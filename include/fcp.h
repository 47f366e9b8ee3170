/*
 * fcp.h — C ABI of the B200 flash-checkpoint data path (libfcp.so).
 *
 * The reference (DLRover flash checkpoint) has NO native code on this path:
 * its "kernel" is a Python loop issuing one pageable torch copy_ per tensor.
 * Each entry point below therefore names the reference *Python* interface it
 * replaces (paths relative to the reference tree):
 *
 *   SAVER = dlrover/python/elastic_agent/torch/ckpt_saver.py
 *   FSDPE = dlrover/trainer/torch/flash_checkpoint/fsdp_engine.py
 *   REPL  = dlrover/trainer/torch/flash_checkpoint/replica.py
 *
 * Conventions: plain pointers and sizes, no torch types; every function
 * returns 0 on success or a negative errno-style code and never throws;
 * fcp_last_error() returns a thread-local message for the last failure.
 * All byte movement is asynchronous on the stream passed in unless the
 * function name ends in _wait or the comment says otherwise.
 * `stream` arguments are cudaStream_t passed as void* (0 = legacy default).
 *
 * Packed image layout (== the reference shm layout, SAVER:255-270): tensor i
 * occupies bytes [dst_off_i, dst_off_i + nbytes_i) of one dense buffer of
 * `total_bytes`; offsets are a running byte sum with NO padding, so they are
 * in general not aligned to anything.  Bytes not covered by any descriptor
 * are written as zero.
 *
 * Checksums (new; the reference has none): standard CRC-32C
 * (Castagnoli, reflected 0x82F63B78, init/xorout 0xFFFFFFFF) of each window
 * [i*chunk_bytes, min((i+1)*chunk_bytes, total_bytes)) of the packed image.
 */
#ifndef FCP_H_
#define FCP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FCP_VERSION 1

/* work granularity of the kernels: one warp owns one 16 KiB region (16 rows of
 * 1 KiB, a lane moves one 32-byte word per row); chunk_bytes must be a
 * multiple of it.                                                           */
#define FCP_REGION_BYTES 16384u
#define FCP_DEFAULT_CHUNK_BYTES (2u << 20)

/* checksum selector */
#define FCP_SUM_NONE 0
#define FCP_SUM_CRC32C 1

/* transfer modes of fcp_save_async / fcp_restore */
#define FCP_MODE_STAGED 0 /* kernel <-> HBM staging, copy engine <-> host   */
#define FCP_MODE_DIRECT 1 /* kernel reads/writes host-mapped memory itself  */
#define FCP_MODE_STAGED_KERNEL 2 /* save: pack into HBM staging (short blocking
                                  * time, as STAGED), then a small copy KERNEL
                                  * on the plan's side stream stores the image
                                  * into the host-mapped segment; restore: as
                                  * STAGED.  Kernel-issued host writes kept
                                  * scaling across GPUs where copy-engine
                                  * writes did not.             */

/* One tensor of the state dict.  Replaces the (value, TensorMeta) pair that
 * _write_shared_memory / _read_tensor_from_buf receive (SAVER:197-206,
 * SAVER:120-137): `ptr` is the tensor's data pointer (device memory, or host
 * memory that the device can address), `dst_off`/`nbytes` are
 * TensorMeta.offset and numel*element_size.                                 */
typedef struct fcp_desc {
  const void* ptr;
  uint64_t dst_off;
  uint64_t nbytes;
} fcp_desc;

typedef struct fcp_plan fcp_plan;

/* ---- library ----------------------------------------------------------- */

int fcp_version(void);
const char* fcp_last_error(void);
/* Upload the CRC tables to `device` (idempotent; done lazily otherwise).    */
int fcp_init(int device);

/* ---- plan: the descriptor table of one rank's shard -------------------- */

/* Replaces the meta-tree walk of SharedMemoryHandler.save_state_dict
 * (SAVER:272-290, _traverse_copy_to_shm SAVER:174-194): the caller flattens
 * the state dict once into `n` descriptors sorted by dst_off, non-overlapping,
 * nbytes > 0.  The plan owns a device copy of the table plus per-region /
 * per-chunk checksum buffers, one copy stream and its events.               */
int fcp_plan_create(const fcp_desc* descs, uint64_t n, uint64_t total_bytes,
                    uint32_t chunk_bytes, int device, fcp_plan** out);
/* Refresh the tensor pointers only (same layout), e.g. after the optimizer
 * re-allocated its state; replaces re-walking the dict on later saves
 * (SAVER:281-282).  Host-side only: the device copy of the table is brought
 * up to date by the next call that launches kernels for this plan, as an
 * asynchronous copy from pinned memory on THAT call's stream, so it is
 * ordered before the kernels on any kind of stream.                         */
int fcp_plan_update_ptrs(fcp_plan* plan, const void* const* ptrs, uint64_t n);
void fcp_plan_destroy(fcp_plan* plan);
uint64_t fcp_plan_total_bytes(const fcp_plan* plan);
uint64_t fcp_plan_num_chunks(const fcp_plan* plan);
uint64_t fcp_plan_num_regions(const fcp_plan* plan);

/* ---- host memory -------------------------------------------------------- */

/* Pin + map the shm segment that SharedMemoryHandler.init_shared_memory
 * created (SAVER:313-317, _create_shared_memory SAVER:140-171) so the copy
 * engines and kernels can address it.  The library never creates or unlinks
 * the segment.  *dev_alias receives the device-side address of `base`.      */
int fcp_host_register(void* base, uint64_t bytes, void** dev_alias);
int fcp_host_unregister(void* base);

/* Standard CRC-32C of every window [i*chunk_bytes, min((i+1)*chunk_bytes,
 * total_bytes)) of a packed image in HOST memory, out[i] per window; the same
 * values the kernels record (fcp_save_wait, fcp_read_checksums).  Pure host
 * code — no CUDA call, usable in a process that has no GPU: the agent-side
 * saver re-checks the shm image with it before persisting (the reference
 * persists whatever the shm holds, SAVER:914-921), and raw image files get
 * their per-file sums from it.  SSE4.2 crc32 where the CPU has it, table
 * slicing otherwise; n_threads <= 0 picks a small default.                  */
int fcp_host_chunk_crc32c(const void* image, uint64_t total_bytes,
                          uint32_t chunk_bytes, uint32_t* out, int n_threads);

/* ---- kernels only (packed image already device-addressable) ------------ */

/* Gather every tensor into the packed image at `packed` (device memory or a
 * host-mapped alias) with one launch, and — if checksum != FCP_SUM_NONE —
 * compute the per-chunk CRC32C in the same pass.  Replaces the per-tensor
 * loop _traverse_copy_to_shm -> _write_shared_memory (SAVER:174-206) and the
 * FSDP twin _write_memory_from_list -> _write_item (FSDPE:85-155).          */
int fcp_pack(fcp_plan* plan, void* packed, int checksum, void* stream);

/* Inverse: read the packed image, recompute the per-chunk checksums, and
 * write every tensor's bytes straight into the live tensor at desc.ptr.
 * Replaces _read_tensor_from_buf + the caller's per-tensor H2D copy_
 * (SAVER:120-137, FSDPE:239-297).                                           */
int fcp_scatter(fcp_plan* plan, const void* packed, int checksum,
                void* stream);

/* Synchronise `stream` and return the checksums of the last fcp_pack /
 * fcp_scatter as standard CRC-32C values, one per chunk.                    */
int fcp_read_checksums(fcp_plan* plan, uint32_t* out, void* stream);

/* Stateless launch used by the two calls above, for callers that own all
 * memory themselves: `descs_dev` is the descriptor table in device memory,
 * `region_sums` is scratch of ceil(total_bytes/FCP_REGION_BYTES) uint32,
 * `chunk_sums` receives one raw (un-finalised) value per chunk; only regions
 * [region_begin, region_end) — a whole number of chunks — are processed.
 * direction 0 = pack, 1 = scatter.                                          */
int fcp_launch(const fcp_desc* descs_dev, uint64_t n, uint64_t total_bytes,
               uint32_t chunk_bytes, uint64_t region_begin,
               uint64_t region_end, void* packed, uint32_t* region_sums,
               uint32_t* chunk_sums, int direction, int checksum,
               void* stream);
/* Convert raw per-chunk values produced by fcp_launch into standard CRC-32C
 * (host-side arithmetic only).                                              */
int fcp_finish_checksums(const uint32_t* raw, uint64_t n_chunks,
                         uint64_t total_bytes, uint32_t chunk_bytes,
                         uint32_t* out);

/* Second checksum algorithm (the reference has none): standard
 * XXH64, seed 0, of every chunk of a DEVICE-resident packed image (the HBM
 * staging after fcp_pack, or after the host->device copy of a restore);
 * out[i] (device memory) receives the hash of chunk i.  XXH64 is not linear,
 * so unlike CRC-32C it cannot be fused into the pack kernel: it is a second
 * pass over the image, four threads per chunk, all chunks in flight.        */
int fcp_chunk_xxh64(const void* packed, uint64_t total_bytes, uint32_t chunk_bytes,
                    uint64_t* out, void* stream);

/* Same work as fcp_launch over the WHOLE image, for callers that must not
 * allocate device memory.  No counterpart in the reference.
 *
 * Who needs it: a trainer that has handed all of its HBM to a caching
 * allocator or runs under a hard per-process memory cap, where a raw
 * cudaMalloc for the per-region scratch can fail or fragment the pool; and
 * last-gasp callers such as an out-of-memory or breakpoint handler that wants
 * the state packed before the process dies and by definition has no memory
 * left to ask for.  cudaMalloc/cudaFree also synchronise the whole device,
 * which a caller inside a latency-critical step may not afford.
 *
 * The per-region checksum scratch lives in a fixed array inside the library
 * (FCP_NOALLOC_MAX_REGIONS uint32 = 256 KiB, resident on every GPU that loads
 * the library until the process exits), so total_bytes may not exceed
 * FCP_NOALLOC_MAX_BYTES (1 GiB); larger images are rejected with -E2BIG.
 * The caller passes everything else: `descs_dev` (device memory), `packed`,
 * and `chunk_sums` (device memory, one raw value per chunk, convert with
 * fcp_finish_checksums).  This call performs no cudaMalloc/cudaFree and calls
 * no allocating library routine.
 *
 * The first call on each device looks up the device address of the scratch
 * array (and uploads the CRC tables if no other call has yet); make that
 * first call before starting any stream capture.  All calls on a device share
 * the one array: every slot a call reads it has written first, so nothing
 * needs clearing between calls, but two calls whose kernels overlap on the
 * GPU (different streams, or different threads) race on it — serialise them.
 * direction 0 = pack, 1 = scatter.                                          */
#define FCP_NOALLOC_MAX_REGIONS 65536u
#define FCP_NOALLOC_MAX_BYTES ((uint64_t)FCP_NOALLOC_MAX_REGIONS * FCP_REGION_BYTES)
int fcp_launch_noalloc(const fcp_desc* descs_dev, uint64_t n,
                       uint64_t total_bytes, uint32_t chunk_bytes,
                       void* packed, uint32_t* chunk_sums, int direction,
                       int checksum, void* stream);

/* ---- transfer engine ---------------------------------------------------- */

/* Blocking part of SharedMemoryHandler.save_state_dict (SAVER:272-290):
 * enqueue pack (+checksum) on `stream`, then the device->host transfer into
 * `host_dst` (registered with fcp_host_register) on the plan's copy stream.
 * Returns as soon as everything is enqueued.  FCP_MODE_STAGED needs
 * `staging` = total_bytes rounded up to FCP_REGION_BYTES of device memory
 * and leaves the live tensors free to change once the pack kernel is done
 * (fcp_save_packed_wait); FCP_MODE_DIRECT ignores `staging`.                */
int fcp_save_async(fcp_plan* plan, void* staging, void* host_dst, int mode,
                   int checksum, void* stream);
/* Split a staged save in two: with defer != 0, fcp_save_async launches the
 * pack kernels only and fcp_save_release_copies later enqueues the
 * device->host copies (fcp_save_wait refuses to run before that).  Lets the
 * caller take a turn on a shared host: on an 8-GPU node 4 ranks writing at a
 * time move 102 GB/s into host memory, all 8 at once 90 GB/s (measured,
 * profiles/r02_hostlink.md); the training-blocking part of the save is not
 * delayed by the wait for a turn.  No counterpart in the reference (its
 * copies are synchronous, SAVER:197-206).                                   */
int fcp_plan_set_defer_copies(fcp_plan* plan, int defer);
int fcp_save_release_copies(fcp_plan* plan);
/* Wait until the pack kernel has consumed the live tensors (the training-
 * blocking part).                                                           */
int fcp_save_packed_wait(fcp_plan* plan);
/* Wait until every byte is in host memory; `chunk_sums` (may be NULL)
 * receives the standard CRC-32C per chunk.                                  */
int fcp_save_wait(fcp_plan* plan, uint32_t* chunk_sums);

/* SharedMemoryHandler.load_state_dict + the caller's load into the model
 * (SAVER:292-311 and the per-tensor H2D it implies): move the packed image
 * from `host_src` into the live tensors, sliced so the host->device copy of
 * slice k+1 overlaps the scatter of slice k.  If `expected` is non-NULL the
 * recomputed checksums are compared with it and -EBADMSG is returned on
 * mismatch (the tensors are then undefined and the caller falls back to
 * storage, as the reference does when shm holds no valid checkpoint,
 * SAVER:303-308).  Synchronous: complete on return.                         */
int fcp_restore(fcp_plan* plan, const void* host_src, void* staging, int mode,
                int checksum, const uint32_t* expected, void* stream);

/* Same result as fcp_restore in staged mode, for a host image that is NOT
 * registered — the situation of a freshly restarted trainer attaching to the
 * shm segment its predecessor left (SAVER:305-306).  Pinning 19 GB takes
 * ~1.6 s (cudaHostRegister does not parallelise), so instead `n_workers` host
 * threads (<= 0: pick from the core count) copy 4 MiB slices of the image
 * into a small pinned ring owned by the plan and DMA them from there;
 * the scatter kernel then verifies and unpacks the staged image.
 * Synchronous.                                                              */
int fcp_restore_bounced(fcp_plan* plan, const void* host_src, void* staging,
                        int checksum, const uint32_t* expected, int n_workers,
                        void* stream);

/* Peer recovery (replaces the gloo broadcast of shm bytes in
 * FullCkptReplicaManager.gather, REPL:275-332): the packed image is already
 * in device memory (received by NCCL broadcast from a surviving replica, or
 * a peer-mapped pointer); verify and unpack it.  Synchronous.               */
int fcp_restore_from_device(fcp_plan* plan, const void* packed, int checksum,
                            const uint32_t* expected, void* stream);

/* The same for an image whose pieces sit in SEVERAL device buffers — the HBM
 * staging of the ranks of a replicated (DDP) group, mapped here with
 * fcp_ipc_open — so that a restore crosses the host link once per node
 * instead of once per rank (the reference has every local rank read the whole
 * image from shard 0's shm, full_ckpt_engine.py:145-166, and moves replica
 * bytes through gloo on CPU tensors, REPL:275-332).  Slice s = image bytes
 * [s*slice_bytes, (s+1)*slice_bytes) is read through slice_bases[s], which is
 * the address image byte 0 WOULD have in that buffer (absolute offsets apply;
 * the buffer only needs to hold the slices that name it: no region reads
 * outside itself).  slice_bytes: a multiple of FCP_REGION_BYTES.  One scatter
 * launch pulls every slice over NVLink with plain per-thread loads, verifies
 * the CRC-32C of every chunk and unpacks into the live tensors; pass
 * read_prefetch = 0 when any base is peer memory (see
 * fcp_plan_set_read_prefetch).  The launch works through the image starting
 * at slice `first_slice` and wraps around: give every rank of the group a
 * different one (its own index + 1) so that at any moment the ranks read
 * different peers instead of all draining one peer's NVLink egress
 * (measured, 8 ranks, all-to-all: 174 GB/s into each GPU with a common
 * start, 405 GB/s rotated).  Synchronous.                                  */
int fcp_restore_from_slices(fcp_plan* plan, const void* const* slice_bases,
                            uint64_t n_slices, uint64_t slice_bytes,
                            int checksum, const uint32_t* expected,
                            int read_prefetch, uint64_t first_slice,
                            void* stream);

/* Host -> staging copy of image bytes [byte_begin, byte_end) only, no
 * kernel: the host-link leg of a striped restore, where every rank of a
 * replicated group fetches 1/n of the shm image (then see
 * fcp_restore_from_slices).  registered != 0: `host_src` is pinned
 * (fcp_host_register), plain copy-engine transfer; else the bytes go through
 * the plan's pinned ring with `n_workers` host threads as in
 * fcp_restore_bounced.  Synchronous.                                        */
int fcp_stage_range(fcp_plan* plan, const void* host_src, void* staging,
                    uint64_t byte_begin, uint64_t byte_end, int registered,
                    int n_workers, void* stream);

/* Restrict the following fcp_save_async calls of this plan to chunks
 * [chunk_begin, chunk_end): only those are packed, checksummed and copied to
 * the host image (staged modes).  A replicated group saves ONE image with
 * every rank writing its own stripe of it over its own host link (the
 * reference lets local rank 0 write all of it, full_ckpt_engine.py:75-88).
 * fcp_save_wait still fills one value per chunk of the image; only those of
 * the window are meaningful.  (0, UINT64_MAX) = everything (default).       */
int fcp_plan_set_window(fcp_plan* plan, uint64_t chunk_begin, uint64_t chunk_end);

/* Let kernels running on `device` dereference memory of `peer` (NVLink /
 * NVSwitch P2P).  With it, fcp_restore_from_device can be given a pointer to
 * the SURVIVING rank's staging image (mapped into this process with CUDA
 * IPC): the scatter kernel then pulls the image over NVLink, verifies and
 * unpacks it in one pass — transfer and unpack fused, no local copy of the
 * image (replaces REPL:275-332 without even the NCCL broadcast).            */
int fcp_enable_peer(int device, int peer);
/* Export / map / unmap a device buffer between the processes of one node
 * (CUDA IPC).  `dev_ptr` may point inside an allocation: the handle names the
 * allocation, *offset is dev_ptr's distance from its base.  fcp_ipc_open is
 * called with the receiving rank's device current and enables peer access to
 * the owner lazily; every open must be paired with fcp_ipc_close before the
 * owner frees the buffer.                                                   */
int fcp_ipc_export(const void* dev_ptr, uint8_t* handle64, uint64_t* offset);
int fcp_ipc_open(const uint8_t* handle64, void** base);
int fcp_ipc_close(void* base);

/* Device-side milliseconds of the phases of the last save / restore:
 * out[0] pack or scatter kernels, out[1] host-link copies, out[2] total.    */
int fcp_plan_timings(fcp_plan* plan, float* out3);
/* Tell the plan that the host image is pinned in separate pieces of
 * `piece_bytes` (counted from its base): staged copies are then split so
 * that none straddles two registrations (CUDA rejects those).  0 = one
 * registration covers the image (default).  Direct mode needs the whole
 * image in ONE registration.                                               */
int fcp_plan_set_copy_piece_bytes(fcp_plan* plan, uint64_t piece_bytes);
/* The kernels pull the next batch of their read side into L2 with one bulk
 * prefetch per warp.  Switch that off (enable = 0) when the read side is PEER
 * memory: remote lines never live in the local L2, and the prefetch made the
 * NVLink scatter 80x slower (6.7 -> 563 GB/s measured without it).          */
int fcp_plan_set_read_prefetch(fcp_plan* plan, int enable);
/* Number of kernels the last fcp_save_async / fcp_restore launched.          */
uint64_t fcp_plan_last_launches(const fcp_plan* plan);
/* Smallest slice of the staged pipelines (multiple of chunk_bytes; default
 * 1 GiB).  Slices grow 4x per step up to 16x this value; a save runs the
 * schedule small-to-large, a restore large-to-small.                        */
int fcp_plan_set_slice_bytes(fcp_plan* plan, uint64_t slice_bytes);

#ifdef __cplusplus
}
#endif
#endif /* FCP_H_ */
