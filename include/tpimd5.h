/*
 * tpimd5.h -- C-ABI of the B200-native workdir-checksum path for terraform-provider-iterative.
 *
 * This is the boundary a cgo package (`cuda/`) binds.  It replaces the CPU
 * hashing that TPI reaches through rclone v1.57.0 (go.mod:32) and Go crypto/md5 (go.mod:3) from
 *   /root/reference/task/common/machine/storage.go:123-159  machine.Transfer -> sync.CopyDir
 * The reference has no FFI today (CGO_ENABLED=0, Makefile:22,29); each entry point below names
 * the Go/rclone interface it stands in for.  Citations "[rclone]" are to the un-vendored
 * module, by package and function name.
 *
 * Conventions
 *   - plain C types only; every function returns 0 on success or a negative TPIMD5_E* code;
 *     tpimd5_last_error() gives the text for the calling thread's last failure.
 *   - digests are 16 raw bytes (A,B,C,D little-endian), i.e. Go's [md5.Size]byte.  rclone's
 *     hash strings are the lowercase hex of these bytes.
 *   - the caller owns every pointer it passes; nothing is retained after a call returns
 *     (cgo pointer rule).  The library owns staging memory, streams and device buffers.
 *   - there is NO CPU fallback: without a usable CUDA device tpimd5_init() fails.
 *   - a context may be used from many threads at once; calls that share a device overlap
 *     (each leases staging buffers, a table and a stream; nothing is locked for a call's length).
 */
#ifndef TPIMD5_H
#define TPIMD5_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define TPIMD5_API __attribute__((visibility("default")))
#else
#define TPIMD5_API
#endif

#define TPIMD5_OK        0
#define TPIMD5_EINVAL   -1   /* bad argument                                              */
#define TPIMD5_ENODEV   -2   /* no CUDA device / device id out of range                   */
#define TPIMD5_ECUDA    -3   /* a CUDA call failed; sticky for the context                */
#define TPIMD5_EIO      -4   /* open/read failed or a file's size changed; see last_error */
#define TPIMD5_ENOMEM   -5   /* host or device allocation failed                          */
#define TPIMD5_ERANGE   -6   /* size beyond what the entry point supports                 */

typedef struct tpimd5_ctx tpimd5_ctx;
typedef struct tpimd5_stream tpimd5_stream;

/* ---- lifecycle ---------------------------------------------------------------------------
 * No counterpart in the reference (rclone needs no set-up to hash).  `device_ids` == NULL
 * means devices 0..n_devices-1.  `staging_bytes_per_device` sizes each of the pinned-host /
 * device staging buffers of the ingest pipeline (0 = default 256 MiB; up to twelve per device,
 * allocated on first use.  Ingest of large files is bounded by bytes in flight divided by the
 * time of the longest single MD5 chain: twelve 256 MiB buffers against 53 ms for an 8 MiB S3
 * part leave the file reads as the limit -- an 8 GiB file on tmpfs runs at 35 GB/s end to end). */
TPIMD5_API int tpimd5_device_count(void);
TPIMD5_API int tpimd5_init(const int *device_ids, int n_devices, size_t staging_bytes_per_device, tpimd5_ctx **out);
TPIMD5_API void tpimd5_destroy(tpimd5_ctx *ctx);
TPIMD5_API int tpimd5_ctx_devices(const tpimd5_ctx *ctx);
/* Progress of the context's host-pointer calls since tpimd5_init, for a reporter like the reference's
 * `progress` goroutine (storage.go:188-211: a stats line every 10 s while a Transfer runs; its counters are
 * process-global in rclone, here they belong to the context).  Monotonic; poll from any thread at any time:
 * bytes_staged grows batch by batch while a call is running (bytes handed to the GPU), bytes_done and
 * units_done when a call returns its digests. */
typedef struct tpimd5_progress {
    uint64_t calls_started, calls_finished;
    uint64_t units_done;
    uint64_t bytes_staged, bytes_done;
} tpimd5_progress;
TPIMD5_API int tpimd5_get_progress(const tpimd5_ctx *ctx, tpimd5_progress *out);
TPIMD5_API const char *tpimd5_last_error(void);
TPIMD5_API const char *tpimd5_version(void);

/* ---- batch / manifest --------------------------------------------------------------------
 * Replaces, for a whole listing at once, [rclone] backend/local Object.Hash(ctx, hash.MD5) ->
 * fs/hash StreamTypes -> io.Copy -> crypto/md5 digest.Write/Sum, which rclone's 8 checker
 * goroutines call one file at a time under operations.CheckHashes (reached from
 * storage.go:158).  Unit i is bytes [offsets[i], offsets[i]+lengths[i]) of `base`
 * (hash_buffers) or the file paths[i] (hash_files; sizes may be NULL = fstat).  Units are
 * spread byte-balanced over the context's devices.  digests_out: 16*n bytes.
 * Read pattern of hash_files: files are read longest first, with open/pread/close by the library's reader
 * threads; files above half a staging buffer -- and, when many equal long files or parts follow one another,
 * files from 4 MiB -- are read in slices (>= 1/384 of a staging buffer each), up to 384 files side by side, not
 * front to back one after the other.  A file is open only while one of its ranges is being read. */
TPIMD5_API int tpimd5_hash_buffers(tpimd5_ctx *ctx, const uint8_t *base, const uint64_t *offsets, const uint64_t *lengths,
                        size_t n, uint8_t *digests_out);
TPIMD5_API int tpimd5_hash_files(tpimd5_ctx *ctx, const char *const *paths, const uint64_t *sizes, size_t n,
                      uint8_t *digests_out, size_t *failed_index_out);
/* Helper for bindings (no reference counterpart): the end of the furthest unit, max over i of offsets[i] +
 * lengths[i] (0 for n == 0), computed on the context's worker pool -- what a cgo / ctypes wrapper compares with the
 * length of the slice it was handed before it passes `base` to tpimd5_hash_buffers, which cannot know it.
 * TPIMD5_ERANGE if an offset + length overflows 64 bits. */
/* Diagnostic (no reference counterpart): the host-ingest ceiling of tpimd5_hash_files and tpimd5_s3_etag_file on this
 * box.  Plans the same batches and lets the same reader threads fill the same pinned staging buffers -- bytes
 * [file_offsets[i], + sizes[i]) of paths[i]; file_offsets may be NULL = whole files -- and stops there: nothing is
 * copied to the device, nothing is hashed, no digests come back.  bench.py quotes the end-to-end rates against it. */
TPIMD5_API int tpimd5_ingest_probe(tpimd5_ctx *ctx, const char *const *paths, const uint64_t *file_offsets, const uint64_t *sizes,
                                   size_t n, size_t *failed_index_out);
TPIMD5_API int tpimd5_units_extent(tpimd5_ctx *ctx, const uint64_t *offsets, const uint64_t *lengths, size_t n, uint64_t *extent_out);

/* ---- S3 multipart ------------------------------------------------------------------------
 * Replaces [rclone] backend/s3 uploadMultipart's per-chunk md5.Sum(buf) and the final
 * md5.Sum(md5s) whose hex + "-N" is the expected ETag (backends registered at
 * storage.go:21).  tpimd5_s3_part_size is rclone's chunk-size growth rule (chunk_size unless
 * size/chunk_size >= max_parts, then whole MiB above size/max_parts).  part_digests_out may be
 * NULL; otherwise 16*nparts bytes (each base64-encodes to a part's Content-MD5). */
TPIMD5_API uint64_t tpimd5_s3_part_size(uint64_t size, uint64_t chunk_size, uint32_t max_parts);
TPIMD5_API uint32_t tpimd5_s3_nparts(uint64_t size, uint64_t part_size);
TPIMD5_API int tpimd5_s3_etag_buffer(tpimd5_ctx *ctx, const uint8_t *data, uint64_t size, uint64_t part_size,
                          uint8_t *part_digests_out, uint8_t etag_md5_out[16], uint32_t *nparts_out);
TPIMD5_API int tpimd5_s3_etag_file(tpimd5_ctx *ctx, const char *path, uint64_t size, uint64_t part_size,
                        uint8_t *part_digests_out, uint8_t etag_md5_out[16], uint32_t *nparts_out);

/* ---- S3 upload plan -------------------------------------------------------------------------
 * Everything [rclone] backend/s3 Object.Update decides and computes about one local file before it
 * uploads it (call path storage.go:21 -> sync.CopyDir storage.go:158 -> operations.Copy): below
 * upload_cutoff a single PutObject with Content-MD5 = base64 of the whole-file MD5 (the ETag S3 then
 * answers with is that MD5 in hex); at or above it a multipart upload with the part size of
 * tpimd5_s3_part_size, one base64 Content-MD5 per UploadPart, the whole-file MD5 kept as metadata
 * X-Amz-Meta-Md5chksum (base64), and the ETag expected back: hex(md5(part digests)) + "-" + nparts.
 * Zero for chunk_size / upload_cutoff / max_parts selects rclone's defaults (5 MiB, 200 MiB, 10,000).
 * TPIMD5_PLAN_NO_WHOLE_MD5 is rclone's --s3-disable-checksum: skip the whole-file MD5 of a multipart
 * object (one serial chain: by far the slowest unit of a large file on a GPU).
 * tpimd5_s3_plan_parts is the pure decision (no I/O): the part count to size the output arrays with.
 * part_md5_out: 16*nparts bytes or NULL; part_content_md5_out: 25*nparts chars (NUL-terminated base64
 * strings) or NULL.  The rule is restated from knowledge of rclone v1.57.0, not verifiable here. */
#define TPIMD5_PLAN_NO_WHOLE_MD5 1u
typedef struct tpimd5_s3_plan {
    int32_t multipart;        /* 0 = single PutObject, 1 = multipart upload                      */
    int32_t has_md5;          /* whole-file MD5 computed (always for a single PutObject)         */
    uint64_t part_size;       /* multipart: bytes per part (the last may be shorter); else size  */
    uint32_t nparts;          /* multipart: parts; else 1                                        */
    uint8_t md5[16];          /* whole-file MD5                                                  */
    char content_md5[25];     /* its base64: Content-MD5 (single) / Md5chksum metadata (multi)   */
    char etag[48];            /* ETag to expect back: 32 hex, or 32 hex + "-" + nparts           */
} tpimd5_s3_plan;
TPIMD5_API uint32_t tpimd5_s3_plan_parts(uint64_t size, uint64_t chunk_size, uint64_t upload_cutoff, uint32_t max_parts,
                                         int *multipart_out, uint64_t *part_size_out);
TPIMD5_API int tpimd5_s3_upload_plan(tpimd5_ctx *ctx, const char *path, uint64_t size, uint64_t chunk_size,
                                     uint64_t upload_cutoff, uint32_t max_parts, uint32_t flags, tpimd5_s3_plan *plan_out,
                                     uint8_t *part_md5_out, char *part_content_md5_out);

/* ---- digest strings -----------------------------------------------------------------------
 * The textual forms the back-ends exchange ([rclone] backend/s3, azureblob, googlecloudstorage;
 * registered at storage.go:18-21): lowercase hex (rclone hash strings, plain S3 ETags), base64
 * (Content-MD5 per part / per object, X-Amz-Meta-Md5chksum, Azure Content-MD5, GCS md5Hash), the
 * multipart ETag "<hex>-<nparts>", and rclone's test whether an ETag is an MD5 at all (after
 * lowercasing and stripping quotes: exactly 32 hex digits; returns 1 and the bytes, else 0).
 * Every "[rclone]" rule in this header is restated from knowledge of v1.57.0; the module's source
 * is not on this machine, so fidelity to it cannot be verified here. */
TPIMD5_API void tpimd5_hex(const uint8_t md5[16], char out33[33]);
TPIMD5_API void tpimd5_base64(const uint8_t md5[16], char out25[25]);
TPIMD5_API int tpimd5_md5_from_etag(const char *etag, uint8_t md5_out[16]);
TPIMD5_API int tpimd5_s3_etag_text(const uint8_t etag_md5[16], uint32_t nparts, char *out, size_t out_len);

/* ---- streaming hasher --------------------------------------------------------------------
 * Go hash.Hash as crypto/md5 implements it: Write (any chunking, never fails for lack of
 * data), Sum (finalises a copy, the stream continues), Reset, Size()==16, BlockSize()==64.
 * This is what a whole-file Azure/GCS content-MD5 goes through ([rclone] azureblob/gcs
 * Object.Hash compare against it).  One stream is one MD5 chain, i.e. one GPU lane: use
 * tpimd5_stream_write_many to advance many streams in a single launch. */
TPIMD5_API int tpimd5_stream_open(tpimd5_ctx *ctx, tpimd5_stream **out);
TPIMD5_API int tpimd5_stream_write(tpimd5_stream *s, const void *p, size_t n);
TPIMD5_API int tpimd5_stream_write_many(tpimd5_stream *const *streams, const void *const *ptrs, const size_t *lens,
                             size_t n_streams);
TPIMD5_API int tpimd5_stream_sum(tpimd5_stream *s, uint8_t out[16]);
TPIMD5_API int tpimd5_stream_reset(tpimd5_stream *s);
TPIMD5_API void tpimd5_stream_close(tpimd5_stream *s);

/* ---- manifest diff -----------------------------------------------------------------------
 * Replaces the per-pair `srcHash != dstHash` of [rclone] fs/operations checkHashes for two
 * whole manifests: indices (ascending) of units whose digests differ. */
TPIMD5_API int tpimd5_diff(tpimd5_ctx *ctx, const uint8_t *old16, const uint8_t *new16, size_t n, uint32_t *changed_idx_out,
                size_t *n_changed_out);

/* ---- file set, manifest, transfer plan ---------------------------------------------------
 * The host logic either side of the hash.
 *
 * tpimd5_filter_*: the set of files one machine.Transfer moves.  Reproduces the rule assembly of
 * /root/reference/task/common/machine/storage.go:37-41,126-138 (built-in excludes "- /main.tf",
 * "- /terraform.tfstate*", "- /.terraform**"; a rule without "+ "/"- " prefix becomes
 * "- " + filepath.Join("/", rule)) and, when limit_subdir is non-NULL, LimitTransfer
 * (storage.go:267-280: include that directory and everything beneath it, exclude the rest), with [rclone] fs/filter glob semantics
 * (first matching rule wins, no match = include; a glob with a leading "/" matches from the root,
 * any other glob after any "/"; "*" stays within a path segment, "**" does not).  Directories
 * follow rclone's derived directory rules (_include_dir): a glob with "**" or a trailing "/" is a
 * directory rule too, an include rule also includes the directories leading to it, so a nested
 * DirectoryOut (include rules for a/b placed before the final exclude-everything) is still reached.  Pinned by the
 * expectations of storage_test.go:55-101.  Pure host code: works without a GPU.
 *
 * tpimd5_manifest_*: {relative path, size, mtime_ns, md5} per object, sorted by path.  The
 * reference persists nothing (every sync re-lists and re-hashes); this is
 * the persisted form the checkpoint loop of machine-script.sh.tpl:108-124 lacks.  _scan walks a
 * local tree like [rclone] backend/local (regular files only, symlinks skipped), applies the
 * filter and hashes every file on the GPU (tpimd5_hash_files).  _write_md5sum emits the
 * "<32 hex>  <path>\n" lines of `md5sum` as used at machine-script.sh.tpl:111.
 *
 * tpimd5_plan_transfer: per source object, what [rclone] fs/march + fs/operations NeedTransfer/
 * equal() decide for a copy src -> dst (sync.CopyDir, storage.go:158).  mode 0 = default
 * (size, then modification-time window, then hash), 1 = --checksum, 2 = --size-only.  Digest
 * comparisons run as one tpimd5_diff over all pairs that need one; ctx may be NULL only if none
 * does.  actions_out: one TPIMD5_ACT_* per source entry, in manifest (path) order. */
typedef struct tpimd5_filter tpimd5_filter;
typedef struct tpimd5_manifest tpimd5_manifest;
#define TPIMD5_ACT_SKIP          0   /* equal: nothing to do                                         */
#define TPIMD5_ACT_TRANSFER_NEW  1   /* not at the destination                                       */
#define TPIMD5_ACT_TRANSFER      2   /* present but different (size or digest)                       */
#define TPIMD5_ACT_FIX_MTIME     3   /* same content, mtime outside the window: rclone updates mtime  */
TPIMD5_API int tpimd5_filter_new(const char *const *exclude, size_t n_exclude, const char *limit_subdir, tpimd5_filter **out);
TPIMD5_API void tpimd5_filter_free(tpimd5_filter *f);
TPIMD5_API int tpimd5_filter_include(const tpimd5_filter *f, const char *rel_path);
TPIMD5_API int tpimd5_filter_include_dir(const tpimd5_filter *f, const char *rel_dir);
TPIMD5_API size_t tpimd5_filter_rule_count(const tpimd5_filter *f);
TPIMD5_API const char *tpimd5_filter_rule(const tpimd5_filter *f, size_t i);
TPIMD5_API int tpimd5_manifest_scan(tpimd5_ctx *ctx, const char *root, const tpimd5_filter *filter, tpimd5_manifest **out);
TPIMD5_API int tpimd5_manifest_from_arrays(const char *const *paths, const uint64_t *sizes, const int64_t *mtimes_ns,
                                           const uint8_t *md5s, const uint8_t *has_md5, size_t n, tpimd5_manifest **out);
TPIMD5_API void tpimd5_manifest_free(tpimd5_manifest *m);
TPIMD5_API size_t tpimd5_manifest_count(const tpimd5_manifest *m);
TPIMD5_API int tpimd5_manifest_entry(const tpimd5_manifest *m, size_t i, const char **path, uint64_t *size, int64_t *mtime_ns,
                                     uint8_t md5_out[16], int *has_md5);
TPIMD5_API int tpimd5_manifest_save(const tpimd5_manifest *m, const char *file);
TPIMD5_API int tpimd5_manifest_load(const char *file, tpimd5_manifest **out);
TPIMD5_API int tpimd5_manifest_write_md5sum(const tpimd5_manifest *m, const char *file);
TPIMD5_API int tpimd5_plan_transfer(tpimd5_ctx *ctx, const tpimd5_manifest *src, const tpimd5_manifest *dst, int mode,
                                    int64_t window_ns, uint8_t *actions_out, size_t *n_transfer_out, size_t *n_dst_only_out);

/* machine.Transfer(ctx, source, destination, exclude) for two LOCAL directories
 * (/root/reference/task/common/machine/storage.go:123-159 -> sync.CopyDir(dst, src, true)): the
 * case storage_test.go:55-101 runs and the tail of the Kubernetes Pull (task/k8s/task.go:294-299).
 * Same filter on both sides, digests only where the mode consults them, equal() per object,
 * directories created (also empty ones), modification times carried over, every copy verified by
 * digest ("corrupted on transfer" removes the copy and fails).  All digests come from the GPU.
 * Cloud remotes (":s3,...:bucket") remain rclone's job and are refused here.
 * limit_subdir: NULL, or the DirectoryOut of a Pull (LimitTransfer).  mode as in plan_transfer. */
typedef struct tpimd5_transfer_stats {
    uint64_t files_listed, bytes_listed;        /* what operations.Count logs ("Transferring %s (%d files)...") */
    uint64_t transferred, bytes_transferred;
    uint64_t skipped, mtime_fixed, dst_only;
    uint64_t bytes_hashed;                      /* bytes that went through the GPU (compare + verify) */
} tpimd5_transfer_stats;
TPIMD5_API int tpimd5_transfer_local(tpimd5_ctx *ctx, const char *source, const char *destination, const char *const *exclude,
                                     size_t n_exclude, const char *limit_subdir, int mode, tpimd5_transfer_stats *stats_out);

/* ---- device-resident variants ------------------------------------------------------------
 * Same arithmetic on data that already sits in HBM of the context's device `slot`; every
 * pointer is a device pointer and `cuda_stream` is a cudaStream_t (NULL = default stream).
 * The calls only enqueue work.  These are what kernel-only throughput is measured through;
 * the reference has no counterpart (its data is never on a device).
 *   offsets == NULL: unit i is bytes [i*stride, min((i+1)*stride, base_bytes)) (fixed-size
 *   files / S3 parts), with lengths ignored; a 16-byte aligned base and stride take the
 *   TMA-fed kernels (throughput kernel for launches that fill the SMs, latency kernel below
 *   ~57,000 units) and may be of any length: a unit longer than 2 GiB is hashed as a run of
 *   1 GiB launches on the caller's stream.  Unaligned strided units and units given by
 *   offset/length arrays may hold at most 31 GiB each (TPIMD5_ERANGE / undefined beyond for
 *   device arrays the host cannot see).  A caller that knows its listing passes large files
 *   as single strided units on streams of their own and the rest as one ragged batch
 *   (bench.py run_c4); the host-pointer calls have no limit, they chain long units themselves.
 *   Scratch: hash_buffers with offset arrays (>= 4,096 units are bucketed by length with a
 *   device radix sort), s3_etag with d_part_digests == NULL, and diff (tile counts) keep their
 *   temporaries in one per-slot buffer that grows on demand (cudaMalloc on first use; an outgrown
 *   buffer stays allocated until tpimd5_destroy, so a call never launches on freed memory).  Calls
 *   that use it on the same slot must not overlap on the GPU: keep them on one stream, or pass
 *   your own part-digest buffer, or use one slot (device) per stream.  The host-pointer calls
 *   (tpimd5_diff, tpimd5_plan_transfer, tpimd5_transfer_local, ...) never touch this buffer.
 *   Every tpimd5_dev_* call makes the slot's device current while it runs and restores the calling
 *   thread's previous device before it returns. */
TPIMD5_API int tpimd5_dev_hash_buffers(tpimd5_ctx *ctx, int slot, const void *d_base, uint64_t base_bytes,
                            const uint64_t *d_offsets, const uint64_t *d_lengths, uint64_t stride, size_t n,
                            void *d_digests_out, void *cuda_stream);
TPIMD5_API int tpimd5_dev_s3_etag(tpimd5_ctx *ctx, int slot, const void *d_data, uint64_t size, uint64_t part_size,
                       void *d_part_digests /* 16*nparts, NULL = library scratch */, void *d_etag_md5_out,
                       void *cuda_stream);
TPIMD5_API int tpimd5_dev_diff(tpimd5_ctx *ctx, int slot, const void *d_old16, const void *d_new16, size_t n,
                    uint32_t *d_changed_idx_out, uint64_t *d_n_changed_out, void *cuda_stream);
/* Synthetic workdir content generated in HBM: file f (f = first_file_index..) occupies
 * [ (f-first)*file_len, +file_len ) of d_out; bytes equal the CPU generator's. */
TPIMD5_API int tpimd5_dev_synth_fill(tpimd5_ctx *ctx, int slot, void *d_out, uint64_t seed, uint64_t first_file_index,
                          uint64_t file_len, uint64_t n_files, void *cuda_stream);

/* ---- S3 ETag for callers that must not allocate ---------------------------------------------
 * This entry point has no counterpart in the reference.
 *
 * Who it is for.  TPI tasks are mostly ML jobs, and a checkpoint hook that wants an ETag often runs
 * inside the training process itself.  There the framework's memory pool has usually claimed
 * nearly all of HBM up front, so a cudaMalloc from a library can simply fail; and cudaMalloc /
 * cudaFree stall every stream of the device while they run, which a step loop cannot afford.
 * Host programs that run under a fixed device-memory budget (the budget is charged at start-up
 * and later allocations are refused) are in the same position.  Such callers already own the
 * input buffer and a 16-byte output; what they cannot supply is the kernel's scratch.
 *
 * What it does.  Same result as tpimd5_dev_s3_etag with d_part_digests == NULL -- per-part MD5s,
 * then the MD5 of their concatenation -- but the part digests go to a fixed array that is part of
 * the library image (static __device__, 16 * TPIMD5_NOALLOC_MAX_PARTS = 160,000 bytes, present on
 * every GPU that loads the library, until the process exits).  It takes no context, creates
 * nothing, and calls no allocator of any kind.  d_data, d_etag_md5_out: device pointers on
 * `device`; d_etag_md5_out 16-byte aligned; cuda_stream: a cudaStream_t of that device (NULL =
 * default stream).  The call only enqueues two kernels.
 *
 * Limits and rules.
 *   - at most TPIMD5_NOALLOC_MAX_PARTS parts (S3's own ceiling of 10,000): more is refused with
 *     TPIMD5_ERANGE before anything is launched.  Pick part_size with tpimd5_s3_part_size.
 *   - one array per GPU, shared by every call: two calls whose kernels can run at the same time on
 *     one GPU (different streams, or different host threads) overwrite each other's part digests
 *     and both ETags are garbage.  Keep such calls on one stream or order them with events.
 *   - the array keeps the previous call's contents.  Nothing here needs it to start at zero: the
 *     second kernel reads exactly the 16 * nparts bytes the first kernel of the same call wrote.
 *   - the first call on each device (whichever thread makes it) looks up the array's device address
 *     and sets the shared-memory attributes of every kernel form the call can launch (the two
 *     TMA-fed shapes for 16-byte aligned data and part size, the general kernel for any other
 *     alignment); later calls on that device, from any thread, only enqueue.  Make that first
 *     call before any stream capture begins.
 *   - `device` is made current for the duration of the call and the thread's previous device is
 *     restored before returning, also on errors (as in every tpimd5_dev_* call). */
#define TPIMD5_NOALLOC_MAX_PARTS 10000
TPIMD5_API int tpimd5_dev_s3_etag_noalloc(int device, const void *d_data, uint64_t size, uint64_t part_size,
                                          void *d_etag_md5_out, void *cuda_stream);

/* ---- sharding ----------------------------------------------------------------------------
 * Byte-balanced assignment of units to `n_shards` GPUs (largest first, each to the least
 * loaded shard).  The library uses it inside one context; a one-process-per-GPU launcher calls
 * it to decide which rank hashes what.  rclone's counterpart is its fixed pool of 8 checker
 * goroutines pulling pairs off a channel ([rclone] fs/march, fs/sync pairChecker). */
TPIMD5_API int tpimd5_shard_lpt(const uint64_t *lengths, size_t n, int n_shards, uint32_t *shard_of_out);

/* ---- tuning / introspection --------------------------------------------------------------
 * Kernel launches issued by this context so far (bench.py's gpu_launches), and the kernel
 * variant knob used for sweeps (0 = library default). */
/* Register-only run of the 64-step MD5 body (no memory traffic): `iters` blocks per thread on
 * 128*ctas_per_sm threads per SM, d_out needs 16 bytes per thread.  Its blocks/s * 64 B is the
 * integer-issue ceiling quoted next to HBM bandwidth.  mode = step formulation. */
TPIMD5_API int tpimd5_dev_int_probe(tpimd5_ctx *ctx, int slot, int mode, int ctas_per_sm, uint32_t iters, void *d_out,
                                    uint64_t *blocks_out, void *cuda_stream);
TPIMD5_API uint64_t tpimd5_launch_count(const tpimd5_ctx *ctx);
TPIMD5_API int tpimd5_set_variant(tpimd5_ctx *ctx, int variant);

#ifdef __cplusplus
}
#endif
#endif /* TPIMD5_H */
