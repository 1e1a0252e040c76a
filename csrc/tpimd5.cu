// tpimd5.cu -- C-ABI (include/tpimd5.h) and host runtime of the B200 checksum path.
//
// Layering (bottom up):
//   md5_device.cuh / md5_lanes.cuh / aux_kernels.cuh   sm_100a kernels
//   launch_lanes(), tpimd5_dev_*                       kernel launches on device-resident data
//   Pipeline                                           host ingest: units -> staging -> H2D -> kernel
//   tpimd5_hash_* / s3 / stream / diff                 the entry points a cgo package binds
//
// The host side mirrors what [rclone] does around crypto/md5 for TPI's machine.Transfer
// (/root/reference/task/common/machine/storage.go:123-159): many independent objects, each
// read once, one digest each.  There is no CPU hashing anywhere in this file.
#include "../include/tpimd5.h"

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <chrono>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <deque>
#include <fcntl.h>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <sched.h>
#include <sys/stat.h>
#include <sys/syscall.h>
#include <thread>
#include <unistd.h>
#include <vector>

#include <cub/device/device_radix_sort.cuh>

#include "aux_kernels.cuh"
#include "md5_lanes.cuh"
#include "md5_strided_tma.cuh"
#include "md5_chain_tma.cuh"
#include "md5_ragged_chain.cuh"

using namespace tpimd5;

// ------------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------------
static thread_local std::string g_err;

static int fail(int code, const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CU(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess) return fail(TPIMD5_ECUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                                           __FILE__, __LINE__);                                          \
    } while (0)

// Makes `dev` current for the calling thread and puts the previous device back on scope exit: a caller
// (a training process, a Go thread that cgo pinned) must not find its thread switched to another GPU.
class DeviceGuard {
  public:
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev_) != cudaSuccess) prev_ = -1;
        err_ = prev_ == dev ? cudaSuccess : cudaSetDevice(dev);
        switched_ = err_ == cudaSuccess && prev_ != dev && prev_ >= 0;
    }
    ~DeviceGuard() {
        if (switched_) cudaSetDevice(prev_);
    }
    cudaError_t error() const { return err_; }
    DeviceGuard(const DeviceGuard &) = delete;
    DeviceGuard &operator=(const DeviceGuard &) = delete;

  private:
    int prev_ = -1;
    bool switched_ = false;
    cudaError_t err_ = cudaSuccess;
};
#define USE_DEVICE(dev)        \
    DeviceGuard dev_guard_(dev); \
    CU(dev_guard_.error())

extern "C" const char *tpimd5_last_error(void) { return g_err.c_str(); }
extern "C" void tpimd5_set_error_(const char *msg) { g_err = msg ? msg : ""; }   // for the other translation units
extern "C" const char *tpimd5_version(void) { return "tpimd5 0.1 (sm_100a)"; }

// ------------------------------------------------------------------------------------------------
// a small worker pool: packs units into pinned staging / reads files, in parallel
// ------------------------------------------------------------------------------------------------
class Pool {
  public:
    explicit Pool(int n) {
        for (int i = 0; i < n; ++i) th_.emplace_back([this] { run(); });
    }
    ~Pool() {
        {
            std::lock_guard<std::mutex> g(mu_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto &t : th_) t.join();
    }
    int size() const { return (int)th_.size(); }
    // fn(i) for i in [0, n), dynamically chunked; returns when all are done.  Callers may overlap: each call
    // waits for its own ITEMS only (not for the helpers it asked for -- those may still be busy with another
    // caller's work and find nothing left when they get here), and requests are served first come, first served.
    // `tick` (optional) is called by the CALLING thread between its own items and every 100 us while it waits for the
    // others: the caller can act on what has been finished so far (the ingest pipeline starts copying the filled
    // part of a staging buffer to the device while the rest is still being read).
    void parallel_for(size_t n, size_t grain, const std::function<void(size_t)> &fn, const std::function<void()> *tick = nullptr) {
        if (n == 0) return;
        if (th_.empty() || n <= grain) {
            for (size_t i = 0; i < n; ++i) { fn(i); if (tick) (*tick)(); }
            return;
        }
        auto job = std::make_shared<Job>();
        job->n = n; job->grain = grain; job->fn = &fn;
        size_t helpers = std::min<size_t>(th_.size(), (n + grain - 1) / grain);
        {
            std::lock_guard<std::mutex> g(mu_);
            for (size_t i = 0; i < helpers; ++i) q_.push_back(job);
        }
        cv_.notify_all();
        work(*job, tick);                // the caller helps too
        std::unique_lock<std::mutex> lk(job->mu);
        if (!tick) {
            job->cv.wait(lk, [&] { return job->done == job->n; });
            return;
        }
        while (job->done != job->n) {
            job->cv.wait_for(lk, std::chrono::microseconds(100));
            lk.unlock();
            (*tick)();
            lk.lock();
        }
    }

  private:
    struct Job {
        size_t n = 0, grain = 1;
        std::atomic<size_t> next{0};
        const std::function<void(size_t)> *fn = nullptr;   // the caller's; only dereferenced for an item < n, i.e. before the caller returns
        std::mutex mu;
        std::condition_variable cv;
        size_t done = 0;                 // items finished
    };
    static void work(Job &j, const std::function<void()> *tick = nullptr) {
        for (;;) {
            size_t lo = j.next.fetch_add(j.grain);
            if (lo >= j.n) break;
            size_t hi = std::min(j.n, lo + j.grain);
            for (size_t i = lo; i < hi; ++i) (*j.fn)(i);
            {
                std::lock_guard<std::mutex> g(j.mu);
                j.done += hi - lo;
                if (j.done == j.n) j.cv.notify_all();
            }
            if (tick) (*tick)();
        }
    }
    // Reader threads get a file-descriptor table of their own.  open() and close() serialise on a lock of the
    // table, which all threads of a process normally share: 16 threads reading 4 KiB files reach 1.5 M files/s
    // with the shared table and well beyond with private ones (tools/smallfile_bench.cc).  unshare(CLONE_FILES)
    // gives the thread a COPY of the table, so every inherited descriptor is closed in that copy at once --
    // otherwise the copies would keep the caller's sockets and deleted files alive.  These threads only ever
    // open, read and close the files they are given (and memcpy); they never call CUDA.
    static void private_fd_table() {
        const char *off = getenv("TPIMD5_SHARED_FDS");
        if (off && off[0] == '1') return;
        if (syscall(SYS_close_range, 1u, 0u, 0u) != -1 || errno != EINVAL) return;   // first > last: EINVAL iff the call exists
        if (unshare(CLONE_FILES) != 0) return;
        syscall(SYS_close_range, 0u, ~0u, 0u);
    }
    void run() {
        private_fd_table();
        for (;;) {
            std::shared_ptr<Job> job;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return stop_ || !q_.empty(); });
                if (stop_ && q_.empty()) return;
                job = q_.front();
                q_.pop_front();
            }
            work(*job);
        }
    }
    std::vector<std::thread> th_;
    std::deque<std::shared_ptr<Job>> q_;
    std::mutex mu_;
    std::condition_variable cv_;
    bool stop_ = false;
};

// Copy into pinned staging with non-temporal stores.  The staging buffer is written once by the CPU and read once by
// the DMA engine; an ordinary store first READS each destination line into the cache (3 bytes of DRAM traffic per
// byte packed, which is what bounds packing pageable memory), a streaming store does not.  Small copies and CPUs
// without AVX2 take memcpy.  The caller issues one sfence per work item (stage_fence) before the buffer is handed on.
#if defined(__x86_64__)
#include <immintrin.h>
__attribute__((target("avx2"))) static void copy_stream_avx2(uint8_t *dst, const uint8_t *src, size_t n) {
    const size_t head = (32 - ((uintptr_t)dst & 31)) & 31;
    if (head) { memcpy(dst, src, head); dst += head; src += head; n -= head; }
    size_t i = 0;
    for (; i + 128 <= n; i += 128) {
        const __m256i a = _mm256_loadu_si256((const __m256i *)(src + i)), b = _mm256_loadu_si256((const __m256i *)(src + i + 32));
        const __m256i c = _mm256_loadu_si256((const __m256i *)(src + i + 64)), d = _mm256_loadu_si256((const __m256i *)(src + i + 96));
        _mm256_stream_si256((__m256i *)(dst + i), a);
        _mm256_stream_si256((__m256i *)(dst + i + 32), b);
        _mm256_stream_si256((__m256i *)(dst + i + 64), c);
        _mm256_stream_si256((__m256i *)(dst + i + 96), d);
    }
    for (; i + 32 <= n; i += 32) _mm256_stream_si256((__m256i *)(dst + i), _mm256_loadu_si256((const __m256i *)(src + i)));
    if (i < n) memcpy(dst + i, src + i, n - i);
}
static bool have_avx2() {
    static const bool ok = __builtin_cpu_supports("avx2");
    return ok;
}
static void stage_copy(uint8_t *dst, const uint8_t *src, size_t n) {
    if (n >= 512 && have_avx2()) copy_stream_avx2(dst, src, n);
    else memcpy(dst, src, n);
}
static void stage_fence() { _mm_sfence(); }
#else
static void stage_copy(uint8_t *dst, const uint8_t *src, size_t n) { memcpy(dst, src, n); }
static void stage_fence() {}
#endif

// ------------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------------
constexpr int NBUF = 12;                      // staging buffers (and batch kernels) in flight per device, allocated as needed:
                                              // bytes in flight / time of the longest chain bounds ingest of large files
constexpr size_t MAX_BATCH_LANES = 1u << 20;  // descriptor rows per staging buffer
constexpr size_t DEFAULT_STAGING = 256u << 20;   // measured best for C1/C2 end to end; 512 MiB is better for multi-GiB files (see tpimd5_init)
constexpr uint64_t MAX_LANE_BYTES = (uint64_t)31 << 30;  // one unit's bytes per launch (the kernels count its 16-byte granules in an int32)

struct LaneDesc {  // device / pinned mirrors, structure of arrays
    uint64_t *off = nullptr, *len = nullptr, *prior = nullptr;
    uint32_t *flags = nullptr, *oidx = nullptr;
};

// A digest / chaining-value table of one call: device rows and their pinned host mirror.
struct TableBuf {
    uint4 *d = nullptr;
    uint8_t *h = nullptr;
    size_t cap = 0;   // rows
};
// A plain block of device memory of one call (the host-pointer diff's inputs and outputs).
struct DevBlock {
    uint8_t *d = nullptr;
    size_t cap = 0;   // bytes
};

// Host threads that belong to one device of a multi-device context: they run that device's share of a call
// (planning, staging, launches), so that a call does not start and join a thread per device.  Two per
// device, so two calls can be in flight on a device at once; bound to the CPUs the kernel lists as local to
// the GPU (/sys/bus/pci/devices/<id>/local_cpulist), where that is a proper subset of the allowed CPUs.
class DevWorkers {
  public:
    DevWorkers(int dev, int n) {
        std::vector<int> cpus = local_cpus(dev);
        for (int i = 0; i < n; ++i)
            th_.emplace_back([this, cpus] {
                if (!cpus.empty()) {
                    cpu_set_t set;
                    CPU_ZERO(&set);
                    for (int c : cpus) CPU_SET(c, &set);
                    sched_setaffinity(0, sizeof set, &set);
                }
                run();
            });
    }
    ~DevWorkers() {
        {
            std::lock_guard<std::mutex> g(mu_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto &t : th_) t.join();
    }
    void post(std::function<void()> fn) {
        {
            std::lock_guard<std::mutex> g(mu_);
            q_.push_back(std::move(fn));
        }
        cv_.notify_one();
    }

  private:
    static std::vector<int> local_cpus(int dev) {
        std::vector<int> out;
        char id[32] = {0};
        if (cudaDeviceGetPCIBusId(id, sizeof id, dev) != cudaSuccess) return out;
        for (char *c = id; *c; ++c) *c = (char)tolower(*c);
        const std::string path = std::string("/sys/bus/pci/devices/") + id + "/local_cpulist";
        FILE *f = fopen(path.c_str(), "r");
        if (!f) return out;
        char line[4096] = {0};
        const bool ok = fgets(line, sizeof line, f) != nullptr;
        fclose(f);
        if (!ok) return out;
        cpu_set_t allowed;
        CPU_ZERO(&allowed);
        if (sched_getaffinity(0, sizeof allowed, &allowed) != 0) return out;
        for (char *p = line; *p && *p != '\n';) {
            char *end;
            long a = strtol(p, &end, 10), b = a;
            if (end == p) break;
            if (*end == '-') { p = end + 1; b = strtol(p, &end, 10); }
            for (long c = a; c <= b && c < CPU_SETSIZE; ++c)
                if (CPU_ISSET((int)c, &allowed)) out.push_back((int)c);
            p = *end == ',' ? end + 1 : end;
        }
        if ((int)out.size() >= CPU_COUNT(&allowed) || out.size() < 2) out.clear();   // not a proper subset: nothing to gain
        return out;
    }
    void run() {
        for (;;) {
            std::function<void()> fn;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return stop_ || !q_.empty(); });
                if (q_.empty()) return;
                fn = std::move(q_.front());
                q_.pop_front();
            }
            fn();
        }
    }
    std::vector<std::thread> th_;
    std::deque<std::function<void()>> q_;
    std::mutex mu_;
    std::condition_variable cv_;
    bool stop_ = false;
};

// One device of a context.  Nothing here is held for the length of a call: callers LEASE what they need --
// staging buffers one at a time out of the slot's NBUF, a table, a stream -- so that several calls (Terraform
// runs ~10 resource operations in one plugin process, rclone calls from 12 goroutines) overlap on one GPU.
// `mu` only guards the free lists.
struct Slot {
    int dev = -1;
    std::mutex mu;
    std::condition_variable cv;          // a staging buffer came back
    cudaStream_t s_comp = nullptr;
    cudaStream_t s_batch[NBUF] = {};     // one per staging buffer (copy, then kernel): successive batches overlap, which
                                         // is what keeps many chains in flight when single units take tens of ms
    // ingest staging (allocated on first use)
    size_t stage_bytes = 0;
    uint8_t *h_stage[NBUF] = {}, *d_stage[NBUF] = {};
    uint8_t *h_desc_raw[NBUF] = {}, *d_desc_raw[NBUF] = {};
    cudaEvent_t ev_done[NBUF] = {};
    // a batch that carries slices of long units next to whole units hashes the two as two kernels: the slices on
    // s_batch (the next batch's slices wait for ev_slice only), the whole units beside them on s_aux
    cudaStream_t s_aux[NBUF] = {};
    cudaEvent_t ev_copied[NBUF] = {}, ev_slice[NBUF] = {}, ev_aux[NBUF] = {};
    bool buf_free[NBUF];
    int n_free = NBUF;
    int active_calls = 0;                // host calls that hold a lease on this slot right now
    int starving = 0;                    // ... of which are waiting for a first staging buffer
    // free lists, kept until tpimd5_destroy
    std::vector<TableBuf> free_tables;
    std::vector<DevBlock> free_blocks;
    std::vector<cudaStream_t> free_streams;
    // scratch of the tpimd5_dev_* calls: S3 part digests, diff tile counts, bucketing temporaries
    std::mutex scratch_mu;
    uint8_t *d_scratch = nullptr;
    size_t scratch_cap = 0;
    std::vector<uint8_t *> retired;   // outgrown scratch buffers: a call that captured the old pointer may still be
                                      // launching on it, so they are only freed in tpimd5_destroy
    std::unique_ptr<DevWorkers> workers;  // contexts over several devices: this device's host threads
    Slot() { for (bool &b : buf_free) b = true; }
};

struct tpimd5_ctx {
    std::vector<std::unique_ptr<Slot>> slots;
    std::unique_ptr<Pool> pool;
    size_t staging = DEFAULT_STAGING;
    std::atomic<uint64_t> launches{0};
    std::atomic<int> variant{0};
    // what tpimd5_get_progress reports (monotonic since tpimd5_init)
    std::atomic<uint64_t> calls_started{0}, calls_finished{0}, units_done{0}, bytes_staged{0}, bytes_done{0};
    // measured rate at which packed batches fill (bytes per second into pinned staging; a running mean over
    // batches of >= 16 MiB), from files and from caller memory: what the slice planner weighs chains against.
    // Starts at the page-cache rate of the boxes this was tuned on; a slow disk brings it down within a call.
    std::atomic<double> feed_files{40e9}, feed_mem{40e9};
    // longest tensor-map row, and bytes of a longer unit hashed per launch (tests shrink both through the
    // environment to run the segment logic on small data)
    uint64_t row_max = (uint64_t)1 << 31, chain_seg = (uint64_t)1 << 30;
};

static constexpr size_t desc_bytes_per_lane = 3 * sizeof(uint64_t) + 2 * sizeof(uint32_t);

static void carve_desc(uint8_t *raw, size_t cap, LaneDesc &d) {
    d.off = (uint64_t *)raw;
    d.len = d.off + cap;
    d.prior = d.len + cap;
    d.flags = (uint32_t *)(d.prior + cap);
    d.oidx = d.flags + cap;
}

// ---- leases ----------------------------------------------------------------------------------------
static int take_table(Slot &s, size_t rows, TableBuf &out) {
    {
        std::lock_guard<std::mutex> g(s.mu);
        int best = -1;
        for (int i = 0; i < (int)s.free_tables.size(); ++i)
            if (s.free_tables[i].cap >= rows && (best < 0 || s.free_tables[i].cap < s.free_tables[best].cap)) best = i;
        if (best >= 0) {
            out = s.free_tables[best];
            s.free_tables.erase(s.free_tables.begin() + best);
            return 0;
        }
        if (!s.free_tables.empty()) {            // all too small: let the largest of them go instead of piling up
            int big = 0;
            for (int i = 1; i < (int)s.free_tables.size(); ++i)
                if (s.free_tables[i].cap > s.free_tables[big].cap) big = i;
            out = s.free_tables[big];
            s.free_tables.erase(s.free_tables.begin() + big);
        }
    }
    if (out.d) cudaFree(out.d);
    if (out.h) cudaFreeHost(out.h);
    out = TableBuf{};
    size_t cap = (std::max<size_t>(rows, 4096) + 4095) & ~(size_t)4095;
    CU(cudaMalloc(&out.d, cap * 16));
    if (cudaHostAlloc(&out.h, cap * 16, cudaHostAllocDefault) != cudaSuccess) {
        cudaFree(out.d);
        out = TableBuf{};
        return fail(TPIMD5_ENOMEM, "pinned host memory for %zu digests", cap);
    }
    out.cap = cap;
    return 0;
}
static void give_table(Slot &s, TableBuf &t) {
    if (!t.d) return;
    std::lock_guard<std::mutex> g(s.mu);
    s.free_tables.push_back(t);
    t = TableBuf{};
}

static int take_block(Slot &s, size_t bytes, DevBlock &out) {
    {
        std::lock_guard<std::mutex> g(s.mu);
        int best = -1;
        for (int i = 0; i < (int)s.free_blocks.size(); ++i)
            if (s.free_blocks[i].cap >= bytes && (best < 0 || s.free_blocks[i].cap < s.free_blocks[best].cap)) best = i;
        if (best >= 0) {
            out = s.free_blocks[best];
            s.free_blocks.erase(s.free_blocks.begin() + best);
            return 0;
        }
        if (!s.free_blocks.empty()) {
            out = s.free_blocks.back();
            s.free_blocks.pop_back();
        }
    }
    if (out.d) cudaFree(out.d);
    out = DevBlock{};
    const size_t cap = (std::max<size_t>(bytes, 1u << 20) + 255) & ~(size_t)255;
    CU(cudaMalloc(&out.d, cap));
    out.cap = cap;
    return 0;
}
static void give_block(Slot &s, DevBlock &b) {
    if (!b.d) return;
    std::lock_guard<std::mutex> g(s.mu);
    s.free_blocks.push_back(b);
    b = DevBlock{};
}

static int take_stream(Slot &s, cudaStream_t &out) {
    {
        std::lock_guard<std::mutex> g(s.mu);
        if (!s.free_streams.empty()) {
            out = s.free_streams.back();
            s.free_streams.pop_back();
            return 0;
        }
    }
    CU(cudaStreamCreateWithFlags(&out, cudaStreamNonBlocking));
    return 0;
}
static void give_stream(Slot &s, cudaStream_t &st) {
    if (!st) return;
    std::lock_guard<std::mutex> g(s.mu);
    s.free_streams.push_back(st);
    st = nullptr;
}

static int ensure_scratch(Slot &s, size_t bytes) {
    if (bytes <= s.scratch_cap) return 0;
    // grow geometrically so that outgrown buffers (kept until destroy, see Slot::retired) stay few
    size_t cap = (std::max<size_t>(std::max<size_t>(bytes, 2 * s.scratch_cap), 1u << 20) + 255) & ~(size_t)255;
    uint8_t *fresh = nullptr;
    CU(cudaMalloc(&fresh, cap));
    if (s.d_scratch) s.retired.push_back(s.d_scratch);
    s.d_scratch = fresh;
    s.scratch_cap = cap;
    return 0;
}

// Staging buffer i of a slot: device side, descriptors and events on first use; the pinned host side only
// when something has to be packed (direct copies out of pinned caller memory never need it).
static int ensure_buffer(tpimd5_ctx *ctx, Slot &s, int i, bool need_host) {
    if (!s.stage_bytes) s.stage_bytes = ctx->staging;
    if (!s.d_stage[i]) {
        CU(cudaMalloc(&s.d_stage[i], s.stage_bytes + 256));
        CU(cudaHostAlloc(&s.h_desc_raw[i], MAX_BATCH_LANES * desc_bytes_per_lane, cudaHostAllocDefault));
        CU(cudaMalloc(&s.d_desc_raw[i], MAX_BATCH_LANES * desc_bytes_per_lane));
        CU(cudaEventCreateWithFlags(&s.ev_done[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&s.ev_copied[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&s.ev_slice[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&s.ev_aux[i], cudaEventDisableTiming));
    }
    if (need_host && !s.h_stage[i]) CU(cudaHostAlloc(&s.h_stage[i], s.stage_bytes + 256, cudaHostAllocDefault));
    return 0;
}

extern "C" int tpimd5_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

extern "C" int tpimd5_init(const int *device_ids, int n_devices, size_t staging_bytes_per_device, tpimd5_ctx **out) {
    if (!out || n_devices < 1) return fail(TPIMD5_EINVAL, "tpimd5_init: need out and n_devices >= 1");
    *out = nullptr;
    int have = 0;
    cudaError_t e = cudaGetDeviceCount(&have);
    if (e != cudaSuccess || have < 1)
        return fail(TPIMD5_ENODEV, "tpimd5_init: no CUDA device (%s); this library has no CPU fallback",
                    e == cudaSuccess ? "count is 0" : cudaGetErrorString(e));
    std::unique_ptr<tpimd5_ctx> ctx(new tpimd5_ctx);
    if (staging_bytes_per_device) {
        if (staging_bytes_per_device < (1u << 20)) return fail(TPIMD5_EINVAL, "tpimd5_init: staging below 1 MiB");
        ctx->staging = (staging_bytes_per_device + 255) & ~(size_t)255;
    }
    for (int i = 0; i < n_devices; ++i) {
        int dev = device_ids ? device_ids[i] : i;
        if (dev < 0 || dev >= have) return fail(TPIMD5_ENODEV, "tpimd5_init: device %d not present (%d visible)", dev, have);
        cudaDeviceProp prop;
        CU(cudaGetDeviceProperties(&prop, dev));
        if (prop.major != 10)
            return fail(TPIMD5_ENODEV, "tpimd5_init: device %d is sm_%d%d; this build carries sm_100a code only", dev,
                        prop.major, prop.minor);
        std::unique_ptr<Slot> s(new Slot);
        s->dev = dev;
        s->stage_bytes = ctx->staging;
        USE_DEVICE(dev);
        CU(cudaStreamCreateWithFlags(&s->s_comp, cudaStreamNonBlocking));
        for (int b = 0; b < NBUF; ++b) CU(cudaStreamCreateWithFlags(&s->s_batch[b], cudaStreamNonBlocking));
        for (int b = 0; b < NBUF; ++b) CU(cudaStreamCreateWithFlags(&s->s_aux[b], cudaStreamNonBlocking));
        ctx->slots.push_back(std::move(s));
    }
    if (const char *e = getenv("TPIMD5_ROW_MAX")) {
        const uint64_t v = strtoull(e, nullptr, 0);
        if (v >= 64 && v <= ((uint64_t)1 << 31)) ctx->row_max = v;
    }
    if (const char *e = getenv("TPIMD5_CHAIN_SEG")) {
        const uint64_t v = strtoull(e, nullptr, 0);
        if (v >= 64 && v <= ((uint64_t)1 << 30) && v % 64 == 0) ctx->chain_seg = v;
    }
    if (ctx->slots.size() > 1)
        for (auto &sp : ctx->slots) sp->workers.reset(new DevWorkers(sp->dev, 2));
    unsigned hc = std::thread::hardware_concurrency();
    int workers = (int)std::min<unsigned>(hc ? hc : 4, 32);
    ctx->pool.reset(new Pool(workers > 1 ? workers - 1 : 0));
    *out = ctx.release();
    return 0;
}

extern "C" void tpimd5_destroy(tpimd5_ctx *ctx) {
    if (!ctx) return;
    for (auto &sp : ctx->slots) sp->workers.reset();
    for (auto &sp : ctx->slots) {
        Slot &s = *sp;
        DeviceGuard guard(s.dev);
        if (s.s_comp) cudaStreamSynchronize(s.s_comp);
        for (int b = 0; b < NBUF; ++b)
            if (s.s_batch[b]) { cudaStreamSynchronize(s.s_batch[b]); cudaStreamDestroy(s.s_batch[b]); }
        for (int b = 0; b < NBUF; ++b)
            if (s.s_aux[b]) { cudaStreamSynchronize(s.s_aux[b]); cudaStreamDestroy(s.s_aux[b]); }
        for (int i = 0; i < NBUF; ++i) {
            if (s.h_stage[i]) cudaFreeHost(s.h_stage[i]);
            if (s.d_stage[i]) cudaFree(s.d_stage[i]);
            if (s.h_desc_raw[i]) cudaFreeHost(s.h_desc_raw[i]);
            if (s.d_desc_raw[i]) cudaFree(s.d_desc_raw[i]);
            if (s.ev_done[i]) cudaEventDestroy(s.ev_done[i]);
            if (s.ev_copied[i]) cudaEventDestroy(s.ev_copied[i]);
            if (s.ev_slice[i]) cudaEventDestroy(s.ev_slice[i]);
            if (s.ev_aux[i]) cudaEventDestroy(s.ev_aux[i]);
        }
        for (TableBuf &t : s.free_tables) { cudaFree(t.d); cudaFreeHost(t.h); }
        for (DevBlock &b : s.free_blocks) cudaFree(b.d);
        for (cudaStream_t st : s.free_streams) { cudaStreamSynchronize(st); cudaStreamDestroy(st); }
        if (s.d_scratch) cudaFree(s.d_scratch);
        for (uint8_t *p : s.retired) cudaFree(p);
        if (s.s_comp) cudaStreamDestroy(s.s_comp);
    }
    delete ctx;
}

extern "C" int tpimd5_ctx_devices(const tpimd5_ctx *ctx) { return ctx ? (int)ctx->slots.size() : 0; }
extern "C" int tpimd5_get_progress(const tpimd5_ctx *ctx, tpimd5_progress *out) {
    if (!ctx || !out) return fail(TPIMD5_EINVAL, "get_progress: null pointer");
    out->calls_started = ctx->calls_started.load();
    out->calls_finished = ctx->calls_finished.load();
    out->units_done = ctx->units_done.load();
    out->bytes_staged = ctx->bytes_staged.load();
    out->bytes_done = ctx->bytes_done.load();
    return 0;
}
extern "C" uint64_t tpimd5_launch_count(const tpimd5_ctx *ctx) { return ctx ? ctx->launches.load() : 0; }
extern "C" int tpimd5_set_variant(tpimd5_ctx *ctx, int variant) {
    if (!ctx || variant < 0 || variant > 599) return fail(TPIMD5_EINVAL, "tpimd5_set_variant: 100*loader(0 cp.async, 2 tensor map for strided units, 3 / 4 latency kernel for strided units, 2 / 4 blocks per iteration, 5 ragged latency kernel) + 10*mode(0,4,6) + shape");
    ctx->variant = variant;
    return 0;
}

// ------------------------------------------------------------------------------------------------
// kernel launch
// ------------------------------------------------------------------------------------------------
// Shared-memory attributes of one kernel instantiation, for the device that is current.
template <int CH, bool ALIGNED, int WARPS, int MINB, int MODE>
static int configure_lanes() {
    using Cfg = LaneCfg<CH, ALIGNED, WARPS>;
    auto kern = md5_lanes_kernel<CH, ALIGNED, WARPS, MINB, MODE>;
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    return 0;
}

// The launch alone (the instantiation must have been configured on this device).
template <int CH, bool ALIGNED, int WARPS, int MINB, int MODE>
static int enqueue_lanes(const LaneArgs &a, cudaStream_t st) {
    using Cfg = LaneCfg<CH, ALIGNED, WARPS>;
    const uint64_t per_cta = WARPS * 32;
    const uint64_t grid = (a.n + per_cta - 1) / per_cta;
    if (grid == 0) return 0;
    if (grid > 0x7fffffffull) return fail(TPIMD5_ERANGE, "too many units for one launch (%llu)", (unsigned long long)a.n);
    md5_lanes_kernel<CH, ALIGNED, WARPS, MINB, MODE><<<(unsigned)grid, WARPS * 32, Cfg::SMEM_BYTES, st>>>(a);
    CU(cudaGetLastError());
    return 0;
}

// Devices on which an instantiation has been configured: one bit per device, process-wide (the
// attribute belongs to the device's primary context, not to the calling thread).
struct ConfiguredSet {
    std::atomic<uint64_t> bits{0};
    bool has(int dev) const { return dev >= 0 && dev < 64 && (bits.load(std::memory_order_acquire) >> dev) & 1; }
    void add(int dev) { if (dev >= 0 && dev < 64) bits.fetch_or(1ull << dev, std::memory_order_release); }
};

template <int CH, bool ALIGNED, int WARPS, int MINB, int MODE>
static int launch_cfg(const LaneArgs &a, cudaStream_t st) {
    static ConfiguredSet configured;                 // per template instantiation
    int dev = 0;
    CU(cudaGetDevice(&dev));
    if (!configured.has(dev)) {
        int rc = configure_lanes<CH, ALIGNED, WARPS, MINB, MODE>();   // idempotent: a race only repeats it
        if (rc) return rc;
        configured.add(dev);
    }
    return enqueue_lanes<CH, ALIGNED, WARPS, MINB, MODE>(a, st);
}

template <bool ALIGNED, int MODE>
static int launch_variant(int variant, const LaneArgs &a, cudaStream_t st) {
    switch (variant) {
        case 1: return launch_cfg<128, ALIGNED, 4, 4, MODE>(a, st);   // 16 warps/SM, 144 KiB smem/SM
        case 2: return launch_cfg<256, ALIGNED, 4, 3, MODE>(a, st);   // 12 warps/SM, 204 KiB
        case 3: return launch_cfg<64, ALIGNED, 4, 5, MODE>(a, st);    // 20 warps/SM, 100 KiB
        default:
        case 4: return launch_cfg<128, ALIGNED, 4, 5, MODE>(a, st);   // 20 warps/SM, 180 KiB (default)
        case 7: return launch_cfg<64, ALIGNED, 4, 8, MODE>(a, st);    // 32 warps/SM, 64 regs, 160 KiB
        case 9: return launch_cfg<128, ALIGNED, 4, 6, MODE>(a, st);   // 24 warps/SM, 80 regs, 216 KiB
    }
}

// ---- TMA path for equal-sized units at a fixed stride (md5_strided_tma.cuh) -----------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
    static std::once_flag once;
    static EncodeTiledFn fn = nullptr;
    std::call_once(once, [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    });
    return fn;
}

// Can this strided layout be described to the TMA unit?  (A row longer than 2^31 bytes cannot; such units
// are hashed in segments, see launch_long_units.)
static bool tma_layout_ok(const void *base, uint64_t stride, uint64_t n) {
    return ((uintptr_t)base & 15) == 0 && (stride & 15) == 0 && stride >= 16 && stride < ((uint64_t)1 << 40) && n >= 1 &&
           n <= ((uint64_t)1 << 31) && encode_tiled_fn() != nullptr;
}
template <int WARPS, int MINB, int NSTAGE, int MODE, int TILES = 1>
static int launch_tma_cfg(const CUtensorMap &tmap, uint64_t n, uint32_t unit_len, uint4 *table, cudaStream_t st) {
    using Cfg = StridedCfg<WARPS, NSTAGE, TILES>;
    auto kern = md5_strided_tma_kernel<WARPS, MINB, NSTAGE, MODE, TILES>;
    static ConfiguredSet configured;
    int dev = 0;
    CU(cudaGetDevice(&dev));
    if (!configured.has(dev)) {
        CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
        CU(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        configured.add(dev);
    }
    const uint64_t per_cta = WARPS * 32;
    const uint64_t grid = (n + per_cta - 1) / per_cta;
    kern<<<(unsigned)grid, WARPS * 32, Cfg::SMEM_BYTES, st>>>(tmap, n, unit_len, table);
    CU(cudaGetLastError());
    return 0;
}

// Shared-memory attributes of one instantiation of the producer-warp kernel, for the device that is current.
template <int WARPS, int MINB, int NSTAGE, int TILES, int MODE, int U>
static int configure_chain() {
    using Cfg = ChainCfg<WARPS, NSTAGE, TILES>;
    auto kern = md5_chain_tma_kernel<WARPS, MINB, NSTAGE, TILES, MODE, U>;
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    return 0;
}

// The launch alone (the instantiation must have been configured on this device).
template <int WARPS, int MINB, int NSTAGE, int TILES, int MODE, int U>
static int enqueue_chain(const CUtensorMap &tmap, uint64_t n, uint32_t unit_len, uint4 *table, cudaStream_t st, uint64_t prior,
                         uint32_t flags, const CUtensorMap *tail = nullptr, uint32_t tail_len = 0) {
    using Cfg = ChainCfg<WARPS, NSTAGE, TILES>;
    const uint64_t per_cta = WARPS * 32;
    const uint64_t grid = (n + per_cta - 1) / per_cta + (tail && tail_len ? 1 : 0);   // + one CTA for a ragged last unit (table[n])
    md5_chain_tma_kernel<WARPS, MINB, NSTAGE, TILES, MODE, U><<<(unsigned)grid, Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(
        tmap, n, unit_len, table, prior, flags, tail && tail_len ? *tail : tmap, tail ? tail_len : 0u);
    CU(cudaGetLastError());
    return 0;
}

template <int WARPS, int MINB, int NSTAGE, int TILES, int MODE, int U = 2>
static int launch_chain_cfg(const CUtensorMap &tmap, uint64_t n, uint32_t unit_len, uint4 *table, cudaStream_t st,
                            uint64_t prior = 0, uint32_t flags = LANE_FINAL, const CUtensorMap *tail = nullptr, uint32_t tail_len = 0) {
    static ConfiguredSet configured;
    int dev = 0;
    CU(cudaGetDevice(&dev));
    if (!configured.has(dev)) {
        int rc = configure_chain<WARPS, MINB, NSTAGE, TILES, MODE, U>();
        if (rc) return rc;
        configured.add(dev);
    }
    return enqueue_chain<WARPS, MINB, NSTAGE, TILES, MODE, U>(tmap, n, unit_len, table, st, prior, flags, tail, tail_len);
}

// The latency kernel (md5_chain_tma.cuh) in one of its shapes; mode 4 unless 0 or 6 is asked for; U blocks
// per loop iteration.
template <int U>
static int launch_chain(const CUtensorMap &tmap, uint64_t n, uint32_t unit_len, uint4 *table, int shape, int mode, cudaStream_t st) {
#define TPI_CHAIN(M)                                                                              \
    switch (shape) {                                                                              \
        default:                                                                                  \
        case 1: return launch_chain_cfg<4, 1, 2, 4, M, U>(tmap, n, unit_len, table, st);          \
        case 2: return launch_chain_cfg<4, 1, 4, 2, M, U>(tmap, n, unit_len, table, st);          \
        case 3: return launch_chain_cfg<4, 3, 2, 2, M, U>(tmap, n, unit_len, table, st);          \
        case 4: return launch_chain_cfg<4, 5, 2, 1, M, 2>(tmap, n, unit_len, table, st);          \
        case 5: return launch_chain_cfg<4, 6, 2, 1, M, 2>(tmap, n, unit_len, table, st);          \
        case 6: return launch_chain_cfg<2, 2, 4, 2, M, U>(tmap, n, unit_len, table, st);          \
        case 7: return launch_chain_cfg<8, 3, 2, 1, M, 2>(tmap, n, unit_len, table, st);          \
        case 8: return launch_chain_cfg<2, 4, 2, 2, M, U>(tmap, n, unit_len, table, st);          \
        case 9: return launch_chain_cfg<3, 1, 2, 4, M, U>(tmap, n, unit_len, table, st);          \
    }
    if (mode == 0) { TPI_CHAIN(0) }
    if (mode == 6) { TPI_CHAIN(6) }
    TPI_CHAIN(4)
#undef TPI_CHAIN
}

// Does a launch of n units fit the latency kernel in one wave (at most 12 hashing warps per SM)?
static bool chain_launch_fits(uint64_t n) { return (n + 31) / 32 <= (uint64_t)148 * 12; }

// The latency kernel in the shape that suits the launch's warp count (see launch_strided_tma).
struct RowsDesc;
static bool encode_rows3(CUtensorMap *tmap, const RowsDesc &d, uint32_t tiles);

static int launch_chain_default(const CUtensorMap &tmap, uint64_t n, uint32_t ul, uint4 *table, cudaStream_t st, uint64_t prior,
                                uint32_t flags, const CUtensorMap *tail = nullptr, uint32_t tail_len = 0, const RowsDesc *rows = nullptr) {
    const uint64_t warps = (n + 31) / 32;
    CUtensorMap t3;
    // latency shapes with several boxes per stage: one 3-D request per stage where the rows allow it
    const uint32_t tiles = warps <= 148 ? 2 : warps <= 888 ? 4 : 2;
    const bool box3d = warps <= (uint64_t)148 * 12 && rows && encode_rows3(&t3, *rows, tiles);
    const CUtensorMap &tm = box3d ? t3 : tmap;
    if (box3d) flags |= LANE_BOX3D;
    if (warps <= 148) return launch_chain_cfg<2, 2, 4, 2, 4, 4>(tm, n, ul, table, st, prior, flags, tail, tail_len);
    if (warps <= 592) return launch_chain_cfg<4, 1, 2, 4, 4, 4>(tm, n, ul, table, st, prior, flags, tail, tail_len);
    if (warps <= 888) return launch_chain_cfg<3, 1, 2, 4, 4, 4>(tm, n, ul, table, st, prior, flags, tail, tail_len);
    if (warps <= (uint64_t)148 * 12) return launch_chain_cfg<4, 3, 2, 2, 6, 4>(tm, n, ul, table, st, prior, flags, tail, tail_len);
    // enough chains to fill the SMs: the integer pipes are the limit -- 4 hashing warps + the producer per CTA, six
    // CTAs per SM (24 hashing warps), one box per stage, step mode 6 (the additions spread over both integer pipes)
    return launch_chain_cfg<4, 6, 2, 1, 6, 2>(tmap, n, ul, table, st, prior, flags, tail, tail_len);
}

static int encode_rows(CUtensorMap *tmap, const void *base, uint64_t row_bytes, uint64_t rows, uint64_t pitch) {
    const cuuint64_t gdim[2] = {row_bytes, rows};
    const cuuint64_t gstride[1] = {pitch};
    const cuuint32_t box[2] = {128, 32};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = encode_tiled_fn()(tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(base), gdim, gstride, box, estr,
                                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(TPIMD5_ECUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
    return 0;
}

// The 3-D view of the same rows for the latency shapes: (byte within a 128-byte chunk, row, chunk), box 128 x 32 x
// `tiles`, so that a stage of `tiles` boxes is ONE TMA request.  Only when a row is a whole number of chunks (the view
// would otherwise reach up to 127 bytes past the last row's end).
struct RowsDesc {
    const void *base;
    uint64_t row_bytes, rows, pitch;
};
static bool encode_rows3(CUtensorMap *tmap, const RowsDesc &d, uint32_t tiles) {
    if (tiles < 2 || d.row_bytes % 128 != 0 || d.row_bytes == 0 || encode_tiled_fn() == nullptr) return false;
    const cuuint64_t gdim[3] = {128, d.rows, d.row_bytes / 128};
    const cuuint64_t gstride[2] = {d.pitch, 128};
    const cuuint32_t box[3] = {128, 32, tiles};
    const cuuint32_t estr[3] = {1, 1, 1};
    return encode_tiled_fn()(tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void *>(d.base), gdim, gstride, box, estr,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// Units too long for one tensor-map row (large files: C4's 32 GiB object): CHAIN_SEG bytes of every unit
// per launch, the chaining values carried from launch to launch in the digest table.
static int launch_long_units(tpimd5_ctx *ctx, const void *base, uint64_t stride, uint64_t unit_len, uint64_t n, uint4 *table,
                             cudaStream_t st) {
    for (uint64_t done = 0; done < unit_len; done += ctx->chain_seg) {
        const uint64_t seg = std::min(ctx->chain_seg, unit_len - done);
        CUtensorMap tmap;
        int rc = encode_rows(&tmap, (const uint8_t *)base + done, seg, n, stride);
        if (rc) return rc;
        const uint32_t flags = (done ? LANE_CONT : 0u) | (done + seg == unit_len ? LANE_FINAL : 0u);
        const RowsDesc rows{(const uint8_t *)base + done, seg, n, stride};
        rc = launch_chain_default(tmap, n, (uint32_t)seg, table, st, done, flags, nullptr, 0, &rows);
        if (rc) return rc;
        ctx->launches.fetch_add(1);
    }
    return 0;
}

// n units of unit_len bytes, unit u at base + u*stride.  shape: 1..4, mode 0 or 4.
// tail_len != 0: the unit after the n whole ones, cut short by the buffer's end (it starts at base + n*stride), goes
// into the same launch and its digest to table[n] -- default kernel choice only.
static int launch_strided_tma(tpimd5_ctx *ctx, const void *base, uint64_t stride, uint64_t unit_len, uint64_t n, uint4 *table,
                              int loader, int shape, int mode, cudaStream_t st, uint64_t tail_len = 0) {
    CUtensorMap tmap, tail;
    if (int rc0 = encode_rows(&tmap, base, unit_len, n, stride)) return rc0;
    if (tail_len)
        if (int rc0 = encode_rows(&tail, (const uint8_t *)base + n * stride, tail_len, 1, (tail_len + 15) & ~(uint64_t)15)) return rc0;
    int rc;
    const RowsDesc rows{base, unit_len, n, stride};
    if (tail_len) {
        rc = launch_chain_default(tmap, n, (uint32_t)unit_len, table, st, 0, LANE_FINAL, &tail, (uint32_t)tail_len, &rows);
        if (rc == 0) ctx->launches.fetch_add(1);
        return rc;
    }
    if (loader >= 3) {
        rc = loader == 3 ? launch_chain<2>(tmap, n, (uint32_t)unit_len, table, shape, mode, st)
                         : launch_chain<4>(tmap, n, (uint32_t)unit_len, table, shape, mode, st);
        if (rc == 0) ctx->launches.fetch_add(1);
        return rc;
    }
    // Too few chains to fill the SMs: every chain runs at its dependent-op latency, so the launch goes to
    // the latency kernel (md5_chain_tma.cuh; mode 4 = LOP3 -> IADD3 -> LEA.HI, the shortest chain), in the
    // shape that spreads its warps thinnest over the SM sub-partitions (measured, profiles/r2_chain_sweep.md):
    //   <= 148 warps             2 hashing warps per CTA, <= 74 CTAs: one warp per sub-partition, and two
    //                            such launches still fit side by side
    //   <= 592 warps             4 hashing warps per CTA, one CTA per SM (128 KiB of stages each); two
    //                            launches of one 64 GiB S3 object each (2 x 64 CTAs) still get an SM per CTA
    //   <= 888 warps             3 per CTA (+ the producer: one warp per sub-partition), up to two CTAs per SM
    //   <= 1,776 warps           4 per CTA, up to three CTAs per SM; three warps share a sub-partition, where
    //                            the integer pipes start to bind: the step formulation that splits the adds (6)
    // Beyond that (12 warps per SM would no longer hold the launch in one wave) the throughput kernel below.
    // All with four blocks per loop iteration.
    // Beyond that the same kernel in its throughput shape (see launch_chain_default).
    if (shape == 0) {
        rc = launch_chain_default(tmap, n, (uint32_t)unit_len, table, st, 0, LANE_FINAL, nullptr, 0, &rows);
        if (rc == 0) ctx->launches.fetch_add(1);
        return rc;
    }
    // explicit shapes of round 1's self-issuing kernel (md5_strided_tma.cuh), kept for comparison sweeps
#define TPI_TMA(M)                                                                          \
    switch (shape) {                                                                        \
        case 1: rc = launch_tma_cfg<4, 4, 3, M>(tmap, n, (uint32_t)unit_len, table, st); break;  \
        default:                                                                                 \
        case 2: rc = launch_tma_cfg<4, 5, 2, M>(tmap, n, (uint32_t)unit_len, table, st); break;  \
        case 3: rc = launch_tma_cfg<4, 3, 4, M>(tmap, n, (uint32_t)unit_len, table, st); break;  \
        case 4: rc = launch_tma_cfg<4, 6, 2, M>(tmap, n, (uint32_t)unit_len, table, st); break;  \
        case 5: rc = launch_tma_cfg<4, 3, 2, M, 2>(tmap, n, (uint32_t)unit_len, table, st); break; \
        case 6: rc = launch_tma_cfg<4, 1, 2, M, 4>(tmap, n, (uint32_t)unit_len, table, st); break; \
        case 7: rc = launch_tma_cfg<4, 2, 3, M, 2>(tmap, n, (uint32_t)unit_len, table, st); break; \
    }
    if (mode == 0) {
        TPI_TMA(0)
    } else if (mode == 6) {
        TPI_TMA(6)
    } else {
        TPI_TMA(4)
    }
#undef TPI_TMA
    if (rc == 0) ctx->launches.fetch_add(1);
    return rc;
}

// Kernel choice.  variant 0 = library default.  Otherwise 100*loader + 10*mode + shape, where loader
// 0 forces the cp.async kernel everywhere and 2 allows the tensor-map kernel for strided units.
static void decode_variant(int v, int &loader, int &mode, int &shape) {
    if (v == 0) { loader = 2; mode = 6; shape = 0; return; }
    loader = v / 100; mode = (v / 10) % 10; shape = v % 10;
}

template <int WARPS, int MINB, int NSTAGE, int SB, int MODE>
static int launch_ragged_chain_cfg(const LaneArgs &a, cudaStream_t st) {
    using Cfg = RaggedCfg<WARPS, NSTAGE, SB>;
    auto kern = md5_ragged_chain_kernel<WARPS, MINB, NSTAGE, SB, MODE, 4>;
    static ConfiguredSet configured;
    int dev = 0;
    CU(cudaGetDevice(&dev));
    if (!configured.has(dev)) {
        CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
        CU(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        configured.add(dev);
    }
    const uint64_t per_cta = WARPS * 32;
    const uint64_t grid = (a.n + per_cta - 1) / per_cta;
    kern<<<(unsigned)grid, Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(a);
    CU(cudaGetLastError());
    return 0;
}

// `aligned`: every unit starts on a 16-byte boundary of device memory.
static int launch_lanes(tpimd5_ctx *ctx, const LaneArgs &a0, bool aligned, cudaStream_t st) {
    if (a0.n == 0) return 0;
    LaneArgs a = a0;
    int loader, mode, shape;
    decode_variant(ctx->variant.load(), loader, mode, shape);
    // A ragged batch with too few chains to fill the SMs (a workdir of a few hundred files of all sizes): the
    // ragged latency kernel (md5_ragged_chain.cuh).  It needs per-unit offset/length arrays and 16-byte aligned
    // units, which is what the host pipeline's packer produces.  Variant 500 + x forces it for tests.
    const uint64_t warps = (a.n + 31) / 32;
    if ((ctx->variant.load() == 0 || loader == 5) && aligned && a.off && a.len && warps <= (uint64_t)148 * 12 &&
        a.base_bytes < ((uint64_t)1 << 32)) {      // (the kernel keeps a unit's length in 32 bits: one staging buffer's worth)
        int rc;
        if (loader == 5 && shape == 2) rc = launch_ragged_chain_cfg<4, 3, 2, 256, 4>(a, st);
        else if (loader == 5 && shape == 3) rc = launch_ragged_chain_cfg<4, 3, 2, 256, 6>(a, st);
        else if (loader == 5 || warps <= 592) rc = launch_ragged_chain_cfg<4, 1, 2, 512, 4>(a, st);
        else if (warps <= 888) rc = launch_ragged_chain_cfg<4, 3, 2, 256, 4>(a, st);
        else rc = launch_ragged_chain_cfg<4, 3, 2, 256, 6>(a, st);
        if (rc == 0) ctx->launches.fetch_add(1);
        return rc;
    }
    // library default: too few chains to fill the SMs -> the formulation with the shortest dependent
    // chain (mode 4); enough of them -> the one that spreads the additions over both integer pipes (6)
    if (ctx->variant.load() == 0 && a.n < (uint64_t)148 * 16 * 32) { mode = 4; shape = 2; }   // + 256-byte stages: half the per-stage work
    int rc;
    if (aligned) {
        switch (mode) {
            case 0: rc = launch_variant<true, 0>(shape, a, st); break;
            case 4: rc = launch_variant<true, 4>(shape, a, st); break;
            default: rc = launch_variant<true, 6>(shape, a, st); break;
        }
    } else {
        switch (mode) {
            case 0: rc = launch_variant<false, 0>(shape, a, st); break;
            case 4: rc = launch_variant<false, 4>(shape, a, st); break;
            default: rc = launch_variant<false, 6>(shape, a, st); break;
        }
    }
    if (rc == 0) ctx->launches.fetch_add(1);
    return rc;
}

// Units of `stride` bytes back to back: unit i = [i*stride, min((i+1)*stride, base_bytes)).  Whole
// units go through the tensor-map kernel when the layout allows it, a ragged last unit (and any
// layout TMA cannot describe) through the general kernel.
static int launch_strided(tpimd5_ctx *ctx, const void *base, uint64_t base_bytes, uint64_t stride, uint64_t n, uint4 *table,
                          cudaStream_t st) {
    if (n == 0) return 0;
    LaneArgs a{};
    a.base = (const uint8_t *)base; a.base_bytes = base_bytes; a.stride = stride; a.table = table; a.n = n;
    const bool aligned = ((uintptr_t)base & 15) == 0 && (stride & 15) == 0;
    int loader, mode, shape;
    decode_variant(ctx->variant.load(), loader, mode, shape);
    const uint64_t n_full = std::min<uint64_t>(n, base_bytes / stride);
    if (loader >= 2 && n_full && tma_layout_ok(base, stride, n_full) && (stride <= ctx->row_max || chain_launch_fits(n_full))) {
        const uint64_t tail = base_bytes - n_full * stride;       // bytes of the unit the buffer's end cuts short
        if (ctx->variant.load() == 0 && stride <= ctx->row_max && n == n_full + 1 && tail >= 1)
            return launch_strided_tma(ctx, base, stride, stride, n_full, table, loader, shape, mode, st, tail);   // one launch for both
        int rc = stride <= ctx->row_max ? launch_strided_tma(ctx, base, stride, stride, n_full, table, loader, shape, mode, st)
                                       : launch_long_units(ctx, base, stride, stride, n_full, table, st);
        if (rc || n_full == n) return rc;
        a.base += n_full * stride; a.base_bytes -= n_full * stride; a.table += n_full; a.n = n - n_full;
    }
    // what is left is units that the buffer's end cuts short: one ragged unit (then possibly empty ones)
    if (loader >= 2 && a.n == 1 && a.base_bytes >= 16 && tma_layout_ok(a.base, stride, 1)) {
        const uint64_t pitch = (a.base_bytes + 15) & ~(uint64_t)15;       // one row: the pitch is never applied
        return a.base_bytes <= ctx->row_max ? launch_strided_tma(ctx, a.base, pitch, a.base_bytes, 1, a.table, loader, shape, mode, st)
                                           : launch_long_units(ctx, a.base, pitch, a.base_bytes, 1, a.table, st);
    }
    if (a.base_bytes > MAX_LANE_BYTES && a.stride > MAX_LANE_BYTES)
        return fail(TPIMD5_ERANGE, "a unit above %llu bytes must start on a 16-byte boundary and have a stride that is a multiple of 16",
                    (unsigned long long)MAX_LANE_BYTES);
    return launch_lanes(ctx, a, aligned, st);
}

static Slot *get_slot(tpimd5_ctx *ctx, int slot) {
    if (!ctx || slot < 0 || slot >= (int)ctx->slots.size()) return nullptr;
    return ctx->slots[slot].get();
}

// ------------------------------------------------------------------------------------------------
// device-resident entry points
// ------------------------------------------------------------------------------------------------
extern "C" int tpimd5_dev_hash_buffers(tpimd5_ctx *ctx, int slot, const void *d_base, uint64_t base_bytes,
                                       const uint64_t *d_offsets, const uint64_t *d_lengths, uint64_t stride, size_t n,
                                       void *d_digests_out, void *cuda_stream) {
    Slot *s = get_slot(ctx, slot);
    if (!s) return fail(TPIMD5_EINVAL, "dev_hash_buffers: bad context or slot");
    if (n == 0) return 0;
    if (!d_base || !d_digests_out) return fail(TPIMD5_EINVAL, "dev_hash_buffers: null pointer");
    if (((uintptr_t)d_digests_out & 15) != 0) return fail(TPIMD5_EINVAL, "dev_hash_buffers: digests_out must be 16-byte aligned");
    bool aligned;
    if (d_offsets) {
        if (!d_lengths) return fail(TPIMD5_EINVAL, "dev_hash_buffers: offsets without lengths");
        aligned = false;  // offsets live on the device; the general kernel takes any alignment
    } else {
        if (stride == 0) return fail(TPIMD5_EINVAL, "dev_hash_buffers: stride 0");
        if ((n - 1) > base_bytes / stride) return fail(TPIMD5_EINVAL, "dev_hash_buffers: n*stride runs past base_bytes");
        aligned = ((uintptr_t)d_base & 15) == 0 && (stride & 15) == 0;
    }
    USE_DEVICE(s->dev);
    LaneArgs a{};
    a.base = (const uint8_t *)d_base; a.base_bytes = base_bytes;
    a.off = d_offsets; a.len = d_lengths; a.stride = stride;
    a.table = (uint4 *)d_digests_out; a.n = n;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    if (!d_offsets) return launch_strided(ctx, d_base, base_bytes, stride, n, (uint4 *)d_digests_out, st);
    if (n < 4096 || n > 0xffffffffull) return launch_lanes(ctx, a, aligned, st);
    // Many units of arbitrary length: bucket them by block count first (aux_kernels.cuh) so that a
    // warp's 32 units end together.  All temporaries live in the slot's scratch.
    size_t cub_bytes = 0;
    CU(cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, (const uint32_t *)nullptr, (uint32_t *)nullptr, (const uint32_t *)nullptr,
                                       (uint32_t *)nullptr, (int)n, 0, 32, st));
    const size_t n4 = (n * 4 + 255) & ~(size_t)255, n8 = (n * 8 + 255) & ~(size_t)255;
    uint8_t *base_scratch;
    {
        std::lock_guard<std::mutex> g(s->scratch_mu);      // the scratch is shared by calls on this slot
        int rc = ensure_scratch(*s, 4 * n4 + 2 * n8 + cub_bytes + 256);
        if (rc) return rc;
        base_scratch = s->d_scratch;
    }
    uint32_t *keys_in = (uint32_t *)base_scratch, *keys_out = (uint32_t *)(base_scratch + n4);
    uint32_t *idx_in = (uint32_t *)(base_scratch + 2 * n4), *perm = (uint32_t *)(base_scratch + 3 * n4);
    uint64_t *off_s = (uint64_t *)(base_scratch + 4 * n4), *len_s = (uint64_t *)(base_scratch + 4 * n4 + n8);
    void *cub_tmp = base_scratch + 4 * n4 + 2 * n8;
    const unsigned grid = (unsigned)((n + 255) / 256);
    bucket_keys_kernel<<<grid, 256, 0, st>>>(d_lengths, n, keys_in, idx_in);
    CU(cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, keys_in, keys_out, idx_in, perm, (int)n, 0, 32, st));
    bucket_gather_kernel<<<grid, 256, 0, st>>>(d_offsets, d_lengths, perm, n, off_s, len_s);
    CU(cudaGetLastError());
    ctx->launches.fetch_add(3);
    a.off = off_s; a.len = len_s; a.out_idx = perm;
    return launch_lanes(ctx, a, aligned, st);
}

extern "C" uint64_t tpimd5_s3_part_size(uint64_t size, uint64_t chunk_size, uint32_t max_parts) {
    // [rclone] backend/s3 uploadMultipart: keep chunk_size unless that needs max_parts or more parts
    if (chunk_size == 0 || max_parts == 0) return 0;
    uint64_t part = chunk_size;
    if (size / part >= max_parts) part = (((size / max_parts) >> 20) + 1) << 20;
    return part;
}

extern "C" uint32_t tpimd5_s3_nparts(uint64_t size, uint64_t part_size) {
    if (part_size == 0) return 0;
    if (size == 0) return 1;
    uint64_t n = (size + part_size - 1) / part_size;
    return n > 0xffffffffull ? 0 : (uint32_t)n;
}

extern "C" int tpimd5_dev_s3_etag(tpimd5_ctx *ctx, int slot, const void *d_data, uint64_t size, uint64_t part_size,
                                  void *d_part_digests, void *d_etag_md5_out, void *cuda_stream) {
    Slot *s = get_slot(ctx, slot);
    if (!s) return fail(TPIMD5_EINVAL, "dev_s3_etag: bad context or slot");
    if (!d_etag_md5_out || (!d_data && size)) return fail(TPIMD5_EINVAL, "dev_s3_etag: null pointer");
    if (part_size == 0) return fail(TPIMD5_EINVAL, "dev_s3_etag: part_size 0");
    const uint32_t np = tpimd5_s3_nparts(size, part_size);
    if (!np) return fail(TPIMD5_ERANGE, "dev_s3_etag: too many parts");
    USE_DEVICE(s->dev);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    uint8_t *parts = (uint8_t *)d_part_digests;
    if (!parts) {
        // library scratch: shared by the slot, so serialise callers that rely on it
        std::lock_guard<std::mutex> g(s->scratch_mu);
        int rc = ensure_scratch(*s, (size_t)np * 16);
        if (rc) return rc;
        parts = s->d_scratch;
    }
    if (((uintptr_t)parts & 15) || ((uintptr_t)d_etag_md5_out & 15)) return fail(TPIMD5_EINVAL, "dev_s3_etag: outputs must be 16-byte aligned");
    // K2: one lane per part
    int rc = launch_strided(ctx, d_data, size, part_size, np, (uint4 *)parts, st);
    if (rc) return rc;
    // K2b: digest of digests -- the part digests are one more message of 16*np bytes
    // (one chain: the latency kernel, fed by TMA like the parts)
    return launch_strided(ctx, parts, (uint64_t)np * 16, (uint64_t)np * 16, 1, (uint4 *)d_etag_md5_out, st);
}

// three launches: per-tile counts, exclusive scan of the counts, ordered scatter
static int diff_launch(tpimd5_ctx *ctx, const void *d_old16, const void *d_new16, size_t n, uint32_t *d_idx,
                       uint64_t *d_count, uint64_t *tile_counts, cudaStream_t st) {
    const uint64_t tiles = (n + DIFF_TILE - 1) / DIFF_TILE;
    uint32_t *bitmap = diff_bitmap(tile_counts, tiles);
    diff_tile_kernel<false><<<(unsigned)tiles, DIFF_THREADS, 0, st>>>((const uint4 *)d_old16, (const uint4 *)d_new16, n, tile_counts, bitmap, nullptr);
    diff_scan_kernel<<<1, 1024, 0, st>>>(tile_counts, tiles, d_count);
    diff_tile_kernel<true><<<(unsigned)tiles, DIFF_THREADS, 0, st>>>(nullptr, nullptr, n, tile_counts, bitmap, d_idx);
    CU(cudaGetLastError());
    ctx->launches.fetch_add(3);
    return 0;
}

extern "C" int tpimd5_dev_diff(tpimd5_ctx *ctx, int slot, const void *d_old16, const void *d_new16, size_t n,
                               uint32_t *d_changed_idx_out, uint64_t *d_n_changed_out, void *cuda_stream) {
    Slot *s = get_slot(ctx, slot);
    if (!s) return fail(TPIMD5_EINVAL, "dev_diff: bad context or slot");
    if (!d_n_changed_out) return fail(TPIMD5_EINVAL, "dev_diff: null count pointer");
    if (n > 0xffffffffull) return fail(TPIMD5_ERANGE, "dev_diff: more than 2^32 units");
    USE_DEVICE(s->dev);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    if (n == 0) {
        CU(cudaMemsetAsync(d_n_changed_out, 0, sizeof(uint64_t), st));
        return 0;
    }
    if (!d_old16 || !d_new16 || !d_changed_idx_out) return fail(TPIMD5_EINVAL, "dev_diff: null pointer");
    if (((uintptr_t)d_old16 | (uintptr_t)d_new16) & 15) return fail(TPIMD5_EINVAL, "dev_diff: manifests must be 16-byte aligned");
    uint64_t *counts;
    {
        // tile counts live in the slot's scratch, which calls on this slot share
        std::lock_guard<std::mutex> g(s->scratch_mu);
        int rc = ensure_scratch(*s, ((n + DIFF_TILE - 1) / DIFF_TILE) * DIFF_SCRATCH_PER_TILE);
        if (rc) return rc;
        counts = (uint64_t *)s->d_scratch;
    }
    return diff_launch(ctx, d_old16, d_new16, n, d_changed_idx_out, d_n_changed_out, counts, st);
}

extern "C" int tpimd5_dev_synth_fill(tpimd5_ctx *ctx, int slot, void *d_out, uint64_t seed, uint64_t first_file_index,
                                     uint64_t file_len, uint64_t n_files, void *cuda_stream) {
    Slot *s = get_slot(ctx, slot);
    if (!s) return fail(TPIMD5_EINVAL, "dev_synth_fill: bad context or slot");
    if (!d_out || ((uintptr_t)d_out & 15)) return fail(TPIMD5_EINVAL, "dev_synth_fill: output must be 16-byte aligned");
    if (file_len == 0 || (file_len & 15)) return fail(TPIMD5_EINVAL, "dev_synth_fill: file_len must be a positive multiple of 16");
    if (n_files == 0) return 0;
    USE_DEVICE(s->dev);
    const uint64_t total16 = (file_len >> 4) * n_files;
    const uint64_t want = (total16 + 255) / 256;
    const unsigned grid = (unsigned)std::min<uint64_t>(want, 148u * 64u);
    synth_fill_kernel<<<grid, 256, 0, (cudaStream_t)cuda_stream>>>((uint4 *)d_out, seed, first_file_index, file_len, n_files);
    CU(cudaGetLastError());
    ctx->launches.fetch_add(1);
    return 0;
}

extern "C" int tpimd5_dev_int_probe(tpimd5_ctx *ctx, int slot, int mode, int ctas_per_sm, uint32_t iters, void *d_out,
                                    uint64_t *blocks_out, void *cuda_stream) {
    Slot *s = get_slot(ctx, slot);
    if (!s || !d_out || ctas_per_sm < 1 || ctas_per_sm > 16 || (mode != 0 && (mode < 4 || mode > 8)))
        return fail(TPIMD5_EINVAL, "dev_int_probe: bad argument");
    USE_DEVICE(s->dev);
    int sms = 0;
    CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->dev));
    const unsigned grid = (unsigned)(sms * ctas_per_sm);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    switch (mode) {
        case 0: int_probe_kernel<0><<<grid, 128, 0, st>>>(iters, (uint4 *)d_out); break;
        case 4: int_probe_kernel<4><<<grid, 128, 0, st>>>(iters, (uint4 *)d_out); break;
        case 5: int_probe_kernel<5><<<grid, 128, 0, st>>>(iters, (uint4 *)d_out); break;
        case 7: int_probe_kernel<7><<<grid, 128, 0, st>>>(iters, (uint4 *)d_out); break;
        case 8: int_probe_kernel<8><<<grid, 128, 0, st>>>(iters, (uint4 *)d_out); break;
        default: int_probe_kernel<6><<<grid, 128, 0, st>>>(iters, (uint4 *)d_out); break;
    }
    CU(cudaGetLastError());
    ctx->launches.fetch_add(1);
    if (blocks_out) *blocks_out = (uint64_t)grid * 128u * iters;
    return 0;
}

// ------------------------------------------------------------------------------------------------
// host ingest pipeline
// ------------------------------------------------------------------------------------------------
// A job is a list of independent units.  Unit bytes come from memory (base+off) or from a file.
// Optional per-unit chaining input (streams): starting state, bytes hashed before, and whether
// to finalise.  Results (digest or chaining value) land in out[16*i].
struct Job {
    size_t n = 0;
    const uint8_t *base = nullptr;
    const uint64_t *off = nullptr;       // may be null when paths are given
    const uint64_t *len = nullptr;
    const char *const *paths = nullptr;
    const uint64_t *file_off = nullptr;    // with paths: unit i is bytes [file_off[i], +len[i]) of its file (S3 parts)
    // streams: unit i is prefix[i][0..prefix_len[i]) followed by base+off[i]; len[i] counts both
    const uint8_t *const *prefix = nullptr;
    const uint64_t *prefix_len = nullptr;
    const uint8_t *init_states = nullptr;  // 16*n or null
    const uint64_t *prior = nullptr;       // or null
    const uint8_t *nonfinal = nullptr;     // per unit: 1 = leave as chaining value; or null
    uint8_t *out = nullptr;
    size_t failed_index = (size_t)-1;
    bool ingest_only = false;              // tpimd5_ingest_probe: plan and fill the staging buffers, copy and hash nothing
};

struct Piece {
    uint32_t unit;      // index into the device's unit list
    uint64_t src_off;   // bytes of the unit before this piece
    uint64_t len;
    uint64_t dst;       // byte offset in the staging buffer
    bool last;
};

// Byte-balanced split of units over devices: largest first, each to the least-loaded device
// (greedy LPT).  Returns per-device unit lists in descending size order, so that a device's
// longest chain starts first (its latency bounds the device's finish time).
static void shard_lpt(const uint64_t *len, size_t n, int ndev, std::vector<std::vector<uint32_t>> &out) {
    out.assign(ndev, {});
    if (ndev == 1) {
        out[0].resize(n);
        for (size_t i = 0; i < n; ++i) out[0][i] = (uint32_t)i;
        return;
    }
    std::vector<uint32_t> order(n);
    for (size_t i = 0; i < n; ++i) order[i] = (uint32_t)i;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return len[a] > len[b]; });
    std::vector<uint64_t> load(ndev, 0);
    for (uint32_t u : order) {
        int best = 0;
        for (int d = 1; d < ndev; ++d)
            if (load[d] < load[best]) best = d;
        out[best].push_back(u);
        load[best] += len[u] + 72;  // padding blocks cost too: keeps runs of empty files spread out
    }
}

extern "C" int tpimd5_shard_lpt(const uint64_t *lengths, size_t n, int n_shards, uint32_t *shard_of_out) {
    if (!lengths || !shard_of_out || n_shards < 1) return fail(TPIMD5_EINVAL, "shard_lpt: bad argument");
    if (n > 0xffffffffull) return fail(TPIMD5_ERANGE, "shard_lpt: more than 2^32 units");
    std::vector<std::vector<uint32_t>> sh;
    shard_lpt(lengths, n, n_shards, sh);
    for (int d = 0; d < n_shards; ++d)
        for (uint32_t u : sh[d]) shard_of_out[u] = (uint32_t)d;
    return 0;
}

// Reads exactly n bytes at `off`.  With `probe`, asks for one byte more: -3 if the file has it (it is
// longer than listed), and end-of-file after exactly n bytes is the normal outcome.  -2: shorter.
static int read_fully(int fd, uint8_t *dst, uint64_t off, uint64_t n, bool probe = false) {
    uint64_t want = n + (probe ? 1 : 0), got = 0;
    while (got < want) {
        const uint64_t left = want - got;
        ssize_t r = pread(fd, dst + got, left > (1u << 30) ? (1u << 30) : (size_t)left, (off_t)(off + got));
        if (r < 0) {
            if (errno == EINTR) continue;
            return -1;
        }
        if (r == 0) break;
        got += (uint64_t)r;
    }
    if (got < n) return -2;
    if (got > n) return -3;
    return 0;
}

// What one host call holds on one device while it runs: a table, a stream of its own, and the staging
// buffers it has drawn so far (oldest first).  Giving everything back is the destructor's job, after every
// stream the call used has drained -- whatever went wrong, no copy out of caller memory and no kernel is
// left in flight when the call returns, so the caller may free its buffers right away.
struct Lease {
    Slot &s;
    TableBuf table;
    cudaStream_t stream = nullptr;
    std::vector<int> bufs;      // staging buffers held, in the order they will be reused
    size_t next = 0;            // index in bufs of the oldest batch in flight
    int share_seen = NBUF;      // staging buffers this call may hold, as of its last acquire
    bool counted = false;
    explicit Lease(Slot &slot) : s(slot) {}
    Lease(const Lease &) = delete;
    Lease &operator=(const Lease &) = delete;
    int begin(size_t rows) {
        {
            std::lock_guard<std::mutex> g(s.mu);
            ++s.active_calls;
            counted = true;
        }
        int rc = take_table(s, rows, table);
        if (rc) return rc;
        return take_stream(s, stream);
    }
    // A staging buffer for the next batch.  A call may draw a free buffer while it is within its share (NBUF /
    // calls, at least one) or while more are free than there are calls that hold none at all; otherwise it reuses
    // its own oldest buffer as soon as that batch's kernel has finished (polled, so that a buffer freed by ANOTHER
    // call in the meantime is taken instead of waiting out a long chain of its own), handing it back instead if
    // it is over its share and someone is starving.  A call that holds nothing waits for a buffer to come back.
    int acquire(int &bi_out, bool need_host) {
        std::unique_lock<std::mutex> lk(s.mu);
        bool starving = false;
        for (;;) {
            const int share = std::max(1, NBUF / std::max(1, s.active_calls));
            share_seen = share;
            if (s.n_free > 0 && (bufs.empty() || (int)bufs.size() < share || s.n_free > s.starving)) {
                // prefer a buffer whose memory exists already: allocating 256 MiB of pinned memory takes tens of
                // milliseconds and stalls every other CUDA call of the process while it runs
                int pick = -1;
                for (int pass = 0; pass < 3 && pick < 0; ++pass)
                    for (int i = 0; i < NBUF && pick < 0; ++i)
                        if (s.buf_free[i] && (pass == 2 || (s.d_stage[i] && (pass == 1 || !need_host || s.h_stage[i])))) pick = i;
                s.buf_free[pick] = false;
                --s.n_free;
                if (starving) --s.starving;
                bufs.insert(bufs.begin() + next, pick);        // in front of the oldest: it is the newest now
                ++next;
                if (next >= bufs.size()) next = 0;
                bi_out = pick;
                return 0;
            }
            if (!bufs.empty()) {
                const int bi = bufs[next];
                lk.unlock();
                const cudaError_t q = cudaEventQuery(s.ev_done[bi]);
                lk.lock();
                if (q == cudaSuccess) {
                    if ((int)bufs.size() > share && s.starving > 0) {      // over the share while someone has nothing: hand it back
                        bufs.erase(bufs.begin() + next);
                        if (next >= bufs.size()) next = 0;
                        s.buf_free[bi] = true;
                        ++s.n_free;
                        s.cv.notify_all();
                        continue;
                    }
                    next = (next + 1) % bufs.size();
                    bi_out = bi;
                    return 0;
                }
                if (q != cudaErrorNotReady) {
                    lk.unlock();
                    return fail(TPIMD5_ECUDA, "cudaEventQuery: %s", cudaGetErrorString(q));
                }
            } else if (!starving) {
                starving = true;
                ++s.starving;
            }
            s.cv.wait_for(lk, std::chrono::microseconds(200));
        }
    }
    ~Lease() {
        for (int bi : bufs) {
            if (s.s_aux[bi]) cudaStreamSynchronize(s.s_aux[bi]);
            if (s.s_batch[bi]) cudaStreamSynchronize(s.s_batch[bi]);
        }
        if (stream) cudaStreamSynchronize(stream);
        give_table(s, table);
        give_stream(s, stream);
        {
            std::lock_guard<std::mutex> g(s.mu);
            for (int bi : bufs) { s.buf_free[bi] = true; ++s.n_free; }
            if (counted) --s.active_calls;
        }
        s.cv.notify_all();
    }
};

// TPIMD5_TRACE=1: phase timings of the ingest pipeline on stderr (diagnostics only)
static bool trace_on() {
    static const bool on = [] { const char *e = getenv("TPIMD5_TRACE"); return e && e[0] == '1'; }();
    return on;
}
static double trace_now() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// Runs one device's share of a job.  `units` are indices into the job.
// Units of a job that one device takes: an explicit list of indices, or the range [base, base + n) of the
// caller's order (the common case: no per-unit vector is built for a listing of millions of files).
struct UnitList {
    const uint32_t *idx = nullptr;
    uint32_t base = 0;
    size_t n = 0;
    uint32_t operator[](size_t i) const { return idx ? idx[i] : base + (uint32_t)i; }
    size_t size() const { return n; }
    bool empty() const { return n == 0; }
};

static int run_device_impl(tpimd5_ctx *ctx, Slot &s, Job &job, const UnitList &units_in) {
    if (units_in.empty()) return 0;
    const double tr0 = trace_now();
    double tr_plan = 0, tr_fill = 0, tr_wait = 0;
    // Files: longest first.  One file is one serial chain (~125 MB/s), so the device's finish time is
    // bounded by its longest file; starting those in the first batch lets the short ones fill in behind.
    std::vector<uint32_t> by_size;
    if (job.paths) {
        by_size.resize(units_in.size());
        for (size_t i = 0; i < by_size.size(); ++i) by_size[i] = units_in[i];
        auto longer = [&](uint32_t a, uint32_t b) { return job.len[a] > job.len[b]; };
        if (!std::is_sorted(by_size.begin(), by_size.end(), longer))          // (a listing of equal files: nothing to sort)
            std::stable_sort(by_size.begin(), by_size.end(), longer);
    }
    const UnitList units = job.paths ? UnitList{by_size.data(), 0, by_size.size()} : units_in;
    USE_DEVICE(s.dev);
    Lease lease(s);                       // declared after the device guard: it drains its streams before the device is put back
    int rc = lease.begin(units.size());
    if (rc) return rc;
    uint4 *const d_table = lease.table.d;
    uint8_t *const h_table = lease.table.h;
    const size_t nu = units.size();
    const bool from_files = job.paths != nullptr;

    // chaining values to start from
    if (job.init_states) {
        for (size_t i = 0; i < nu; ++i) memcpy(h_table + 16 * i, job.init_states + 16 * (size_t)units[i], 16);
        CU(cudaMemcpyAsync(d_table, h_table, nu * 16, cudaMemcpyHostToDevice, lease.stream));
        CU(cudaStreamSynchronize(lease.stream));    // in place before any batch's kernel reads it (a few KiB)
    }

    // Direct mode: units already lie close together and in order in caller memory, so each batch is
    // one H2D copy straight from there (full speed when the caller's buffer is pinned).  Otherwise
    // (scattered units, files) the pool packs units 16-byte aligned into pinned staging first.
    bool direct = !from_files && !job.prefix;
    if (direct && job.base) {
        // Pageable caller memory (a Go slice, a numpy array): the driver would stage a direct copy through
        // its own bounce buffers on one thread.  The worker pool packing into our pinned staging is faster.
        cudaPointerAttributes attr;
        const bool pinned = cudaPointerGetAttributes(&attr, job.base + job.off[units[0]]) == cudaSuccess &&
                            (attr.type == cudaMemoryTypeHost || attr.type == cudaMemoryTypeManaged);
        cudaGetLastError();
        const char *force = getenv("TPIMD5_PAGEABLE_DIRECT");
        if (!pinned && !(force && force[0] == '1')) direct = false;
    }
    if (direct) {
        // (millions of units: checked by the pool, 64 Ki units per item)
        constexpr size_t CHUNK = 1u << 16;
        std::atomic<uint64_t> a_bytes{0};
        std::atomic<bool> in_order{true};
        std::function<void(size_t)> check = [&](size_t c) {
            uint64_t t = 0;
            bool ok = true;
            const size_t hi = std::min(nu, (c + 1) * CHUNK);
            for (size_t i = c * CHUNK; i < hi; ++i) {
                const uint32_t u = units[i];
                if (i && job.off[u] < job.off[units[i - 1]] + job.len[units[i - 1]]) ok = false;
                t += job.len[u];
            }
            a_bytes.fetch_add(t);
            if (!ok) in_order.store(false);
        };
        ctx->pool->parallel_for((nu + CHUNK - 1) / CHUNK, 1, check);
        const uint64_t bytes = a_bytes.load();
        direct = in_order.load();
        if (direct) {
            const uint64_t span = job.off[units[nu - 1]] + job.len[units[nu - 1]] - job.off[units[0]];
            if (span > bytes + bytes / 4 + (1u << 20)) direct = false;
        }
    }

    int tma_loader, tma_mode, tma_shape;
    decode_variant(ctx->variant.load(), tma_loader, tma_mode, tma_shape);
    std::vector<Piece> pieces;
    std::vector<uint32_t> order;
    size_t ui = 0;            // next unit
    uint64_t consumed = 0;    // bytes of units[ui] already planned
    int batch = 0, prev_bi = -1;
    uint64_t batch_bytes = 0, total_bytes = 0;
    std::atomic<int> io_err{0};
    std::atomic<size_t> io_idx{(size_t)-1};

    // Packed batches (files, scattered or pageable units) take LONG units in slices, many units side by side:
    // a unit is one serial chain (~160 MB/s once its bytes are on the device), so a long unit that arrives
    // whole, after the ones before it, finishes its-length / chain-rate after everything before it was read
    // (the last 8 MiB part of a multipart object: 53 ms after the file's last byte; the tenth 1 GiB file of a
    // workdir: after nine others' chains).  Sliced, every long unit's chain starts in the first batches and
    // keeps pace with the reads; a batch's kernel then runs for slice / chain-rate, which the slice bounds
    // below set to about the time the batch takes to read and copy (256 MiB at ~43 GB/s ~ 1 MiB of chain).
    struct Active { uint32_t ui; uint64_t consumed; int last_bi; };
    std::vector<Active> active;
    // Which units are sliced: always those above half a staging buffer; from 4 MiB when whole units in listing
    // order would end later than slices do by these estimates (reads at the context's measured fill rate, ~40 GB/s
    // from the page cache; one chain ~0.16 GB/s):
    //   whole units:  the latest (everything up to and including unit i read) + (unit i's chain)
    //   slices:       max(everything read, the longest chain) + one slice's chain
    // Whole units win when a few long chains bound the call and come first anyway (a workdir of files up to 8 MiB,
    // longest first); slices win when many equal long units arrive one after another (the parts of one object).
    uint64_t stripe_min = s.stage_bytes / 2;
    // How many units are sliced side by side.  A batch of A slices takes A x slice / 40 GB/s to read and copy and
    // slice / 0.16 GB/s to hash, and each batch's slices wait for the previous batch's: with 256 the kernels
    // just keep up with full-speed reads and a chain gets its bytes at the rate it can hash them -- right when the
    // longest chain bounds the call.  When the reads bound it, 384 thinner slices keep the kernels ahead of them
    // (an 8 GiB object in 8 MiB parts: the kernels finished 31 ms after the last read with 256, 5 ms with 384).
    size_t max_active = 256;
    if (!direct) {
        const uint64_t cand = std::min<uint64_t>((uint64_t)4 << 20, s.stage_bytes / 4);
        constexpr size_t CHUNK = 1u << 16;
        std::atomic<uint64_t> a_total{0}, a_max{0}, a_long{0};
        std::function<void(size_t)> sum = [&](size_t c) {
            uint64_t t = 0, m = 0, nl = 0;
            const size_t hi = std::min(nu, (c + 1) * CHUNK);
            for (size_t i = c * CHUNK; i < hi; ++i) {
                const uint64_t l = job.len[units[i]];
                t += l; m = std::max(m, l); nl += l >= cand;
            }
            a_total.fetch_add(t); a_long.fetch_add(nl);
            uint64_t cur = a_max.load();
            while (m > cur && !a_max.compare_exchange_weak(cur, m)) {}
        };
        ctx->pool->parallel_for((nu + CHUNK - 1) / CHUNK, 1, sum);
        if (a_long.load() > 1 && a_max.load() < stripe_min) {
            const double R = std::max(0.05e9, (from_files ? ctx->feed_files : ctx->feed_mem).load()), r = 0.16e9, total = (double)a_total.load();
            double whole = total / R, pre = 0;
            for (size_t i = 0; i < nu; ++i) {
                const uint64_t l = job.len[units[i]];
                pre += (double)l;
                if (l >= cand) whole = std::max(whole, pre / R + (double)l / r);
                else if (from_files) break;                       // files come longest first
            }
            const double sliced = std::max(total / R, (double)a_max.load() / r) + (double)(s.stage_bytes / 256) / r;
            static const int force = [] { const char *e = getenv("TPIMD5_SLICE"); return e ? atoi(e) : -1; }();   // diagnostics: 1 always, 0 never
            if (force < 0 ? whole > 1.08 * sliced : force == 1) stripe_min = cand;
            if (total / R > (double)a_max.load() / r) {
                // ... in equal groups, so that the last group is not a short one whose chains outlast its reads
                const size_t groups = (a_long.load() + 383) / 384;
                max_active = std::max<size_t>(256, (a_long.load() + groups - 1) / groups);
            }
        } else if (a_long.load() > 1) {
            stripe_min = cand;      // units above half a buffer are sliced anyway: the other long ones go with them
        }
    }
    size_t next_long = 0, next_short = 0;      // scan positions in `units` for the two classes
    auto skip_long = [&] { while (next_long < nu && job.len[units[next_long]] < stripe_min) ++next_long; };
    auto skip_short = [&] { while (next_short < nu && job.len[units[next_short]] >= stripe_min) ++next_short; };
    if (!direct) { skip_long(); skip_short(); }
    auto more = [&] { return direct ? ui < nu : (!active.empty() || next_long < nu || next_short < nu); };

    while (more()) {
        // ---- a staging buffer: a fresh one, or this call's oldest once its batch has been consumed ----
        const double tr_w0 = trace_now();
        int bi = -1;
        rc = lease.acquire(bi, !direct);
        if (rc) return rc;
        tr_wait += trace_now() - tr_w0;
        rc = ensure_buffer(ctx, s, bi, !direct);
        if (rc) return rc;
        const double tr_b0 = trace_now();
        // ---- plan one batch --------------------------------------------------------------------
        // The first batches are short (32, 64, 128 MiB, then whole buffers): units are taken longest first, so
        // the longest chains -- which bound the whole call -- start a few milliseconds after the call does
        // instead of after a full buffer has been read and copied.
        // (Only while the call has the device to itself: with other calls active every call is down to a few
        // buffers, and more, shorter batches would just make it wait for its own kernels more often.)
        const uint64_t cap = lease.share_seen < NBUF ? s.stage_bytes
                                                     : std::min<uint64_t>(s.stage_bytes, ((uint64_t)32 << 20) << std::min(batch, 3));
        pieces.clear();
        uint64_t used = 0;       // bytes of staging consumed
        uint64_t lo = 0;         // direct mode: caller offset of the batch's first byte ...
        uint64_t shift = 0;      // ... which lands at this staging offset (keeps offsets' alignment mod 16)
        bool have_lo = false;
        uint32_t dep_bufs = 0;   // staging buffers whose kernels wrote chaining values this batch continues from
        size_t ns = 0;           // pieces [0, ns) are slices of long units, [ns, np) whole short units
        if (!direct) {
            const bool shorts_left = next_short < nu;
            while (active.size() < max_active && next_long < nu) {
                active.push_back({(uint32_t)next_long, 0, -1});
                ++next_long;
                skip_long();
            }
            if (!active.empty()) {
                // with short units still to come the slice is what a chain hashes while a batch is read and copied
                // (the short units fill the rest of the buffer); once only long units are left, the whole buffer
                const uint64_t lo_s = (cap / max_active - 16) & ~(uint64_t)63;
                const uint64_t slice = shorts_left ? lo_s : std::max(lo_s, (cap / active.size() - 16) & ~(uint64_t)63);
                for (Active &a : active) {
                    if (pieces.size() >= MAX_BATCH_LANES) break;
                    const uint64_t ulen = job.len[units[a.ui]];
                    const uint64_t rem = ulen - a.consumed;
                    const uint64_t spare = from_files ? 1 : 0;              // files: one spare byte for the grew-since-listing probe
                    const bool fin = rem + spare <= slice;
                    uint64_t take = fin ? rem : std::min(rem, slice) & ~(uint64_t)63;
                    if (!fin && take == rem) take -= 64;                    // never a final piece of no bytes
                    const uint64_t extra = fin ? spare : 0;
                    const uint64_t dst = (used + 15) & ~(uint64_t)15;
                    if (dst + take + extra > cap) break;                   // the rest of the active units wait a batch
                    if (a.consumed > 0 && a.last_bi >= 0) dep_bufs |= 1u << a.last_bi;
                    pieces.push_back({a.ui, a.consumed, take, dst, fin});
                    used = dst + take + extra;
                    a.consumed += take;
                    a.last_bi = bi;
                    if (fin) a.last_bi = -2;                                // done: dropped below
                }
                active.erase(std::remove_if(active.begin(), active.end(), [](const Active &a) { return a.last_bi == -2; }), active.end());
            }
            ns = pieces.size();
            while (next_short < nu && pieces.size() < MAX_BATCH_LANES) {
                const uint64_t ulen = job.len[units[next_short]];
                const uint64_t dst = (used + 15) & ~(uint64_t)15;
                // (a first batch is capped below a buffer; a unit under half a buffer still goes in whole)
                if (dst + ulen + (from_files ? 1 : 0) > (pieces.empty() ? s.stage_bytes : cap)) break;
                pieces.push_back({(uint32_t)next_short, 0, ulen, dst, true});
                used = dst + ulen + (from_files ? 1 : 0);
                ++next_short;
                skip_short();
            }
            if (pieces.empty()) return fail(TPIMD5_EINVAL, "internal: a batch without units");
        }
        while (direct && ui < nu && pieces.size() < MAX_BATCH_LANES) {
            const uint32_t u = units[ui];
            const uint64_t ulen = job.len[u];
            const uint64_t rem = ulen - consumed;
            uint64_t dst;
            if (direct) {
                const uint64_t src = job.off[u] + consumed;
                if (!have_lo) { lo = src; shift = src & 15; have_lo = true; }
                dst = src - lo + shift;
            } else {
                dst = (used + 15) & ~(uint64_t)15;
            }
            if (dst > cap) break;
            const uint64_t space = cap - dst;
            if (rem + (from_files ? 1 : 0) <= space) {
                pieces.push_back({(uint32_t)ui, consumed, rem, dst, true});
                used = dst + rem + (from_files ? 1 : 0);      // files: one spare byte for the grew-since-listing probe
                ++ui; consumed = 0;
            } else if (pieces.empty() || space >= (8u << 20)) {
                const uint64_t take = space & ~(uint64_t)63;
                pieces.push_back({(uint32_t)ui, consumed, take, dst, false});
                used = dst + take;
                consumed += take;
                break;
            } else {
                break;
            }
        }
        const size_t np = pieces.size();

        tr_plan += trace_now() - tr_b0;

        // ---- lanes, longest first so that a warp's 32 units finish together ---------------------
        // (slices and whole units are two launches: each group is ordered by itself)
        order.resize(np);
        for (size_t i = 0; i < np; ++i) order[i] = (uint32_t)i;
        bool ragged = false;
        for (size_t g0 = 0, g1 = ns ? ns : np; g0 < np; g0 = g1, g1 = np) {
            bool rg = false;
            for (size_t i = g0 + 1; i < g1 && !rg; ++i) rg = (pieces[i].len >> 6) != (pieces[g0].len >> 6);
            if (rg) std::stable_sort(order.begin() + g0, order.begin() + g1, [&](uint32_t a, uint32_t b) { return pieces[a].len > pieces[b].len; });
            ragged = ragged || rg;
        }
        // the five descriptor arrays are carved for exactly np lanes, so they are one contiguous copy
        LaneDesc h, dv;
        carve_desc(s.h_desc_raw[bi], np, h);
        carve_desc(s.d_desc_raw[bi], np, dv);
        bool aligned = true;
        uint64_t longest = 0;
        for (size_t i = 0; i < np; ++i) {
            aligned = aligned && (pieces[i].dst & 15) == 0;
            longest = std::max(longest, pieces[i].len);
            if (pieces[i].len > MAX_LANE_BYTES) return fail(TPIMD5_ERANGE, "staging above %llu bytes per unit is not supported", (unsigned long long)MAX_LANE_BYTES);
        }
        // A batch of whole, fresh units of one length at one spacing (a workdir of equal small files, the parts of
        // one S3 object) needs no per-unit descriptors at all: it is described to the TMA unit as a 2-D tensor and
        // goes to the strided kernels (throughput or latency kernel by launch size), like device-resident input.
        bool uniform = np >= 1 && ns == 0 && !ragged && !job.init_states && !job.nonfinal && !job.prior && tma_loader >= 2;
        uint64_t ustride = 0;
        if (uniform) {
            const uint64_t ulen = pieces[0].len;
            ustride = np > 1 ? pieces[1].dst - pieces[0].dst : ((ulen + 15) & ~(uint64_t)15);
            uniform = ulen >= 1 && ulen <= ctx->row_max && (pieces[0].dst & 15) == 0 && (ustride & 15) == 0 && ustride >= ulen &&
                      tma_layout_ok(s.d_stage[bi] + pieces[0].dst, ustride, np);
            for (size_t i = 0; i < np && uniform; ++i) {
                const Piece &p = pieces[i];
                uniform = p.src_off == 0 && p.last && p.len == ulen && p.unit == pieces[0].unit + i && p.dst == pieces[0].dst + i * ustride;
            }
        }
        // descriptor rows [i0, i1) of this batch; run on the pool next to the packing, or, in direct mode,
        // by this thread while the data copy is already under way
        auto fill_desc = [&](size_t i0, size_t i1) {
            if (uniform) return;
            for (size_t i = i0; i < i1; ++i) {
                const Piece &p = pieces[order[i]];
                const uint32_t u = units[p.unit];
                h.off[i] = p.dst;
                h.len[i] = p.len;
                h.prior[i] = (job.prior ? job.prior[u] : 0) + p.src_off;
                const bool cont = p.src_off > 0 || job.init_states != nullptr;
                const bool fin = p.last && !(job.nonfinal && job.nonfinal[u]);
                h.flags[i] = (fin ? LANE_FINAL : 0u) | (cont ? LANE_CONT : 0u);
                h.oidx[i] = p.unit;
            }
        };

        // ---- bring the bytes to the device -----------------------------------------------------
        if (direct) {
            if (used > shift) CU(cudaMemcpyAsync(s.d_stage[bi] + shift, job.base + lo, used - shift, cudaMemcpyHostToDevice, s.s_batch[bi]));
            fill_desc(0, np);
        } else {
            uint8_t *stage = s.h_stage[bi];
            // Work items are byte ranges, not units: a large file is read by several workers at once
            // (pread at different offsets), a run of tiny ones by a single worker.
            struct Task { uint32_t first, count; uint64_t lo, hi; };      // pieces [first, first+count); byte range of a lone piece
                                                                          // (count 0: descriptor rows [lo, hi) instead)
            constexpr uint64_t CHUNK = 2u << 20;
            std::vector<Task> tasks;
            for (size_t i = 0; i < np;) {
                if (pieces[i].len > CHUNK) {
                    for (uint64_t lo = 0; lo < pieces[i].len; lo += CHUNK)
                        tasks.push_back({(uint32_t)i, 1, lo, std::min<uint64_t>(pieces[i].len, lo + CHUNK)});
                    ++i;
                } else {
                    size_t j = i;
                    uint64_t bytes = 0;
                    while (j < np && pieces[j].len <= CHUNK && bytes < CHUNK && j - i < 64) bytes += pieces[j++].len;
                    tasks.push_back({(uint32_t)i, (uint32_t)(j - i), 0, 0});
                    i = j;
                }
            }
            auto fail_io = [&](int r, uint32_t u) {
                int expect = 0;
                if (io_err.compare_exchange_strong(expect, r == -1 ? (errno ? errno : EIO) : r)) io_idx = u;
            };
            for (size_t i0 = 0; i0 < np; i0 += 8192) tasks.push_back({0, 0, i0, std::min<uint64_t>(np, i0 + 8192)});
            std::function<void(size_t)> fill = [&](size_t t) {
                const Task &tk = tasks[t];
                if (tk.count == 0) { fill_desc((size_t)tk.lo, (size_t)tk.hi); return; }
                for (uint32_t i = tk.first; i < tk.first + tk.count; ++i) {
                    const Piece &p = pieces[i];
                    const uint32_t u = units[p.unit];
                    const bool whole = tk.count > 1 || tk.hi == 0;               // the piece in one go
                    const uint64_t lo = whole ? 0 : tk.lo, hi = whole ? p.len : tk.hi;
                    if (!from_files) {
                        // logical bytes [a, b) of the unit: first its prefix (a stream's buffered bytes), then base+off
                        const uint64_t P = job.prefix_len ? job.prefix_len[u] : 0;
                        uint64_t a = p.src_off + lo;
                        const uint64_t b = p.src_off + hi;
                        uint8_t *dst = stage + p.dst + lo;
                        if (a < P) {
                            const uint64_t k = std::min(b, P) - a;
                            stage_copy(dst, job.prefix[u] + a, k);
                            dst += k; a += k;
                        }
                        if (b > a) stage_copy(dst, job.base + job.off[u] + (a - P), b - a);
                        continue;
                    }
                    int fd = open(job.paths[u], O_RDONLY | O_CLOEXEC);
                    if (fd < 0) { fail_io(-1, u); continue; }
                    const uint64_t fo = job.file_off ? job.file_off[u] : 0;
                    // The listing promised exactly job.len[u] bytes.  The end of a whole file is read with
                    // one byte to spare (the packer leaves that byte free): getting it means the file grew.
                    const bool probe = p.last && !job.file_off && hi == p.len;
                    const int r = read_fully(fd, stage + p.dst + lo, fo + p.src_off + lo, hi - lo, probe);
                    close(fd);
                    if (r != 0) fail_io(r, u);
                }
                if (!from_files) stage_fence();             // streaming stores of this item are visible before the item counts as done
            };
            // Copy what has been filled while the rest is still being read: tasks are in staging order, so the
            // finished PREFIX of the task list is a finished prefix of the buffer.  This thread (the only one that
            // calls CUDA) sends it off in slices of >= 16 MiB between its own work items.
            const size_t n_data = tasks.size() - (np + 8191) / 8192;      // the descriptor tasks come last
            std::vector<uint64_t> task_end(n_data);
            for (size_t t = 0; t < n_data; ++t) {
                const Task &tk = tasks[t];
                const Piece &last = pieces[tk.first + tk.count - 1];
                task_end[t] = (tk.count > 1 || tk.hi == 0) ? last.dst + last.len + (from_files ? 1 : 0) : last.dst + tk.hi;
            }
            std::unique_ptr<std::atomic<uint8_t>[]> task_done(new std::atomic<uint8_t>[tasks.size()]);
            for (size_t t = 0; t < tasks.size(); ++t) task_done[t].store(0, std::memory_order_relaxed);
            std::function<void(size_t)> fill_mark = [&](size_t t) {
                fill(t);
                task_done[t].store(1, std::memory_order_release);
            };
            size_t cursor = 0;
            uint64_t sent = 0;
            cudaError_t copy_err = cudaSuccess;
            std::function<void()> tick = [&] {
                while (cursor < n_data && task_done[cursor].load(std::memory_order_acquire)) ++cursor;
                if (!cursor || copy_err != cudaSuccess || io_err.load() != 0) return;
                const uint64_t upto = std::min<uint64_t>(task_end[cursor - 1], used) & ~(uint64_t)15;
                if (upto >= sent + (16u << 20) && cursor < n_data && !job.ingest_only) {
                    copy_err = cudaMemcpyAsync(s.d_stage[bi] + sent, stage + sent, upto - sent, cudaMemcpyHostToDevice, s.s_batch[bi]);
                    sent = upto;
                }
            };
            const double tr_f0 = trace_now();
            // (Only for a call's first batches, and only when they hold units of a MiB or more: units come longest first, so that is where the chains that bound the
            // call start; once the pipeline streams, the extra copies just take memory bandwidth from the readers:
            // an 8 GiB file went 250 -> 262 ms with slices on every batch, C1 72 -> 70 ms either way.)
            // When the reads are the slow side anyway (small files: open/close bound, ~17 GB/s against 55 for the copy)
            // every batch is copied as it fills: the copy of a whole buffer no longer trails its last read.
            const bool slow_reads = from_files && ctx->feed_files.load() < 25e9;
            ctx->pool->parallel_for(tasks.size(), 1, fill_mark,
                                    (used >= (64u << 20) && ((batch < 6 && longest >= (1u << 20)) || slow_reads)) ? &tick : nullptr);
            const double fill_s = trace_now() - tr_f0;
            tr_fill += fill_s;
            if (used >= (16u << 20) && fill_s > 0) {
                std::atomic<double> &rate = from_files ? ctx->feed_files : ctx->feed_mem;
                rate.store(0.5 * rate.load() + 0.5 * ((double)used / fill_s));
            }
            if (io_err.load() != 0) break;
            if (copy_err != cudaSuccess) return fail(TPIMD5_ECUDA, "cudaMemcpyAsync (staging): %s", cudaGetErrorString(copy_err));
            if (used > sent && !job.ingest_only) CU(cudaMemcpyAsync(s.d_stage[bi] + sent, stage + sent, used - sent, cudaMemcpyHostToDevice, s.s_batch[bi]));
        }
        // descriptors: five arrays, copied as stored (their device mirror has the same carving)
        if (np && !uniform && !job.ingest_only) CU(cudaMemcpyAsync(s.d_desc_raw[bi], s.h_desc_raw[bi], np * desc_bytes_per_lane, cudaMemcpyHostToDevice, s.s_batch[bi]));

        // ---- hash (same stream: the kernel follows its batch's copies) -----------------------------
        cudaStream_t sk = s.s_batch[bi];
        // a unit continued from the previous batch reads the chaining value that batch's kernel writes
        if (direct && batch > 0 && np && pieces[0].src_off > 0) CU(cudaStreamWaitEvent(sk, s.ev_done[prev_bi], 0));
        auto lanes_of = [&](size_t g0, size_t g1) {
            LaneArgs a{};
            a.base = s.d_stage[bi]; a.base_bytes = used;
            a.off = dv.off + g0; a.len = dv.len + g0; a.prior = dv.prior + g0;
            a.flags = dv.flags + g0; a.out_idx = dv.oidx + g0;
            a.table = d_table; a.n = g1 - g0;
            return a;
        };
        if (job.ingest_only) {
            rc = 0;
        } else if (uniform) {
            rc = launch_strided_tma(ctx, s.d_stage[bi] + pieces[0].dst, ustride, pieces[0].len, np, d_table + pieces[0].unit,
                                    tma_loader, tma_shape, tma_mode, sk);
        } else if (ns == 0) {
            rc = launch_lanes(ctx, lanes_of(0, np), aligned, sk);
        } else {
            // whole units beside the slices, on the buffer's second stream, once the batch's bytes are in place
            if (np > ns) {
                CU(cudaEventRecord(s.ev_copied[bi], sk));
                CU(cudaStreamWaitEvent(s.s_aux[bi], s.ev_copied[bi], 0));
                rc = launch_lanes(ctx, lanes_of(ns, np), aligned, s.s_aux[bi]);
                if (rc) return rc;
                CU(cudaEventRecord(s.ev_aux[bi], s.s_aux[bi]));
            }
            // the slices: after the kernels that wrote the chaining values they continue from
            for (int b = 0; b < NBUF; ++b)
                if ((dep_bufs >> b & 1) && b != bi) CU(cudaStreamWaitEvent(sk, s.ev_slice[b], 0));
            rc = launch_lanes(ctx, lanes_of(0, ns), aligned, sk);
            if (rc) return rc;
            CU(cudaEventRecord(s.ev_slice[bi], sk));
            if (np > ns) CU(cudaStreamWaitEvent(sk, s.ev_aux[bi], 0));
        }
        if (rc) return rc;
        CU(cudaEventRecord(s.ev_done[bi], sk));
        for (size_t i = 0; i < np; ++i) batch_bytes += pieces[i].len;
        if (trace_on())
            fprintf(stderr, "[tpimd5]   batch %d: %zu units, %.1f MiB, longest %.2f MiB, launched at %.1f ms%s\n", batch, np,
                    batch_bytes / 1048576.0, np ? pieces[order.empty() ? 0 : order[0]].len / 1048576.0 : 0.0, 1e3 * (trace_now() - tr0),
                    uniform ? " (uniform)" : "");
        ctx->bytes_staged.fetch_add(batch_bytes);
        total_bytes += batch_bytes;
        batch_bytes = 0;
        prev_bi = bi;
        ++batch;
    }

    if (io_err.load() != 0) {
        const size_t u = io_idx.load();
        job.failed_index = u;
        const int e = io_err.load();
        if (e == -2) return fail(TPIMD5_EIO, "unit %zu (%s): file is shorter than its listed size", u, job.paths[u]);
        if (e == -3) return fail(TPIMD5_EIO, "unit %zu (%s): file is longer than its listed size", u, job.paths[u]);
        return fail(TPIMD5_EIO, "unit %zu (%s): %s", u, job.paths[u], strerror(e));
    }

    // ---- digests home -------------------------------------------------------------------------
    for (int b : lease.bufs) CU(cudaStreamWaitEvent(lease.stream, s.ev_done[b], 0));
    if (job.ingest_only) {
        CU(cudaStreamSynchronize(lease.stream));
        return 0;
    }
    CU(cudaMemcpyAsync(h_table, d_table, nu * 16, cudaMemcpyDeviceToHost, lease.stream));
    const double tr_s0 = trace_now();
    CU(cudaStreamSynchronize(lease.stream));
    if (trace_on())
        fprintf(stderr, "[tpimd5] device %d: %zu units, %d batches, total %.1f ms: plan %.1f, buffer wait %.1f, fill %.1f, final sync %.1f\n",
                s.dev, nu, batch, 1e3 * (trace_now() - tr0), 1e3 * tr_plan, 1e3 * tr_wait, 1e3 * tr_fill, 1e3 * (trace_now() - tr_s0));
    // table row i belongs to unit units[i]; runs of consecutive units move as one copy (16 MB of digests for a
    // million files: the pool moves them, 64 Ki rows per item)
    {
        constexpr size_t CHUNK = 1u << 16;
        std::function<void(size_t)> home = [&](size_t c) {
            const size_t hi = std::min(nu, (c + 1) * CHUNK);
            if (!units.idx) {
                memcpy(job.out + 16 * ((size_t)units.base + c * CHUNK), h_table + 16 * c * CHUNK, 16 * (hi - c * CHUNK));
                return;
            }
            for (size_t i = c * CHUNK; i < hi;) {
                size_t j = i + 1;
                while (j < hi && units[j] == units[j - 1] + 1) ++j;
                memcpy(job.out + 16 * (size_t)units[i], h_table + 16 * i, 16 * (j - i));
                i = j;
            }
        };
        ctx->pool->parallel_for((nu + CHUNK - 1) / CHUNK, 1, home);
    }
    ctx->units_done.fetch_add(nu);
    ctx->bytes_done.fetch_add(total_bytes);
    return 0;
}

// Runs one device's share.  (The lease inside run_device_impl drains and returns everything on every path.)
static int run_device(tpimd5_ctx *ctx, Slot &s, Job &job, const UnitList &units) {
    return run_device_impl(ctx, s, job, units);
}

static int run_job(tpimd5_ctx *ctx, Job &job) {
    if (job.n == 0) return 0;
    if (job.n > 0xffffffffull) return fail(TPIMD5_ERANGE, "more than 2^32 units in one call");
    const int ndev = (int)ctx->slots.size();
    struct Count {
        tpimd5_ctx *c;
        explicit Count(tpimd5_ctx *x) : c(x) { c->calls_started.fetch_add(1); }
        ~Count() { c->calls_finished.fetch_add(1); }
    } count(ctx);
    if (ndev == 1 || job.init_states)
        // chained units (streams) stay on device 0: their state lives there between calls
        return run_device(ctx, *ctx->slots[0], job, UnitList{nullptr, 0, job.n});
    // Unless the sizes are heavy-tailed, cut the caller's order into ndev runs of equal bytes: every device
    // then gets one dense stretch of the caller's buffer (direct copies, no packing, no index list) and the
    // imbalance is at most one unit.  A unit bigger than a quarter of a device's share means the tail matters:
    // greedy LPT.
    uint64_t total = 0, biggest = 0;
    {
        // (millions of units: summed by the pool, 64 Ki units per item)
        constexpr size_t CHUNK = 1u << 16;
        std::atomic<uint64_t> a_total{0}, a_big{0};
        std::function<void(size_t)> sum = [&](size_t c) {
            uint64_t t = 0, b = 0;
            const size_t hi = std::min(job.n, (c + 1) * CHUNK);
            for (size_t i = c * CHUNK; i < hi; ++i) { t += job.len[i] + 72; b = std::max(b, job.len[i]); }
            a_total.fetch_add(t);
            uint64_t cur = a_big.load();
            while (b > cur && !a_big.compare_exchange_weak(cur, b)) {}
        };
        ctx->pool->parallel_for((job.n + CHUNK - 1) / CHUNK, 1, sum);
        total = a_total.load(); biggest = a_big.load();
    }
    std::vector<UnitList> shards(ndev);
    std::vector<std::vector<uint32_t>> lists;
    if (biggest <= total / (4 * (uint64_t)ndev)) {
        uint64_t acc = 0;
        size_t first = 0;
        int d = 0;
        for (size_t i = 0; i < job.n; ++i) {
            if (d + 1 < ndev && acc >= (total / ndev) * (uint64_t)(d + 1)) {
                shards[d] = UnitList{nullptr, (uint32_t)first, i - first};
                first = i;
                ++d;
            }
            acc += job.len[i] + 72;
        }
        shards[d] = UnitList{nullptr, (uint32_t)first, job.n - first};
    } else {
        shard_lpt(job.len, job.n, ndev, lists);
        // direct mode wants ascending addresses; LPT order is by size.  Re-sort each shard by the
        // caller's order (for files that is listing order, which is also friendliest to the page cache).
        for (int d = 0; d < ndev; ++d) {
            std::sort(lists[d].begin(), lists[d].end());
            shards[d] = UnitList{lists[d].data(), 0, lists[d].size()};
        }
    }
    std::vector<int> rcs(ndev, 0);
    std::vector<std::string> errs(ndev);
    std::mutex done_mu;
    std::condition_variable done_cv;
    int pending = ndev;
    for (int d = 0; d < ndev; ++d)
        ctx->slots[d]->workers->post([&, d] {
            rcs[d] = shards[d].empty() ? 0 : run_device(ctx, *ctx->slots[d], job, shards[d]);
            if (rcs[d]) errs[d] = g_err;
            std::lock_guard<std::mutex> g(done_mu);
            if (--pending == 0) done_cv.notify_all();
        });
    {
        std::unique_lock<std::mutex> lk(done_mu);
        done_cv.wait(lk, [&] { return pending == 0; });
    }
    for (int d = 0; d < ndev; ++d)
        if (rcs[d]) {
            g_err = errs[d];
            return rcs[d];
        }
    return 0;
}

// ------------------------------------------------------------------------------------------------
// host entry points
// ------------------------------------------------------------------------------------------------
extern "C" int tpimd5_hash_buffers(tpimd5_ctx *ctx, const uint8_t *base, const uint64_t *offsets, const uint64_t *lengths,
                                   size_t n, uint8_t *digests_out) {
    if (!ctx) return fail(TPIMD5_EINVAL, "hash_buffers: null context");
    if (n == 0) return 0;
    if (!offsets || !lengths || !digests_out) return fail(TPIMD5_EINVAL, "hash_buffers: null pointer");
    Job job;
    job.n = n; job.base = base; job.off = offsets; job.len = lengths; job.out = digests_out;
    return run_job(ctx, job);
}

extern "C" int tpimd5_ingest_probe(tpimd5_ctx *ctx, const char *const *paths, const uint64_t *file_offsets, const uint64_t *sizes, size_t n,
                                   size_t *failed_index_out) {
    if (failed_index_out) *failed_index_out = (size_t)-1;
    if (!ctx) return fail(TPIMD5_EINVAL, "ingest_probe: null context");
    if (n == 0) return 0;
    if (!paths || !sizes) return fail(TPIMD5_EINVAL, "ingest_probe: null pointer");
    Job job;
    job.n = n; job.len = sizes; job.paths = paths; job.file_off = file_offsets; job.ingest_only = true;
    int rc = run_job(ctx, job);
    if (rc && failed_index_out) *failed_index_out = job.failed_index;
    return rc;
}

extern "C" int tpimd5_units_extent(tpimd5_ctx *ctx, const uint64_t *offsets, const uint64_t *lengths, size_t n, uint64_t *extent_out) {
    if (!ctx || !extent_out) return fail(TPIMD5_EINVAL, "units_extent: null pointer");
    *extent_out = 0;
    if (n == 0) return 0;
    if (!offsets || !lengths) return fail(TPIMD5_EINVAL, "units_extent: null pointer");
    constexpr size_t CHUNK = 1u << 16;
    std::atomic<uint64_t> a_end{0};
    std::atomic<bool> wrapped{false};
    std::function<void(size_t)> scan = [&](size_t c) {
        uint64_t e = 0;
        bool w = false;
        const size_t hi = std::min(n, (c + 1) * CHUNK);
        for (size_t i = c * CHUNK; i < hi; ++i) {
            const uint64_t end = offsets[i] + lengths[i];
            w |= end < offsets[i];
            e = std::max(e, end);
        }
        if (w) wrapped.store(true);
        uint64_t cur = a_end.load();
        while (e > cur && !a_end.compare_exchange_weak(cur, e)) {}
    };
    ctx->pool->parallel_for((n + CHUNK - 1) / CHUNK, 1, scan);
    if (wrapped.load()) return fail(TPIMD5_ERANGE, "units_extent: an offset + length overflows 64 bits");
    *extent_out = a_end.load();
    return 0;
}

extern "C" int tpimd5_hash_files(tpimd5_ctx *ctx, const char *const *paths, const uint64_t *sizes, size_t n,
                                 uint8_t *digests_out, size_t *failed_index_out) {
    if (failed_index_out) *failed_index_out = (size_t)-1;
    if (!ctx) return fail(TPIMD5_EINVAL, "hash_files: null context");
    if (n == 0) return 0;
    if (!paths || !digests_out) return fail(TPIMD5_EINVAL, "hash_files: null pointer");
    std::vector<uint64_t> st_sizes;
    if (!sizes) {
        // (a listing of 200,000 files is 0.1 s of stat calls on one thread: the pool makes them, 4 Ki files per item;
        // the error reported is the first one in listing order)
        st_sizes.resize(n);
        constexpr size_t CHUNK = 4096;
        std::atomic<size_t> first_bad{(size_t)-1};
        std::vector<int> bad_errno((n + CHUNK - 1) / CHUNK, 0);
        std::function<void(size_t)> stat_chunk = [&](size_t c) {
            const size_t hi = std::min(n, (c + 1) * CHUNK);
            for (size_t i = c * CHUNK; i < hi; ++i) {
                struct stat sb;
                if (stat(paths[i], &sb) != 0) {
                    bad_errno[c] = errno;
                    size_t cur = first_bad.load();
                    while (i < cur && !first_bad.compare_exchange_weak(cur, i)) {}
                    return;
                }
                st_sizes[i] = (uint64_t)sb.st_size;
            }
        };
        ctx->pool->parallel_for((n + CHUNK - 1) / CHUNK, 1, stat_chunk);
        if (first_bad.load() != (size_t)-1) {
            const size_t i = first_bad.load();
            if (failed_index_out) *failed_index_out = i;
            return fail(TPIMD5_EIO, "unit %zu (%s): %s", i, paths[i], strerror(bad_errno[i / CHUNK]));
        }
        sizes = st_sizes.data();
    }
    Job job;
    job.n = n; job.paths = paths; job.len = sizes; job.out = digests_out;
    int rc = run_job(ctx, job);
    if (rc && failed_index_out) *failed_index_out = job.failed_index;
    return rc;
}

static int s3_common(tpimd5_ctx *ctx, const uint8_t *data, const char *path, uint64_t size, uint64_t part_size,
                     uint8_t *part_digests_out, uint8_t etag_md5_out[16], uint32_t *nparts_out) {
    if (!ctx || !etag_md5_out) return fail(TPIMD5_EINVAL, "s3_etag: null pointer");
    if (part_size == 0) return fail(TPIMD5_EINVAL, "s3_etag: part_size 0");
    const uint32_t np = tpimd5_s3_nparts(size, part_size);
    if (!np) return fail(TPIMD5_ERANGE, "s3_etag: too many parts");
    std::vector<uint8_t> own;
    uint8_t *pd = part_digests_out;
    if (!pd) { own.resize((size_t)np * 16); pd = own.data(); }
    {
        std::vector<uint64_t> off(np), len(np);
        for (uint32_t i = 0; i < np; ++i) {
            off[i] = (uint64_t)i * part_size;
            len[i] = std::min(part_size, size - off[i]);
        }
        Job job;
        job.n = np; job.len = len.data(); job.out = pd;
        std::vector<const char *> same;
        if (path) {
            // parts of one file: np units that share a path and differ in file offset
            struct stat sb;
            if (stat(path, &sb) != 0) return fail(TPIMD5_EIO, "%s: %s", path, strerror(errno));
            if ((uint64_t)sb.st_size != size)
                return fail(TPIMD5_EIO, "%s: size is %llu, not the listed %llu bytes", path, (unsigned long long)sb.st_size,
                            (unsigned long long)size);
            same.assign(np, path);
            job.paths = same.data(); job.file_off = off.data();
        } else {
            job.base = data; job.off = off.data();
        }
        int rc = run_job(ctx, job);
        if (rc) return rc;
    }
    // digest of digests: one more unit, hashed on the GPU as well
    const uint64_t zero = 0, dl = (uint64_t)np * 16;
    Job j2;
    j2.n = 1; j2.base = pd; j2.off = &zero; j2.len = &dl; j2.out = etag_md5_out;
    int rc = run_job(ctx, j2);
    if (rc) return rc;
    if (nparts_out) *nparts_out = np;
    return 0;
}

extern "C" int tpimd5_s3_etag_buffer(tpimd5_ctx *ctx, const uint8_t *data, uint64_t size, uint64_t part_size,
                                     uint8_t *part_digests_out, uint8_t etag_md5_out[16], uint32_t *nparts_out) {
    if (!data && size) return fail(TPIMD5_EINVAL, "s3_etag_buffer: null data");
    return s3_common(ctx, data, nullptr, size, part_size, part_digests_out, etag_md5_out, nparts_out);
}

extern "C" int tpimd5_s3_etag_file(tpimd5_ctx *ctx, const char *path, uint64_t size, uint64_t part_size,
                                   uint8_t *part_digests_out, uint8_t etag_md5_out[16], uint32_t *nparts_out) {
    if (!path) return fail(TPIMD5_EINVAL, "s3_etag_file: null path");
    return s3_common(ctx, nullptr, path, size, part_size, part_digests_out, etag_md5_out, nparts_out);
}

// ------------------------------------------------------------------------------------------------
// S3 upload plan: what [rclone v1.57.0] backend/s3 Object.Update needs to know about one local file
// before it sends a byte (rule restated from memory of that release)
// ------------------------------------------------------------------------------------------------
extern "C" uint32_t tpimd5_s3_plan_parts(uint64_t size, uint64_t chunk_size, uint64_t upload_cutoff, uint32_t max_parts,
                                         int *multipart_out, uint64_t *part_size_out) {
    if (!chunk_size) chunk_size = (uint64_t)5 << 20;          // rclone: chunk_size default (and minimum) 5 MiB
    if (!upload_cutoff) upload_cutoff = (uint64_t)200 << 20;  // upload_cutoff default 200 MiB
    if (!max_parts) max_parts = 10000;                        // max_upload_parts default
    const bool multi = size >= upload_cutoff;                 // "multipart := size < 0 || size >= upload_cutoff"
    uint64_t part = size;
    uint32_t np = 1;
    if (multi) {
        part = tpimd5_s3_part_size(size, chunk_size, max_parts);
        np = tpimd5_s3_nparts(size, part);
    }
    if (multipart_out) *multipart_out = multi ? 1 : 0;
    if (part_size_out) *part_size_out = part;
    return np;
}

extern "C" int tpimd5_s3_upload_plan(tpimd5_ctx *ctx, const char *path, uint64_t size, uint64_t chunk_size, uint64_t upload_cutoff,
                                     uint32_t max_parts, uint32_t flags, tpimd5_s3_plan *plan_out, uint8_t *part_md5_out,
                                     char *part_content_md5_out) {
    if (!ctx || !path || !plan_out) return fail(TPIMD5_EINVAL, "s3_upload_plan: null pointer");
    memset(plan_out, 0, sizeof *plan_out);
    int multi = 0;
    uint64_t part = 0;
    const uint32_t np = tpimd5_s3_plan_parts(size, chunk_size, upload_cutoff, max_parts, &multi, &part);
    if (!np) return fail(TPIMD5_ERANGE, "s3_upload_plan: too many parts");
    const bool whole = !multi || !(flags & TPIMD5_PLAN_NO_WHOLE_MD5);
    struct stat sb;
    if (stat(path, &sb) != 0) return fail(TPIMD5_EIO, "%s: %s", path, strerror(errno));
    if ((uint64_t)sb.st_size != size)
        return fail(TPIMD5_EIO, "%s: size is %llu, not the listed %llu bytes", path, (unsigned long long)sb.st_size,
                    (unsigned long long)size);
    // one job: the parts (multipart only) and, unless switched off, the whole file as one more unit --
    // both read the same bytes, so the second read comes out of the page cache
    const size_t nu = (multi ? np : 0) + (whole ? 1 : 0);
    std::vector<uint64_t> off(nu), len(nu);
    std::vector<const char *> same(nu, path);
    std::vector<uint8_t> dig(nu * 16);
    size_t k = 0;
    if (whole) { off[k] = 0; len[k] = size; ++k; }             // first: it is the longest chain by far
    if (multi)
        for (uint32_t i = 0; i < np; ++i, ++k) {
            off[k] = (uint64_t)i * part;
            len[k] = std::min(part, size - off[k]);
        }
    Job job;
    job.n = nu; job.len = len.data(); job.out = dig.data(); job.paths = same.data(); job.file_off = off.data();
    int rc = run_job(ctx, job);
    if (rc) return rc;
    plan_out->multipart = multi;
    plan_out->part_size = part;
    plan_out->nparts = np;
    plan_out->has_md5 = whole ? 1 : 0;
    const uint8_t *parts = dig.data() + (whole ? 16 : 0);
    if (whole) {
        memcpy(plan_out->md5, dig.data(), 16);
        tpimd5_base64(plan_out->md5, plan_out->content_md5);   // Content-MD5 (single) / X-Amz-Meta-Md5chksum (multipart)
    }
    if (!multi) {
        tpimd5_hex(plan_out->md5, plan_out->etag);              // a plain PutObject answers with the MD5 as its ETag
        if (part_md5_out) memcpy(part_md5_out, plan_out->md5, 16);
        if (part_content_md5_out) memcpy(part_content_md5_out, plan_out->content_md5, 25);
        return 0;
    }
    if (part_md5_out) memcpy(part_md5_out, parts, (size_t)np * 16);
    if (part_content_md5_out)
        for (uint32_t i = 0; i < np; ++i) tpimd5_base64(parts + 16 * (size_t)i, part_content_md5_out + 25 * (size_t)i);
    // expected ETag: MD5 of the concatenated part digests, hashed on the GPU like everything else
    uint8_t etag_md5[16];
    const uint64_t zero = 0, dl = (uint64_t)np * 16;
    Job j2;
    j2.n = 1; j2.base = parts; j2.off = &zero; j2.len = &dl; j2.out = etag_md5;
    rc = run_job(ctx, j2);
    if (rc) return rc;
    return tpimd5_s3_etag_text(etag_md5, np, plan_out->etag, sizeof plan_out->etag);
}

extern "C" int tpimd5_diff(tpimd5_ctx *ctx, const uint8_t *old16, const uint8_t *new16, size_t n, uint32_t *changed_idx_out,
                           size_t *n_changed_out) {
    if (!ctx || !n_changed_out) return fail(TPIMD5_EINVAL, "diff: null pointer");
    *n_changed_out = 0;
    if (n == 0) return 0;
    if (!old16 || !new16 || !changed_idx_out) return fail(TPIMD5_EINVAL, "diff: null pointer");
    if (n > 0xffffffffull) return fail(TPIMD5_ERANGE, "diff: more than 2^32 units");
    Slot &s = *ctx->slots[0];
    USE_DEVICE(s.dev);
    // One block of device memory out of the slot's free list and a stream of the call's own: no allocation
    // in the steady state, and concurrent diffs (or a diff next to a running hash job) do not wait for each other.
    const size_t tiles = (n + DIFF_TILE - 1) / DIFF_TILE;
    const size_t b_old = 0, b_new = n * 16, b_idx = 2 * n * 16, b_cnt = (b_idx + n * 4 + 255) & ~(size_t)255;
    struct Held {
        Slot &s;
        DevBlock blk;
        cudaStream_t st = nullptr;
        explicit Held(Slot &slot) : s(slot) {}
        ~Held() {
            if (st) cudaStreamSynchronize(st);
            give_block(s, blk);
            give_stream(s, st);
        }
    } held(s);
    int rc = take_block(s, b_cnt + 8 + tiles * DIFF_SCRATCH_PER_TILE, held.blk);
    if (rc == 0) rc = take_stream(s, held.st);
    if (rc) return rc;
    uint8_t *const base = held.blk.d;
    uint64_t *const d_cnt = (uint64_t *)(base + b_cnt);
    uint64_t cnt = 0;
    CU(cudaMemcpyAsync(base + b_old, old16, n * 16, cudaMemcpyHostToDevice, held.st));
    CU(cudaMemcpyAsync(base + b_new, new16, n * 16, cudaMemcpyHostToDevice, held.st));
    rc = diff_launch(ctx, base + b_old, base + b_new, n, (uint32_t *)(base + b_idx), d_cnt, d_cnt + 1, held.st);
    if (rc) return rc;
    CU(cudaMemcpyAsync(&cnt, d_cnt, 8, cudaMemcpyDeviceToHost, held.st));
    CU(cudaStreamSynchronize(held.st));
    if (cnt) {
        CU(cudaMemcpyAsync(changed_idx_out, base + b_idx, cnt * 4, cudaMemcpyDeviceToHost, held.st));
        CU(cudaStreamSynchronize(held.st));
    }
    *n_changed_out = (size_t)cnt;
    return 0;
}

// ------------------------------------------------------------------------------------------------
// streaming hasher (Go hash.Hash semantics)
// ------------------------------------------------------------------------------------------------
struct tpimd5_stream {
    tpimd5_ctx *ctx = nullptr;
    uint8_t state[16];          // chaining value after `hashed` bytes
    uint64_t hashed = 0;        // multiple of 64
    std::vector<uint8_t> pend;  // bytes written but not hashed yet
    std::mutex mu;
};

static const uint8_t MD5_IV_BYTES[16] = {0x01, 0x23, 0x45, 0x67, 0x89, 0xab, 0xcd, 0xef,
                                         0xfe, 0xdc, 0xba, 0x98, 0x76, 0x54, 0x32, 0x10};
constexpr size_t STREAM_FLUSH = 4u << 20;  // pending bytes that trigger a launch

extern "C" int tpimd5_stream_open(tpimd5_ctx *ctx, tpimd5_stream **out) {
    if (!ctx || !out) return fail(TPIMD5_EINVAL, "stream_open: null pointer");
    auto *s = new tpimd5_stream;
    s->ctx = ctx;
    memcpy(s->state, MD5_IV_BYTES, 16);
    *out = s;
    return 0;
}

extern "C" void tpimd5_stream_close(tpimd5_stream *s) { delete s; }

extern "C" int tpimd5_stream_reset(tpimd5_stream *s) {
    if (!s) return fail(TPIMD5_EINVAL, "stream_reset: null stream");
    std::lock_guard<std::mutex> g(s->mu);
    memcpy(s->state, MD5_IV_BYTES, 16);
    s->hashed = 0;
    s->pend.clear();
    return 0;
}

// Advance streams together, one lane each, one launch: stream i hashes the whole 64-byte blocks of
// (its buffered bytes || ptrs[i][0..lens[i])) straight from where those bytes lie -- the buffered part is
// the unit's prefix, the new data is not copied into the stream first -- and keeps the < 64 bytes left
// over.  ptrs may be null (flush only).
static int advance_streams(tpimd5_stream *const *ss, const void *const *ptrs, const size_t *lens, size_t n) {
    if (n == 0) return 0;
    tpimd5_ctx *ctx = ss[0]->ctx;
    std::vector<uint64_t> off, len, prior, plen;
    std::vector<const uint8_t *> pfx;
    std::vector<uint8_t> init, nonfinal, out;
    std::vector<size_t> who;
    static const uint8_t dummy = 0;
    for (size_t i = 0; i < n; ++i) {
        tpimd5_stream *s = ss[i];
        const uint64_t add = ptrs ? lens[i] : 0;
        const uint64_t total = s->pend.size() + add;
        const uint64_t full = total & ~(uint64_t)63;
        if (!full || (ptrs && add < 64 && s->pend.size() + add < STREAM_FLUSH)) {
            // nothing worth a lane yet: buffer it (also keeps the left-over logic below simple: the tail of
            // a hashed unit always lies inside the new data)
            if (add) s->pend.insert(s->pend.end(), (const uint8_t *)ptrs[i], (const uint8_t *)ptrs[i] + add);
            continue;
        }
        who.push_back(i);
        pfx.push_back(s->pend.empty() ? &dummy : s->pend.data());
        plen.push_back(s->pend.size());
        off.push_back(add ? (uint64_t)(uintptr_t)ptrs[i] : (uint64_t)(uintptr_t)&dummy);
        len.push_back(full);
        prior.push_back(s->hashed);
        init.insert(init.end(), s->state, s->state + 16);
        nonfinal.push_back(1);
    }
    if (who.empty()) return 0;
    out.resize(who.size() * 16);
    Job job;
    job.n = who.size(); job.base = nullptr; job.off = off.data(); job.len = len.data();
    job.prefix = pfx.data(); job.prefix_len = plen.data();
    job.init_states = init.data(); job.prior = prior.data(); job.nonfinal = nonfinal.data(); job.out = out.data();
    int rc = run_job(ctx, job);
    if (rc) return rc;
    for (size_t k = 0; k < who.size(); ++k) {
        const size_t i = who[k];
        tpimd5_stream *s = ss[i];
        const uint64_t add = ptrs ? lens[i] : 0, P = plen[k], total = P + add, left = total - len[k];
        memcpy(s->state, out.data() + 16 * k, 16);
        s->hashed += len[k];
        // the unhashed tail: logical bytes [len[k], total) of (pend || new data)
        std::vector<uint8_t> tail;
        tail.reserve(64);
        for (uint64_t x = len[k]; x < total; ++x) tail.push_back(x < P ? s->pend[x] : ((const uint8_t *)ptrs[i])[x - P]);
        (void)left;
        s->pend.swap(tail);
    }
    return 0;
}

extern "C" int tpimd5_stream_write(tpimd5_stream *s, const void *p, size_t n) {
    if (!s || (!p && n)) return fail(TPIMD5_EINVAL, "stream_write: null pointer");
    std::lock_guard<std::mutex> g(s->mu);
    // Small writes (io.Copy hands over 32 KiB at a time) are buffered up to STREAM_FLUSH so that a launch
    // always has megabytes to chew on; a large write goes to the GPU from where it lies.
    if (s->pend.size() + n < STREAM_FLUSH) {
        s->pend.insert(s->pend.end(), (const uint8_t *)p, (const uint8_t *)p + n);
        return 0;
    }
    tpimd5_stream *one = s;
    const void *ptr = p;
    return advance_streams(&one, &ptr, &n, 1);
}

extern "C" int tpimd5_stream_write_many(tpimd5_stream *const *streams, const void *const *ptrs, const size_t *lens,
                                        size_t n_streams) {
    if (n_streams == 0) return 0;
    if (!streams || !ptrs || !lens) return fail(TPIMD5_EINVAL, "stream_write_many: null pointer");
    {
        std::vector<const tpimd5_stream *> seen(streams, streams + n_streams);
        std::sort(seen.begin(), seen.end());
        if (std::adjacent_find(seen.begin(), seen.end()) != seen.end()) return fail(TPIMD5_EINVAL, "stream_write_many: stream listed twice");
    }
    for (size_t i = 0; i < n_streams; ++i) {
        if (!streams[i] || (!ptrs[i] && lens[i])) return fail(TPIMD5_EINVAL, "stream_write_many: null entry %zu", i);
        if (streams[i]->ctx != streams[0]->ctx) return fail(TPIMD5_EINVAL, "stream_write_many: streams of different contexts");
    }
    return advance_streams(streams, ptrs, lens, n_streams);
}

extern "C" int tpimd5_stream_sum(tpimd5_stream *s, uint8_t out[16]) {
    if (!s || !out) return fail(TPIMD5_EINVAL, "stream_sum: null pointer");
    std::lock_guard<std::mutex> g(s->mu);
    // finalise a copy: the pending bytes are hashed with padding from the stream's chaining value,
    // the stream itself keeps state and pending bytes untouched
    const uint64_t off = (uint64_t)(uintptr_t)s->pend.data(), len = s->pend.size(), prior = s->hashed;
    static const uint8_t dummy = 0;
    const uint64_t off_safe = len ? off : (uint64_t)(uintptr_t)&dummy;
    Job job;
    job.n = 1; job.base = nullptr; job.off = &off_safe; job.len = &len;
    job.init_states = s->state; job.prior = &prior; job.out = out;
    return run_job(s->ctx, job);
}

// ------------------------------------------------------------------------------------------------
// S3 ETag with no allocation at all (see include/tpimd5.h for who needs this and why)
// ------------------------------------------------------------------------------------------------
// Scratch for the part digests, carried in the library image: one copy per GPU that loads this
// module.  In host code the name below is only a handle for the runtime; the kernels need the
// array's address on the current device, which cudaGetSymbolAddress supplies (once per device).
static __device__ __align__(256) uint8_t g_noalloc_part_digests[(size_t)TPIMD5_NOALLOC_MAX_PARTS * 16];

constexpr int NOALLOC_MAX_DEVICES = 64;
static uint8_t *g_noalloc_dev_ptr[NOALLOC_MAX_DEVICES];  // [device] -> device address, null until set up
static std::mutex g_noalloc_mu;

extern "C" int tpimd5_dev_s3_etag_noalloc(int device, const void *d_data, uint64_t size, uint64_t part_size,
                                          void *d_etag_md5_out, void *cuda_stream) {
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return fail(TPIMD5_ENODEV, "s3_etag_noalloc: device %d out of range", device);
    if (!d_etag_md5_out || (!d_data && size)) return fail(TPIMD5_EINVAL, "s3_etag_noalloc: null pointer");
    if ((uintptr_t)d_etag_md5_out & 15) return fail(TPIMD5_EINVAL, "s3_etag_noalloc: d_etag_md5_out must be 16-byte aligned");
    if (part_size == 0) return fail(TPIMD5_EINVAL, "s3_etag_noalloc: part_size 0");
    if (part_size > MAX_LANE_BYTES) return fail(TPIMD5_ERANGE, "s3_etag_noalloc: part above %llu bytes", (unsigned long long)MAX_LANE_BYTES);
    const uint64_t np = size == 0 ? 1 : (size + part_size - 1) / part_size;
    if (np > TPIMD5_NOALLOC_MAX_PARTS)
        return fail(TPIMD5_ERANGE,
                    "s3_etag_noalloc: %llu parts, but the built-in scratch holds %d part digests; use a larger part_size "
                    "(tpimd5_s3_part_size) or tpimd5_dev_s3_etag",
                    (unsigned long long)np, TPIMD5_NOALLOC_MAX_PARTS);
    USE_DEVICE(device);

    // set-up, first call on this device (any thread): where does the array live on *this* GPU, and the
    // shared-memory attributes of the two kernel instantiations this entry point can launch
    uint8_t *parts;
    {
        std::lock_guard<std::mutex> g(g_noalloc_mu);
        if (!g_noalloc_dev_ptr[device]) {
            uint8_t *ptr = nullptr;
            CU(cudaGetSymbolAddress((void **)&ptr, g_noalloc_part_digests));
            if (!ptr) return fail(TPIMD5_ECUDA, "s3_etag_noalloc: no device address for the scratch array on device %d", device);
            int rc = configure_lanes<256, true, 4, 3, 4>();
            if (rc == 0) rc = configure_lanes<256, false, 4, 3, 4>();
            if (rc == 0) rc = configure_chain<2, 2, 4, 2, 4, 4>();
            if (rc == 0) rc = configure_chain<4, 1, 2, 4, 4, 4>();
            if (rc) return rc;
            g_noalloc_dev_ptr[device] = ptr;
        }
        parts = g_noalloc_dev_ptr[device];
    }

    cudaStream_t st = (cudaStream_t)cuda_stream;
    // At most 10,000 chains: far too few to fill the SMs, so every chain runs at its dependent-op latency.
    // Parts that the TMA unit can address (16-byte aligned data and part size, at most 2 GiB per part) take the
    // latency kernel in the two shapes tpimd5_dev_s3_etag would pick for this many parts; anything else the
    // general kernel with 256-byte stages and the shortest step formulation.  Encoding a tensor map is
    // host-side arithmetic into a struct on this stack: nothing is allocated.
    // per-part MD5s into the built-in array (one lane per part) ...
    const bool aligned = ((uintptr_t)d_data & 15) == 0 && (part_size & 15) == 0;
    const uint64_t n_full = size / part_size;               // whole parts; a ragged last part follows (or the one empty part)
    int rc = 0;
    uint64_t done_parts = 0;
    if (aligned && n_full && part_size <= ((uint64_t)1 << 31) && tma_layout_ok(d_data, part_size, n_full)) {
        CUtensorMap tmap, tail;
        rc = encode_rows(&tmap, d_data, part_size, n_full, part_size);
        if (rc) return rc;
        const uint64_t tail_len = size - n_full * part_size;   // the ragged last part rides in the same launch
        if (tail_len) {
            rc = encode_rows(&tail, (const uint8_t *)d_data + n_full * part_size, tail_len, 1, (tail_len + 15) & ~(uint64_t)15);
            if (rc) return rc;
        }
        rc = (n_full + 31) / 32 <= 148
                 ? enqueue_chain<2, 2, 4, 2, 4, 4>(tmap, n_full, (uint32_t)part_size, (uint4 *)parts, st, 0, LANE_FINAL, &tail, (uint32_t)tail_len)
                 : enqueue_chain<4, 1, 2, 4, 4, 4>(tmap, n_full, (uint32_t)part_size, (uint4 *)parts, st, 0, LANE_FINAL, &tail, (uint32_t)tail_len);
        if (rc) return rc;
        done_parts = np;
    }
    if (done_parts < np) {
        LaneArgs a{};
        a.base = (const uint8_t *)d_data + done_parts * part_size; a.base_bytes = size - done_parts * part_size; a.stride = part_size;
        a.table = (uint4 *)parts + done_parts; a.n = np - done_parts;
        rc = aligned ? enqueue_lanes<256, true, 4, 3, 4>(a, st) : enqueue_lanes<256, false, 4, 3, 4>(a, st);
        if (rc) return rc;
    }
    // ... then the MD5 of those 16*np bytes, read back from the same array (one chain)
    if (encode_tiled_fn() == nullptr) {
        LaneArgs b{};
        b.base = parts; b.base_bytes = np * 16; b.stride = np * 16;
        b.table = (uint4 *)d_etag_md5_out; b.n = 1;
        return enqueue_lanes<256, true, 4, 3, 4>(b, st);
    }
    CUtensorMap tmap2;
    rc = encode_rows(&tmap2, parts, np * 16, 1, np * 16);
    if (rc) return rc;
    return enqueue_chain<2, 2, 4, 2, 4, 4>(tmap2, 1, (uint32_t)(np * 16), (uint4 *)d_etag_md5_out, st, 0, LANE_FINAL);
}
