// smallfile_bench.cc -- host-only: how fast can N x 4 KiB files be brought into one buffer?
//
// The ingest of BASELINE config 2 from real files is bound by open/read/close on the host.
// This tool times the candidates on the box it runs on: full-path open, openat relative to a cached directory
// descriptor, O_NOATIME, reader threads beyond the core count, and io_uring with linked openat -> read -> close
// requests (raw system calls; prints whether the kernel / seccomp profile allows io_uring at all).
//     g++ -O2 -std=c++17 -o build/smallfile_bench tools/smallfile_bench.cc -lpthread && build/smallfile_bench [files]
#include <atomic>
#include <chrono>
#include <cerrno>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fcntl.h>
#include <linux/io_uring.h>
#include <sched.h>
#include <string>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/syscall.h>
#include <thread>
#include <unistd.h>
#include <vector>

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

struct Files {
    std::string root;
    std::vector<std::string> paths;      // full paths, listing order (directory by directory)
    std::vector<uint32_t> dir_of;        // index into dirs
    std::vector<std::string> dirs, names;
};

static Files make_files(size_t n) {
    Files f;
    char tmpl[] = "/dev/shm/tpi_sfb_XXXXXX";
    f.root = mkdtemp(tmpl);
    const size_t ndirs = 1024, per = (n + ndirs - 1) / ndirs;
    std::vector<char> buf(4096, 5);
    for (size_t d = 0; d < ndirs; ++d) {
        char name[64];
        snprintf(name, sizeof name, "%s/%04zu", f.root.c_str(), d);
        mkdir(name, 0755);
        f.dirs.push_back(name);
    }
    for (size_t i = 0; i < n; ++i) {
        char name[32];
        snprintf(name, sizeof name, "%07zu", i);
        f.names.push_back(name);
        f.dir_of.push_back((uint32_t)(i / per));
        f.paths.push_back(f.dirs[i / per] + "/" + name);
        int fd = open(f.paths.back().c_str(), O_CREAT | O_WRONLY | O_TRUNC, 0644);
        memcpy(buf.data(), &i, sizeof i);
        if (write(fd, buf.data(), 4096) != 4096) perror("write");
        close(fd);
    }
    return f;
}

template <class Fn>
static double run_threads(int nth, size_t n, size_t grain, Fn fn) {
    std::atomic<size_t> next{0};
    const double t0 = now();
    std::vector<std::thread> th;
    for (int t = 0; t < nth; ++t)
        th.emplace_back([&] {
            for (;;) {
                const size_t lo = next.fetch_add(grain);
                if (lo >= n) break;
                fn(lo, std::min(n, lo + grain));
            }
        });
    for (auto &t : th) t.join();
    return now() - t0;
}

// ---- minimal io_uring (no liburing on this image) ----------------------------------------------------------
struct Ring {
    int fd = -1;
    io_uring_params p{};
    uint8_t *sq_ptr = nullptr, *cq_ptr = nullptr;
    io_uring_sqe *sqes = nullptr;
    unsigned *sq_head, *sq_tail, *sq_mask, *sq_array, *cq_head, *cq_tail, *cq_mask;
    io_uring_cqe *cqes;
    size_t sq_len = 0, cq_len = 0;
    int init(unsigned entries) {
        memset(&p, 0, sizeof p);
        fd = (int)syscall(__NR_io_uring_setup, entries, &p);
        if (fd < 0) return -errno;
        sq_len = p.sq_off.array + p.sq_entries * sizeof(unsigned);
        cq_len = p.cq_off.cqes + p.cq_entries * sizeof(io_uring_cqe);
        sq_ptr = (uint8_t *)mmap(nullptr, sq_len, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_POPULATE, fd, IORING_OFF_SQ_RING);
        cq_ptr = (uint8_t *)mmap(nullptr, cq_len, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_POPULATE, fd, IORING_OFF_CQ_RING);
        sqes = (io_uring_sqe *)mmap(nullptr, p.sq_entries * sizeof(io_uring_sqe), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_POPULATE, fd,
                                    IORING_OFF_SQES);
        if (sq_ptr == MAP_FAILED || cq_ptr == MAP_FAILED || sqes == MAP_FAILED) return -ENOMEM;
        sq_head = (unsigned *)(sq_ptr + p.sq_off.head); sq_tail = (unsigned *)(sq_ptr + p.sq_off.tail);
        sq_mask = (unsigned *)(sq_ptr + p.sq_off.ring_mask); sq_array = (unsigned *)(sq_ptr + p.sq_off.array);
        cq_head = (unsigned *)(cq_ptr + p.cq_off.head); cq_tail = (unsigned *)(cq_ptr + p.cq_off.tail);
        cq_mask = (unsigned *)(cq_ptr + p.cq_off.ring_mask); cqes = (io_uring_cqe *)(cq_ptr + p.cq_off.cqes);
        return 0;
    }
    void close_ring() { if (fd >= 0) close(fd); }
};

int main(int argc, char **argv) {
    const size_t n = argc > 1 ? strtoull(argv[1], nullptr, 10) : 200000;
    const unsigned hc = std::thread::hardware_concurrency();
    printf("{\"files\": %zu, \"host_threads\": %u}\n", n, hc);
    Files f = make_files(n);
    std::vector<uint8_t> dst(n * 4096 + 64);
    std::vector<const char *> cpaths(n);
    for (size_t i = 0; i < n; ++i) cpaths[i] = f.paths[i].c_str();

    for (int nth : {(int)hc, (int)hc * 2, (int)hc * 4}) {
        for (int method = 0; method < 4; ++method) {
            std::atomic<size_t> bad{0};
            const double s = run_threads(nth, n, 64, [&](size_t lo, size_t hi) {
                int dfd = -1;
                uint32_t cur = UINT32_MAX;
                for (size_t i = lo; i < hi; ++i) {
                    int fd;
                    if (method == 0) fd = open(cpaths[i], O_RDONLY | O_CLOEXEC);
                    else if (method == 1) fd = open(cpaths[i], O_RDONLY | O_CLOEXEC | O_NOATIME);
                    else {
                        if (f.dir_of[i] != cur) {
                            if (dfd >= 0) close(dfd);
                            dfd = open(f.dirs[f.dir_of[i]].c_str(), O_RDONLY | O_DIRECTORY | O_CLOEXEC);
                            cur = f.dir_of[i];
                        }
                        fd = openat(dfd, f.names[i].c_str(), O_RDONLY | O_CLOEXEC | (method == 3 ? O_NOATIME : 0));
                    }
                    if (fd < 0) { ++bad; continue; }
                    if (pread(fd, dst.data() + i * 4096, 4097, 0) != 4096) ++bad;
                    close(fd);
                }
                if (dfd >= 0) close(dfd);
            });
            const char *names[] = {"open(full path)", "open(full path, O_NOATIME)", "openat(dirfd)", "openat(dirfd, O_NOATIME)"};
            printf("{\"method\": \"%s\", \"threads\": %d, \"s\": %.4f, \"files_per_s\": %.0f, \"bad\": %zu}\n", names[method], nth, s, n / s,
                   bad.load());
        }
    }
    // every reader thread with a file-descriptor table of its own (unshare(CLONE_FILES)): open/close then take no
    // lock that the process's other threads share
    for (int nth : {(int)hc, (int)hc * 2}) {
        std::atomic<size_t> bad{0};
        std::atomic<size_t> next{0};
        std::atomic<int> unshare_errno{0};
        const double t0 = now();
        std::vector<std::thread> th;
        for (int t = 0; t < nth; ++t)
            th.emplace_back([&] {
                if (unshare(CLONE_FILES) != 0) unshare_errno = errno;
                for (;;) {
                    const size_t lo = next.fetch_add(64);
                    if (lo >= n) break;
                    for (size_t i = lo; i < std::min(n, lo + 64); ++i) {
                        int fd = open(cpaths[i], O_RDONLY | O_CLOEXEC);
                        if (fd < 0) { ++bad; continue; }
                        if (pread(fd, dst.data() + i * 4096, 4097, 0) != 4096) ++bad;
                        close(fd);
                    }
                }
            });
        for (auto &t : th) t.join();
        const double s = now() - t0;
        printf("{\"method\": \"open(full path), private fd table per thread\", \"threads\": %d, \"s\": %.4f, \"files_per_s\": %.0f, "
               "\"bad\": %zu, \"unshare_errno\": %d}\n", nth, s, n / s, bad.load(), unshare_errno.load());
    }
    // stat-free baseline of the syscalls alone: open + close, no read
    {
        const double s = run_threads((int)hc, n, 64, [&](size_t lo, size_t hi) {
            for (size_t i = lo; i < hi; ++i) { int fd = open(cpaths[i], O_RDONLY | O_CLOEXEC); if (fd >= 0) close(fd); }
        });
        printf("{\"method\": \"open+close only\", \"threads\": %u, \"s\": %.4f, \"files_per_s\": %.0f}\n", hc, s, n / s);
    }
    // io_uring: per thread one ring; per file three linked requests (openat into a fixed-file slot is avoided:
    // plain openat returns the fd in the completion, so read/close cannot be linked to it without direct
    // descriptors; use direct descriptors: IORING_FILE_INDEX_ALLOC is too new to rely on, so slots are assigned by hand)
    {
        Ring probe;
        const int rc = probe.init(8);
        if (rc < 0) {
            printf("{\"method\": \"io_uring\", \"available\": false, \"errno\": %d, \"error\": \"%s\"}\n", -rc, strerror(-rc));
        } else {
            probe.close_ring();
            for (int nth : {(int)hc, (int)hc / 2 > 0 ? (int)hc / 2 : 1}) {
                std::atomic<size_t> bad{0};
                std::atomic<int> setup_err{0};
                const double s = run_threads(nth, n, 4096, [&](size_t lo, size_t hi) {
                    constexpr unsigned DEPTH = 64;                     // files in flight per ring
                    Ring r;
                    if (r.init(DEPTH * 4) < 0) { setup_err = errno; return; }
                    std::vector<int> slots(DEPTH, -1);
                    if (syscall(__NR_io_uring_register, r.fd, IORING_REGISTER_FILES, slots.data(), DEPTH) < 0) { setup_err = errno; r.close_ring(); return; }
                    size_t next = lo, inflight = 0;
                    std::vector<unsigned> free_slots;
                    for (unsigned k = 0; k < DEPTH; ++k) free_slots.push_back(k);
                    while (next < hi || inflight) {
                        unsigned tail = *r.sq_tail, queued = 0;
                        while (next < hi && !free_slots.empty()) {
                            const unsigned slot = free_slots.back();
                            free_slots.pop_back();
                            const size_t i = next++;
                            io_uring_sqe *e;
                            e = &r.sqes[tail & *r.sq_mask]; memset(e, 0, sizeof *e);
                            e->opcode = IORING_OP_OPENAT; e->fd = AT_FDCWD; e->addr = (uint64_t)(uintptr_t)cpaths[i]; e->open_flags = O_RDONLY;
                            e->file_index = slot + 1; e->flags = IOSQE_IO_LINK; e->user_data = (i << 2) | 0;
                            r.sq_array[tail & *r.sq_mask] = tail & *r.sq_mask; ++tail;
                            e = &r.sqes[tail & *r.sq_mask]; memset(e, 0, sizeof *e);
                            e->opcode = IORING_OP_READ; e->fd = (int)slot; e->flags = IOSQE_FIXED_FILE | IOSQE_IO_LINK; e->addr = (uint64_t)(uintptr_t)(dst.data() + i * 4096);
                            e->len = 4096; e->off = 0; e->user_data = (i << 2) | 1;
                            r.sq_array[tail & *r.sq_mask] = tail & *r.sq_mask; ++tail;
                            e = &r.sqes[tail & *r.sq_mask]; memset(e, 0, sizeof *e);
                            e->opcode = IORING_OP_CLOSE; e->file_index = slot + 1; e->user_data = ((uint64_t)slot << 34) | (i << 2) | 2;
                            r.sq_array[tail & *r.sq_mask] = tail & *r.sq_mask; ++tail;
                            queued += 3; ++inflight;
                        }
                        __atomic_store_n(r.sq_tail, tail, __ATOMIC_RELEASE);
                        const int got = (int)syscall(__NR_io_uring_enter, r.fd, queued, 1, IORING_ENTER_GETEVENTS, nullptr, 0);
                        if (got < 0 && errno != EINTR) { setup_err = errno; break; }
                        unsigned head = *r.cq_head;
                        const unsigned ctail = __atomic_load_n(r.cq_tail, __ATOMIC_ACQUIRE);
                        while (head != ctail) {
                            const io_uring_cqe &c = r.cqes[head & *r.cq_mask];
                            const unsigned kind = c.user_data & 3;
                            if (kind == 1 && c.res != 4096) ++bad;
                            if (kind == 0 && c.res < 0) ++bad;
                            if (kind == 2) { free_slots.push_back((unsigned)(c.user_data >> 34)); --inflight; }
                            ++head;
                        }
                        __atomic_store_n(r.cq_head, head, __ATOMIC_RELEASE);
                    }
                    r.close_ring();
                });
                printf("{\"method\": \"io_uring linked openat->read->close (direct descriptors)\", \"available\": true, \"threads\": %d, \"s\": %.4f, "
                       "\"files_per_s\": %.0f, \"bad\": %zu, \"setup_errno\": %d}\n", nth, s, n / s, bad.load(), setup_err.load());
            }
        }
    }
    std::string cmd = "rm -rf " + f.root;
    if (system(cmd.c_str()) != 0) perror("rm");
    return 0;
}
