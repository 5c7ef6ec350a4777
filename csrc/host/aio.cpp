// Asynchronous tensor <-> file I/O engine for ZeRO-Infinity (NVMe tier).
// Kernel AIO through raw syscalls (io_setup / io_submit / io_getevents; no libaio dependency), O_DIRECT with
// automatic fallback to buffered I/O when the file system refuses it, one io_context per worker thread, each thread
// owning the contiguous slice [tid * bytes / T, (tid + 1) * bytes / T) of the transfer, block_size-sized requests
// with at most queue_depth in flight, sequential (single_submit) or batched submission, optional overlap of
// submission and completion reaping. Plus a page-aligned pinned-style allocator helper and a parallel memcpy.
// Same capabilities as reference csrc/aio/py_lib/* and csrc/aio/common/* (deepspeed_py_aio_handle.cpp:15-241,
// deepspeed_aio_thread.cpp:12-84, deepspeed_aio_common.cpp:69-265, deepspeed_py_copy.cpp:23-133).
#include <errno.h>
#include <fcntl.h>
#include <linux/aio_abi.h>
#include <omp.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/syscall.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

namespace {

inline int sys_io_setup(unsigned nr, aio_context_t* ctx) { return (int)syscall(SYS_io_setup, nr, ctx); }
inline int sys_io_destroy(aio_context_t ctx) { return (int)syscall(SYS_io_destroy, ctx); }
inline int sys_io_submit(aio_context_t ctx, long n, struct iocb** iocbs) { return (int)syscall(SYS_io_submit, ctx, n, iocbs); }
inline int sys_io_getevents(aio_context_t ctx, long min_nr, long nr, struct io_event* ev, struct timespec* ts) {
    return (int)syscall(SYS_io_getevents, ctx, min_nr, nr, ev, ts);
}

struct AioConfig {
    long block_size;
    int queue_depth;
    bool single_submit;
    bool overlap_events;
};

double now_sec() {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + ts.tv_nsec * 1e-9;
}

int open_file(const char* path, bool read, bool* direct) {
    int flags = read ? O_RDONLY : (O_WRONLY | O_CREAT);
    int fd = open(path, flags | O_DIRECT, 0644);
    *direct = true;
    if (fd < 0) {
        fd = open(path, flags, 0644);
        *direct = false;
    }
    return fd;
}

// Synchronous fallback for unaligned tails / file systems without O_DIRECT+AIO support.
long rw_fallback(int fd, bool read, char* buf, long nbytes, long offset) {
    long done = 0;
    while (done < nbytes) {
        ssize_t r = read ? pread(fd, buf + done, nbytes - done, offset + done) : pwrite(fd, buf + done, nbytes - done, offset + done);
        if (r < 0) {
            if (errno == EINTR) continue;
            return -errno;
        }
        if (r == 0) break;
        done += r;
    }
    return done;
}

// Transfer [offset, offset + nbytes) of the file <-> buf with kernel AIO; returns bytes moved or -errno.
long do_aio(int fd, bool read, bool direct, char* buf, long nbytes, long offset, const AioConfig& cfg) {
    const long align = 512;
    const bool aligned = ((uintptr_t)buf % align == 0) && (offset % align == 0);
    long aio_bytes = (direct && !aligned) ? 0 : (direct ? nbytes / align * align : nbytes);
    long moved = 0;
    if (aio_bytes > 0) {
        aio_context_t ctx = 0;
        if (sys_io_setup(cfg.queue_depth, &ctx) < 0) {
            aio_bytes = 0;
        } else {
            const long nblocks = (aio_bytes + cfg.block_size - 1) / cfg.block_size;
            std::vector<struct iocb> cbs(cfg.queue_depth);
            std::vector<struct io_event> events(cfg.queue_depth);
            long next = 0, inflight = 0, completed = 0;
            bool failed = false;
            std::vector<int> free_slots;
            for (int i = cfg.queue_depth - 1; i >= 0; --i) free_slots.push_back(i);
            while (completed < nblocks && !failed) {
                // fill the queue
                std::vector<struct iocb*> batch;
                while (next < nblocks && !free_slots.empty()) {
                    int slot = free_slots.back();
                    free_slots.pop_back();
                    struct iocb* cb = &cbs[slot];
                    memset(cb, 0, sizeof(*cb));
                    const long boff = next * cfg.block_size;
                    const long blen = (boff + cfg.block_size <= aio_bytes) ? cfg.block_size : aio_bytes - boff;
                    cb->aio_fildes = fd;
                    cb->aio_lio_opcode = read ? IOCB_CMD_PREAD : IOCB_CMD_PWRITE;
                    cb->aio_buf = (uint64_t)(uintptr_t)(buf + boff);
                    cb->aio_nbytes = blen;
                    cb->aio_offset = offset + boff;
                    cb->aio_data = (uint64_t)slot;
                    batch.push_back(cb);
                    ++next;
                    if (cfg.single_submit) {
                        struct iocb* one = cb;
                        if (sys_io_submit(ctx, 1, &one) != 1) { failed = true; break; }
                        batch.pop_back();
                        ++inflight;
                    }
                }
                if (!failed && !batch.empty()) {
                    long sub = 0;
                    while (sub < (long)batch.size()) {
                        int r = sys_io_submit(ctx, batch.size() - sub, batch.data() + sub);
                        if (r <= 0) { failed = true; break; }
                        sub += r;
                        inflight += r;
                    }
                }
                if (failed) break;
                // reap: with overlap_events take whatever is ready (at least one), otherwise drain the whole queue
                const long want = cfg.overlap_events ? 1 : inflight;
                int got = sys_io_getevents(ctx, want, inflight, events.data(), nullptr);
                if (got < 0) {
                    if (errno == EINTR) continue;
                    failed = true;
                    break;
                }
                for (int i = 0; i < got; ++i) {
                    const int slot = (int)events[i].data;
                    // a short transfer (EOF, ENOSPC part way) is a failure of the block, not a success with fewer bytes
                    if ((long)events[i].res < 0 || (uint64_t)events[i].res != cbs[slot].aio_nbytes) failed = true;
                    else moved += events[i].res;
                    free_slots.push_back(slot);
                }
                inflight -= got;
                completed += got;
            }
            sys_io_destroy(ctx);
            if (failed) {
                // the file system may not support AIO for this file: redo everything synchronously
                moved = 0;
                aio_bytes = 0;
            }
        }
    }
    if (aio_bytes < nbytes) {
        // tail (or everything): O_DIRECT cannot take unaligned lengths -> use a buffered descriptor
        long r = rw_fallback(fd, read, buf + aio_bytes, nbytes - aio_bytes, offset + aio_bytes);
        if (r < 0 && direct) {
            char linkpath[64], real[4096];
            snprintf(linkpath, sizeof(linkpath), "/proc/self/fd/%d", fd);
            ssize_t n = readlink(linkpath, real, sizeof(real) - 1);
            if (n > 0) {
                real[n] = 0;
                int fd2 = open(real, read ? O_RDONLY : O_WRONLY);
                if (fd2 >= 0) {
                    r = rw_fallback(fd2, read, buf + aio_bytes, nbytes - aio_bytes, offset + aio_bytes);
                    close(fd2);
                }
            }
        }
        if (r < 0) return r;
        moved += r;
    }
    if (moved != nbytes) return -EIO;       // never report a partial transfer as done
    return moved;
}

struct Op {
    bool read;
    char* buf;
    long nbytes;
    char path[4096];
    bool validate;
};

struct Handle {
    AioConfig cfg;
    int num_threads;
    std::vector<std::thread> workers;
    std::mutex mu;
    std::condition_variable cv_work, cv_done;
    std::deque<Op> queue;       // every queued op is executed by ALL threads (each its slice)
    long op_seq = 0;            // ops pushed
    std::vector<long> thread_seq;
    long ops_completed = 0;     // ops finished by all threads
    std::vector<int> pending_per_op;
    std::deque<long> errors;
    bool stop = false;
    long inflight_ops = 0;
    double last_gbps = 0.0;

    void worker(int tid) {
        long my_next = 0;
        for (;;) {
            Op op;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv_work.wait(lk, [&] { return stop || my_next < op_seq; });
                if (stop && my_next >= op_seq) return;
                op = queue[my_next - (op_seq - (long)queue.size())];
            }
            long err = run_slice(op, tid);
            {
                std::unique_lock<std::mutex> lk(mu);
                if (err < 0) errors.push_back(err);
                const long qi = my_next - (op_seq - (long)queue.size());
                if (--pending_per_op[qi] == 0) {
                    // ops complete in order because every thread walks the queue in order
                    while (!queue.empty() && pending_per_op.front() == 0) {
                        queue.pop_front();
                        pending_per_op.erase(pending_per_op.begin());
                        ++ops_completed;
                        --inflight_ops;
                    }
                    cv_done.notify_all();
                }
                ++my_next;
            }
        }
    }

    long run_slice(const Op& op, int tid) {
        bool direct;
        int fd = open_file(op.path, op.read, &direct);
        if (fd < 0) return -errno;
        // slices are 512-byte aligned so O_DIRECT stays usable inside every thread's range
        long per = (op.nbytes / num_threads) / 512 * 512;
        long start = (long)tid * per;
        long len = (tid == num_threads - 1) ? op.nbytes - start : per;
        long r = 0;
        if (len > 0) r = do_aio(fd, op.read, direct, op.buf + start, len, start, cfg);
        close(fd);
        if (r < 0) return r;
        return (r == len) ? 0 : -EIO;
    }
};

}  // namespace

extern "C" {

void* dsb_aio_handle_create(long block_size, int queue_depth, int single_submit, int overlap_events, int num_threads) {
    Handle* h = new Handle();
    h->cfg.block_size = block_size > 0 ? block_size : (1 << 20);
    h->cfg.queue_depth = queue_depth > 0 ? queue_depth : 8;
    h->cfg.single_submit = single_submit != 0;
    h->cfg.overlap_events = overlap_events != 0;
    h->num_threads = num_threads > 0 ? num_threads : 1;
    for (int t = 0; t < h->num_threads; ++t) h->workers.emplace_back(&Handle::worker, h, t);
    return h;
}

void dsb_aio_handle_destroy(void* hp) {
    Handle* h = reinterpret_cast<Handle*>(hp);
    {
        std::unique_lock<std::mutex> lk(h->mu);
        h->stop = true;
    }
    h->cv_work.notify_all();
    for (auto& w : h->workers) w.join();
    delete h;
}

// Queue a read (file -> buf) or write (buf -> file) of nbytes; returns immediately. For writes the file is truncated
// to nbytes first so the parallel slices land in a file of the right size.
int dsb_aio_submit(void* hp, int is_read, void* buf, long nbytes, const char* path) {
    Handle* h = reinterpret_cast<Handle*>(hp);
    if (!is_read) {
        int fd = open(path, O_WRONLY | O_CREAT, 0644);
        if (fd < 0) return -errno;
        if (ftruncate(fd, nbytes) != 0) { int e = errno; close(fd); return -e; }
        close(fd);
    } else {
        struct stat st;
        if (stat(path, &st) != 0) return -errno;
        if (st.st_size < nbytes) return -EINVAL;
    }
    Op op;
    op.read = is_read != 0;
    op.buf = reinterpret_cast<char*>(buf);
    op.nbytes = nbytes;
    strncpy(op.path, path, sizeof(op.path) - 1);
    op.path[sizeof(op.path) - 1] = 0;
    op.validate = false;
    {
        std::unique_lock<std::mutex> lk(h->mu);
        h->queue.push_back(op);
        h->pending_per_op.push_back(h->num_threads);
        ++h->op_seq;
        ++h->inflight_ops;
    }
    h->cv_work.notify_all();
    return 0;
}

// Block until every queued operation has finished; returns the number completed since the last wait, or -errno.
long dsb_aio_wait(void* hp) {
    Handle* h = reinterpret_cast<Handle*>(hp);
    std::unique_lock<std::mutex> lk(h->mu);
    h->cv_done.wait(lk, [&] { return h->inflight_ops == 0; });
    long done = h->ops_completed;
    h->ops_completed = 0;
    if (!h->errors.empty()) {
        long e = h->errors.front();
        h->errors.clear();
        return e;
    }
    return done;
}

long dsb_aio_pending(void* hp) {
    Handle* h = reinterpret_cast<Handle*>(hp);
    std::unique_lock<std::mutex> lk(h->mu);
    return h->inflight_ops;
}

// One-shot synchronous helpers (reference aio_read / aio_write free functions).
long dsb_aio_sync_rw(int is_read, void* buf, long nbytes, const char* path, long block_size, int queue_depth,
                     int single_submit, int overlap_events) {
    void* h = dsb_aio_handle_create(block_size, queue_depth, single_submit, overlap_events, 1);
    int rc = dsb_aio_submit(h, is_read, buf, nbytes, path);
    long r = rc < 0 ? rc : dsb_aio_wait(h);
    dsb_aio_handle_destroy(h);
    return r;
}

// Page-aligned, optionally mlock'ed host buffer (staging for O_DIRECT).
void* dsb_aio_alloc_aligned(long nbytes, int lock_pages) {
    void* p = nullptr;
    const long sz = (nbytes + 4095) / 4096 * 4096;
    if (posix_memalign(&p, 4096, sz) != 0) return nullptr;
    if (lock_pages) mlock(p, sz);
    return p;
}

void dsb_aio_free_aligned(void* p, long nbytes, int locked) {
    if (!p) return;
    if (locked) munlock(p, (nbytes + 4095) / 4096 * 4096);
    free(p);
}

// OpenMP host memcpy (reference deepspeed_memcpy): large copies are bandwidth-bound on one core otherwise.
int dsb_parallel_memcpy(void* dst, const void* src, long nbytes) {
    const long chunk = 1L << 22;
    const long n = (nbytes + chunk - 1) / chunk;
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i) {
        const long off = i * chunk;
        const long len = off + chunk <= nbytes ? chunk : nbytes - off;
        memcpy(static_cast<char*>(dst) + off, static_cast<const char*>(src) + off, len);
    }
    return 0;
}

}  // extern "C"
