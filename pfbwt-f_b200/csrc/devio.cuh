// devio.cuh -- device buffer -> file (the writers of src/pfbwt-f.cpp:229-243,298-328 and
// include/pfbwt_io.hpp:234-249 hand std::vectors to fwrite; here the bytes live in HBM).
//
// A GB-sized output copied into a pageable std::vector and then fwrite()n pays for page faults twice and runs
// on one core: 0.8 s for the 1.1 GB .dict of config 2, more than the whole device pipeline.  write_dev_file
// keeps up to eight workers busy, each with its own pinned 16 MB buffer and copy stream: D2H of chunk i overlaps
// the pwrite() of the chunks before it, and the page-cache copies run on several cores.  read_file_to_dev is the
// same thing in the other direction (pread -> H2D) for the FASTA reader of ingest.cuh.
#pragma once
#include "prim.cuh"
#include <fcntl.h>
#include <unistd.h>
#include <sys/stat.h>
#include <mutex>
#include <thread>
#include <atomic>

namespace pfb {

#ifndef PFB_EMU
struct DevIoPool {
    static const int MAXW = 8;
    static const size_t CHUNK = 16u << 20;
    u8 *buf[MAXW] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaStream_t st[MAXW];
    int ready = 0;
    std::mutex mu;                  // one file at a time per process
    bool ensure(int nw) {
        while (ready < nw) {
            if (cudaHostAlloc((void **)&buf[ready], CHUNK, cudaHostAllocDefault) != cudaSuccess) return false;
            if (cudaStreamCreateWithFlags(&st[ready], cudaStreamNonBlocking) != cudaSuccess) return false;
            ++ready;
        }
        return true;
    }
};
// one pool per thread: the background writer of the parse files (engine.hpp) and the pipeline thread copy at the same time
inline DevIoPool &devio_pool() { static thread_local DevIoPool p; return p; }
#endif

// Writes `bytes` bytes of device memory at `off` of the file.  whole = true: the file is created/truncated and
// holds exactly these bytes; false: the file is sized to final_size and other writers (ranks of a multi-GPU job)
// fill the rest.  The producing stream `st` is synchronised first.
inline int write_dev_file(const std::string &path, const void *dptr, size_t bytes, u64 off, u64 final_size, bool whole, int device,
                          pfbStream_t st, std::string &err, bool sync_producer = true) {
    if (sync_producer) stream_sync(st);
    int fd = open(path.c_str(), whole ? (O_CREAT | O_WRONLY | O_TRUNC) : (O_CREAT | O_WRONLY), 0644);
    if (fd < 0) { err = "could not open " + path + " for writing"; return -1; }
    if (!whole && ftruncate(fd, (off_t)final_size) != 0) { close(fd); err = "could not size " + path; return -1; }
    auto put = [&](const u8 *p, size_t len, u64 at) -> bool {
        while (len) {
            ssize_t k = pwrite(fd, p, len, (off_t)at);
            if (k <= 0) return false;
            p += k; at += (u64)k; len -= (size_t)k;
        }
        return true;
    };
    bool ok = true;
#ifdef PFB_EMU
    (void)device;
    ok = put((const u8 *)dptr, bytes, off);
#else
    DevIoPool &pool = devio_pool();
    std::lock_guard<std::mutex> lock(pool.mu);
    const size_t CH = DevIoPool::CHUNK;
    const u64 nchunk = (bytes + CH - 1) / CH;
    int nw = nchunk < (u64)DevIoPool::MAXW ? (int)nchunk : DevIoPool::MAXW;
    if (nw < 1) nw = 1;
    if (!pool.ensure(nw)) { close(fd); err = "could not allocate pinned staging memory"; return -1; }
    std::atomic<int> bad(0);
    auto worker = [&](int k) {
        if (cudaSetDevice(device) != cudaSuccess) { bad = 1; return; }
        for (u64 c = (u64)k; c < nchunk && !bad; c += (u64)nw) {
            const size_t o = (size_t)c * CH, len = bytes - o < CH ? bytes - o : CH;
            if (cudaMemcpyAsync(pool.buf[k], (const u8 *)dptr + o, len, cudaMemcpyDeviceToHost, pool.st[k]) != cudaSuccess ||
                cudaStreamSynchronize(pool.st[k]) != cudaSuccess) { bad = 2; return; }
            if (!put(pool.buf[k], len, off + o)) { bad = 3; return; }
        }
    };
    if (nw == 1) worker(0);
    else {
        std::vector<std::thread> th;
        for (int k = 1; k < nw; ++k) th.emplace_back(worker, k);
        worker(0);
        for (auto &t : th) t.join();
    }
    if (bad == 1 || bad == 2) { close(fd); err = "device to host copy failed while writing " + path; return -1; }
    ok = bad == 0;
#endif
    if (close(fd) != 0) ok = false;
    if (!ok) { err = "could not write file " + path; return -1; }
    return 0;
}

#ifndef PFB_EMU
// Reads bytes [foff, foff + m) of fd into device memory.  first(buf, len) sees the first chunk on the host (the FASTA reader
// looks for the first header there).  Returns 0, 1 when the file turned out shorter than m, -1 on errors.
template <class F>
inline int read_file_to_dev(int fd, u64 m, void *dptr, int device, F first, std::string &err, u64 foff = 0) {
    DevIoPool &pool = devio_pool();
    std::lock_guard<std::mutex> lock(pool.mu);
    const size_t CH = DevIoPool::CHUNK;
    const u64 nchunk = (m + CH - 1) / CH;
    int nw = nchunk < (u64)DevIoPool::MAXW ? (int)nchunk : DevIoPool::MAXW;
    if (nw < 1) nw = 1;
    if (!pool.ensure(nw)) { err = "could not allocate pinned staging memory"; return -1; }
    std::atomic<int> bad(0);
    auto worker = [&](int k) {
        if (cudaSetDevice(device) != cudaSuccess) { bad = 2; return; }
        for (u64 c = (u64)k; c < nchunk && !bad; c += (u64)nw) {
            const u64 o = c * CH;
            const size_t len = (size_t)(m - o < CH ? m - o : CH);
            size_t have = 0;
            while (have < len) {
                ssize_t got = pread(fd, pool.buf[k] + have, len - have, (off_t)(foff + o + have));
                if (got < 0) { bad = 2; return; }
                if (got == 0) { bad = 1; return; }
                have += (size_t)got;
            }
            if (c == 0) first(pool.buf[k], len);
            if (cudaMemcpyAsync((u8 *)dptr + o, pool.buf[k], len, cudaMemcpyHostToDevice, pool.st[k]) != cudaSuccess ||
                cudaStreamSynchronize(pool.st[k]) != cudaSuccess) { bad = 2; return; }
        }
    };
    if (nw == 1) worker(0);
    else {
        std::vector<std::thread> th;
        for (int k = 1; k < nw; ++k) th.emplace_back(worker, k);
        worker(0);
        for (auto &t : th) t.join();
    }
    if (bad == 2) { err = "error reading input"; return -1; }
    return bad == 1 ? 1 : 0;
}
#endif

// device array -> a stream that must be written in order (--stdout bwt|sa, src/pfbwt-f.cpp:113-127): two pinned buffers,
// the D2H of the next chunk runs while the current one is written.  No host copy of the whole array (a 64 GB .sa
// would not fit next to the page cache).  0 ok, -1 with err set.
inline int write_dev_stream(FILE *f, const void *dptr, size_t bytes, int device, pfbStream_t st, std::string &err) {
    stream_sync(st);
#ifdef PFB_EMU
    (void)device;
    if (bytes && fwrite(dptr, 1, bytes, f) != bytes) { err = "could not write to the output stream"; return -1; }
    fflush(f);
    return 0;
#else
    DevIoPool &pool = devio_pool();
    std::lock_guard<std::mutex> lock(pool.mu);
    if (!pool.ensure(2)) { err = "could not allocate pinned staging memory"; return -1; }
    if (cudaSetDevice(device) != cudaSuccess) { err = "cudaSetDevice failed"; return -1; }
    const size_t CH = DevIoPool::CHUNK;
    const u64 nchunk = (bytes + CH - 1) / CH;
    auto start = [&](u64 c) {
        const size_t o = (size_t)c * CH, len = bytes - o < CH ? bytes - o : CH;
        return cudaMemcpyAsync(pool.buf[c & 1], (const u8 *)dptr + o, len, cudaMemcpyDeviceToHost, pool.st[c & 1]) == cudaSuccess;
    };
    if (nchunk && !start(0)) { err = "device to host copy failed"; return -1; }
    for (u64 c = 0; c < nchunk; ++c) {
        const size_t o = (size_t)c * CH, len = bytes - o < CH ? bytes - o : CH;
        if (cudaStreamSynchronize(pool.st[c & 1]) != cudaSuccess) { err = "device to host copy failed"; return -1; }
        if (c + 1 < nchunk && !start(c + 1)) { err = "device to host copy failed"; return -1; }
        if (fwrite(pool.buf[c & 1], 1, len, f) != len) { err = "could not write to the output stream"; return -1; }
    }
    fflush(f);
    return 0;
#endif
}

// whole file -> device buffer (allocated here, `slack` spare bytes behind the data).  0 ok, -1 with err set.
inline int read_whole_file_to_dev(const std::string &path, DevBuf<u8> &buf, u64 &size, size_t slack, int device, pfbStream_t st, std::string &err) {
    int fd = open(path.c_str(), O_RDONLY);
    if (fd < 0) { err = "error opening " + path; return -1; }
    struct stat sb;
    if (fstat(fd, &sb) != 0) { close(fd); err = "error opening " + path; return -1; }
    size = (u64)sb.st_size;
    buf.alloc(size + slack + 1);
    int rc = 0;
#ifdef PFB_EMU
    (void)device; (void)st;
    u64 off = 0;
    while (off < size) {
        ssize_t got = read(fd, buf.get() + off, (size_t)(size - off < (1u << 26) ? size - off : (1u << 26)));
        if (got <= 0) { rc = -1; break; }
        off += (u64)got;
    }
#else
    stream_sync(st);
    std::string e;
    if (size) rc = read_file_to_dev(fd, size, buf.get(), device, [](const u8 *, size_t) {}, e) ? -1 : 0;
#endif
    close(fd);
    if (rc) { err = "error reading " + path; return -1; }
    return 0;
}

}  // namespace pfb
