// comm.hpp -- exchange layer of the single-job multi-GPU path.
//
// One process per GPU; a job of `world` ranks exchanges data only through all-gathers of
// device buffers (fixed-size or variable-size).  Two transports:
//   * NcclComm      NCCL over NVLink/NVSwitch.  libnccl.so.2 is dlopen'ed (the torch-bundled
//                   copy is already resident in a torchrun process); the unique id is made on
//                   rank 0 and handed to the other ranks by the launcher (torch.distributed).
//   * HostCallComm  the caller supplies a host-level all-gather (e.g. torch.distributed/gloo);
//                   device buffers are staged through host memory.  This is what the CPU tests
//                   (world_size 2, gloo) and the single-GPU multi-process GPU tests use.
// Replaces nothing in the reference one-to-one: the reference's only parallel step is the
// thread pool of merge_pfp (src/merge_pfp.cpp:97-150).
#pragma once
#include "prim.cuh"
#ifndef PFB_EMU
#include <dlfcn.h>
#endif

extern "C" {
// host-level all-gather-v: every rank passes `send` (counts[rank] bytes) and receives the
// concatenation of all ranks' buffers, in rank order, in `recv` (sum(counts) bytes).  0 = ok.
typedef int (*pfb_allgatherv_fn)(void *user, const void *send, void *recv, const uint64_t *counts, int world);
}

namespace pfb {

struct Comm {
    int rank = 0, world = 1;
    uint64_t bytes_moved = 0;      // bytes received from other ranks so far
    virtual ~Comm() {}
    // device buffers; recv holds sum(counts) bytes, rank r's part at offset sum(counts[0..r))
    virtual void all_gather_v(const void *d_send, void *d_recv, const u64 *counts, pfbStream_t st) = 0;

    // all-to-all-v of device buffers.  cnt[s * world + d] = bytes rank s sends to rank d (the whole matrix, known on every
    // rank); d_send holds this rank's row back to back (part for rank 0 first), d_recv receives column `rank` back to back
    // (part from rank 0 first).
    virtual void all_to_all_v(const void *d_send, void *d_recv, const u64 *cnt, pfbStream_t st) = 0;

    // Peer windows: every rank passes k device pointers into its own memory; out[r * k + i] = rank r's i-th buffer mapped
    // into this process, so that a kernel of this rank can store straight into it over NVLink.  Collective; false (on every
    // rank alike) when the transport has no peer windows - the caller then stages and uses all_to_all_v.
    virtual bool peer_pointers(void *const *mine, int k, std::vector<void *> &out, pfbStream_t st) {
        (void)mine; (void)k; (void)out; (void)st;
        return false;
    }
    void barrier(pfbStream_t st) { u64 z = 0; gather_u64(&z, 1, st); }

    void all_gather(const void *d_send, void *d_recv, u64 bytes, pfbStream_t st) {
        std::vector<u64> c((size_t)world, bytes);
        all_gather_v(d_send, d_recv, c.data(), st);
    }
    // K 64-bit values per rank, host in / host out (out[r*K + i])
    std::vector<u64> gather_u64(const u64 *vals, int K, pfbStream_t st) {
        DevBuf<u64> s((size_t)K), r((size_t)K * world);
        h2d(s.get(), vals, (size_t)K * 8, st);
        all_gather(s.get(), r.get(), (u64)K * 8, st);
        std::vector<u64> out((size_t)K * world);
        d2h(out.data(), r.get(), out.size() * 8, st);
        stream_sync(st);
        return out;
    }
};

struct HostCallComm : Comm {
    pfb_allgatherv_fn fn = nullptr;
    void *user = nullptr;
    void all_gather_v(const void *d_send, void *d_recv, const u64 *counts, pfbStream_t st) override {
        u64 tot = 0;
        for (int r = 0; r < world; ++r) tot += counts[r];
        std::vector<uint8_t> hs(counts[rank] ? counts[rank] : 1), hr(tot ? tot : 1);
        if (counts[rank]) d2h(hs.data(), d_send, counts[rank], st);
        stream_sync(st);
        if (fn(user, hs.data(), hr.data(), counts, world)) throw Error(-8, "host all-gather callback failed");
        if (tot) h2d(d_recv, hr.data(), tot, st);
        stream_sync(st);
        bytes_moved += tot - counts[rank];
    }
    // test transport: every rank's whole row travels through the host all-gather, the receiver picks its column
    void all_to_all_v(const void *d_send, void *d_recv, const u64 *cnt, pfbStream_t st) override {
        std::vector<u64> row((size_t)world, 0);
        u64 tot = 0, mine = 0;
        for (int s = 0; s < world; ++s) { for (int d = 0; d < world; ++d) row[(size_t)s] += cnt[(size_t)s * world + d]; tot += row[(size_t)s]; }
        for (int s = 0; s < world; ++s) mine += cnt[(size_t)s * world + rank];
        std::vector<uint8_t> hs(row[(size_t)rank] ? row[(size_t)rank] : 1), hr(tot ? tot : 1), out(mine ? mine : 1);
        if (row[(size_t)rank]) d2h(hs.data(), d_send, row[(size_t)rank], st);
        stream_sync(st);
        if (fn(user, hs.data(), hr.data(), row.data(), world)) throw Error(-8, "host all-gather callback failed");
        u64 base = 0, o = 0;
        for (int s = 0; s < world; ++s) {
            u64 off = base;
            for (int d = 0; d < rank; ++d) off += cnt[(size_t)s * world + d];
            const u64 c = cnt[(size_t)s * world + rank];
            if (c) memcpy(out.data() + o, hr.data() + off, c);
            o += c;
            base += row[(size_t)s];
        }
        if (mine) h2d(d_recv, out.data(), mine, st);
        stream_sync(st);
        bytes_moved += mine - cnt[(size_t)rank * world + rank];
    }
};

#ifndef PFB_EMU
// ---- NCCL through dlopen: only the handful of entry points the path needs ----------------
struct NcclApi {
    typedef struct ncclComm *comm_t;
    struct uid { char b[128]; };
    int (*GetUniqueId)(uid *) = nullptr;
    int (*CommInitRank)(comm_t *, int, uid, int) = nullptr;
    int (*CommDestroy)(comm_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, comm_t, cudaStream_t) = nullptr;
    int (*Broadcast)(const void *, void *, size_t, int, int, comm_t, cudaStream_t) = nullptr;
    int (*Send)(const void *, size_t, int, int, comm_t, cudaStream_t) = nullptr;
    int (*Recv)(void *, size_t, int, int, comm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    void *h = nullptr;
    std::string err;
    bool load() {
        if (h) return true;
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char *n : names) { h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
        if (!h) { err = std::string("cannot load NCCL: ") + dlerror(); return false; }
#define PFB_NCCL_SYM(f) *(void **)(&f) = dlsym(h, "nccl" #f); if (!f) { err = "NCCL symbol nccl" #f " missing"; return false; }
        PFB_NCCL_SYM(GetUniqueId) PFB_NCCL_SYM(CommInitRank) PFB_NCCL_SYM(CommDestroy) PFB_NCCL_SYM(AllGather)
        PFB_NCCL_SYM(Broadcast) PFB_NCCL_SYM(Send) PFB_NCCL_SYM(Recv) PFB_NCCL_SYM(GroupStart) PFB_NCCL_SYM(GroupEnd) PFB_NCCL_SYM(GetErrorString)
#undef PFB_NCCL_SYM
        return true;
    }
};
inline NcclApi &nccl_api() { static NcclApi a; return a; }

struct NcclComm : Comm {
    NcclApi::comm_t comm = nullptr;
    DevBuf<u8> stage;              // padded slots of the variable-size all-gather
    void check(int rc, const char *what) {
        if (rc != 0) throw Error(-5, std::string("NCCL ") + what + ": " + nccl_api().GetErrorString(rc));
    }
    void init(int rank_, int world_, const void *uid128) {
        NcclApi &a = nccl_api();
        if (!a.load()) throw Error(-5, a.err);
        rank = rank_; world = world_;
        NcclApi::uid id;
        memcpy(id.b, uid128, 128);
        check(a.CommInitRank(&comm, world, id, rank), "CommInitRank");
    }
    ~NcclComm() override {
        for (auto &o : opened) cudaIpcCloseMemHandle(o.second);
        if (comm) nccl_api().CommDestroy(comm);
    }
    // CUDA IPC: the slab a buffer lives in is exported once, opened once per peer and kept open (slabs live as long as
    // the process); what travels per call is (handle, offset) per buffer.  PFB_PEER=0 keeps everything on NCCL.
    std::map<std::string, void *> opened;          // 64-byte handle -> where the peer's slab is mapped here
    bool peer_ok = !(getenv("PFB_PEER") && !strcmp(getenv("PFB_PEER"), "0"));
    bool peer_pointers(void *const *mine, int k, std::vector<void *> &out, pfbStream_t st) override {
        const int W = 10;                          // u64 words per buffer: 8 handle, offset, ok
        std::vector<u64> msg((size_t)k * W, 0);
        bool ok = peer_ok;
        for (int i = 0; i < k && ok; ++i) {
            char *base = nullptr; size_t off = 0;
            cudaIpcMemHandle_t h;
            if (!slab_pool().locate(mine[i], base, off) || cudaIpcGetMemHandle(&h, base) != cudaSuccess) { cudaGetLastError(); ok = false; break; }
            static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
            memcpy(&msg[(size_t)i * W], &h, 64);
            msg[(size_t)i * W + 8] = off;
        }
        for (int i = 0; i < k; ++i) msg[(size_t)i * W + 9] = ok ? 1 : 0;
        std::vector<u64> all = gather_u64(msg.data(), k * W, st);
        for (int r = 0; r < world; ++r) for (int i = 0; i < k; ++i) ok = ok && all[((size_t)r * k + i) * W + 9] != 0;
        out.assign((size_t)world * k, nullptr);
        for (int r = 0; r < world && ok; ++r)
            for (int i = 0; i < k && ok; ++i) {
                if (r == rank) { out[(size_t)r * k + i] = mine[i]; continue; }
                const u64 *m = &all[((size_t)r * k + i) * W];
                std::string key((const char *)m, 64);
                auto it = opened.find(key);
                void *base = nullptr;
                if (it != opened.end()) base = it->second;
                else {
                    cudaIpcMemHandle_t h;
                    memcpy(&h, m, 64);
                    if (cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = false; break; }
                    opened[key] = base;
                }
                out[(size_t)r * k + i] = (char *)base + m[8];
            }
        // a rank that could not map a peer makes every rank fall back
        u64 f = ok ? 1 : 0;
        std::vector<u64> fl = gather_u64(&f, 1, st);
        for (int r = 0; r < world; ++r) ok = ok && fl[(size_t)r] != 0;
        if (!ok) peer_ok = false;
        return ok;
    }
    void all_gather_v(const void *d_send, void *d_recv, const u64 *counts, pfbStream_t st) override {
        NcclApi &a = nccl_api();
        bool equal = true;
        u64 tot = 0, mx = 0;
        for (int r = 0; r < world; ++r) { equal &= counts[r] == counts[0]; tot += counts[r]; if (counts[r] > mx) mx = counts[r]; }
        if (equal) {
            if (counts[0]) check(a.AllGather(d_send, d_recv, counts[0], /*ncclChar*/ 0, comm, st), "AllGather");
        } else if (tot >= (8u << 20) && mx * (u64)world <= tot + tot / 4) {
            // large and nearly equal parts (the ranks' dictionaries / parses): one padded in-place all-gather at full
            // NVLink rate, then the parts are packed.  Eight separate broadcasts reach about half of that.
            const u64 slot = (mx + 15) & ~15ull;
            if (stage.size() < slot * (u64)world) stage.alloc(slot * (u64)world);
            u8 *sg = stage.get();
            if (counts[rank]) PFB_CUDA_CHECK(cudaMemcpyAsync(sg + slot * (u64)rank, d_send, counts[rank], cudaMemcpyDeviceToDevice, st));
            check(a.AllGather(sg + slot * (u64)rank, sg, slot, /*ncclChar*/ 0, comm, st), "AllGather");
            u64 off = 0;
            for (int r = 0; r < world; ++r) {
                if (counts[r]) PFB_CUDA_CHECK(cudaMemcpyAsync((char *)d_recv + off, sg + slot * (u64)r, counts[r], cudaMemcpyDeviceToDevice, st));
                off += counts[r];
            }
        } else {
            // all-gather-v = one broadcast per root inside a group (fused by NCCL into one launch)
            check(a.GroupStart(), "GroupStart");
            u64 off = 0;
            int rcb = 0;
            for (int r = 0; r < world && !rcb; ++r) {
                if (counts[r]) rcb = a.Broadcast(d_send, (char *)d_recv + off, counts[r], 0, r, comm, st);
                off += counts[r];
            }
            const int rce = a.GroupEnd();        // the group is closed on the error path too
            check(rcb, "Broadcast");
            check(rce, "GroupEnd");
        }
        bytes_moved += tot - counts[rank];
    }
    // grouped point-to-point sends / receives over NVLink (one fused launch); the part a rank keeps is a device copy
    void all_to_all_v(const void *d_send, void *d_recv, const u64 *cnt, pfbStream_t st) override {
        NcclApi &a = nccl_api();
        std::vector<u64> soff((size_t)world + 1, 0), roff((size_t)world + 1, 0);
        for (int r = 0; r < world; ++r) {
            soff[(size_t)r + 1] = soff[(size_t)r] + cnt[(size_t)rank * world + r];
            roff[(size_t)r + 1] = roff[(size_t)r] + cnt[(size_t)r * world + rank];
        }
        const u64 self = cnt[(size_t)rank * world + rank];
        if (self) PFB_CUDA_CHECK(cudaMemcpyAsync((char *)d_recv + roff[(size_t)rank], (const char *)d_send + soff[(size_t)rank], self, cudaMemcpyDeviceToDevice, st));
        const u64 CH = 1ull << 30;               // a send / receive pair moves at most 1 GiB per call
        int rc = a.GroupStart(), rc2 = 0;
        for (int r = 0; r < world && !rc; ++r) {
            if (r == rank) continue;
            const u64 sc = cnt[(size_t)rank * world + r], rcv = cnt[(size_t)r * world + rank];
            for (u64 o = 0; o < sc && !rc; o += CH) rc = a.Send((const char *)d_send + soff[(size_t)r] + o, (size_t)(sc - o < CH ? sc - o : CH), /*ncclChar*/ 0, r, comm, st);
            for (u64 o = 0; o < rcv && !rc; o += CH) rc = a.Recv((char *)d_recv + roff[(size_t)r] + o, (size_t)(rcv - o < CH ? rcv - o : CH), /*ncclChar*/ 0, r, comm, st);
        }
        rc2 = a.GroupEnd();                      // always closed, also on the error path
        check(rc, "Send/Recv");
        check(rc2, "GroupEnd");
        bytes_moved += roff[(size_t)world] - self;
    }
};
#endif

}  // namespace pfb
