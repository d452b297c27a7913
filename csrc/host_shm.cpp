// CPU shared-memory transport: see include/ccmpi_host.h for the contract.
//
// Collectives follow one discipline so that the per-rank staging areas can be shared by every
// communicator a rank belongs to:
//     write own IN area -> barrier -> read others' IN, write own RES area -> barrier ->
//     read others' RES
// and any phase that reads IN is followed by a barrier before the call returns.
// Reductions run in group-rank order 0..n-1, i.e. the order of the root loop in the
// reference's myAllreduce (mpi_wrapper/comm.py:85-93), so results are deterministic.
#include "../include/ccmpi_host.h"

#include <errno.h>
#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <deque>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../include/ccmpi_b200.h"  // dtype / op enums

namespace {

thread_local std::string g_err;
void set_error(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
}

constexpr uint64_t MAGIC = 0x63636d70695f7631ull;  // "ccmpi_v1"

struct alignas(64) CommSlot {
    std::atomic<uint32_t> count;
    std::atomic<uint32_t> gen;
};

struct alignas(64) RingCtl {
    std::atomic<uint64_t> head;  // bytes produced
    char pad0[56];
    std::atomic<uint64_t> tail;   // bytes consumed
    std::atomic<uint64_t> acked;  // out-of-band messages the consumer has finished with (see ccmpi_host_ack_post)
    char pad1[48];
};
static_assert(sizeof(RingCtl) == 128, "ring control block layout");

struct Header {
    std::atomic<uint64_t> magic;
    uint32_t nranks;
    std::atomic<uint32_t> comm_ids;  // communicator ids handed out so far, job-wide (ccmpi_host_reserve_comm_ids)
    uint64_t stage_bytes;
    uint64_t ring_bytes;
    std::atomic<uint32_t> attached;
    std::atomic<uint32_t> failed;  // any rank that errors out raises this so peers stop spinning
};

struct MsgHdr {
    int32_t comm_id;
    int32_t tag;
    uint64_t bytes;
    uint64_t check;
};

struct Req {
    bool is_send;
    int peer;
    MsgHdr hdr;
    char* buf;
    size_t bytes;
    size_t hdr_done = 0;
    size_t done = 0;
    bool complete = false;
    int error = 0;
};

struct Handle {
    std::string name;
    int rank = 0, nranks = 0;
    size_t stage_bytes = 0, ring_bytes = 0, total = 0;
    double timeout_s = 120.0;
    char* base = nullptr;
    Header* hdr = nullptr;
    CommSlot* slots = nullptr;
    RingCtl* rings = nullptr;
    char* ring_data = nullptr;
    char* stage = nullptr;
    // pending p2p requests, FIFO per (direction, peer)
    std::map<int64_t, Req*> reqs;
    std::vector<std::deque<Req*>> sendq, recvq;
    int64_t next_req = 1;

    RingCtl& ring(int src, int dst) { return rings[(size_t)src * nranks + dst]; }
    char* rdata(int src, int dst) { return ring_data + ((size_t)src * nranks + dst) * ring_bytes; }
    char* in_area(int r) { return stage + (size_t)r * 2 * stage_bytes; }
    char* res_area(int r) { return stage + (size_t)r * 2 * stage_bytes + stage_bytes; }
};

std::mutex g_mu;
std::map<int64_t, Handle*> g_handles;
int64_t g_next = 1;

Handle* get(int64_t h) {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_handles.find(h);
    return it == g_handles.end() ? nullptr : it->second;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct Backoff {
    uint32_t spins = 0;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    // returns false on timeout
    bool pause(double timeout_s) {
        ++spins;
        if (spins < 64) {
#if defined(__x86_64__)
            __builtin_ia32_pause();
#endif
        } else {
            sched_yield();
            if ((spins & 0x3ff) == 0) {
                double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
                if (s > timeout_s) return false;
            }
        }
        return true;
    }
};

int barrier(Handle* H, int slot, int n) {
    if (n <= 1) return 0;
    CommSlot& s = H->slots[slot];
    const uint32_t g = s.gen.load(std::memory_order_acquire);
    if (s.count.fetch_add(1, std::memory_order_acq_rel) + 1 == (uint32_t)n) {
        s.count.store(0, std::memory_order_relaxed);
        s.gen.store(g + 1, std::memory_order_release);
        return 0;
    }
    Backoff b;
    while (s.gen.load(std::memory_order_acquire) == g) {
        if (H->hdr->failed.load(std::memory_order_relaxed)) {
            set_error("a peer rank failed while this rank waited in a barrier (slot %d)", slot);
            return -4;
        }
        if (!b.pause(H->timeout_s)) {
            set_error("barrier timeout after %.0f s (slot %d, %d members)", H->timeout_s, slot, n);
            H->hdr->failed.store(1);
            return -4;
        }
    }
    return 0;
}

// ---- typed reductions ----
inline float bf16_to_f32(uint16_t v) {
    uint32_t u = (uint32_t)v << 16;
    float f;
    memcpy(&f, &u, 4);
    return f;
}
inline uint16_t f32_to_bf16(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffff) > 0x7f800000) return (uint16_t)((u >> 16) | 0x40);  // NaN
    u += 0x7fff + ((u >> 16) & 1);                                          // round to nearest even
    return (uint16_t)(u >> 16);
}

template <typename T>
void fold(T* acc, const T* x, size_t n, int op) {
    switch (op) {
        case CCMPI_SUM: for (size_t i = 0; i < n; ++i) acc[i] = acc[i] + x[i]; break;
        case CCMPI_MIN: for (size_t i = 0; i < n; ++i) acc[i] = x[i] < acc[i] ? x[i] : acc[i]; break;
        default:        for (size_t i = 0; i < n; ++i) acc[i] = x[i] > acc[i] ? x[i] : acc[i]; break;
    }
}

size_t dtype_size(int dtype) {
    switch (dtype) {
        case CCMPI_F32: case CCMPI_I32: return 4;
        case CCMPI_BF16: case CCMPI_F16: return 2;
        case CCMPI_F64: case CCMPI_I64: return 8;
        case CCMPI_U8: return 1;
    }
    return 0;
}

// out[0:n] = reduce over j of srcs[j][0:n], folded in order j = 0..k-1
int reduce_many(void* out, const char* const* srcs, int k, size_t n, int dtype, int op) {
    const size_t es = dtype_size(dtype);
    if (dtype == CCMPI_BF16) {
        // accumulate in fp32, round once
        std::vector<float> acc(n);
        const uint16_t* s0 = (const uint16_t*)srcs[0];
        for (size_t i = 0; i < n; ++i) acc[i] = bf16_to_f32(s0[i]);
        std::vector<float> tmp(n);
        for (int j = 1; j < k; ++j) {
            const uint16_t* s = (const uint16_t*)srcs[j];
            for (size_t i = 0; i < n; ++i) tmp[i] = bf16_to_f32(s[i]);
            fold<float>(acc.data(), tmp.data(), n, op);
        }
        uint16_t* o = (uint16_t*)out;
        for (size_t i = 0; i < n; ++i) o[i] = f32_to_bf16(acc[i]);
        return 0;
    }
    if (out != (const void*)srcs[0]) memcpy(out, srcs[0], n * es);
    for (int j = 1; j < k; ++j) {
        switch (dtype) {
            case CCMPI_F32: fold<float>((float*)out, (const float*)srcs[j], n, op); break;
            case CCMPI_F64: fold<double>((double*)out, (const double*)srcs[j], n, op); break;
            case CCMPI_I32: fold<int32_t>((int32_t*)out, (const int32_t*)srcs[j], n, op); break;
            case CCMPI_I64: fold<int64_t>((int64_t*)out, (const int64_t*)srcs[j], n, op); break;
            case CCMPI_U8: fold<uint8_t>((uint8_t*)out, (const uint8_t*)srcs[j], n, op); break;
            default:
                set_error("host reduction: unsupported dtype %d", dtype);
                return -3;
        }
    }
    return 0;
}

int check_group(Handle* H, int slot, const int* members, int n, int me) {
    if (!H) { set_error("bad host transport handle"); return -2; }
    if (slot < 0 || slot >= CCMPI_HOST_MAX_COMMS) { set_error("barrier slot %d out of range", slot); return -2; }
    if (n < 1 || n > H->nranks || me < 0 || me >= n || !members) { set_error("bad group description (n=%d me=%d)", n, me); return -2; }
    if (members[me] != H->rank) { set_error("group member %d is world rank %d, not this rank %d", me, members[me], H->rank); return -2; }
    return 0;
}

// ---- p2p progress engine ----
// copy `len` bytes into the ring at logical position `pos`
void ring_put(char* data, size_t cap, uint64_t pos, const char* src, size_t len) {
    size_t off = pos % cap, first = std::min(len, cap - off);
    memcpy(data + off, src, first);
    if (len > first) memcpy(data, src + first, len - first);
}
void ring_get(const char* data, size_t cap, uint64_t pos, char* dst, size_t len) {
    size_t off = pos % cap, first = std::min(len, cap - off);
    memcpy(dst, data + off, first);
    if (len > first) memcpy(dst + first, data, len - first);
}

bool progress_send(Handle* H, Req* r) {
    RingCtl& rc = H->ring(H->rank, r->peer);
    char* data = H->rdata(H->rank, r->peer);
    bool moved = false;
    for (;;) {
        const uint64_t head = rc.head.load(std::memory_order_relaxed);
        const uint64_t tail = rc.tail.load(std::memory_order_acquire);
        size_t space = H->ring_bytes - (size_t)(head - tail);
        if (space == 0) break;
        if (r->hdr_done < sizeof(MsgHdr)) {
            size_t n = std::min(space, sizeof(MsgHdr) - r->hdr_done);
            ring_put(data, H->ring_bytes, head, (const char*)&r->hdr + r->hdr_done, n);
            r->hdr_done += n;
            rc.head.store(head + n, std::memory_order_release);
            moved = true;
            continue;
        }
        if (r->done < r->bytes) {
            size_t n = std::min(space, r->bytes - r->done);
            ring_put(data, H->ring_bytes, head, r->buf + r->done, n);
            r->done += n;
            rc.head.store(head + n, std::memory_order_release);
            moved = true;
            continue;
        }
        break;
    }
    if (r->hdr_done == sizeof(MsgHdr) && r->done == r->bytes) r->complete = true;
    return moved;
}

bool progress_recv(Handle* H, Req* r) {
    RingCtl& rc = H->ring(r->peer, H->rank);
    const char* data = H->rdata(r->peer, H->rank);
    bool moved = false;
    for (;;) {
        const uint64_t tail = rc.tail.load(std::memory_order_relaxed);
        const uint64_t head = rc.head.load(std::memory_order_acquire);
        size_t avail = (size_t)(head - tail);
        if (avail == 0) break;
        if (r->hdr_done < sizeof(MsgHdr)) {
            if (avail < sizeof(MsgHdr)) break;  // headers are read whole
            MsgHdr h;
            ring_get(data, H->ring_bytes, tail, (char*)&h, sizeof(h));
            if (h.check != (MAGIC ^ h.bytes) || h.comm_id != r->hdr.comm_id || h.bytes > r->bytes ||
                (r->hdr.tag >= 0 && h.tag != r->hdr.tag)) {
                set_error("recv from rank %d: message mismatch (comm %d/%d tag %d/%d bytes %llu, buffer %zu)", r->peer,
                          h.comm_id, r->hdr.comm_id, h.tag, r->hdr.tag, (unsigned long long)h.bytes, r->bytes);
                r->error = -2;
                r->complete = true;
                return true;
            }
            r->bytes = (size_t)h.bytes;  // shorter messages are allowed, like MPI
            r->hdr_done = sizeof(MsgHdr);
            rc.tail.store(tail + sizeof(MsgHdr), std::memory_order_release);
            moved = true;
            continue;
        }
        if (r->done < r->bytes) {
            size_t n = std::min(avail, r->bytes - r->done);
            ring_get(data, H->ring_bytes, tail, r->buf + r->done, n);
            r->done += n;
            rc.tail.store(tail + n, std::memory_order_release);
            moved = true;
            continue;
        }
        break;
    }
    if (r->hdr_done == sizeof(MsgHdr) && r->done == r->bytes) r->complete = true;
    return moved;
}

// advance the head of every queue; completed requests leave their queue
bool progress_all(Handle* H) {
    bool moved = false;
    for (int p = 0; p < H->nranks; ++p) {
        auto& sq = H->sendq[p];
        while (!sq.empty()) {
            moved |= progress_send(H, sq.front());
            if (!sq.front()->complete) break;
            sq.pop_front();
        }
        auto& rq = H->recvq[p];
        while (!rq.empty()) {
            moved |= progress_recv(H, rq.front());
            if (!rq.front()->complete) break;
            rq.pop_front();
        }
    }
    return moved;
}

int wait_reqs(Handle* H, const int64_t* ids, int n) {
    Backoff b;
    for (;;) {
        bool all = true;
        int err = 0;
        for (int i = 0; i < n; ++i) {
            auto it = H->reqs.find(ids[i]);
            if (it == H->reqs.end()) continue;  // already reaped
            if (!it->second->complete) all = false;
            else if (it->second->error) err = it->second->error;
        }
        if (all) {
            for (int i = 0; i < n; ++i) {
                auto it = H->reqs.find(ids[i]);
                if (it != H->reqs.end()) {
                    delete it->second;
                    H->reqs.erase(it);
                }
            }
            return err;
        }
        if (progress_all(H)) {
            b = Backoff();
        } else {
            if (H->hdr->failed.load(std::memory_order_relaxed)) {
                set_error("a peer rank failed while this rank waited on point-to-point traffic");
                return -4;
            }
            if (!b.pause(H->timeout_s)) {
                set_error("point-to-point wait timed out after %.0f s", H->timeout_s);
                H->hdr->failed.store(1);
                return -4;
            }
        }
    }
}

int64_t post(Handle* H, bool is_send, int peer, int comm_id, int tag, void* buf, size_t bytes) {
    if (!H) { set_error("bad host transport handle"); return -2; }
    if (peer < 0 || peer >= H->nranks || peer == H->rank) { set_error("p2p peer %d invalid for rank %d", peer, H->rank); return -2; }
    Req* r = new Req();
    r->is_send = is_send;
    r->peer = peer;
    r->hdr.comm_id = comm_id;
    r->hdr.tag = tag;
    r->hdr.bytes = bytes;
    r->hdr.check = MAGIC ^ (uint64_t)bytes;
    r->buf = (char*)buf;
    r->bytes = bytes;
    int64_t id = H->next_req++;
    H->reqs[id] = r;
    (is_send ? H->sendq : H->recvq)[peer].push_back(r);
    progress_all(H);
    return id;
}

}  // namespace

extern "C" const char* ccmpi_host_last_error(void) { return g_err.c_str(); }

extern "C" int64_t ccmpi_host_attach(const char* job, int rank, int nranks, size_t stage_bytes, size_t ring_bytes,
                                     double timeout_s) {
    if (nranks < 1 || nranks > CCMPI_HOST_MAX_RANKS || rank < 0 || rank >= nranks) {
        set_error("attach: rank %d / nranks %d out of range", rank, nranks);
        return -2;
    }
    stage_bytes = align_up(stage_bytes ? stage_bytes : (4u << 20), 4096);
    ring_bytes = align_up(ring_bytes ? ring_bytes : (256u << 10), 4096);
    Handle* H = new Handle();
    H->name = std::string("/ccmpi-") + job;
    H->rank = rank;
    H->nranks = nranks;
    H->stage_bytes = stage_bytes;
    H->ring_bytes = ring_bytes;
    H->timeout_s = timeout_s > 0 ? timeout_s : 120.0;
    size_t off = align_up(sizeof(Header), 4096);
    const size_t off_slots = off;
    off += align_up(sizeof(CommSlot) * CCMPI_HOST_MAX_COMMS, 4096);
    const size_t off_rings = off;
    off += align_up(sizeof(RingCtl) * nranks * nranks, 4096);
    const size_t off_rdata = off;
    off += ring_bytes * nranks * nranks;
    const size_t off_stage = off;
    off += 2 * stage_bytes * nranks;
    H->total = off;

    int fd = shm_open(H->name.c_str(), O_CREAT | O_RDWR, 0600);
    if (fd < 0) {
        set_error("shm_open(%s): %s", H->name.c_str(), strerror(errno));
        delete H;
        return -4;
    }
    if (ftruncate(fd, (off_t)H->total) != 0) {
        set_error("ftruncate(%s, %zu): %s", H->name.c_str(), H->total, strerror(errno));
        close(fd);
        delete H;
        return -4;
    }
    void* p = mmap(nullptr, H->total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) {
        set_error("mmap(%s, %zu): %s", H->name.c_str(), H->total, strerror(errno));
        delete H;
        return -4;
    }
    H->base = (char*)p;
    H->hdr = (Header*)H->base;
    H->slots = (CommSlot*)(H->base + off_slots);
    H->rings = (RingCtl*)(H->base + off_rings);
    H->ring_data = H->base + off_rdata;
    H->stage = H->base + off_stage;
    H->sendq.resize(nranks);
    H->recvq.resize(nranks);
    // a fresh segment is zero-filled (ftruncate), which is a valid initial state for every
    // counter; rank 0 stamps the geometry, everyone else waits for the stamp and checks it
    if (rank == 0) {
        H->hdr->nranks = nranks;
        H->hdr->stage_bytes = stage_bytes;
        H->hdr->ring_bytes = ring_bytes;
        H->hdr->magic.store(MAGIC, std::memory_order_release);
    } else {
        Backoff b;
        while (H->hdr->magic.load(std::memory_order_acquire) != MAGIC) {
            if (!b.pause(H->timeout_s)) {
                set_error("attach: rank 0 never initialised segment %s", H->name.c_str());
                munmap(H->base, H->total);
                delete H;
                return -4;
            }
        }
        if (H->hdr->nranks != (uint32_t)nranks || H->hdr->stage_bytes != stage_bytes || H->hdr->ring_bytes != ring_bytes) {
            set_error("attach: segment %s has a different geometry (stale job name?)", H->name.c_str());
            munmap(H->base, H->total);
            delete H;
            return -2;
        }
    }
    H->hdr->attached.fetch_add(1);
    std::lock_guard<std::mutex> lk(g_mu);
    int64_t id = g_next++;
    g_handles[id] = H;
    return id;
}

extern "C" int ccmpi_host_detach(int64_t h, int unlink_segment) {
    Handle* H;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        auto it = g_handles.find(h);
        if (it == g_handles.end()) return 0;
        H = it->second;
        g_handles.erase(it);
    }
    if (unlink_segment) shm_unlink(H->name.c_str());
    for (auto& kv : H->reqs) delete kv.second;
    munmap(H->base, H->total);
    delete H;
    return 0;
}

extern "C" int ccmpi_host_barrier(int64_t h, int slot, const int* members, int n, int me) {
    Handle* H = get(h);
    int rc = check_group(H, slot, members, n, me);
    if (rc) return rc;
    return barrier(H, slot, n);
}

extern "C" int ccmpi_host_allreduce(int64_t h, int slot, const int* members, int n, int me, const void* src, void* dst,
                                    size_t count, int dtype, int op) {
    Handle* H = get(h);
    int rc = check_group(H, slot, members, n, me);
    if (rc) return rc;
    const size_t es = dtype_size(dtype);
    if (!es || op < CCMPI_SUM || op > CCMPI_MAX) { set_error("allreduce: unsupported dtype %d / op %d", dtype, op); return -3; }
    if (n == 1) { if (dst != src) memmove(dst, src, count * es); return 0; }
    const size_t chunk = H->stage_bytes / es;
    const char* srcs[CCMPI_HOST_MAX_RANKS];
    for (size_t base = 0; base < count; base += chunk) {
        const size_t len = std::min(chunk, count - base);
        memcpy(H->in_area(H->rank), (const char*)src + base * es, len * es);
        if ((rc = barrier(H, slot, n))) return rc;
        if (len * es <= 32 * 1024) {
            // small: every member folds all inputs itself, straight into dst -- two barriers, no
            // result area
            for (int j = 0; j < n; ++j) srcs[j] = H->in_area(members[j]);
            if ((rc = reduce_many((char*)dst + base * es, srcs, n, len, dtype, op))) return rc;
            if ((rc = barrier(H, slot, n))) return rc;
            continue;
        }
        const size_t per = (len + n - 1) / n;
        const size_t lo = std::min(len, per * me), hi = std::min(len, lo + per);
        if (hi > lo) {
            for (int j = 0; j < n; ++j) srcs[j] = H->in_area(members[j]) + lo * es;
            if ((rc = reduce_many(H->res_area(H->rank) + lo * es, srcs, n, hi - lo, dtype, op))) return rc;
        }
        if ((rc = barrier(H, slot, n))) return rc;
        for (int j = 0; j < n; ++j) {
            const size_t jlo = std::min(len, per * j), jhi = std::min(len, jlo + per);
            if (jhi > jlo) memcpy((char*)dst + (base + jlo) * es, H->res_area(members[j]) + jlo * es, (jhi - jlo) * es);
        }
        // Nobody leaves while a member is still reading result areas: the staging areas belong to
        // the RANK, not to the communicator, and a member that ran ahead into a collective of
        // another communicator (whose barrier does not include the slow reader) would overwrite
        // its result area under that reader.
        if ((rc = barrier(H, slot, n))) return rc;
    }
    return 0;
}

extern "C" int ccmpi_host_allgather(int64_t h, int slot, const int* members, int n, int me, const void* src, void* dst,
                                    size_t bytes) {
    Handle* H = get(h);
    int rc = check_group(H, slot, members, n, me);
    if (rc) return rc;
    if (n == 1) { if (dst != src) memmove(dst, src, bytes); return 0; }
    for (size_t base = 0; base < bytes; base += H->stage_bytes) {
        const size_t len = std::min(H->stage_bytes, bytes - base);
        memcpy(H->in_area(H->rank), (const char*)src + base, len);
        if ((rc = barrier(H, slot, n))) return rc;
        for (int j = 0; j < n; ++j) memcpy((char*)dst + (size_t)j * bytes + base, H->in_area(members[j]), len);
        if ((rc = barrier(H, slot, n))) return rc;
    }
    return 0;
}

extern "C" int ccmpi_host_reduce_scatter(int64_t h, int slot, const int* members, int n, int me, const void* src,
                                         void* dst, size_t count, int dtype, int op) {
    Handle* H = get(h);
    int rc = check_group(H, slot, members, n, me);
    if (rc) return rc;
    const size_t es = dtype_size(dtype);
    if (!es || op < CCMPI_SUM || op > CCMPI_MAX) { set_error("reduce_scatter: unsupported dtype %d / op %d", dtype, op); return -3; }
    if (n == 1) { if (dst != src) memmove(dst, src, count * es); return 0; }
    // block b of every rank travels through that rank's IN area, `chunk` elements of each
    // block at a time
    const size_t chunk = H->stage_bytes / es / n;
    if (chunk == 0) { set_error("reduce_scatter: staging area too small for %d ranks", n); return -5; }
    const char* srcs[CCMPI_HOST_MAX_RANKS];
    for (size_t base = 0; base < count; base += chunk) {
        const size_t len = std::min(chunk, count - base);
        for (int b = 0; b < n; ++b)
            memcpy(H->in_area(H->rank) + (size_t)b * len * es, (const char*)src + ((size_t)b * count + base) * es, len * es);
        if ((rc = barrier(H, slot, n))) return rc;
        for (int j = 0; j < n; ++j) srcs[j] = H->in_area(members[j]) + (size_t)me * len * es;
        if ((rc = reduce_many((char*)dst + base * es, srcs, n, len, dtype, op))) return rc;
        if ((rc = barrier(H, slot, n))) return rc;
    }
    return 0;
}

extern "C" int ccmpi_host_alltoall(int64_t h, int slot, const int* members, int n, int me, const void* src, void* dst,
                                   size_t seg) {
    Handle* H = get(h);
    int rc = check_group(H, slot, members, n, me);
    if (rc) return rc;
    if (n == 1) { if (dst != src) memmove(dst, src, seg); return 0; }
    const size_t chunk = H->stage_bytes / n;
    if (chunk == 0) { set_error("alltoall: staging area too small for %d ranks", n); return -5; }
    for (size_t base = 0; base < seg; base += chunk) {
        const size_t len = std::min(chunk, seg - base);
        for (int b = 0; b < n; ++b) memcpy(H->in_area(H->rank) + (size_t)b * len, (const char*)src + (size_t)b * seg + base, len);
        if ((rc = barrier(H, slot, n))) return rc;
        for (int j = 0; j < n; ++j) memcpy((char*)dst + (size_t)j * seg + base, H->in_area(members[j]) + (size_t)me * len, len);
        if ((rc = barrier(H, slot, n))) return rc;
    }
    return 0;
}

extern "C" int64_t ccmpi_host_isend(int64_t h, int dst, int comm_id, int tag, const void* buf, size_t bytes) {
    return post(get(h), true, dst, comm_id, tag, const_cast<void*>(buf), bytes);
}
extern "C" int64_t ccmpi_host_irecv(int64_t h, int src, int comm_id, int tag, void* buf, size_t bytes) {
    return post(get(h), false, src, comm_id, tag, buf, bytes);
}
extern "C" int ccmpi_host_test(int64_t h, int64_t req) {
    Handle* H = get(h);
    if (!H) { set_error("bad host transport handle"); return -2; }
    progress_all(H);
    auto it = H->reqs.find(req);
    if (it == H->reqs.end()) return 1;
    if (!it->second->complete) return 0;
    int err = it->second->error;
    delete it->second;
    H->reqs.erase(it);
    return err ? err : 1;
}
extern "C" int ccmpi_host_waitall(int64_t h, const int64_t* reqs, int n) {
    Handle* H = get(h);
    if (!H) { set_error("bad host transport handle"); return -2; }
    for (int i = 0; i < n; ++i)
        if (reqs[i] < 0) { set_error("waitall: request %d is invalid", i); return -2; }
    return wait_reqs(H, reqs, n);
}
extern "C" int ccmpi_host_send(int64_t h, int dst, int comm_id, int tag, const void* buf, size_t bytes) {
    int64_t r = ccmpi_host_isend(h, dst, comm_id, tag, buf, bytes);
    if (r < 0) return (int)r;
    return ccmpi_host_waitall(h, &r, 1);
}
extern "C" int ccmpi_host_recv(int64_t h, int src, int comm_id, int tag, void* buf, size_t bytes) {
    int64_t r = ccmpi_host_irecv(h, src, comm_id, tag, buf, bytes);
    if (r < 0) return (int)r;
    return ccmpi_host_waitall(h, &r, 1);
}
// Out-of-band completion counters, one per ordered pair, next to that pair's ring. The device
// point-to-point path sends a small descriptor through the ring while the payload stays in the
// sender's symmetric heap; when the receiver has copied the payload out (over NVLink) it posts an
// ack here and the sender may reuse the staging block. Counters only grow; messages of one pair
// are consumed in send order, so a count identifies which blocks are free.
// Communicator ids come from ONE job-wide counter: two groups that share no rank (the two halves of a
// split, say) still split again at the same time, and ids picked from their own members' state alone
// would collide -- and with them the barrier slot and the signal pad both index by id.
extern "C" int ccmpi_host_reserve_comm_ids(int64_t h, int count) {
    Handle* H = get(h);
    if (!H) { set_error("bad handle"); return -1; }
    if (count < 0) { set_error("reserve_comm_ids: count %d", count); return -2; }
    // id 0 is the world; the counter stores ids in use beyond it
    uint32_t base = 1 + H->hdr->comm_ids.fetch_add((uint32_t)count, std::memory_order_relaxed);
    if (base + (uint32_t)count > CCMPI_HOST_MAX_COMMS) {
        set_error("out of communicator ids (%d): reuse communicators instead of re-splitting", CCMPI_HOST_MAX_COMMS);
        return -3;
    }
    return (int)base;
}

extern "C" int ccmpi_host_ack_post(int64_t h, int src) {
    Handle* H = get(h);
    if (!H || src < 0 || src >= H->nranks || src == H->rank) { set_error("ack_post: bad handle or peer %d", src); return -2; }
    H->ring(src, H->rank).acked.fetch_add(1, std::memory_order_release);
    return 0;
}

extern "C" int64_t ccmpi_host_ack_count(int64_t h, int dst) {
    Handle* H = get(h);
    if (!H || dst < 0 || dst >= H->nranks || dst == H->rank) { set_error("ack_count: bad handle or peer %d", dst); return -2; }
    return (int64_t)H->ring(H->rank, dst).acked.load(std::memory_order_acquire);
}

// Non-blocking look at the next message queued from `src`: 1 and its header fields when a whole header
// is at the head of the ring and no posted receive is ahead of it, 0 otherwise. What MPI_Iprobe and
// receives from MPI_ANY_SOURCE are built from (mpi/core.py); ordering per pair stays strict FIFO.
extern "C" int ccmpi_host_iprobe(int64_t h, int src, int* comm_id, int* tag, int64_t* bytes) {
    Handle* H = get(h);
    if (!H) { set_error("bad host transport handle"); return -2; }
    if (src < 0 || src >= H->nranks || src == H->rank) { set_error("iprobe: bad source %d", src); return -2; }
    if (H->hdr->failed.load(std::memory_order_relaxed)) { set_error("a peer rank failed"); return -4; }
    progress_all(H);
    if (!H->recvq[src].empty()) return 0;
    RingCtl& rc = H->ring(src, H->rank);
    const uint64_t tail = rc.tail.load(std::memory_order_relaxed);
    const uint64_t head = rc.head.load(std::memory_order_acquire);
    if (head - tail < sizeof(MsgHdr)) return 0;
    MsgHdr m;
    ring_get(H->rdata(src, H->rank), H->ring_bytes, tail, (char*)&m, sizeof(m));
    if (m.check != (MAGIC ^ m.bytes)) { set_error("iprobe: corrupt message header from rank %d", src); return -2; }
    if (comm_id) *comm_id = m.comm_id;
    if (tag) *tag = m.tag;
    if (bytes) *bytes = (int64_t)m.bytes;
    return 1;
}

extern "C" int64_t ccmpi_host_probe(int64_t h, int src) {
    Handle* H = get(h);
    if (!H) { set_error("bad host transport handle"); return -2; }
    if (src < 0 || src >= H->nranks || src == H->rank) { set_error("probe: bad source %d", src); return -2; }
    RingCtl& rc = H->ring(src, H->rank);
    Backoff b;
    for (;;) {
        progress_all(H);
        if (H->recvq[src].empty()) {
            const uint64_t tail = rc.tail.load(std::memory_order_relaxed);
            const uint64_t head = rc.head.load(std::memory_order_acquire);
            if (head - tail >= sizeof(MsgHdr)) {
                MsgHdr m;
                ring_get(H->rdata(src, H->rank), H->ring_bytes, tail, (char*)&m, sizeof(m));
                return (int64_t)m.bytes;
            }
        }
        if (!b.pause(H->timeout_s)) {
            set_error("probe: no message from rank %d within %.0f s", src, H->timeout_s);
            return -4;
        }
    }
}
