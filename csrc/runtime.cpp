// N0 device runtime: symmetric heap over the CUDA virtual-memory-management API, segment
// sharing between the one-process-per-GPU ranks (POSIX fds passed over abstract unix sockets
// with SCM_RIGHTS), and NVLS multicast objects per (sub-)communicator.
//
// This is what stands where the reference has "mpi4py + libmpi" (requirements.txt:4,
// README.md:11-24): MPI.COMM_WORLD becomes a table of peer-mapped segment pointers, and
// MPI_Comm_split (comm.py:38-39, func_impl.py:58,62) becomes a subset of that table plus a
// multicast object bound to the member GPUs. No payload ever travels through this file: it
// only creates mappings. The python layer (runtime/world.py) orders the phases with its
// control-plane barrier.
#include <errno.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <map>
#include <mutex>
#include <thread>
#include <vector>

#include "host_common.h"

namespace {

using namespace ccmpi;

// ---- driver API through cudart entry points (no link-time libcuda) ----
typedef CUresult (*PFN_MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
typedef CUresult (*PFN_MemRelease)(CUmemGenericAllocationHandle);
typedef CUresult (*PFN_MemExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
typedef CUresult (*PFN_MemImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
typedef CUresult (*PFN_MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
typedef CUresult (*PFN_MemAddressFree)(CUdeviceptr, size_t);
typedef CUresult (*PFN_MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
typedef CUresult (*PFN_MemUnmap)(CUdeviceptr, size_t);
typedef CUresult (*PFN_MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
typedef CUresult (*PFN_MemGetGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
typedef CUresult (*PFN_DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
typedef CUresult (*PFN_McCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
typedef CUresult (*PFN_McAddDevice)(CUmemGenericAllocationHandle, CUdevice);
typedef CUresult (*PFN_McBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
typedef CUresult (*PFN_McGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
typedef CUresult (*PFN_McUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);

struct Drv {
    PFN_MemCreate MemCreate;
    PFN_MemRelease MemRelease;
    PFN_MemExport MemExport;
    PFN_MemImport MemImport;
    PFN_MemAddressReserve MemAddressReserve;
    PFN_MemAddressFree MemAddressFree;
    PFN_MemMap MemMap;
    PFN_MemUnmap MemUnmap;
    PFN_MemSetAccess MemSetAccess;
    PFN_MemGetGranularity MemGetGranularity;
    PFN_DeviceGetAttribute DeviceGetAttribute;
    PFN_McCreate McCreate;
    PFN_McAddDevice McAddDevice;
    PFN_McBindMem McBindMem;
    PFN_McGetGranularity McGetGranularity;
    PFN_McUnbind McUnbind;
    bool ok = false;
    bool mc_ok = false;
};

Drv& D() {
    static Drv d;
    static std::once_flag once;
    std::call_once(once, [] {
        d.MemCreate = drv<PFN_MemCreate>("cuMemCreate");
        d.MemRelease = drv<PFN_MemRelease>("cuMemRelease");
        d.MemExport = drv<PFN_MemExport>("cuMemExportToShareableHandle");
        d.MemImport = drv<PFN_MemImport>("cuMemImportFromShareableHandle");
        d.MemAddressReserve = drv<PFN_MemAddressReserve>("cuMemAddressReserve");
        d.MemAddressFree = drv<PFN_MemAddressFree>("cuMemAddressFree");
        d.MemMap = drv<PFN_MemMap>("cuMemMap");
        d.MemUnmap = drv<PFN_MemUnmap>("cuMemUnmap");
        d.MemSetAccess = drv<PFN_MemSetAccess>("cuMemSetAccess");
        d.MemGetGranularity = drv<PFN_MemGetGranularity>("cuMemGetAllocationGranularity");
        d.DeviceGetAttribute = drv<PFN_DeviceGetAttribute>("cuDeviceGetAttribute");
        d.ok = d.MemCreate && d.MemRelease && d.MemExport && d.MemImport && d.MemAddressReserve && d.MemAddressFree &&
               d.MemMap && d.MemUnmap && d.MemSetAccess && d.MemGetGranularity && d.DeviceGetAttribute;
        d.McCreate = drv<PFN_McCreate>("cuMulticastCreate");
        d.McAddDevice = drv<PFN_McAddDevice>("cuMulticastAddDevice");
        d.McBindMem = drv<PFN_McBindMem>("cuMulticastBindMem");
        d.McGetGranularity = drv<PFN_McGetGranularity>("cuMulticastGetGranularity");
        d.McUnbind = drv<PFN_McUnbind>("cuMulticastUnbind");
        d.mc_ok = d.McCreate && d.McAddDevice && d.McBindMem && d.McGetGranularity;
    });
    return d;
}

struct McGroup {
    int tag = -1;
    int nmembers = 0;
    bool leader = false;
    CUmemGenericAllocationHandle handle = 0;
    int export_fd = -1;  // leader only
    CUdeviceptr va = 0;
    size_t size = 0;
    bool bound = false;
};

struct Heap {
    int device = 0;
    size_t bytes = 0;
    bool emulated = false;
    int nranks = 1;
    int rank = 0;
    CUmemGenericAllocationHandle local_handle = 0;
    int local_fd = -1;
    CUdeviceptr ptr[CCMPI_MAX_RANKS] = {0};
    CUmemGenericAllocationHandle peer_handle[CCMPI_MAX_RANKS] = {0};
    // fd server
    int listen_fd = -1;
    std::thread server;
    std::atomic<bool> stop{false};
    std::mutex mu;  // guards mc
    std::vector<McGroup> mc;
};

std::mutex g_mu;
std::map<int64_t, Heap*> g_heaps;
int64_t g_next = 1;

Heap* get(int64_t h) {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_heaps.find(h);
    return it == g_heaps.end() ? nullptr : it->second;
}

// ---- unix-socket plumbing ----
struct Req {
    int32_t kind;  // 0 = heap segment fd, 1 = multicast fd
    int32_t tag;
    int32_t from;
    int32_t pad;
};

socklen_t make_addr(sockaddr_un& a, const char* job, int rank) {
    memset(&a, 0, sizeof(a));
    a.sun_family = AF_UNIX;
    // abstract namespace: leading NUL, no filesystem entry to clean up
    int n = snprintf(a.sun_path + 1, sizeof(a.sun_path) - 1, "ccmpi-%s-%d", job, rank);
    return (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
}

int send_fd(int sock, int fd, int32_t status) {
    msghdr msg;
    memset(&msg, 0, sizeof(msg));
    iovec io{&status, sizeof(status)};
    msg.msg_iov = &io;
    msg.msg_iovlen = 1;
    char ctrl[CMSG_SPACE(sizeof(int))];
    if (fd >= 0) {
        memset(ctrl, 0, sizeof(ctrl));
        msg.msg_control = ctrl;
        msg.msg_controllen = sizeof(ctrl);
        cmsghdr* c = CMSG_FIRSTHDR(&msg);
        c->cmsg_level = SOL_SOCKET;
        c->cmsg_type = SCM_RIGHTS;
        c->cmsg_len = CMSG_LEN(sizeof(int));
        memcpy(CMSG_DATA(c), &fd, sizeof(int));
    }
    return sendmsg(sock, &msg, 0) == (ssize_t)sizeof(status) ? 0 : -1;
}

int recv_fd(int sock, int* fd_out, int32_t* status) {
    msghdr msg;
    memset(&msg, 0, sizeof(msg));
    iovec io{status, sizeof(*status)};
    msg.msg_iov = &io;
    msg.msg_iovlen = 1;
    char ctrl[CMSG_SPACE(sizeof(int))];
    memset(ctrl, 0, sizeof(ctrl));
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof(ctrl);
    ssize_t n = recvmsg(sock, &msg, 0);
    if (n != (ssize_t)sizeof(*status)) return -1;
    *fd_out = -1;
    for (cmsghdr* c = CMSG_FIRSTHDR(&msg); c; c = CMSG_NXTHDR(&msg, c)) {
        if (c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS) memcpy(fd_out, CMSG_DATA(c), sizeof(int));
    }
    return 0;
}

void serve(Heap* h) {
    while (!h->stop.load()) {
        pollfd pf{h->listen_fd, POLLIN, 0};
        int pr = poll(&pf, 1, 100);
        if (pr <= 0) continue;
        int s = accept(h->listen_fd, nullptr, nullptr);
        if (s < 0) continue;
        Req rq;
        if (recv(s, &rq, sizeof(rq), MSG_WAITALL) == (ssize_t)sizeof(rq)) {
            if (rq.kind == 0) {
                send_fd(s, h->local_fd, 0);
            } else {
                // the requester may be ahead of the leader's cuMulticastCreate: wait for the tag
                int fd = -1;
                for (int tries = 0; tries < 600 && fd < 0 && !h->stop.load(); ++tries) {
                    {
                        std::lock_guard<std::mutex> lk(h->mu);
                        for (auto& g : h->mc)
                            if (g.tag == rq.tag && g.leader) fd = g.export_fd;
                    }
                    if (fd < 0) std::this_thread::sleep_for(std::chrono::milliseconds(50));
                }
                send_fd(s, fd, fd >= 0 ? 0 : -1);
            }
        }
        close(s);
    }
}

int request_fd(const char* job, int peer, const Req& rq, int* fd_out, int timeout_ms) {
    sockaddr_un addr;
    socklen_t len = make_addr(addr, job, peer);
    const auto t0 = std::chrono::steady_clock::now();
    for (;;) {
        int s = socket(AF_UNIX, SOCK_STREAM, 0);
        if (s < 0) {
            set_error("socket(): %s", strerror(errno));
            return CCMPI_ERR_SYS;
        }
        if (connect(s, (sockaddr*)&addr, len) == 0) {
            int32_t status = -1;
            if (send(s, &rq, sizeof(rq), 0) == (ssize_t)sizeof(rq) && recv_fd(s, fd_out, &status) == 0 && status == 0 &&
                *fd_out >= 0) {
                close(s);
                return CCMPI_OK;
            }
            close(s);
            set_error("peer %d refused handle request kind %d tag %d", peer, rq.kind, rq.tag);
            return CCMPI_ERR_SYS;
        }
        close(s);
        auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
        if (ms > timeout_ms) {
            set_error("timed out connecting to peer %d (job %s): %s", peer, job, strerror(errno));
            return CCMPI_ERR_SYS;
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(20));
    }
}

CUmemAllocationProp alloc_prop(int device) {
    CUmemAllocationProp prop;
    memset(&prop, 0, sizeof(prop));
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = device;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return prop;
}

int map_handle(CUmemGenericAllocationHandle handle, size_t bytes, size_t align, int device, CUdeviceptr* out) {
    Drv& d = D();
    CUdeviceptr va = 0;
    CCMPI_DRV_CHECK(d.MemAddressReserve(&va, bytes, align, 0, 0));
    CCMPI_DRV_CHECK(d.MemMap(va, bytes, 0, handle, 0));
    CUmemAccessDesc acc;
    memset(&acc, 0, sizeof(acc));
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = device;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    CCMPI_DRV_CHECK(d.MemSetAccess(va, bytes, &acc, 1));
    *out = va;
    return CCMPI_OK;
}

size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

extern "C" int ccmpi_probe(int device, int* out8) {
    CCMPI_REQUIRE(out8 != nullptr, "probe: null output");
    memset(out8, 0, 8 * sizeof(int));
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        (void)cudaGetLastError();
        n = 0;
    }
    out8[0] = n;
    if (n == 0 || device >= n) return CCMPI_OK;
    cudaDeviceProp p;
    CCMPI_CUDA_CHECK(cudaGetDeviceProperties(&p, device));
    out8[1] = p.major;
    out8[2] = p.minor;
    out8[3] = p.multiProcessorCount;
    CCMPI_CUDA_CHECK(cudaSetDevice(device));
    CCMPI_CUDA_CHECK(cudaFree(0));
    Drv& d = D();
    if (!d.ok) return CCMPI_OK;
    int v = 0;
    if (d.mc_ok && d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, device) == CUDA_SUCCESS) out8[4] = v;
    if (d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, device) == CUDA_SUCCESS) out8[5] = v;
    if (d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, device) == CUDA_SUCCESS) out8[6] = v;
    CUmemAllocationProp prop = alloc_prop(device);
    size_t g = 0;
    if (d.MemGetGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS) out8[7] = (int)(g & 0x7fffffff);
    return CCMPI_OK;
}

extern "C" int64_t ccmpi_heap_create(int device, size_t bytes) {
    if (cudaSetDevice(device) != cudaSuccess || cudaFree(0) != cudaSuccess) {
        set_error("cannot initialise CUDA device %d: %s", device, cudaGetErrorString(cudaGetLastError()));
        return CCMPI_ERR_CUDA;
    }
    Drv& d = D();
    if (!d.ok) {
        set_error("CUDA virtual memory management entry points are unavailable");
        return CCMPI_ERR_UNSUPPORTED;
    }
    CUmemAllocationProp prop = alloc_prop(device);
    size_t gran = 0;
    if (d.MemGetGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || gran == 0) gran = 2u << 20;
    // multicast binding wants its own (usually larger) granularity: honour it up front
    if (d.mc_ok) {
        CUmulticastObjectProp mp;
        memset(&mp, 0, sizeof(mp));
        mp.numDevices = 2;
        mp.size = round_up(bytes, gran);
        mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        size_t mg = 0;
        if (d.McGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > gran) gran = mg;
    }
    bytes = round_up(bytes, gran);
    Heap* h = new Heap();
    h->device = device;
    h->bytes = bytes;
    CUresult r = d.MemCreate(&h->local_handle, bytes, &prop, 0);
    if (r != CUDA_SUCCESS) {
        set_error("cuMemCreate(%zu bytes) failed: CUresult %d", bytes, (int)r);
        delete h;
        return CCMPI_ERR_CUDA;
    }
    if (map_handle(h->local_handle, bytes, gran, device, &h->ptr[0]) != CCMPI_OK) {
        delete h;
        return CCMPI_ERR_CUDA;
    }
    r = d.MemExport(&h->local_fd, h->local_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
        set_error("cuMemExportToShareableHandle failed: CUresult %d", (int)r);
        delete h;
        return CCMPI_ERR_CUDA;
    }
    if (cudaMemset((void*)h->ptr[0], 0, bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
        set_error("zeroing the heap failed: %s", cudaGetErrorString(cudaGetLastError()));
        delete h;
        return CCMPI_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    int64_t id = g_next++;
    g_heaps[id] = h;
    return id;
}

extern "C" int64_t ccmpi_heap_create_emulated(int device, size_t bytes, int nranks) {
    if (nranks < 1 || nranks > CCMPI_MAX_RANKS) {
        set_error("emulated heap: nranks %d out of range", nranks);
        return CCMPI_ERR_ARG;
    }
    if (cudaSetDevice(device) != cudaSuccess) {
        set_error("cannot select CUDA device %d", device);
        return CCMPI_ERR_CUDA;
    }
    Heap* h = new Heap();
    h->device = device;
    h->bytes = round_up(bytes, 4096);
    h->emulated = true;
    h->nranks = nranks;
    for (int i = 0; i < nranks; ++i) {
        void* p = nullptr;
        if (cudaMalloc(&p, h->bytes) != cudaSuccess || cudaMemset(p, 0, h->bytes) != cudaSuccess) {
            set_error("emulated heap: cudaMalloc(%zu) failed: %s", h->bytes, cudaGetErrorString(cudaGetLastError()));
            for (int j = 0; j < i; ++j) cudaFree((void*)h->ptr[j]);
            delete h;
            return CCMPI_ERR_CUDA;
        }
        h->ptr[i] = (CUdeviceptr)p;
    }
    cudaDeviceSynchronize();
    std::lock_guard<std::mutex> lk(g_mu);
    int64_t id = g_next++;
    g_heaps[id] = h;
    return id;
}

extern "C" size_t ccmpi_heap_bytes(int64_t heap) {
    Heap* h = get(heap);
    return h ? h->bytes : 0;
}

extern "C" void* ccmpi_heap_local_ptr(int64_t heap) {
    Heap* h = get(heap);
    if (!h) return nullptr;
    return (void*)h->ptr[h->emulated ? 0 : h->rank];
}

extern "C" int ccmpi_heap_listen(int64_t heap, const char* job, int rank) {
    Heap* h = get(heap);
    CCMPI_REQUIRE(h && !h->emulated, "listen: bad heap handle");
    CCMPI_REQUIRE(rank >= 0 && rank < CCMPI_MAX_RANKS, "listen: rank %d out of range", rank);
    // the local segment moves to its rank slot
    if (rank != 0) {
        h->ptr[rank] = h->ptr[0];
        h->ptr[0] = 0;
    }
    h->rank = rank;
    int s = socket(AF_UNIX, SOCK_STREAM, 0);
    if (s < 0) {
        set_error("socket(): %s", strerror(errno));
        return CCMPI_ERR_SYS;
    }
    sockaddr_un addr;
    socklen_t len = make_addr(addr, job, rank);
    if (bind(s, (sockaddr*)&addr, len) != 0 || listen(s, 64) != 0) {
        set_error("bind/listen on ccmpi-%s-%d: %s", job, rank, strerror(errno));
        close(s);
        return CCMPI_ERR_SYS;
    }
    h->listen_fd = s;
    h->server = std::thread(serve, h);
    return CCMPI_OK;
}

extern "C" int ccmpi_heap_connect_all(int64_t heap, const char* job, int rank, int nranks) {
    Heap* h = get(heap);
    CCMPI_REQUIRE(h && !h->emulated, "connect: bad heap handle");
    CCMPI_REQUIRE(nranks >= 1 && nranks <= CCMPI_MAX_RANKS && rank == h->rank, "connect: rank/nranks mismatch");
    Drv& d = D();
    h->nranks = nranks;
    CCMPI_CUDA_CHECK(cudaSetDevice(h->device));
    for (int i = 1; i < nranks; ++i) {
        const int peer = (rank + i) % nranks;
        int fd = -1;
        Req rq{0, 0, rank, 0};
        int rc = request_fd(job, peer, rq, &fd, 60000);
        if (rc) return rc;
        CUresult r = d.MemImport(&h->peer_handle[peer], (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
        close(fd);
        if (r != CUDA_SUCCESS) {
            set_error("cuMemImportFromShareableHandle(peer %d) failed: CUresult %d", peer, (int)r);
            return CCMPI_ERR_CUDA;
        }
        rc = map_handle(h->peer_handle[peer], h->bytes, 2u << 20, h->device, &h->ptr[peer]);
        if (rc) {
            std::string prev = ccmpi_last_error();
            set_error("mapping peer %d segment failed (is P2P available between the GPUs?): %s", peer, prev.c_str());
            return rc;
        }
    }
    return CCMPI_OK;
}

extern "C" void* ccmpi_heap_peer_ptr(int64_t heap, int peer) {
    Heap* h = get(heap);
    if (!h || peer < 0 || peer >= CCMPI_MAX_RANKS) return nullptr;
    return (void*)h->ptr[peer];
}

extern "C" int ccmpi_mc_create(int64_t heap, const char* job, int rank, const int* members, int nmembers, int group_tag) {
    Heap* h = get(heap);
    CCMPI_REQUIRE(h && !h->emulated, "multicast: bad heap handle");
    CCMPI_REQUIRE(members && nmembers >= 2 && nmembers <= CCMPI_MAX_RANKS, "multicast: need 2..%d members", CCMPI_MAX_RANKS);
    Drv& d = D();
    if (!d.mc_ok) {
        set_error("multicast entry points are unavailable in this driver");
        return CCMPI_ERR_UNSUPPORTED;
    }
    int sup = 0;
    if (d.DeviceGetAttribute(&sup, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, h->device) != CUDA_SUCCESS || !sup) {
        set_error("device %d does not support multicast (no NVSwitch / NVLS)", h->device);
        return CCMPI_ERR_UNSUPPORTED;
    }
    CCMPI_CUDA_CHECK(cudaSetDevice(h->device));
    McGroup g;
    g.tag = group_tag;
    g.nmembers = nmembers;
    g.leader = (members[0] == rank);
    g.size = h->bytes;
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = nmembers;
    mp.size = h->bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    if (g.leader) {
        CUresult r = d.McCreate(&g.handle, &mp);
        if (r != CUDA_SUCCESS) {
            set_error("cuMulticastCreate(%d devices, %zu bytes) failed: CUresult %d", nmembers, h->bytes, (int)r);
            return r == CUDA_ERROR_NOT_SUPPORTED ? CCMPI_ERR_UNSUPPORTED : CCMPI_ERR_CUDA;
        }
        CCMPI_DRV_CHECK(d.MemExport(&g.export_fd, g.handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
        // publish before anyone asks for it
        std::lock_guard<std::mutex> lk(h->mu);
        h->mc.push_back(g);
    } else {
        int fd = -1;
        Req rq{1, group_tag, rank, 0};
        int rc = request_fd(job, members[0], rq, &fd, 60000);
        if (rc) return rc;
        CUresult r = d.MemImport(&g.handle, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
        close(fd);
        if (r != CUDA_SUCCESS) {
            set_error("importing the multicast handle failed: CUresult %d", (int)r);
            return CCMPI_ERR_CUDA;
        }
        std::lock_guard<std::mutex> lk(h->mu);
        h->mc.push_back(g);
    }
    int idx;
    CUmemGenericAllocationHandle handle;
    {
        std::lock_guard<std::mutex> lk(h->mu);
        idx = (int)h->mc.size() - 1;
        handle = h->mc[idx].handle;
    }
    CCMPI_DRV_CHECK(d.McAddDevice(handle, h->device));
    return idx;
}

extern "C" int ccmpi_mc_bind(int64_t heap, int group) {
    Heap* h = get(heap);
    CCMPI_REQUIRE(h && !h->emulated, "multicast bind: bad heap handle");
    Drv& d = D();
    CUmemGenericAllocationHandle handle;
    size_t size;
    {
        std::lock_guard<std::mutex> lk(h->mu);
        CCMPI_REQUIRE(group >= 0 && group < (int)h->mc.size(), "multicast bind: bad group %d", group);
        handle = h->mc[group].handle;
        size = h->mc[group].size;
    }
    CCMPI_CUDA_CHECK(cudaSetDevice(h->device));
    CCMPI_DRV_CHECK(d.McBindMem(handle, 0, h->local_handle, 0, size, 0));
    CUdeviceptr va = 0;
    int rc = map_handle(handle, size, 2u << 20, h->device, &va);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(h->mu);
    h->mc[group].va = va;
    h->mc[group].bound = true;
    return CCMPI_OK;
}

extern "C" void* ccmpi_mc_ptr(int64_t heap, int group) {
    Heap* h = get(heap);
    if (!h) return nullptr;
    std::lock_guard<std::mutex> lk(h->mu);
    if (group < 0 || group >= (int)h->mc.size() || !h->mc[group].bound) return nullptr;
    return (void*)h->mc[group].va;
}

extern "C" int ccmpi_heap_destroy(int64_t heap) {
    Heap* h;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        auto it = g_heaps.find(heap);
        if (it == g_heaps.end()) return CCMPI_OK;
        h = it->second;
        g_heaps.erase(it);
    }
    h->stop.store(true);
    if (h->server.joinable()) h->server.join();
    if (h->listen_fd >= 0) close(h->listen_fd);
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    if (h->emulated) {
        for (int i = 0; i < h->nranks; ++i)
            if (h->ptr[i]) cudaFree((void*)h->ptr[i]);
        delete h;
        return CCMPI_OK;
    }
    Drv& d = D();
    for (auto& g : h->mc) {
        if (g.va) {
            d.MemUnmap(g.va, g.size);
            d.MemAddressFree(g.va, g.size);
        }
        if (g.bound && d.McUnbind) d.McUnbind(g.handle, h->device, 0, g.size);
        if (g.export_fd >= 0) close(g.export_fd);
        if (g.handle) d.MemRelease(g.handle);
    }
    for (int i = 0; i < CCMPI_MAX_RANKS; ++i) {
        if (h->ptr[i]) {
            d.MemUnmap(h->ptr[i], h->bytes);
            d.MemAddressFree(h->ptr[i], h->bytes);
        }
        if (h->peer_handle[i]) d.MemRelease(h->peer_handle[i]);
    }
    if (h->local_fd >= 0) close(h->local_fd);
    if (h->local_handle) d.MemRelease(h->local_handle);
    delete h;
    return CCMPI_OK;
}
