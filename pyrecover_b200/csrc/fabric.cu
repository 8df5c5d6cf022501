// Set-up code for NVSwitch multicast regions (see fabric.cuh).  Host only; no kernels here.
#include "fabric.cuh"

#include <errno.h>
#include <poll.h>
#include <stddef.h>
#include <stdlib.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <mutex>

namespace {

// ---- driver entry points (resolved once, through the runtime: no libcuda link dependency) ----
struct Drv {
    bool ok = false;
    CUresult (*GetErrorString)(CUresult, const char **) = nullptr;
    CUresult (*DeviceGet)(CUdevice *, int) = nullptr;
    CUresult (*DeviceGetAttribute)(int *, CUdevice_attribute, CUdevice) = nullptr;
    CUresult (*MemGetAllocationGranularity)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags) = nullptr;
    CUresult (*MemCreate)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long) = nullptr;
    CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*MemAddressReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t) = nullptr;
    CUresult (*MemExportToShareableHandle)(void *, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
    CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle *, void *, CUmemAllocationHandleType) = nullptr;
    CUresult (*MulticastCreate)(CUmemGenericAllocationHandle *, const CUmulticastObjectProp *) = nullptr;
    CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
    CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long) = nullptr;
    CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
    CUresult (*MulticastGetGranularity)(size_t *, const CUmulticastObjectProp *, CUmulticastGranularity_flags) = nullptr;
};

Drv g_drv;
std::once_flag g_drv_once;

template <typename F>
bool resolve(const char *name, F *&fn, unsigned version) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
    if (cudaGetDriverEntryPointByVersion(name, &p, version, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess || p == nullptr) {
        cudaGetLastError();
        return false;
    }
    fn = reinterpret_cast<F *>(p);
    return true;
}

const Drv *drv() {
    std::call_once(g_drv_once, [] {
        Drv d;
        bool ok = cudaFree(nullptr) == cudaSuccess;     // make sure a context exists
        ok = ok && resolve("cuGetErrorString", d.GetErrorString, 6000);
        ok = ok && resolve("cuDeviceGet", d.DeviceGet, 2000);
        ok = ok && resolve("cuDeviceGetAttribute", d.DeviceGetAttribute, 2000);
        ok = ok && resolve("cuMemGetAllocationGranularity", d.MemGetAllocationGranularity, 10020);
        ok = ok && resolve("cuMemCreate", d.MemCreate, 10020);
        ok = ok && resolve("cuMemRelease", d.MemRelease, 10020);
        ok = ok && resolve("cuMemAddressReserve", d.MemAddressReserve, 10020);
        ok = ok && resolve("cuMemAddressFree", d.MemAddressFree, 10020);
        ok = ok && resolve("cuMemMap", d.MemMap, 10020);
        ok = ok && resolve("cuMemUnmap", d.MemUnmap, 10020);
        ok = ok && resolve("cuMemSetAccess", d.MemSetAccess, 10020);
        ok = ok && resolve("cuMemExportToShareableHandle", d.MemExportToShareableHandle, 10020);
        ok = ok && resolve("cuMemImportFromShareableHandle", d.MemImportFromShareableHandle, 10020);
        ok = ok && resolve("cuMulticastCreate", d.MulticastCreate, 12010);
        ok = ok && resolve("cuMulticastAddDevice", d.MulticastAddDevice, 12010);
        ok = ok && resolve("cuMulticastBindMem", d.MulticastBindMem, 12010);
        ok = ok && resolve("cuMulticastUnbind", d.MulticastUnbind, 12010);
        ok = ok && resolve("cuMulticastGetGranularity", d.MulticastGetGranularity, 12010);
        d.ok = ok;
        g_drv = d;
    });
    return g_drv.ok ? &g_drv : nullptr;
}

const char *cu_text(CUresult r) {
    const char *s = nullptr;
    if (g_drv.GetErrorString && g_drv.GetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
    return "unknown CUDA driver error";
}

#define PR_CU(call)                                                                          \
    do {                                                                                     \
        CUresult _r = (call);                                                                \
        if (_r != CUDA_SUCCESS) {                                                            \
            pr_set_error("%s failed at %s:%d: %s (%d)", #call, __FILE__, __LINE__, cu_text(_r), (int)_r); \
            return PR_ERR_CUDA;                                                              \
        }                                                                                    \
    } while (0)

CUmemAllocationProp alloc_prop(int device) {
    CUmemAllocationProp p;
    memset(&p, 0, sizeof(p));
    p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    p.location.id = device;
    return p;
}

int map_rw(const Drv *d, int device, CUmemGenericAllocationHandle h, size_t size, size_t align, CUdeviceptr *va_out) {
    CUdeviceptr va = 0;
    PR_CU(d->MemAddressReserve(&va, size, align, 0, 0));
    CUresult r = d->MemMap(va, size, 0, h, 0);
    if (r != CUDA_SUCCESS) {
        d->MemAddressFree(va, size);
        pr_set_error("cuMemMap failed: %s (%d)", cu_text(r), (int)r);
        return PR_ERR_CUDA;
    }
    CUmemAccessDesc acc;
    memset(&acc, 0, sizeof(acc));
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = device;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r = d->MemSetAccess(va, size, &acc, 1);
    if (r != CUDA_SUCCESS) {
        d->MemUnmap(va, size);
        d->MemAddressFree(va, size);
        pr_set_error("cuMemSetAccess failed: %s (%d)", cu_text(r), (int)r);
        return PR_ERR_CUDA;
    }
    *va_out = va;
    return PR_OK;
}

size_t lcm_pow2ish(size_t a, size_t b) {   // granularities are powers of two in practice; stay safe otherwise
    if (a == 0) return b;
    if (b == 0) return a;
    size_t x = a, y = b;
    while (y) { size_t t = x % y; x = y; y = t; }
    return a / x * b;
}

}  // namespace

PrFabricCaps pr_fabric_caps(int device, int world) {
    PrFabricCaps c;
    const Drv *d = drv();
    if (!d) return c;
    c.driver = true;
    CUdevice dev;
    if (d->DeviceGet(&dev, device) != CUDA_SUCCESS) return c;
    int v = 0;
    if (d->DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev) == CUDA_SUCCESS)
        c.posix_fd = v != 0;
    v = 0;
    if (d->DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) == CUDA_SUCCESS)
        c.multicast = v != 0;
    CUmemAllocationProp p = alloc_prop(device);
    size_t g = 0;
    if (d->MemGetAllocationGranularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS) c.alloc_gran = g;
    if (c.alloc_gran == 0) c.posix_fd = false;
    if (c.multicast) {
        CUmulticastObjectProp mp;
        memset(&mp, 0, sizeof(mp));
        mp.numDevices = (unsigned)world;
        mp.size = c.alloc_gran ? c.alloc_gran : (2u << 20);
        mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        size_t gmin = 0, grec = 0;
        if (d->MulticastGetGranularity(&gmin, &mp, CU_MULTICAST_GRANULARITY_MINIMUM) != CUDA_SUCCESS ||
            d->MulticastGetGranularity(&grec, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || gmin == 0) {
            c.multicast = false;
        } else {
            c.mc_gran_min = gmin;
            c.mc_gran_rec = grec ? grec : gmin;
        }
    }
    return c;
}

size_t pr_fabric_round(const PrFabricCaps &c, size_t nbytes) {
    size_t g = c.alloc_gran ? c.alloc_gran : (2u << 20);
    if (c.multicast) {
        g = lcm_pow2ish(g, c.mc_gran_min);
        // the recommended multicast granularity is honoured while it costs < 1/8 of the region
        if (c.mc_gran_rec > g && c.mc_gran_rec <= nbytes / 8) g = lcm_pow2ish(g, c.mc_gran_rec);
    }
    return (nbytes + g - 1) / g * g;
}

int pr_vmm_alloc(int device, size_t size, PrVmmBlock *out) {
    const Drv *d = drv();
    PR_REQUIRE(d != nullptr, "pr_vmm_alloc: CUDA driver entry points unavailable");
    CUmemAllocationProp p = alloc_prop(device);
    PrVmmBlock b;
    PR_CU(d->MemCreate(&b.handle, size, &p, 0));
    int rc = map_rw(d, device, b.handle, size, 0, &b.va);
    if (rc) {
        d->MemRelease(b.handle);
        return rc;
    }
    b.size = size;
    b.mapped = true;
    *out = b;
    return PR_OK;
}

int pr_vmm_export_fd(const PrVmmBlock &b, int *fd_out) {
    const Drv *d = drv();
    PR_REQUIRE(d != nullptr, "pr_vmm_export_fd: CUDA driver entry points unavailable");
    int fd = -1;
    PR_CU(d->MemExportToShareableHandle(&fd, b.handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    *fd_out = fd;
    return PR_OK;
}

int pr_vmm_import_fd(int device, int fd, size_t size, PrVmmBlock *out) {
    const Drv *d = drv();
    PR_REQUIRE(d != nullptr, "pr_vmm_import_fd: CUDA driver entry points unavailable");
    PrVmmBlock b;
    PR_CU(d->MemImportFromShareableHandle(&b.handle, (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    int rc = map_rw(d, device, b.handle, size, 0, &b.va);
    if (rc) {
        d->MemRelease(b.handle);
        return rc;
    }
    b.size = size;
    b.mapped = true;
    *out = b;
    return PR_OK;
}

void pr_vmm_free(PrVmmBlock *b) {
    const Drv *d = drv();
    if (!d || !b) return;
    if (b->mapped) {
        d->MemUnmap(b->va, b->size);
        d->MemAddressFree(b->va, b->size);
    }
    if (b->handle) d->MemRelease(b->handle);
    *b = PrVmmBlock();
}

int pr_mc_create(int world, size_t size, CUmemGenericAllocationHandle *mc_out, int *fd_out) {
    const Drv *d = drv();
    PR_REQUIRE(d != nullptr, "pr_mc_create: CUDA driver entry points unavailable");
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = (unsigned)world;
    mp.size = size;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CUmemGenericAllocationHandle mc = 0;
    PR_CU(d->MulticastCreate(&mc, &mp));
    int fd = -1;
    CUresult r = d->MemExportToShareableHandle(&fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
        d->MemRelease(mc);
        pr_set_error("exporting the multicast object failed: %s (%d)", cu_text(r), (int)r);
        return PR_ERR_CUDA;
    }
    *mc_out = mc;
    *fd_out = fd;
    return PR_OK;
}

int pr_mc_import(int fd, CUmemGenericAllocationHandle *mc_out) {
    const Drv *d = drv();
    PR_REQUIRE(d != nullptr, "pr_mc_import: CUDA driver entry points unavailable");
    PR_CU(d->MemImportFromShareableHandle(mc_out, (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    return PR_OK;
}

int pr_mc_attach(CUmemGenericAllocationHandle mc, int device, const PrVmmBlock &mem, CUdeviceptr *mc_va_out) {
    const Drv *d = drv();
    PR_REQUIRE(d != nullptr, "pr_mc_attach: CUDA driver entry points unavailable");
    CUdevice dev;
    PR_CU(d->DeviceGet(&dev, device));
    PR_CU(d->MulticastAddDevice(mc, dev));
    PR_CU(d->MulticastBindMem(mc, 0, mem.handle, 0, mem.size, 0));   // blocks until every device is added
    CUdeviceptr va = 0;
    int rc = map_rw(d, device, mc, mem.size, 0, &va);
    if (rc) {
        d->MulticastUnbind(mc, dev, 0, mem.size);
        return rc;
    }
    *mc_va_out = va;
    return PR_OK;
}

void pr_mc_detach(CUmemGenericAllocationHandle mc, int device, CUdeviceptr mc_va, size_t size, bool bound) {
    const Drv *d = drv();
    if (!d) return;
    if (mc_va) {
        d->MemUnmap(mc_va, size);
        d->MemAddressFree(mc_va, size);
    }
    CUdevice dev;
    if (bound && d->DeviceGet(&dev, device) == CUDA_SUCCESS) d->MulticastUnbind(mc, dev, 0, size);
    if (mc) d->MemRelease(mc);
}

// ---- descriptor exchange ---------------------------------------------------------------------
static socklen_t abstract_addr(const char *name, sockaddr_un *sa) {
    memset(sa, 0, sizeof(*sa));
    sa->sun_family = AF_UNIX;
    size_t n = strlen(name);
    if (n > sizeof(sa->sun_path) - 2) n = sizeof(sa->sun_path) - 2;
    memcpy(sa->sun_path + 1, name, n);            // sun_path[0] = 0: abstract namespace, no file to clean up
    return (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
}

int pr_sock_open(PrSock *s, int rank) {
    static int counter = 0;
    int fd = socket(AF_UNIX, SOCK_DGRAM | SOCK_CLOEXEC, 0);
    if (fd < 0) {
        pr_set_error("socket(AF_UNIX) failed: %s", strerror(errno));
        return PR_ERR_IO;
    }
    snprintf(s->name, sizeof(s->name), "pr-b200-%d-%d-%d", (int)getpid(), rank, counter++);
    sockaddr_un sa;
    socklen_t len = abstract_addr(s->name, &sa);
    if (bind(fd, (sockaddr *)&sa, len) != 0) {
        pr_set_error("bind(@%s) failed: %s", s->name, strerror(errno));
        close(fd);
        return PR_ERR_IO;
    }
    s->fd = fd;
    return PR_OK;
}

void pr_sock_close(PrSock *s) {
    if (s && s->fd >= 0) close(s->fd);
    if (s) s->fd = -1;
}

int pr_sock_send(const PrSock &s, const char *peer, const void *payload, size_t nbytes, const int *fds, int nfds) {
    sockaddr_un sa;
    socklen_t len = abstract_addr(peer, &sa);
    iovec iov;
    iov.iov_base = const_cast<void *>(payload);
    iov.iov_len = nbytes;
    alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int) * 16)];
    msghdr msg;
    memset(&msg, 0, sizeof(msg));
    msg.msg_name = &sa;
    msg.msg_namelen = len;
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    if (nfds > 0) {
        PR_REQUIRE(nfds <= 16, "pr_sock_send: too many descriptors (%d)", nfds);
        memset(ctrl, 0, sizeof(ctrl));
        msg.msg_control = ctrl;
        msg.msg_controllen = CMSG_SPACE(sizeof(int) * nfds);
        cmsghdr *cm = CMSG_FIRSTHDR(&msg);
        cm->cmsg_level = SOL_SOCKET;
        cm->cmsg_type = SCM_RIGHTS;
        cm->cmsg_len = CMSG_LEN(sizeof(int) * nfds);
        memcpy(CMSG_DATA(cm), fds, sizeof(int) * nfds);
    }
    for (int attempt = 0; attempt < 200; ++attempt) {
        if (sendmsg(s.fd, &msg, 0) >= 0) return PR_OK;
        if (errno == EINTR) continue;
        if (errno == ECONNREFUSED || errno == ENOENT || errno == EAGAIN || errno == ENOBUFS) {
            usleep(50 * 1000);    // the peer has not bound its socket yet / its queue is full
            continue;
        }
        break;
    }
    pr_set_error("sendmsg(@%s) failed: %s", peer, strerror(errno));
    return PR_ERR_IO;
}

int pr_sock_recv(const PrSock &s, void *payload, size_t nbytes, int *fds, int max_fds, int *nfds, int timeout_ms) {
    pollfd pf;
    pf.fd = s.fd;
    pf.events = POLLIN;
    pf.revents = 0;
    int pr = 0;
    do {
        pr = poll(&pf, 1, timeout_ms);
    } while (pr < 0 && errno == EINTR);
    if (pr <= 0) {
        pr_set_error("timed out after %d ms waiting for a peer's descriptor message", timeout_ms);
        return PR_ERR_STATE;
    }
    iovec iov;
    iov.iov_base = payload;
    iov.iov_len = nbytes;
    alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int) * 16)];
    memset(ctrl, 0, sizeof(ctrl));
    msghdr msg;
    memset(&msg, 0, sizeof(msg));
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof(ctrl);
    ssize_t got;
    do {
        got = recvmsg(s.fd, &msg, MSG_CMSG_CLOEXEC);
    } while (got < 0 && errno == EINTR);
    if (got < 0 || (size_t)got != nbytes) {
        pr_set_error("recvmsg failed or short message (%zd of %zu): %s", got, nbytes, got < 0 ? strerror(errno) : "short");
        return PR_ERR_IO;
    }
    int n = 0;
    for (cmsghdr *cm = CMSG_FIRSTHDR(&msg); cm; cm = CMSG_NXTHDR(&msg, cm)) {
        if (cm->cmsg_level == SOL_SOCKET && cm->cmsg_type == SCM_RIGHTS) {
            int cnt = (int)((cm->cmsg_len - CMSG_LEN(0)) / sizeof(int));
            const int *src = reinterpret_cast<const int *>(CMSG_DATA(cm));
            for (int i = 0; i < cnt; ++i) {
                if (n < max_fds) fds[n++] = src[i];
                else close(src[i]);
            }
        }
    }
    *nfds = n;
    return PR_OK;
}
