// Host plumbing for NVSwitch multicast (NVLS) regions: CUDA virtual-memory-management
// allocations shared between the per-GPU processes as POSIX file descriptors, multicast
// objects bound over them, and the tiny unix-socket exchange that carries the descriptors.
//
// Only set-up code lives here; the kernels that use the mappings (multimem.ld_reduce /
// multimem.st) are in allreduce.cu.  The CUDA driver API is reached through
// cudaGetDriverEntryPointByVersion, so the library keeps no link-time dependency on libcuda
// (it must still load, and export its symbols, on a box without a driver).
#pragma once
#include <cuda.h>

#include "common.cuh"

struct PrVmmBlock {
    CUmemGenericAllocationHandle handle = 0;
    CUdeviceptr va = 0;
    size_t size = 0;      // mapped (granularity-rounded) size
    bool mapped = false;
};

struct PrFabricCaps {
    bool driver = false;      // driver entry points resolved
    bool posix_fd = false;    // cuMemCreate allocations can be exported as file descriptors
    bool multicast = false;   // CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED
    size_t alloc_gran = 0, mc_gran_min = 0, mc_gran_rec = 0;
};

// Capabilities of `device` for a team of `world` devices (never fails: all-false on any error).
PrFabricCaps pr_fabric_caps(int device, int world);

// size actually mapped for a request of nbytes (multiple of every granularity in play)
size_t pr_fabric_round(const PrFabricCaps &caps, size_t nbytes);

int pr_vmm_alloc(int device, size_t rounded_bytes, PrVmmBlock *out);          // zero-filled by the caller
int pr_vmm_export_fd(const PrVmmBlock &b, int *fd_out);
int pr_vmm_import_fd(int device, int fd, size_t rounded_bytes, PrVmmBlock *out);
void pr_vmm_free(PrVmmBlock *b);

int pr_mc_create(int world, size_t rounded_bytes, CUmemGenericAllocationHandle *mc_out, int *fd_out);
int pr_mc_import(int fd, CUmemGenericAllocationHandle *mc_out);
// add `device` to the team, bind `mem` (blocks until every rank has added its device) and map
// the multicast object into this process
int pr_mc_attach(CUmemGenericAllocationHandle mc, int device, const PrVmmBlock &mem, CUdeviceptr *mc_va_out);
void pr_mc_detach(CUmemGenericAllocationHandle mc, int device, CUdeviceptr mc_va, size_t size, bool bound);

// ---- descriptor exchange: one abstract-namespace unix datagram socket per rank ---------------
struct PrSock {
    int fd = -1;
    char name[48] = {0};   // abstract socket name (without the leading NUL)
};
int pr_sock_open(PrSock *s, int rank);
void pr_sock_close(PrSock *s);
int pr_sock_send(const PrSock &s, const char *peer_name, const void *payload, size_t nbytes, const int *fds, int nfds);
// one message; fails with PR_ERR_STATE after timeout_ms
int pr_sock_recv(const PrSock &s, void *payload, size_t nbytes, int *fds, int max_fds, int *nfds, int timeout_ms);
