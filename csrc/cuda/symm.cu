// Symmetric memory segments: the same-sized allocation on every rank of a group, mapped into
// every peer's address space (and, where the box supports it, bound to an NVSwitch multicast
// object), so kernels can load/store remote gradient buckets directly over NVLink.
//
// This replaces what the reference gets from NCCL's internal buffers (its only data plane,
// backend='nccl' at launch_dist.py:49): physical memory comes from the CUDA VMM API
// (cuMemCreate), is exported as a POSIX file descriptor, and the descriptors travel between
// the local ranks over abstract Unix-domain sockets (SCM_RIGHTS).  The Python side provides
// the barriers between the phases (begin -> exchange -> bind_multicast) through the store.
#include "common.cuh"

#include <cuda.h>
#include <cerrno>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <string>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>
#include <vector>

static thread_local char g_err[1024] = "";

void pdl_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

PDL_API const char* pdl_last_error() { return g_err; }

long long g_pdl_launch_calls = 0;
// Number of launcher calls (each enqueues at least one kernel of this library) made so far in this process.
PDL_API long long pdl_launch_count() { return g_pdl_launch_calls; }

// Programmatic dependent launch (common.cuh) is on unless PDL_NO_PDL=1.
bool pdl_dependent_launch_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* s = getenv("PDL_NO_PDL");
    v = (s && s[0] == '1') ? 0 : 1;
  }
  return v != 0;
}

// ---- driver entry points fetched at run time (the .so must load on a box without libcuda) ----
namespace {

struct Driver {
  bool ok = false;
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGet)(CUdevice*, int);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*GetErrorString)(CUresult, const char**);
} drv;

template <typename F>
bool load_sym(F& fn, const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
    pdl_set_error("driver entry point %s is not available", name);
    return false;
  }
  fn = reinterpret_cast<F>(p);
  return true;
}

bool load_driver() {
  if (drv.ok) return true;
  bool ok = load_sym(drv.MemCreate, "cuMemCreate") && load_sym(drv.MemRelease, "cuMemRelease") &&
            load_sym(drv.MemGetAllocationGranularity, "cuMemGetAllocationGranularity") &&
            load_sym(drv.MemExportToShareableHandle, "cuMemExportToShareableHandle") &&
            load_sym(drv.MemImportFromShareableHandle, "cuMemImportFromShareableHandle") &&
            load_sym(drv.MemAddressReserve, "cuMemAddressReserve") && load_sym(drv.MemAddressFree, "cuMemAddressFree") &&
            load_sym(drv.MemMap, "cuMemMap") && load_sym(drv.MemUnmap, "cuMemUnmap") && load_sym(drv.MemSetAccess, "cuMemSetAccess") &&
            load_sym(drv.MulticastCreate, "cuMulticastCreate") && load_sym(drv.MulticastAddDevice, "cuMulticastAddDevice") &&
            load_sym(drv.MulticastBindMem, "cuMulticastBindMem") && load_sym(drv.MulticastGetGranularity, "cuMulticastGetGranularity") &&
            load_sym(drv.DeviceGet, "cuDeviceGet") && load_sym(drv.DeviceGetAttribute, "cuDeviceGetAttribute") &&
            load_sym(drv.GetErrorString, "cuGetErrorString");
  drv.ok = ok;
  return ok;
}

#define PDL_CU_CHECK(expr)                                                                      \
  do {                                                                                          \
    CUresult _r = (expr);                                                                       \
    if (_r != CUDA_SUCCESS) {                                                                   \
      const char* _s = "?";                                                                     \
      drv.GetErrorString(_r, &_s);                                                              \
      pdl_set_error("%s failed at %s:%d: %s (%d)", #expr, __FILE__, __LINE__, _s, (int)_r);     \
      return -static_cast<int>(_r) - 1000;                                                      \
    }                                                                                           \
  } while (0)

constexpr int kMaxWorld = 16;

struct SymmSeg {
  int rank = 0, world = 1, device = 0;
  size_t bytes = 0;
  bool want_mc = false, has_mc = false;
  CUmemGenericAllocationHandle handles[kMaxWorld] = {};
  CUdeviceptr ptrs[kMaxWorld] = {};
  CUmemGenericAllocationHandle mc_handle = 0;
  CUdeviceptr mc_ptr = 0;
  int local_fd = -1, mc_fd = -1;
  int listen_fd = -1;
  std::string prefix;
  void** d_ptrs = nullptr;
};

void uds_addr(const std::string& prefix, int rank, sockaddr_un* addr, socklen_t* len) {
  memset(addr, 0, sizeof(*addr));
  addr->sun_family = AF_UNIX;
  std::string name = prefix + "-" + std::to_string(rank);
  // abstract namespace: leading NUL, no filesystem entry to clean up
  size_t n = name.size() < sizeof(addr->sun_path) - 2 ? name.size() : sizeof(addr->sun_path) - 2;
  memcpy(addr->sun_path + 1, name.data(), n);
  *len = static_cast<socklen_t>(offsetof(sockaddr_un, sun_path) + 1 + n);
}

bool send_fds(int sock, int32_t from_rank, const int* fds, int nfds) {
  msghdr msg{};
  iovec iov{&from_rank, sizeof(from_rank)};
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  char ctrl[CMSG_SPACE(sizeof(int) * 2)] = {};
  msg.msg_control = ctrl;
  msg.msg_controllen = CMSG_SPACE(sizeof(int) * nfds);
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int) * nfds);
  memcpy(CMSG_DATA(c), fds, sizeof(int) * nfds);
  return sendmsg(sock, &msg, 0) == static_cast<ssize_t>(sizeof(from_rank));
}

bool recv_fds(int sock, int32_t* from_rank, int* fds, int* nfds) {
  msghdr msg{};
  iovec iov{from_rank, sizeof(*from_rank)};
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  char ctrl[CMSG_SPACE(sizeof(int) * 2)] = {};
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  ssize_t k;
  do { k = recvmsg(sock, &msg, 0); } while (k < 0 && errno == EINTR);
  if (k != static_cast<ssize_t>(sizeof(*from_rank))) return false;
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  if (!c || c->cmsg_level != SOL_SOCKET || c->cmsg_type != SCM_RIGHTS) return false;
  int n = static_cast<int>((c->cmsg_len - CMSG_LEN(0)) / sizeof(int));
  memcpy(fds, CMSG_DATA(c), sizeof(int) * n);
  *nfds = n;
  return true;
}

int map_handle(SymmSeg* s, CUmemGenericAllocationHandle h, CUdeviceptr* out) {
  CUdeviceptr va = 0;
  PDL_CU_CHECK(drv.MemAddressReserve(&va, s->bytes, 0, 0, 0));
  PDL_CU_CHECK(drv.MemMap(va, s->bytes, 0, h, 0));
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = s->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  PDL_CU_CHECK(drv.MemSetAccess(va, s->bytes, &acc, 1));
  *out = va;
  return 0;
}

}  // namespace

PDL_API int pdl_cuda_device_count() {
  int n = 0;
  return cudaGetDeviceCount(&n) == cudaSuccess ? n : 0;
}

// 1 if the device can join an NVSwitch multicast group (NVLS), else 0.
PDL_API int pdl_symm_multicast_supported(int device) {
  if (!load_driver()) return 0;
  if (cudaSetDevice(device) != cudaSuccess || cudaFree(0) != cudaSuccess) return 0;
  CUdevice dev;
  int v = 0;
  if (drv.DeviceGet(&dev, device) != CUDA_SUCCESS) return 0;
  if (drv.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) return 0;
  return v;
}

// Phase 1: allocate local physical memory, export it, start listening for the peers.
PDL_API int pdl_symm_begin(int rank, int world, int device, uint64_t bytes, const char* uds_prefix, int want_multicast, void** out) {
  if (world > kMaxWorld) { pdl_set_error("world %d exceeds the supported maximum %d", world, kMaxWorld); return -1; }
  if (!load_driver()) return -2;
  PDL_CUDA_CHECK(cudaSetDevice(device));
  PDL_CUDA_CHECK(cudaFree(0));
  auto* s = new SymmSeg();
  s->rank = rank; s->world = world; s->device = device; s->prefix = uds_prefix; s->want_mc = want_multicast && world > 1;

  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  PDL_CU_CHECK(drv.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (s->want_mc) {
    CUmulticastObjectProp mp{};
    mp.numDevices = world;
    mp.size = bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mgran = 0;
    if (drv.MulticastGetGranularity(&mgran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mgran > gran) gran = mgran;
  }
  s->bytes = (bytes + gran - 1) / gran * gran;
  PDL_CU_CHECK(drv.MemCreate(&s->handles[rank], s->bytes, &prop, 0));
  PDL_CU_CHECK(drv.MemExportToShareableHandle(&s->local_fd, s->handles[rank], CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  if (int rc = map_handle(s, s->handles[rank], &s->ptrs[rank])) return rc;
  PDL_CUDA_CHECK(cudaMemset(reinterpret_cast<void*>(s->ptrs[rank]), 0, s->bytes));
  PDL_CUDA_CHECK(cudaDeviceSynchronize());

  if (s->want_mc && rank == 0) {
    CUmulticastObjectProp mp{};
    mp.numDevices = world;
    mp.size = s->bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CUresult r = drv.MulticastCreate(&s->mc_handle, &mp);
    if (r == CUDA_SUCCESS) r = drv.MemExportToShareableHandle(&s->mc_fd, s->mc_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) { s->mc_handle = 0; s->mc_fd = -1; }  // peers learn this from a 1-fd message
  }

  if (world > 1) {
    s->listen_fd = socket(AF_UNIX, SOCK_STREAM, 0);
    sockaddr_un addr; socklen_t len;
    uds_addr(s->prefix, rank, &addr, &len);
    if (s->listen_fd < 0 || bind(s->listen_fd, reinterpret_cast<sockaddr*>(&addr), len) != 0 || listen(s->listen_fd, 64) != 0) {
      pdl_set_error("symm: cannot listen on the handle-exchange socket: %s", strerror(errno));
      return -3;
    }
  }
  *out = s;
  return 0;
}

// Phase 2 (after every rank finished phase 1): swap descriptors, import and map the peers.
PDL_API int pdl_symm_exchange(void* h) {
  auto* s = static_cast<SymmSeg*>(h);
  if (s->world == 1) {
    PDL_CUDA_CHECK(cudaMalloc(&s->d_ptrs, sizeof(void*)));
    void* p = reinterpret_cast<void*>(s->ptrs[0]);
    PDL_CUDA_CHECK(cudaMemcpy(s->d_ptrs, &p, sizeof(void*), cudaMemcpyHostToDevice));
    return 0;
  }
  for (int p = 0; p < s->world; ++p) {
    if (p == s->rank) continue;
    int sock = socket(AF_UNIX, SOCK_STREAM, 0);
    sockaddr_un addr; socklen_t len;
    uds_addr(s->prefix, p, &addr, &len);
    if (sock < 0 || connect(sock, reinterpret_cast<sockaddr*>(&addr), len) != 0) {
      pdl_set_error("symm: cannot reach rank %d's handle-exchange socket: %s", p, strerror(errno));
      return -4;
    }
    int fds[2] = {s->local_fd, s->mc_fd};
    int nf = (s->rank == 0 && s->mc_fd >= 0) ? 2 : 1;
    bool ok = send_fds(sock, s->rank, fds, nf);
    close(sock);
    if (!ok) { pdl_set_error("symm: sending the memory handle to rank %d failed: %s", p, strerror(errno)); return -5; }
  }
  bool mc_ok = s->want_mc && (s->rank != 0 || s->mc_handle != 0);
  for (int k = 0; k < s->world - 1; ++k) {
    int c = accept(s->listen_fd, nullptr, nullptr);
    int32_t from = -1; int fds[2] = {-1, -1}, nf = 0;
    if (c < 0 || !recv_fds(c, &from, fds, &nf) || from < 0 || from >= s->world) {
      pdl_set_error("symm: receiving a peer memory handle failed: %s", strerror(errno));
      return -6;
    }
    close(c);
    PDL_CU_CHECK(drv.MemImportFromShareableHandle(&s->handles[from], reinterpret_cast<void*>(static_cast<uintptr_t>(fds[0])),
                                                  CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    close(fds[0]);
    if (int rc = map_handle(s, s->handles[from], &s->ptrs[from])) return rc;
    if (from == 0 && s->want_mc) {
      if (nf == 2) {
        CUresult r = drv.MemImportFromShareableHandle(&s->mc_handle, reinterpret_cast<void*>(static_cast<uintptr_t>(fds[1])),
                                                      CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
        close(fds[1]);
        if (r != CUDA_SUCCESS) mc_ok = false;
      } else {
        mc_ok = false;
      }
    }
  }
  close(s->listen_fd); s->listen_fd = -1;
  close(s->local_fd); s->local_fd = -1;
  if (s->mc_fd >= 0) { close(s->mc_fd); s->mc_fd = -1; }
  if (mc_ok) {
    CUdevice dev;
    mc_ok = drv.DeviceGet(&dev, s->device) == CUDA_SUCCESS && drv.MulticastAddDevice(s->mc_handle, dev) == CUDA_SUCCESS;
  }
  s->has_mc = mc_ok;  // provisional: every rank must report success before bind (Python allreduces this flag)
  void* host_ptrs[kMaxWorld];
  for (int p = 0; p < s->world; ++p) host_ptrs[p] = reinterpret_cast<void*>(s->ptrs[p]);
  PDL_CUDA_CHECK(cudaMalloc(&s->d_ptrs, sizeof(void*) * s->world));
  PDL_CUDA_CHECK(cudaMemcpy(s->d_ptrs, host_ptrs, sizeof(void*) * s->world, cudaMemcpyHostToDevice));
  return 0;
}

PDL_API int pdl_symm_multicast_ready(void* h) { return static_cast<SymmSeg*>(h)->has_mc ? 1 : 0; }

// Phase 3 (after every rank added its device): bind local memory to the multicast object and map it.
// `all_ok` is the AND of pdl_symm_multicast_ready over the group.
PDL_API int pdl_symm_bind_multicast(void* h, int all_ok) {
  auto* s = static_cast<SymmSeg*>(h);
  if (!all_ok || !s->has_mc) { s->has_mc = false; return 0; }
  if (drv.MulticastBindMem(s->mc_handle, 0, s->handles[s->rank], 0, s->bytes, 0) != CUDA_SUCCESS) { s->has_mc = false; return 1; }
  CUdeviceptr va = 0;
  if (map_handle(s, s->mc_handle, &va) != 0) { s->has_mc = false; return 1; }
  s->mc_ptr = va;
  return 0;
}

PDL_API uint64_t pdl_symm_bytes(void* h) { return static_cast<SymmSeg*>(h)->bytes; }
PDL_API void* pdl_symm_ptr(void* h, int rank) { return reinterpret_cast<void*>(static_cast<SymmSeg*>(h)->ptrs[rank]); }
PDL_API void* pdl_symm_mc_ptr(void* h) { auto* s = static_cast<SymmSeg*>(h); return s->has_mc ? reinterpret_cast<void*>(s->mc_ptr) : nullptr; }
PDL_API void** pdl_symm_dev_ptrs(void* h) { return static_cast<SymmSeg*>(h)->d_ptrs; }

PDL_API int pdl_symm_destroy(void* h) {
  auto* s = static_cast<SymmSeg*>(h);
  if (!s) return 0;
  cudaSetDevice(s->device);
  cudaDeviceSynchronize();
  if (s->mc_ptr) { drv.MemUnmap(s->mc_ptr, s->bytes); drv.MemAddressFree(s->mc_ptr, s->bytes); }
  for (int p = 0; p < s->world; ++p) {
    if (s->ptrs[p]) { drv.MemUnmap(s->ptrs[p], s->bytes); drv.MemAddressFree(s->ptrs[p], s->bytes); }
    if (s->handles[p]) drv.MemRelease(s->handles[p]);
  }
  if (s->mc_handle) drv.MemRelease(s->mc_handle);
  if (s->d_ptrs) cudaFree(s->d_ptrs);
  if (s->listen_fd >= 0) close(s->listen_fd);
  if (s->local_fd >= 0) close(s->local_fd);
  delete s;
  return 0;
}
