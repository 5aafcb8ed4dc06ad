// Symmetric device memory across the ranks of one node (C1 in SURVEY.md §2.5).
//
// Every rank allocates the same number of bytes with the CUDA virtual-memory API (cuMemCreate, POSIX
// file-descriptor handles), passes the descriptor to every peer over a unix-domain socket
// (SCM_RIGHTS) and maps all `world` allocations into its own address space. After that, kernels on
// any GPU can load/store any rank's buffer directly over NVLink 5 / NVSwitch. When the platform
// supports it, the allocations are also bound to one multicast object (NVLS): a store to the
// multicast address lands in every rank's buffer, a multimem.ld_reduce sums all of them in the switch.
//
// The reference gets its communication from NCCL through torch DDP (implicit at
// my_ray_module.py:135,159); it has no counterpart of this file.
#include <dlfcn.h>
#include <errno.h>
#include <poll.h>
#include <pthread.h>
#include <stdlib.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

#include <string>
#include <vector>

#include "common.h"
#include "symm.h"

namespace rtdc {

// ---------------------------------------------------------------- driver API, resolved at run time
#define RTDC_DRV_FUNCS(X)                                                                                 \
  X(cuMemCreate) X(cuMemRelease) X(cuMemAddressReserve) X(cuMemAddressFree) X(cuMemMap) X(cuMemUnmap)     \
  X(cuMemSetAccess) X(cuMemExportToShareableHandle) X(cuMemImportFromShareableHandle)                     \
  X(cuMemGetAllocationGranularity) X(cuDeviceGetAttribute) X(cuMulticastCreate) X(cuMulticastAddDevice)   \
  X(cuMulticastBindMem) X(cuMulticastGetGranularity) X(cuMulticastUnbind) X(cuGetErrorString)

struct Drv {
#define X(name) decltype(&::name) name = nullptr;
  RTDC_DRV_FUNCS(X)
#undef X
  bool ok = false;
};

static Drv& drv() {
  static Drv d;
  static bool tried = false;
  if (tried) return d;
  tried = true;
  bool all = true;
#define X(name)                                                                                      \
  {                                                                                                  \
    void* sym = nullptr;                                                                             \
    cudaDriverEntryPointQueryResult q;                                                               \
    if (cudaGetDriverEntryPoint(#name, &sym, cudaEnableDefault, &q) == cudaSuccess &&                \
        q == cudaDriverEntryPointSuccess && sym)                                                     \
      d.name = reinterpret_cast<decltype(&::name)>(sym);                                               \
    else if (std::string(#name).rfind("cuMulticast", 0) != 0)                                        \
      all = false;                                                                                   \
  }
  RTDC_DRV_FUNCS(X)
#undef X
  d.ok = all;
  return d;
}

static const char* cu_err(CUresult r) {
  const char* s = nullptr;
  if (drv().cuGetErrorString && drv().cuGetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
  return "unknown CUresult";
}
#define RTDC_CU_CHECK(expr)                                                                              \
  do {                                                                                                   \
    CUresult _r = (expr);                                                                                \
    if (_r != CUDA_SUCCESS) return fail(RTDC_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cu_err(_r), __FILE__, __LINE__); \
  } while (0)

// ---------------------------------------------------------------- fd passing
static int send_fd(int sock, int fd, int payload) {
  msghdr msg;
  memset(&msg, 0, sizeof(msg));
  char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  iovec io = {&payload, sizeof(payload)};
  msg.msg_iov = &io;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(c), &fd, sizeof(int));
  return sendmsg(sock, &msg, 0) == (ssize_t)sizeof(payload) ? 0 : -1;
}

static int recv_fd(int sock, int* fd, int* payload) {
  msghdr msg;
  memset(&msg, 0, sizeof(msg));
  char ctrl[CMSG_SPACE(sizeof(int))];
  iovec io = {payload, sizeof(*payload)};
  msg.msg_iov = &io;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  if (recvmsg(sock, &msg, 0) != (ssize_t)sizeof(*payload)) return -1;
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  if (!c || c->cmsg_type != SCM_RIGHTS) return -1;
  memcpy(fd, CMSG_DATA(c), sizeof(int));
  return 0;
}

static double now_s() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec + ts.tv_nsec * 1e-9;
}

static std::string sock_path(const std::string& prefix, int rank) { return prefix + "." + std::to_string(rank) + ".sock"; }

struct ServeArgs {
  int listen_fd;
  int n_clients;
  std::vector<int> fds;  // indexed by request kind: 0 = memory handle, 1 = multicast handle (rank 0 only)
  double timeout_s;
  int rc;
};

// Serve `n_clients` requests: each client sends (kind) and gets back the descriptor of that kind.
static void* serve_thread(void* p) {
  ServeArgs* a = static_cast<ServeArgs*>(p);
  a->rc = 0;
  const double deadline = now_s() + a->timeout_s;
  for (int served = 0; served < a->n_clients;) {
    pollfd pf = {a->listen_fd, POLLIN, 0};
    int pr = poll(&pf, 1, 200);
    if (pr < 0 && errno != EINTR) { a->rc = -1; return nullptr; }
    if (pr <= 0) {
      if (now_s() > deadline) { a->rc = -2; return nullptr; }
      continue;
    }
    int c = accept(a->listen_fd, nullptr, nullptr);
    if (c < 0) continue;
    int kind = -1;
    if (read(c, &kind, sizeof(kind)) == (ssize_t)sizeof(kind) && kind >= 0 && kind < (int)a->fds.size() && a->fds[kind] >= 0) {
      if (send_fd(c, a->fds[kind], kind) != 0) a->rc = -3;
    } else {
      a->rc = -4;
    }
    close(c);
    ++served;
  }
  return nullptr;
}

static int request_fd(const std::string& path, int kind, double timeout_s, int* out_fd) {
  const double deadline = now_s() + timeout_s;
  while (true) {
    int s = socket(AF_UNIX, SOCK_STREAM, 0);
    if (s < 0) return -1;
    sockaddr_un addr;
    memset(&addr, 0, sizeof(addr));
    addr.sun_family = AF_UNIX;
    strncpy(addr.sun_path, path.c_str(), sizeof(addr.sun_path) - 1);
    if (connect(s, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) == 0) {
      int got_kind = -1;
      int rc = -1;
      if (write(s, &kind, sizeof(kind)) == (ssize_t)sizeof(kind)) rc = recv_fd(s, out_fd, &got_kind);
      close(s);
      return rc;
    }
    close(s);
    if (now_s() > deadline) return -2;
    usleep(2000);  // peer has not bound its socket yet
  }
}

struct Symm {
  int rank = 0, world = 1, device = 0;
  size_t bytes = 0;        // usable bytes per rank (caller's request, rounded up)
  size_t alloc_bytes = 0;  // physical allocation per rank = bytes + signal pad, granularity aligned
  size_t pad_offset = 0;   // signal pad lives at the end of each rank's allocation
  CUmemGenericAllocationHandle local_handle = 0;
  std::vector<CUmemGenericAllocationHandle> peer_handles;
  CUdeviceptr va_base = 0;  // world consecutive windows of alloc_bytes
  std::vector<void*> peer_ptrs;
  void** peer_table_dev = nullptr;  // device copy of peer_ptrs
  void** pad_table_dev = nullptr;   // device copy of per-rank signal pad pointers
  // multicast
  bool has_mc = false;
  CUmemGenericAllocationHandle mc_handle = 0;
  CUdeviceptr mc_va = 0;
  size_t mc_bytes = 0;
  unsigned int* epochs_dev = nullptr;  // per-block launch counters (local, not shared)
};

}  // namespace rtdc

using namespace rtdc;

RTDC_API int rtdc_symm_create(int rank, int world, int device, size_t bytes, const char* rendezvous_prefix,
                              int enable_multicast, double timeout_s, void** out_handle) {
  if (!out_handle || world < 1 || world > RTDC_SYMM_MAX_WORLD || rank < 0 || rank >= world)
    return fail(RTDC_ERR_ARG, "symm_create: bad rank/world (%d/%d, max world %d)", rank, world, RTDC_SYMM_MAX_WORLD);
  if (!rendezvous_prefix) return fail(RTDC_ERR_ARG, "symm_create: rendezvous prefix required");
  RTDC_CUDA_CHECK(cudaSetDevice(device));
  RTDC_CUDA_CHECK(cudaFree(0));
  Drv& d = drv();
  if (!d.ok) return fail(RTDC_ERR_CUDA, "CUDA driver lacks the virtual memory management API");

  Symm* h = new Symm();
  h->rank = rank; h->world = world; h->device = device;

  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  RTDC_CU_CHECK(d.cuMemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));

  // multicast support?
  int mc_supported = 0;
  if (enable_multicast && world > 1 && d.cuMulticastCreate)
    d.cuDeviceGetAttribute(&mc_supported, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, device);
  CUmulticastObjectProp mcprop;
  memset(&mcprop, 0, sizeof(mcprop));
  size_t mc_gran = 0;
  if (mc_supported) {
    mcprop.numDevices = world;
    mcprop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    mcprop.size = gran;  // placeholder to query granularity
    if (d.cuMulticastGetGranularity(&mc_gran, &mcprop, CU_MULTICAST_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) mc_supported = 0;
    if (mc_gran > gran) gran = mc_gran;
  }

  const size_t data_bytes = (bytes + 4095) / 4096 * 4096;
  h->pad_offset = data_bytes;
  h->alloc_bytes = (data_bytes + RTDC_SYMM_PAD_BYTES + gran - 1) / gran * gran;
  h->bytes = data_bytes;

  RTDC_CU_CHECK(d.cuMemCreate(&h->local_handle, h->alloc_bytes, &prop, 0));
  int my_fd = -1;
  RTDC_CU_CHECK(d.cuMemExportToShareableHandle(&my_fd, h->local_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));

  // multicast object is created by rank 0 and shared like the memory handles
  int mc_fd = -1;
  if (mc_supported && rank == 0) {
    mcprop.size = h->alloc_bytes;
    if (d.cuMulticastCreate(&h->mc_handle, &mcprop) == CUDA_SUCCESS &&
        d.cuMemExportToShareableHandle(&mc_fd, h->mc_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) == CUDA_SUCCESS) {
      h->has_mc = true;
    } else {
      mc_fd = -1;
    }
  }

  // --- exchange descriptors
  h->peer_handles.assign(world, 0);
  h->peer_handles[rank] = h->local_handle;
  const std::string prefix(rendezvous_prefix);
  pthread_t th;
  ServeArgs sa;
  int lfd = -1;
  if (world > 1) {
    lfd = socket(AF_UNIX, SOCK_STREAM, 0);
    if (lfd < 0) return fail(RTDC_ERR_SYSTEM, "socket(): %s", strerror(errno));
    sockaddr_un addr;
    memset(&addr, 0, sizeof(addr));
    addr.sun_family = AF_UNIX;
    const std::string mine = sock_path(prefix, rank);
    if (mine.size() >= sizeof(addr.sun_path)) return fail(RTDC_ERR_ARG, "rendezvous prefix too long for a unix socket path");
    strncpy(addr.sun_path, mine.c_str(), sizeof(addr.sun_path) - 1);
    unlink(mine.c_str());
    if (bind(lfd, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) != 0 || listen(lfd, 64) != 0)
      return fail(RTDC_ERR_SYSTEM, "bind/listen %s: %s", mine.c_str(), strerror(errno));
    sa.listen_fd = lfd;
    // every peer asks me for my memory handle; rank 0 is also asked for the multicast handle
    sa.n_clients = (world - 1) * ((rank == 0 && mc_supported) ? 2 : 1);
    sa.fds = {my_fd, mc_fd};
    sa.timeout_s = timeout_s;
    sa.rc = 0;
    pthread_create(&th, nullptr, serve_thread, &sa);
  }
  int rc_all = RTDC_OK;
  std::vector<int> got(world, -1);
  for (int p = 0; p < world && rc_all == RTDC_OK; ++p) {
    if (p == rank) continue;
    if (request_fd(sock_path(prefix, p), 0, timeout_s, &got[p]) != 0)
      rc_all = fail(RTDC_ERR_SYSTEM, "rank %d: could not obtain the memory handle of rank %d within %.0fs", rank, p, timeout_s);
  }
  int got_mc_fd = -1;
  bool mc_failed = false;
  if (rc_all == RTDC_OK && mc_supported && rank != 0) {
    // rank 0 answers kind 1 with -1 fd impossible; a failed create on rank 0 shows up as a refused request
    if (request_fd(sock_path(prefix, 0), 1, timeout_s, &got_mc_fd) != 0) mc_failed = true;
  }
  if (world > 1) {
    pthread_join(th, nullptr);
    close(lfd);
    unlink(sock_path(prefix, rank).c_str());
    if (rc_all == RTDC_OK && sa.rc != 0 && !(sa.rc == -4 && rank == 0))
      rc_all = fail(RTDC_ERR_SYSTEM, "rank %d: serving memory handles failed (%d)", rank, sa.rc);
  }
  if (rc_all != RTDC_OK) return rc_all;

  for (int p = 0; p < world; ++p) {
    if (p == rank) continue;
    RTDC_CU_CHECK(d.cuMemImportFromShareableHandle(&h->peer_handles[p], reinterpret_cast<void*>(static_cast<intptr_t>(got[p])),
                                                   CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    close(got[p]);
  }
  close(my_fd);

  // --- map everything: one VA range with `world` windows
  RTDC_CU_CHECK(d.cuMemAddressReserve(&h->va_base, h->alloc_bytes * world, gran, 0, 0));
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  h->peer_ptrs.resize(world);
  for (int p = 0; p < world; ++p) {
    CUdeviceptr va = h->va_base + static_cast<CUdeviceptr>(h->alloc_bytes) * p;
    RTDC_CU_CHECK(d.cuMemMap(va, h->alloc_bytes, 0, h->peer_handles[p], 0));
    RTDC_CU_CHECK(d.cuMemSetAccess(va, h->alloc_bytes, &acc, 1));
    h->peer_ptrs[p] = reinterpret_cast<void*>(va);
  }
  RTDC_CUDA_CHECK(cudaMemset(h->peer_ptrs[rank], 0, h->alloc_bytes));

  // --- multicast binding (best effort: any failure just disables NVLS)
  if (mc_supported && !mc_failed) {
    bool okmc = true;
    if (rank != 0) {
      okmc = d.cuMemImportFromShareableHandle(&h->mc_handle, reinterpret_cast<void*>(static_cast<intptr_t>(got_mc_fd)),
                                              CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) == CUDA_SUCCESS;
      close(got_mc_fd);
    } else {
      okmc = h->has_mc;
      if (mc_fd >= 0) close(mc_fd);
    }
    okmc = okmc && d.cuMulticastAddDevice(h->mc_handle, device) == CUDA_SUCCESS;
    h->has_mc = okmc;
    h->mc_bytes = h->alloc_bytes;
    // cuMulticastBindMem must run after every device was added: the caller synchronises the ranks
    // (rtdc_symm_bind_multicast is phase two).
  } else {
    h->has_mc = false;
  }

  // --- device-side tables
  std::vector<void*> pads(world);
  for (int p = 0; p < world; ++p) pads[p] = static_cast<char*>(h->peer_ptrs[p]) + h->pad_offset;
  RTDC_CUDA_CHECK(cudaMalloc(&h->peer_table_dev, sizeof(void*) * world));
  RTDC_CUDA_CHECK(cudaMalloc(&h->pad_table_dev, sizeof(void*) * world));
  RTDC_CUDA_CHECK(cudaMemcpy(h->peer_table_dev, h->peer_ptrs.data(), sizeof(void*) * world, cudaMemcpyHostToDevice));
  RTDC_CUDA_CHECK(cudaMemcpy(h->pad_table_dev, pads.data(), sizeof(void*) * world, cudaMemcpyHostToDevice));
  RTDC_CUDA_CHECK(cudaMalloc(&h->epochs_dev, sizeof(unsigned int) * RTDC_SYMM_MAX_BLOCKS));
  RTDC_CUDA_CHECK(cudaMemset(h->epochs_dev, 0, sizeof(unsigned int) * RTDC_SYMM_MAX_BLOCKS));
  RTDC_CUDA_CHECK(cudaDeviceSynchronize());
  *out_handle = h;
  return RTDC_OK;
}

// Phase two of multicast set-up; call on every rank after all ranks returned from rtdc_symm_create
// (they must all have added their device). Returns 0 and leaves multicast disabled on failure.
RTDC_API int rtdc_symm_bind_multicast(void* handle) {
  Symm* h = static_cast<Symm*>(handle);
  if (!h || !h->has_mc) return RTDC_OK;
  Drv& d = drv();
  bool ok = d.cuMulticastBindMem(h->mc_handle, 0, h->local_handle, 0, h->alloc_bytes, 0) == CUDA_SUCCESS;
  if (ok) {
    size_t gran = 0;
    CUmulticastObjectProp mcprop;
    memset(&mcprop, 0, sizeof(mcprop));
    mcprop.numDevices = h->world;
    mcprop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    mcprop.size = h->alloc_bytes;
    d.cuMulticastGetGranularity(&gran, &mcprop, CU_MULTICAST_GRANULARITY_RECOMMENDED);
    ok = d.cuMemAddressReserve(&h->mc_va, h->alloc_bytes, gran, 0, 0) == CUDA_SUCCESS &&
         d.cuMemMap(h->mc_va, h->alloc_bytes, 0, h->mc_handle, 0) == CUDA_SUCCESS;
    if (ok) {
      CUmemAccessDesc acc;
      memset(&acc, 0, sizeof(acc));
      acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
      acc.location.id = h->device;
      acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
      ok = d.cuMemSetAccess(h->mc_va, h->alloc_bytes, &acc, 1) == CUDA_SUCCESS;
    }
  }
  if (!ok) {
    h->has_mc = false;
    h->mc_va = 0;
  }
  return RTDC_OK;
}

RTDC_API int rtdc_symm_info(void* handle, rtdc_symm_info_t* out) {
  Symm* h = static_cast<Symm*>(handle);
  if (!h || !out) return fail(RTDC_ERR_ARG, "symm_info: null");
  out->rank = h->rank; out->world = h->world; out->device = h->device;
  out->bytes = h->bytes;
  out->local_ptr = h->peer_ptrs[h->rank];
  out->multicast_ptr = h->has_mc ? reinterpret_cast<void*>(h->mc_va) : nullptr;
  out->peer_table_dev = h->peer_table_dev;
  out->pad_table_dev = h->pad_table_dev;
  out->epochs_dev = h->epochs_dev;
  for (int p = 0; p < RTDC_SYMM_MAX_WORLD; ++p) out->peer_ptrs[p] = p < h->world ? h->peer_ptrs[p] : nullptr;
  return RTDC_OK;
}

RTDC_API int rtdc_symm_destroy(void* handle) {
  Symm* h = static_cast<Symm*>(handle);
  if (!h) return RTDC_OK;
  Drv& d = drv();
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  if (h->mc_va) { d.cuMemUnmap(h->mc_va, h->alloc_bytes); d.cuMemAddressFree(h->mc_va, h->alloc_bytes); }
  if (h->has_mc && d.cuMulticastUnbind) d.cuMulticastUnbind(h->mc_handle, h->device, 0, h->alloc_bytes);
  if (h->mc_handle) d.cuMemRelease(h->mc_handle);
  for (int p = 0; p < h->world; ++p) {
    if (h->peer_ptrs.size() > (size_t)p && h->peer_ptrs[p]) d.cuMemUnmap(reinterpret_cast<CUdeviceptr>(h->peer_ptrs[p]), h->alloc_bytes);
    if (h->peer_handles.size() > (size_t)p && h->peer_handles[p]) d.cuMemRelease(h->peer_handles[p]);
  }
  if (h->va_base) d.cuMemAddressFree(h->va_base, h->alloc_bytes * h->world);
  cudaFree(h->peer_table_dev);
  cudaFree(h->pad_table_dev);
  cudaFree(h->epochs_dev);
  delete h;
  return RTDC_OK;
}
