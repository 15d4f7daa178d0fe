// Symmetric buffers bound to an NVSwitch multicast object (NVLS), without any library underneath.
//
// Replaces what ProcessGroupNCCL / NCCL's NVLS transport set up internally for the reference's
// collectives (pytorch_allreduce.py:31, data_parallel_training.py:44, zero_dp_training.py:46-50) —
// SURVEY.md N2 / W2 "alloc, handle exchange, peer table, multicast bind".
//
// Every rank creates physical memory with cuMemCreate (shareable as a POSIX file descriptor), maps
// its own and every peer's allocation (peer table for P2P loads/stores) and binds its allocation to
// ONE multicast object created by rank 0.  The multicast object is mapped too: a store to that
// address lands in every rank's copy (multimem.st), a multimem.ld_reduce makes the switch read all
// copies and return the sum.  File descriptors travel between the rank processes as SCM_RIGHTS
// messages over abstract-namespace unix datagram sockets (one socket per rank, named after its pid);
// pids and the order of the steps are exchanged by the Python side over the bootstrap group:
//
//   dtds_mc_create      every rank   physical allocation + fd, opens its fd socket
//   dtds_mc_root_create rank 0       cuMulticastCreate + fd
//   dtds_mc_send_fds    every rank   own memory fd -> every peer (rank 0: also the multicast fd)
//   dtds_mc_recv_add    every rank   receive fds, import, cuMulticastAddDevice
//   (bootstrap barrier: all devices added)
//   dtds_mc_bind_map    every rank   cuMulticastBindMem, map local / peers / multicast, fill tables
//   (bootstrap barrier: all bound before first use)
#include <cuda.h>
#include <cuda_runtime.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

#include "dtds_b200.h"

namespace {

constexpr int kMaxRanks = 16;

struct Drv {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                               unsigned long long);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*DeviceGet)(CUdevice*, int);
  bool ok = false;
};

template <class F>
bool load_sym(F& fn, const char* name) {
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &sym, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess ||
      sym == nullptr) {
    fprintf(stderr, "[dtds] multicast: driver entry point %s not available\n", name);
    return false;
  }
  fn = reinterpret_cast<F>(sym);
  return true;
}

Drv& drv() {
  static Drv d;
  static std::once_flag once;
  std::call_once(once, [] {
    cudaFree(nullptr);  // make sure the primary context exists
    bool ok = true;
    ok &= load_sym(d.MemCreate, "cuMemCreate");
    ok &= load_sym(d.MemGetAllocationGranularity, "cuMemGetAllocationGranularity");
    ok &= load_sym(d.MemExportToShareableHandle, "cuMemExportToShareableHandle");
    ok &= load_sym(d.MemImportFromShareableHandle, "cuMemImportFromShareableHandle");
    ok &= load_sym(d.MemAddressReserve, "cuMemAddressReserve");
    ok &= load_sym(d.MemAddressFree, "cuMemAddressFree");
    ok &= load_sym(d.MemMap, "cuMemMap");
    ok &= load_sym(d.MemUnmap, "cuMemUnmap");
    ok &= load_sym(d.MemSetAccess, "cuMemSetAccess");
    ok &= load_sym(d.MemRelease, "cuMemRelease");
    ok &= load_sym(d.MulticastCreate, "cuMulticastCreate");
    ok &= load_sym(d.MulticastAddDevice, "cuMulticastAddDevice");
    ok &= load_sym(d.MulticastBindMem, "cuMulticastBindMem");
    ok &= load_sym(d.MulticastGetGranularity, "cuMulticastGetGranularity");
    ok &= load_sym(d.DeviceGetAttribute, "cuDeviceGetAttribute");
    ok &= load_sym(d.DeviceGet, "cuDeviceGet");
    d.ok = ok;
  });
  return d;
}

#define DRV_CHECK(call)                                                                       \
  do {                                                                                        \
    CUresult _r = (call);                                                                     \
    if (_r != CUDA_SUCCESS) {                                                                 \
      fprintf(stderr, "[dtds] multicast: %s failed with CUresult %d (%s:%d)\n", #call, (int)_r, __FILE__, __LINE__); \
      return -10 - (int)_r;                                                                   \
    }                                                                                         \
  } while (0)

struct McBuffer {
  int dev = 0, rank = 0, world = 0;
  size_t bytes = 0;
  CUmemGenericAllocationHandle mem = 0, mc = 0;
  CUmemGenericAllocationHandle peer_mem[kMaxRanks] = {};
  int mem_fd = -1, mc_fd = -1, sock = -1;
  int peer_fd[kMaxRanks];
  CUdeviceptr local_ptr = 0, mc_ptr = 0, peer_ptr[kMaxRanks] = {};
  bool in_use = false;
};

std::mutex g_mu;
std::vector<McBuffer*> g_bufs;

McBuffer* get(int id) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (id < 0 || id >= (int)g_bufs.size() || g_bufs[id] == nullptr) return nullptr;
  return g_bufs[id];
}

void sock_name(sockaddr_un* a, socklen_t* len, int pid, int id) {
  memset(a, 0, sizeof(*a));
  a->sun_family = AF_UNIX;
  // abstract namespace: sun_path[0] == 0, no file system entry to clean up
  const int n = snprintf(a->sun_path + 1, sizeof(a->sun_path) - 1, "dtds_mc_%d_%d", pid, id);
  *len = static_cast<socklen_t>(offsetof(sockaddr_un, sun_path) + 1 + n);
}

int send_fd(int sock, int dst_pid, int id, int fd, int tag) {
  sockaddr_un to;
  socklen_t tolen;
  sock_name(&to, &tolen, dst_pid, id);
  msghdr msg;
  memset(&msg, 0, sizeof(msg));
  iovec iov;
  int payload = tag;
  iov.iov_base = &payload;
  iov.iov_len = sizeof(payload);
  char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  msg.msg_name = &to;
  msg.msg_namelen = tolen;
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(c), &fd, sizeof(int));
  for (int attempt = 0; attempt < 200; ++attempt) {
    if (sendmsg(sock, &msg, 0) >= 0) return 0;
    usleep(10000);  // peer socket not bound yet / queue full
  }
  perror("[dtds] multicast: sendmsg(SCM_RIGHTS)");
  return -1;
}

int recv_fd(int sock, int* fd_out, int* tag_out) {
  msghdr msg;
  memset(&msg, 0, sizeof(msg));
  iovec iov;
  int payload = 0;
  iov.iov_base = &payload;
  iov.iov_len = sizeof(payload);
  char ctrl[CMSG_SPACE(sizeof(int))];
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  if (recvmsg(sock, &msg, 0) < 0) {
    perror("[dtds] multicast: recvmsg");
    return -1;
  }
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  if (c == nullptr || c->cmsg_type != SCM_RIGHTS) {
    fprintf(stderr, "[dtds] multicast: message without a file descriptor\n");
    return -1;
  }
  memcpy(fd_out, CMSG_DATA(c), sizeof(int));
  *tag_out = payload;
  return 0;
}

CUmemAllocationProp alloc_prop(int dev) {
  CUmemAllocationProp p;
  memset(&p, 0, sizeof(p));
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = dev;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

int map_rw(Drv& d, CUdeviceptr* out, CUmemGenericAllocationHandle h, size_t bytes, size_t align, int dev) {
  DRV_CHECK(d.MemAddressReserve(out, bytes, align, 0, 0));
  DRV_CHECK(d.MemMap(*out, bytes, 0, h, 0));
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  DRV_CHECK(d.MemSetAccess(*out, bytes, &acc, 1));
  return 0;
}

}  // namespace

// 1 if the current device can join multicast objects (NVSwitch + driver support), else 0.
extern "C" int dtds_mc_supported(void) {
  Drv& d = drv();
  if (!d.ok) return 0;
  int dev = 0;
  cudaGetDevice(&dev);
  CUdevice cd;
  if (d.DeviceGet(&cd, dev) != CUDA_SUCCESS) return 0;
  int v = 0;
  if (d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cd) != CUDA_SUCCESS) return 0;
  return v != 0;
}

// Physical allocation of at least `bytes` on the current device (rounded up to the multicast
// granularity; *bytes_out is the same on every rank for the same request).  Returns a buffer id.
extern "C" int dtds_mc_create(size_t bytes, int rank, int world, size_t* bytes_out) {
  Drv& d = drv();
  if (!d.ok || world < 2 || world > kMaxRanks) return -1;
  McBuffer* b = new McBuffer();
  cudaGetDevice(&b->dev);
  b->rank = rank;
  b->world = world;
  for (int i = 0; i < kMaxRanks; ++i) b->peer_fd[i] = -1;
  CUmemAllocationProp prop = alloc_prop(b->dev);
  size_t gran = 0, mc_gran = 0;
  DRV_CHECK(d.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = world;
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  DRV_CHECK(d.MulticastGetGranularity(&mc_gran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  const size_t g = gran > mc_gran ? gran : mc_gran;
  b->bytes = (bytes + g - 1) / g * g;
  DRV_CHECK(d.MemCreate(&b->mem, b->bytes, &prop, 0));
  DRV_CHECK(d.MemExportToShareableHandle(&b->mem_fd, b->mem, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  int id;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    id = (int)g_bufs.size();
    g_bufs.push_back(b);
  }
  b->sock = socket(AF_UNIX, SOCK_DGRAM, 0);
  sockaddr_un me;
  socklen_t melen;
  sock_name(&me, &melen, (int)getpid(), id);
  if (b->sock < 0 || bind(b->sock, reinterpret_cast<sockaddr*>(&me), melen) != 0) {
    perror("[dtds] multicast: fd socket");
    return -2;
  }
  timeval tv{60, 0};
  setsockopt(b->sock, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
  *bytes_out = b->bytes;
  return id;
}

// Rank 0: creates the multicast object for `world` devices.
extern "C" int dtds_mc_root_create(int id) {
  Drv& d = drv();
  McBuffer* b = get(id);
  if (b == nullptr) return -1;
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = b->world;
  mp.size = b->bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  DRV_CHECK(d.MulticastCreate(&b->mc, &mp));
  DRV_CHECK(d.MemExportToShareableHandle(&b->mc_fd, b->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  return 0;
}

// Sends this rank's memory fd to every peer (tag = rank); rank 0 also sends the multicast fd
// (tag = -1).  `pids[r]` / `ids[r]` = process id and buffer id of rank r (buffer ids are equal when
// every rank allocates in the same order, but they are exchanged anyway).
extern "C" int dtds_mc_send_fds(int id, const int* pids, const int* ids) {
  McBuffer* b = get(id);
  if (b == nullptr) return -1;
  for (int r = 0; r < b->world; ++r) {
    if (r == b->rank) continue;
    if (send_fd(b->sock, pids[r], ids[r], b->mem_fd, b->rank)) return -2;
    if (b->rank == 0 && send_fd(b->sock, pids[r], ids[r], b->mc_fd, -1)) return -2;
  }
  return 0;
}

// Receives the peers' fds, imports them and adds this device to the multicast object.
extern "C" int dtds_mc_recv_add(int id) {
  Drv& d = drv();
  McBuffer* b = get(id);
  if (b == nullptr) return -1;
  const int expect = (b->world - 1) + (b->rank == 0 ? 0 : 1);
  for (int i = 0; i < expect; ++i) {
    int fd = -1, tag = 0;
    if (recv_fd(b->sock, &fd, &tag)) return -2;
    if (tag == -1) {
      b->mc_fd = fd;
      DRV_CHECK(d.MemImportFromShareableHandle(&b->mc, reinterpret_cast<void*>(static_cast<intptr_t>(fd)),
                                               CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    } else if (tag >= 0 && tag < b->world && tag != b->rank) {
      b->peer_fd[tag] = fd;
      DRV_CHECK(d.MemImportFromShareableHandle(&b->peer_mem[tag], reinterpret_cast<void*>(static_cast<intptr_t>(fd)),
                                               CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    } else {
      fprintf(stderr, "[dtds] multicast: unexpected fd message tag %d\n", tag);
      return -3;
    }
  }
  CUdevice cd;
  DRV_CHECK(d.DeviceGet(&cd, b->dev));
  DRV_CHECK(d.MulticastAddDevice(b->mc, cd));
  return 0;
}

// After every rank has added its device: binds the local memory to the multicast object and maps
// everything.  ptrs_out[world] = address of every rank's buffer in this process (own one included),
// *mc_ptr_out = multicast address.  The memory is zero-filled.
extern "C" int dtds_mc_bind_map(int id, unsigned long long* ptrs_out, unsigned long long* mc_ptr_out) {
  Drv& d = drv();
  McBuffer* b = get(id);
  if (b == nullptr) return -1;
  DRV_CHECK(d.MulticastBindMem(b->mc, 0, b->mem, 0, b->bytes, 0));
  CUmemAllocationProp prop = alloc_prop(b->dev);
  size_t gran = 0;
  DRV_CHECK(d.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  int rc;
  if ((rc = map_rw(d, &b->local_ptr, b->mem, b->bytes, gran, b->dev))) return rc;
  if ((rc = map_rw(d, &b->mc_ptr, b->mc, b->bytes, gran, b->dev))) return rc;
  for (int r = 0; r < b->world; ++r) {
    if (r == b->rank) {
      b->peer_ptr[r] = b->local_ptr;
    } else if ((rc = map_rw(d, &b->peer_ptr[r], b->peer_mem[r], b->bytes, gran, b->dev))) {
      return rc;
    }
    ptrs_out[r] = static_cast<unsigned long long>(b->peer_ptr[r]);
  }
  *mc_ptr_out = static_cast<unsigned long long>(b->mc_ptr);
  if (cudaMemset(reinterpret_cast<void*>(b->local_ptr), 0, b->bytes) != cudaSuccess) return -4;
  if (cudaDeviceSynchronize() != cudaSuccess) return -4;
  // the fds are no longer needed once imported and mapped
  for (int r = 0; r < b->world; ++r)
    if (b->peer_fd[r] >= 0) { close(b->peer_fd[r]); b->peer_fd[r] = -1; }
  if (b->mem_fd >= 0) { close(b->mem_fd); b->mem_fd = -1; }
  if (b->mc_fd >= 0) { close(b->mc_fd); b->mc_fd = -1; }
  if (b->sock >= 0) { close(b->sock); b->sock = -1; }
  b->in_use = true;
  return 0;
}

extern "C" int dtds_mc_free(int id) {
  Drv& d = drv();
  McBuffer* b = get(id);
  if (b == nullptr) return -1;
  cudaDeviceSynchronize();
  if (b->mc_ptr) { d.MemUnmap(b->mc_ptr, b->bytes); d.MemAddressFree(b->mc_ptr, b->bytes); }
  for (int r = 0; r < b->world; ++r) {
    if (b->peer_ptr[r]) { d.MemUnmap(b->peer_ptr[r], b->bytes); d.MemAddressFree(b->peer_ptr[r], b->bytes); }
    if (r != b->rank && b->peer_mem[r]) d.MemRelease(b->peer_mem[r]);
  }
  if (b->mc) d.MemRelease(b->mc);
  if (b->mem) d.MemRelease(b->mem);
  if (b->sock >= 0) close(b->sock);
  {
    std::lock_guard<std::mutex> lk(g_mu);
    g_bufs[id] = nullptr;
  }
  delete b;
  return 0;
}
