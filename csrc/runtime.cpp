// Bring-up of the peer-memory transport: single-node rendezvous, symmetric heap, topology probe.
//
// Replaces what the reference gets from dist.init_process_group('nccl') + NCCL communicator setup
// (ref:main.py:190-193; c10d TCPStore + ProcessGroupNCCL; SURVEY.md N2, N2', N3, W3, W4):
//   * rendezvous   -- an abstract-namespace unix socket named after the run id (no fixed port:
//                     the reference's 127.0.0.1:20080 collides between concurrent runs).  Rank 0
//                     serves; the protocol is all-gather of byte blobs plus all-gather of file
//                     descriptors (SCM_RIGHTS), which is what exporting CUDA VMM handles needs;
//   * symmetric heap -- every rank creates one VMM allocation (cuMemCreate, POSIX fd handle), the
//                     fds are exchanged and every rank maps every peer's allocation, so a buffer is
//                     identified by its OFFSET and kernels address rank w's copy as bases[w] + offset.
//                     Falls back to cudaMalloc + cudaIpc handles if VMM export is refused;
//   * multicast    -- where CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED is set, the same allocations are
//                     bound to one multicast object so kernels can use multimem.ld_reduce / multimem.st
//                     (in-switch reduction, "NVLS");
//   * the first bytes of the heap are the signal pad (flags + epoch counters) used by comm.cu.
#include <cuda.h>
#include <cuda_runtime.h>
#include <errno.h>
#include <fcntl.h>
#include <poll.h>
#include <stddef.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

#include <string>
#include <vector>

#include "../include/pmdist.h"
#include "../include/pmdist_comm.h"

#define PMD_API __attribute__((visibility("default")))

namespace pmd {
void set_error(const std::string& msg);
}
using pmd::set_error;

#define RT_REQUIRE(cond, msg)                                                        \
  do {                                                                               \
    if (!(cond)) {                                                                   \
      set_error(std::string(msg) + " [" #cond "] errno=" + std::to_string(errno));   \
      return -1;                                                                     \
    }                                                                                \
  } while (0)

// ------------------------------------------------------------------------------------ rendezvous
struct PmdRendezvous {
  int rank = 0, world = 1;
  int listen_fd = -1;
  std::vector<int> peers;   // rank 0: socket to each rank (index = rank, [0] unused); others: [0] = socket to rank 0
  double timeout_s = 120.0;
};

namespace {

double now_s() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec + ts.tv_nsec * 1e-9;
}

int make_addr(const std::string& run_id, sockaddr_un* addr, socklen_t* len) {
  memset(addr, 0, sizeof(*addr));
  addr->sun_family = AF_UNIX;
  const std::string name = "pmdist_" + run_id;
  if (name.size() + 1 >= sizeof(addr->sun_path)) return -1;
  addr->sun_path[0] = '\0';   // abstract namespace: vanishes with the last socket, nothing to unlink
  memcpy(addr->sun_path + 1, name.data(), name.size());
  *len = static_cast<socklen_t>(offsetof(sockaddr_un, sun_path) + 1 + name.size());
  return 0;
}

int send_all(int fd, const void* buf, size_t n) {
  const char* p = static_cast<const char*>(buf);
  while (n > 0) {
    ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
    if (k < 0) {
      if (errno == EINTR) continue;
      return -1;
    }
    p += k;
    n -= static_cast<size_t>(k);
  }
  return 0;
}

int recv_all(int fd, void* buf, size_t n, double timeout_s) {
  char* p = static_cast<char*>(buf);
  const double deadline = now_s() + timeout_s;
  while (n > 0) {
    pollfd pfd{fd, POLLIN, 0};
    const int ms = static_cast<int>((deadline - now_s()) * 1000);
    if (ms <= 0) {
      errno = ETIMEDOUT;
      return -1;
    }
    const int pr = ::poll(&pfd, 1, ms > 1000 ? 1000 : ms);
    if (pr < 0 && errno != EINTR) return -1;
    if (pr <= 0) continue;
    ssize_t k = ::recv(fd, p, n, 0);
    if (k == 0) {
      errno = ECONNRESET;   // peer died: surface it instead of hanging (failure detection)
      return -1;
    }
    if (k < 0) {
      if (errno == EINTR || errno == EAGAIN) continue;
      return -1;
    }
    p += k;
    n -= static_cast<size_t>(k);
  }
  return 0;
}

int send_fd(int sock, int fd) {
  char byte = 'F';
  iovec iov{&byte, 1};
  char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  msghdr msg{};
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(c), &fd, sizeof(int));
  for (;;) {
    if (::sendmsg(sock, &msg, MSG_NOSIGNAL) >= 0) return 0;
    if (errno != EINTR) return -1;
  }
}

int recv_fd(int sock, int* fd_out, double timeout_s) {
  const double deadline = now_s() + timeout_s;
  for (;;) {
    pollfd pfd{sock, POLLIN, 0};
    const int ms = static_cast<int>((deadline - now_s()) * 1000);
    if (ms <= 0) {
      errno = ETIMEDOUT;
      return -1;
    }
    const int pr = ::poll(&pfd, 1, ms > 1000 ? 1000 : ms);
    if (pr < 0 && errno != EINTR) return -1;
    if (pr <= 0) continue;
    char byte = 0;
    iovec iov{&byte, 1};
    char ctrl[CMSG_SPACE(sizeof(int))];
    msghdr msg{};
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof(ctrl);
    const ssize_t k = ::recvmsg(sock, &msg, 0);
    if (k == 0) {
      errno = ECONNRESET;
      return -1;
    }
    if (k < 0) {
      if (errno == EINTR || errno == EAGAIN) continue;
      return -1;
    }
    cmsghdr* c = CMSG_FIRSTHDR(&msg);
    if (c == nullptr || c->cmsg_type != SCM_RIGHTS) {
      errno = EPROTO;
      return -1;
    }
    memcpy(fd_out, CMSG_DATA(c), sizeof(int));
    return 0;
  }
}

}  // namespace

extern "C" PMD_API int pmd_rdv_create(int rank, int world, const char* run_id, double timeout_s, PmdRendezvous** out) {
  RT_REQUIRE(world >= 1 && rank >= 0 && rank < world, "bad rank/world");
  PmdRendezvous* r = new PmdRendezvous();
  r->rank = rank;
  r->world = world;
  r->timeout_s = timeout_s > 0 ? timeout_s : 120.0;
  r->peers.assign(world, -1);
  *out = r;
  if (world == 1) return 0;
  sockaddr_un addr;
  socklen_t len;
  RT_REQUIRE(make_addr(run_id, &addr, &len) == 0, "run id too long for a socket name");
  if (rank == 0) {
    r->listen_fd = ::socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
    RT_REQUIRE(r->listen_fd >= 0, "socket() failed");
    RT_REQUIRE(::bind(r->listen_fd, reinterpret_cast<sockaddr*>(&addr), len) == 0,
               "bind() failed (is another job using the same run id?)");
    RT_REQUIRE(::listen(r->listen_fd, world) == 0, "listen() failed");
    const double deadline = now_s() + r->timeout_s;
    for (int got = 0; got < world - 1;) {
      pollfd pfd{r->listen_fd, POLLIN, 0};
      const int pr = ::poll(&pfd, 1, 1000);
      RT_REQUIRE(now_s() < deadline, "rendezvous timed out waiting for peers to connect");
      if (pr <= 0) continue;
      const int fd = ::accept4(r->listen_fd, nullptr, nullptr, SOCK_CLOEXEC);
      if (fd < 0) continue;
      int peer_rank = -1;
      RT_REQUIRE(recv_all(fd, &peer_rank, sizeof(int), r->timeout_s) == 0, "handshake failed");
      RT_REQUIRE(peer_rank > 0 && peer_rank < world && r->peers[peer_rank] < 0, "bad or duplicate peer rank");
      r->peers[peer_rank] = fd;
      ++got;
    }
  } else {
    const double deadline = now_s() + r->timeout_s;
    int fd = -1;
    for (;;) {
      fd = ::socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
      RT_REQUIRE(fd >= 0, "socket() failed");
      if (::connect(fd, reinterpret_cast<sockaddr*>(&addr), len) == 0) break;
      ::close(fd);
      RT_REQUIRE(now_s() < deadline, "rendezvous timed out connecting to rank 0");
      usleep(20000);
    }
    RT_REQUIRE(send_all(fd, &rank, sizeof(int)) == 0, "handshake send failed");
    r->peers[0] = fd;
  }
  return 0;
}

extern "C" PMD_API int pmd_rdv_destroy(PmdRendezvous* r) {
  if (r == nullptr) return 0;
  for (int fd : r->peers)
    if (fd >= 0) ::close(fd);
  if (r->listen_fd >= 0) ::close(r->listen_fd);
  delete r;
  return 0;
}

// out must hold world * bytes; every rank contributes `bytes`.
extern "C" PMD_API int pmd_rdv_allgather(PmdRendezvous* r, const void* in, int bytes, void* out) {
  char* o = static_cast<char*>(out);
  if (r->world == 1) {
    memcpy(o, in, bytes);
    return 0;
  }
  if (r->rank == 0) {
    memcpy(o, in, bytes);
    for (int w = 1; w < r->world; ++w)
      RT_REQUIRE(recv_all(r->peers[w], o + static_cast<size_t>(w) * bytes, bytes, r->timeout_s) == 0,
                 "allgather: a peer went away or timed out");
    for (int w = 1; w < r->world; ++w)
      RT_REQUIRE(send_all(r->peers[w], o, static_cast<size_t>(r->world) * bytes) == 0, "allgather: send failed");
  } else {
    RT_REQUIRE(send_all(r->peers[0], in, bytes) == 0, "allgather: send failed");
    RT_REQUIRE(recv_all(r->peers[0], o, static_cast<size_t>(r->world) * bytes, r->timeout_s) == 0,
               "allgather: rank 0 went away or timed out");
  }
  return 0;
}

extern "C" PMD_API int pmd_rdv_barrier(PmdRendezvous* r) {
  std::vector<char> tmp(r->world);
  char z = 1;
  return pmd_rdv_allgather(r, &z, 1, tmp.data());
}

// fds_out[w] = a descriptor in THIS process referring to rank w's `fd` (fds_out[rank] = dup of own).
extern "C" PMD_API int pmd_rdv_allgather_fds(PmdRendezvous* r, int fd, int* fds_out) {
  if (r->world == 1) {
    fds_out[0] = ::dup(fd);
    return 0;
  }
  if (r->rank == 0) {
    std::vector<int> all(r->world, -1);
    all[0] = fd;
    for (int w = 1; w < r->world; ++w)
      RT_REQUIRE(recv_fd(r->peers[w], &all[w], r->timeout_s) == 0, "fd gather failed");
    for (int w = 1; w < r->world; ++w)
      for (int s = 0; s < r->world; ++s) RT_REQUIRE(send_fd(r->peers[w], all[s]) == 0, "fd scatter failed");
    for (int w = 0; w < r->world; ++w) fds_out[w] = (w == 0) ? ::dup(fd) : all[w];
  } else {
    RT_REQUIRE(send_fd(r->peers[0], fd) == 0, "fd send failed");
    for (int s = 0; s < r->world; ++s) RT_REQUIRE(recv_fd(r->peers[0], &fds_out[s], r->timeout_s) == 0, "fd recv failed");
  }
  return 0;
}

// ------------------------------------------------------------------------------------ driver API
namespace {

struct Driver {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*McCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*McAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*McBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                        unsigned long long);
  CUresult (*McGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  bool ok = false;
};

template <typename F>
bool load_sym(const char* name, F* fn) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess ||
      p == nullptr)
    return false;
  *fn = reinterpret_cast<F>(p);
  return true;
}

Driver& driver() {
  static Driver d;
  static bool tried = false;
  if (!tried) {
    tried = true;
    bool ok = true;
    ok &= load_sym("cuMemCreate", &d.MemCreate);
    ok &= load_sym("cuMemRelease", &d.MemRelease);
    ok &= load_sym("cuMemAddressReserve", &d.MemAddressReserve);
    ok &= load_sym("cuMemAddressFree", &d.MemAddressFree);
    ok &= load_sym("cuMemMap", &d.MemMap);
    ok &= load_sym("cuMemUnmap", &d.MemUnmap);
    ok &= load_sym("cuMemSetAccess", &d.MemSetAccess);
    ok &= load_sym("cuMemExportToShareableHandle", &d.MemExport);
    ok &= load_sym("cuMemImportFromShareableHandle", &d.MemImport);
    ok &= load_sym("cuMemGetAllocationGranularity", &d.MemGranularity);
    ok &= load_sym("cuDeviceGetAttribute", &d.DeviceGetAttribute);
    d.ok = ok;
    // multicast is optional
    if (!load_sym("cuMulticastCreate", &d.McCreate) || !load_sym("cuMulticastAddDevice", &d.McAddDevice) ||
        !load_sym("cuMulticastBindMem", &d.McBindMem) || !load_sym("cuMulticastGetGranularity", &d.McGranularity))
      d.McCreate = nullptr;
  }
  return d;
}

std::string cu_err(const char* what, CUresult r) {
  return std::string(what) + " failed with CUresult " + std::to_string(static_cast<int>(r));
}

}  // namespace

// ------------------------------------------------------------------------------------ communicator
struct PmdComm {
  PmdRendezvous* rdv = nullptr;
  int rank = 0, world = 1, device = 0;
  size_t heap_bytes = 0;          // usable bytes (rounded up to granularity)
  int transport = PMD_TRANSPORT_SINGLE;
  std::vector<void*> bases;       // bases[w]: rank w's heap as mapped here
  std::vector<CUmemGenericAllocationHandle> handles;
  unsigned long long* d_bases = nullptr;   // device copy of bases
  void* mc_base = nullptr;        // multicast mapping (NULL if unavailable)
  CUmemGenericAllocationHandle mc_handle = 0;
  size_t bump = 0;                // next free offset
  int multicast_supported = 0;
  int p2p_ok = 1;
  std::string note;
};

namespace {

int vmm_setup(PmdComm* c, bool want_multicast) {
  Driver& d = driver();
  RT_REQUIRE(d.ok, "CUDA VMM driver entry points are not available");
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = c->device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  CUresult r = d.MemGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
  if (r != CUDA_SUCCESS) {
    set_error(cu_err("cuMemGetAllocationGranularity", r));
    return -1;
  }
  int mc_attr = 0;
  d.DeviceGetAttribute(&mc_attr, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, c->device);
  c->multicast_supported = mc_attr;
  size_t mc_gran = 0;
  CUmulticastObjectProp mprop{};
  const bool try_mc = want_multicast && mc_attr && d.McCreate != nullptr && c->world > 1;
  if (try_mc) {
    mprop.numDevices = c->world;
    mprop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    mprop.size = c->heap_bytes;
    if (d.McGranularity(&mc_gran, &mprop, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mc_gran > gran)
      gran = mc_gran;
  }
  c->heap_bytes = (c->heap_bytes + gran - 1) / gran * gran;

  CUmemGenericAllocationHandle mine;
  r = d.MemCreate(&mine, c->heap_bytes, &prop, 0);
  if (r != CUDA_SUCCESS) {
    set_error(cu_err("cuMemCreate", r));
    return -1;
  }
  int my_fd = -1;
  r = d.MemExport(&my_fd, mine, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    d.MemRelease(mine);
    set_error(cu_err("cuMemExportToShareableHandle", r));
    return -1;
  }
  std::vector<int> fds(c->world, -1);
  if (pmd_rdv_allgather_fds(c->rdv, my_fd, fds.data()) != 0) return -1;
  ::close(my_fd);

  c->handles.assign(c->world, 0);
  c->bases.assign(c->world, nullptr);
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = c->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  for (int w = 0; w < c->world; ++w) {
    CUmemGenericAllocationHandle h = mine;
    if (w != c->rank) {
      r = d.MemImport(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fds[w])),
                      CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
      if (r != CUDA_SUCCESS) {
        set_error(cu_err("cuMemImportFromShareableHandle", r));
        return -1;
      }
    }
    ::close(fds[w]);
    c->handles[w] = h;
    CUdeviceptr va = 0;
    r = d.MemAddressReserve(&va, c->heap_bytes, gran, 0, 0);
    if (r != CUDA_SUCCESS) {
      set_error(cu_err("cuMemAddressReserve", r));
      return -1;
    }
    r = d.MemMap(va, c->heap_bytes, 0, h, 0);
    if (r != CUDA_SUCCESS) {
      set_error(cu_err("cuMemMap", r));
      return -1;
    }
    r = d.MemSetAccess(va, c->heap_bytes, &acc, 1);
    if (r != CUDA_SUCCESS) {
      set_error(cu_err("cuMemSetAccess (peer mapping)", r));
      return -1;
    }
    c->bases[w] = reinterpret_cast<void*>(va);
  }
  c->transport = PMD_TRANSPORT_VMM;

  if (try_mc) {
    // Multicast object: created by rank 0, shared as an fd, every device added, every heap bound.
    bool ok = true;
    CUmemGenericAllocationHandle mc = 0;
    int mc_fd = -1;
    if (c->rank == 0) {
      mprop.size = c->heap_bytes;
      ok = d.McCreate(&mc, &mprop) == CUDA_SUCCESS &&
           d.MemExport(&mc_fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) == CUDA_SUCCESS;
    }
    // agree on success before exchanging descriptors
    std::vector<char> flags(c->world);
    char f = ok ? 1 : 0;
    if (pmd_rdv_allgather(c->rdv, &f, 1, flags.data()) != 0) return -1;
    if (flags[0]) {
      std::vector<int> mfds(c->world, -1);
      int send = c->rank == 0 ? mc_fd : ::open("/dev/null", O_RDONLY);
      if (pmd_rdv_allgather_fds(c->rdv, send, mfds.data()) != 0) return -1;
      ::close(send);
      if (c->rank != 0)
        ok = d.MemImport(&mc, reinterpret_cast<void*>(static_cast<uintptr_t>(mfds[0])),
                         CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) == CUDA_SUCCESS;
      for (int fd : mfds)
        if (fd >= 0) ::close(fd);
      if (ok) ok = d.McAddDevice(mc, c->device) == CUDA_SUCCESS;
      f = ok ? 1 : 0;
      if (pmd_rdv_allgather(c->rdv, &f, 1, flags.data()) != 0) return -1;   // all devices added before any bind
      bool all = true;
      for (char x : flags) all = all && x;
      if (all) {
        ok = d.McBindMem(mc, 0, mine, 0, c->heap_bytes, 0) == CUDA_SUCCESS;
        CUdeviceptr va = 0;
        if (ok) ok = d.MemAddressReserve(&va, c->heap_bytes, gran, 0, 0) == CUDA_SUCCESS;
        if (ok) ok = d.MemMap(va, c->heap_bytes, 0, mc, 0) == CUDA_SUCCESS;
        if (ok) ok = d.MemSetAccess(va, c->heap_bytes, &acc, 1) == CUDA_SUCCESS;
        f = ok ? 1 : 0;
        if (pmd_rdv_allgather(c->rdv, &f, 1, flags.data()) != 0) return -1;
        all = true;
        for (char x : flags) all = all && x;
        if (all) {
          c->mc_base = reinterpret_cast<void*>(va);
          c->mc_handle = mc;
        } else {
          c->note += "multicast bind/map failed on some rank; ";
        }
      } else {
        c->note += "cuMulticastAddDevice failed on some rank; ";
      }
    } else {
      c->note += "cuMulticastCreate failed on rank 0; ";
    }
  }
  return 0;
}

int ipc_setup(PmdComm* c) {
  void* mine = nullptr;
  if (cudaMalloc(&mine, c->heap_bytes) != cudaSuccess) {
    set_error("cudaMalloc for the symmetric heap failed");
    return -1;
  }
  cudaIpcMemHandle_t h;
  if (cudaIpcGetMemHandle(&h, mine) != cudaSuccess) {
    set_error("cudaIpcGetMemHandle failed");
    return -1;
  }
  std::vector<cudaIpcMemHandle_t> all(c->world);
  if (pmd_rdv_allgather(c->rdv, &h, sizeof(h), all.data()) != 0) return -1;
  c->bases.assign(c->world, nullptr);
  for (int w = 0; w < c->world; ++w) {
    if (w == c->rank) {
      c->bases[w] = mine;
      continue;
    }
    if (cudaIpcOpenMemHandle(&c->bases[w], all[w], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
      set_error("cudaIpcOpenMemHandle failed for rank " + std::to_string(w));
      return -1;
    }
  }
  c->transport = PMD_TRANSPORT_IPC;
  return 0;
}

}  // namespace

extern "C" PMD_API int pmd_comm_create(int rank, int world, const char* run_id, int device,
                                       unsigned long long heap_bytes, int flags, PmdComm** out) {
  PmdComm* c = new PmdComm();
  *out = c;
  c->rank = rank;
  c->world = world;
  c->device = device;
  c->heap_bytes = heap_bytes < PMD_SIGNAL_BYTES * 2 ? PMD_SIGNAL_BYTES * 2 : heap_bytes;
  if (cudaSetDevice(device) != cudaSuccess || cudaFree(nullptr) != cudaSuccess) {
    set_error("cudaSetDevice failed for device " + std::to_string(device));
    return -1;
  }
  if (pmd_rdv_create(rank, world, run_id, 180.0, &c->rdv) != 0) return -1;
  // topology sanity: every peer must be P2P reachable (NVLink / NVSwitch)
  if (world > 1) {
    std::vector<int> devs(world);
    if (pmd_rdv_allgather(c->rdv, &device, sizeof(int), devs.data()) != 0) return -1;
    for (int w = 0; w < world; ++w) {
      if (devs[w] == device) continue;
      int can = 0;
      cudaDeviceCanAccessPeer(&can, device, devs[w]);
      if (!can) c->p2p_ok = 0;
    }
  }
  int rc = -1;
  const bool force_ipc = (flags & PMD_COMM_FORCE_IPC) != 0;
  if (!force_ipc) {
    rc = vmm_setup(c, (flags & PMD_COMM_NO_MULTICAST) == 0);
    // every rank must agree on the transport
    std::vector<char> okv(world);
    char ok = rc == 0 ? 1 : 0;
    if (pmd_rdv_allgather(c->rdv, &ok, 1, okv.data()) != 0) return -1;
    for (char x : okv)
      if (!x) rc = -1;
    if (rc != 0) c->note += std::string("VMM path unavailable (") + pmd_last_error() + "), using cudaIpc; ";
  }
  if (rc != 0) {
    if (ipc_setup(c) != 0) return -1;
  }
  if (cudaMalloc(&c->d_bases, sizeof(unsigned long long) * world) != cudaSuccess) {
    set_error("cudaMalloc(d_bases) failed");
    return -1;
  }
  std::vector<unsigned long long> hb(world);
  for (int w = 0; w < world; ++w) hb[w] = reinterpret_cast<unsigned long long>(c->bases[w]);
  cudaMemcpy(c->d_bases, hb.data(), sizeof(unsigned long long) * world, cudaMemcpyHostToDevice);
  cudaMemset(c->bases[rank], 0, PMD_SIGNAL_BYTES);   // flags and epochs start at zero
  cudaDeviceSynchronize();
  c->bump = PMD_SIGNAL_BYTES;
  if (pmd_rdv_barrier(c->rdv) != 0) return -1;        // nobody signals before every pad is cleared
  return 0;
}

extern "C" PMD_API int pmd_comm_destroy(PmdComm* c) {
  if (c == nullptr) return 0;
  cudaDeviceSynchronize();
  if (c->rdv != nullptr && c->world > 1) pmd_rdv_barrier(c->rdv);   // nobody unmaps while a peer may still touch it
  Driver& d = driver();
  if (c->transport == PMD_TRANSPORT_VMM) {
    if (c->mc_base != nullptr) {
      d.MemUnmap(reinterpret_cast<CUdeviceptr>(c->mc_base), c->heap_bytes);
      d.MemAddressFree(reinterpret_cast<CUdeviceptr>(c->mc_base), c->heap_bytes);
      d.MemRelease(c->mc_handle);
    }
    for (int w = 0; w < c->world; ++w) {
      if (c->bases[w] == nullptr) continue;
      d.MemUnmap(reinterpret_cast<CUdeviceptr>(c->bases[w]), c->heap_bytes);
      d.MemAddressFree(reinterpret_cast<CUdeviceptr>(c->bases[w]), c->heap_bytes);
      d.MemRelease(c->handles[w]);
    }
  } else if (c->transport == PMD_TRANSPORT_IPC) {
    for (int w = 0; w < c->world; ++w) {
      if (c->bases[w] == nullptr) continue;
      if (w == c->rank)
        cudaFree(c->bases[w]);
      else
        cudaIpcCloseMemHandle(c->bases[w]);
    }
  }
  if (c->d_bases != nullptr) cudaFree(c->d_bases);
  pmd_rdv_destroy(c->rdv);
  delete c;
  return 0;
}

extern "C" PMD_API int pmd_comm_info(PmdComm* c, PmdCommInfo* info) {
  memset(info, 0, sizeof(*info));
  info->rank = c->rank;
  info->world = c->world;
  info->device = c->device;
  info->transport = c->transport;
  info->multicast_supported = c->multicast_supported;
  info->multicast_mapped = c->mc_base != nullptr;
  info->p2p_ok = c->p2p_ok;
  info->heap_bytes = c->heap_bytes;
  info->local_base = reinterpret_cast<unsigned long long>(c->bases[c->rank]);
  info->mc_base = reinterpret_cast<unsigned long long>(c->mc_base);
  info->d_bases = reinterpret_cast<unsigned long long>(c->d_bases);
  strncpy(info->note, c->note.c_str(), sizeof(info->note) - 1);
  return 0;
}

extern "C" PMD_API unsigned long long pmd_comm_peer_base(PmdComm* c, int w) {
  return reinterpret_cast<unsigned long long>(c->bases[w]);
}

// Bump allocation inside the symmetric heap.  Every rank must issue the same sequence of calls so
// that offsets agree (checked cheaply by the caller through pmd_comm_allgather_host when wanted).
extern "C" PMD_API long long pmd_comm_alloc(PmdComm* c, unsigned long long bytes, unsigned long long align) {
  if (align < 256) align = 256;
  const size_t off = (c->bump + align - 1) / align * align;
  if (off + bytes > c->heap_bytes) {
    set_error("symmetric heap exhausted: need " + std::to_string(off + bytes) + " of " + std::to_string(c->heap_bytes));
    return -1;
  }
  c->bump = off + bytes;
  return static_cast<long long>(off);
}

extern "C" PMD_API int pmd_comm_host_barrier(PmdComm* c) { return pmd_rdv_barrier(c->rdv); }
extern "C" PMD_API int pmd_comm_allgather_host(PmdComm* c, const void* in, int bytes, void* out) {
  return pmd_rdv_allgather(c->rdv, in, bytes, out);
}

// ------------------------------------------------------------------------------------ topology probe
// What NCCL discovers internally (NVML / sysfs topology graph) reduced to what this transport needs to know
// about a single NVSwitch box: for every ordered device pair whether peer access exists, its performance rank,
// whether native atomics work across the link, and per device whether multicast (NVLS) is offered.
extern "C" PMD_API int pmd_topology_probe(int max_devices, int* n_devices, int* can_access /*[max*max]*/,
                                          int* perf_rank /*[max*max]*/, int* native_atomics /*[max*max]*/,
                                          int* multicast /*[max]*/) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    set_error("cudaGetDeviceCount failed (no CUDA device?)");
    return -1;
  }
  if (n > max_devices) n = max_devices;
  *n_devices = n;
  Driver& d = driver();
  for (int a = 0; a < n; ++a) {
    int mc = 0;
    if (d.ok) d.DeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, a);
    multicast[a] = mc;
    for (int b = 0; b < n; ++b) {
      int acc = a == b, rank = 0, atom = a == b;
      if (a != b) {
        cudaDeviceCanAccessPeer(&acc, a, b);
        cudaDeviceGetP2PAttribute(&rank, cudaDevP2PAttrPerformanceRank, a, b);
        cudaDeviceGetP2PAttribute(&atom, cudaDevP2PAttrNativeAtomicSupported, a, b);
      }
      can_access[a * max_devices + b] = acc;
      perf_rank[a * max_devices + b] = rank;
      native_atomics[a * max_devices + b] = atom;
    }
  }
  return 0;
}
