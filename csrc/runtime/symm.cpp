// symm.cpp — symmetric memory over NVLink: every rank allocates one physical block on its GPU,
// exports it as a POSIX file descriptor, passes the descriptor to every peer process over
// Unix-domain sockets (SCM_RIGHTS) and maps all peers' blocks into its own address space.
// After that a kernel on any GPU can load/store any peer's buffer directly through NVSwitch.
// Where the fabric exposes it, the same blocks are additionally bound to an NVLS multicast
// object: one address that reduces on load (multimem.ld_reduce) and broadcasts on store.
//
// One process per GPU, one box; no NCCL, no NVSHMEM, no torch on this path.
// (SURVEY.md section 5.9 / W12.  The reference's transport is gRPC between hosts,
// README.md:395; nothing of it is reused.)
#include <errno.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/time.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../cuda/host_common.h"

using namespace b2d;

namespace {

constexpr int kMaxRanks = 16;

#define DRV(name) reinterpret_cast<decltype(&name)>(driver_fn(#name))

double now_s() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec + ts.tv_nsec * 1e-9;
}

// Abstract-namespace Unix socket address: no filesystem entry, vanishes with the process.
socklen_t make_addr(sockaddr_un* a, const std::string& name) {
  std::memset(a, 0, sizeof(*a));
  a->sun_family = AF_UNIX;
  const size_t n = std::min(name.size(), sizeof(a->sun_path) - 2);
  std::memcpy(a->sun_path + 1, name.data(), n);
  return static_cast<socklen_t>(offsetof(sockaddr_un, sun_path) + 1 + n);
}

int send_fd(int sock, int fd, int32_t rank) {
  msghdr msg{};
  iovec iov{&rank, sizeof(rank)};
  char ctrl[CMSG_SPACE(sizeof(int))] = {};
  msg.msg_iov = &iov, msg.msg_iovlen = 1;
  msg.msg_control = ctrl, msg.msg_controllen = sizeof(ctrl);
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET, c->cmsg_type = SCM_RIGHTS, c->cmsg_len = CMSG_LEN(sizeof(int));
  std::memcpy(CMSG_DATA(c), &fd, sizeof(int));
  return sendmsg(sock, &msg, 0) == (ssize_t)sizeof(rank) ? 0 : -1;
}

int recv_fd(int sock, int* fd, int32_t* rank) {
  msghdr msg{};
  iovec iov{rank, sizeof(*rank)};
  char ctrl[CMSG_SPACE(sizeof(int))] = {};
  msg.msg_iov = &iov, msg.msg_iovlen = 1;
  msg.msg_control = ctrl, msg.msg_controllen = sizeof(ctrl);
  if (recvmsg(sock, &msg, 0) != (ssize_t)sizeof(*rank)) return -1;
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  if (c == nullptr || c->cmsg_type != SCM_RIGHTS) return -1;
  std::memcpy(fd, CMSG_DATA(c), sizeof(int));
  return 0;
}

}  // namespace

struct b2d_symm {
  int rank = 0, world = 1, device = 0;
  size_t bytes = 0;  // mapped size (granularity-rounded)
  CUmemGenericAllocationHandle handle = 0;
  CUmemGenericAllocationHandle peer_handle[kMaxRanks] = {};
  CUdeviceptr va[kMaxRanks] = {};  // va[rank] is the local mapping
  int listen_sock = -1;
  std::string tag;
  double timeout_s = 120.0;
  // multicast (NVLS)
  bool mc_supported = false, mc_bound = false;
  CUmemGenericAllocationHandle mc_handle = 0;
  CUdeviceptr mc_va = 0;
  size_t mc_bytes = 0;
};

// Round-robin descriptor exchange: in round p every other rank connects to rank p and hands
// over its descriptor (root < 0), or only `root` distributes one descriptor to everyone.
static int exchange_fds(b2d_symm* s, const char* phase, int my_fd, int root, int* fds_out) {
  const std::string base = "b2d-" + s->tag + "-" + phase + "-";
  sockaddr_un addr;
  // every rank that receives in this phase listens before anybody connects to it
  const bool i_receive = root < 0 || s->rank != root;
  int lsock = -1;
  if (i_receive) {
    lsock = socket(AF_UNIX, SOCK_STREAM, 0);
    socklen_t len = make_addr(&addr, base + std::to_string(s->rank));
    if (lsock < 0 || bind(lsock, reinterpret_cast<sockaddr*>(&addr), len) != 0 || listen(lsock, kMaxRanks) != 0) {
      set_error("symm: cannot listen on unix socket (%s)", strerror(errno));
      if (lsock >= 0) close(lsock);
      return B2D_ERR_SYS;
    }
  }
  auto connect_and_send = [&](int peer, int fd) -> int {
    const double t0 = now_s();
    while (true) {
      int c = socket(AF_UNIX, SOCK_STREAM, 0);
      socklen_t len = make_addr(&addr, base + std::to_string(peer));
      if (connect(c, reinterpret_cast<sockaddr*>(&addr), len) == 0) {
        const int rc = send_fd(c, fd, s->rank);
        close(c);
        if (rc != 0) set_error("symm: sendmsg to rank %d failed (%s)", peer, strerror(errno));
        return rc == 0 ? B2D_OK : B2D_ERR_SYS;
      }
      close(c);
      if (now_s() - t0 > s->timeout_s) {
        set_error("symm: rank %d never opened its socket (phase %s)", peer, phase);
        return B2D_ERR_SYS;
      }
      usleep(2000);
    }
  };
  int rc = B2D_OK;
  if (root < 0) {
    for (int p = 0; p < s->world && rc == B2D_OK; ++p)
      if (p != s->rank) rc = connect_and_send(p, my_fd);
  } else if (s->rank == root) {
    for (int p = 0; p < s->world && rc == B2D_OK; ++p)
      if (p != root) rc = connect_and_send(p, my_fd);
  }
  if (rc == B2D_OK && i_receive) {
    const int expect = root < 0 ? s->world - 1 : 1;
    const double t0 = now_s();
    for (int k = 0; k < expect; ++k) {
      // a peer that died after the rendezvous must not hang this rank: wait for the connection with the
      // remaining time budget, and bound the receive as well
      pollfd pfd{lsock, POLLIN, 0};
      const double left = s->timeout_s - (now_s() - t0);
      if (left <= 0 || poll(&pfd, 1, static_cast<int>(left * 1000)) <= 0) {
        set_error("symm: timed out after %.0f s waiting for descriptor %d of %d (phase %s)", s->timeout_s, k + 1,
                  expect, phase);
        rc = B2D_ERR_SYS;
        break;
      }
      int c = accept(lsock, nullptr, nullptr);
      int fd = -1;
      int32_t from = -1;
      if (c >= 0) {
        timeval tv{static_cast<time_t>(s->timeout_s), 0};
        setsockopt(c, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
      }
      if (c < 0 || recv_fd(c, &fd, &from) != 0 || from < 0 || from >= s->world) {
        set_error("symm: receiving a descriptor failed (%s)", strerror(errno));
        rc = B2D_ERR_SYS;
        if (c >= 0) close(c);
        break;
      }
      close(c);
      fds_out[from] = fd;
    }
  }
  if (lsock >= 0) close(lsock);
  return rc;
}

static int map_handle(CUmemGenericAllocationHandle h, size_t bytes, size_t gran, int device,
                      CUdeviceptr* out) {
  auto reserve = DRV(cuMemAddressReserve);
  auto map = DRV(cuMemMap);
  auto set_access = DRV(cuMemSetAccess);
  if (!reserve || !map || !set_access) return B2D_ERR_UNSUPPORTED;
  B2D_CU_CHECK(reserve(out, bytes, gran, 0, 0));
  B2D_CU_CHECK(map(*out, bytes, 0, h, 0));
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  B2D_CU_CHECK(set_access(*out, bytes, &acc, 1));
  return B2D_OK;
}

static int symm_create_impl(b2d_symm* s, size_t bytes);

// Collective over all ranks of `tag`: allocate, exchange, map.  `bytes` must be equal everywhere.
// Every failure path releases what was created so far (handle, address ranges, mappings, imported peers).
extern "C" int b2d_symm_create(b2d_symm** out, const char* tag, int rank, int world, int device,
                               size_t bytes, double timeout_s) {
  B2D_REQUIRE(out && tag && world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world && bytes > 0,
              "symm_create: bad arguments (world <= %d)", kMaxRanks);
  B2D_CUDA_CHECK(cudaSetDevice(device));
  B2D_CUDA_CHECK(cudaFree(nullptr));  // make sure the primary context exists
  b2d_symm* s = new (std::nothrow) b2d_symm();
  B2D_REQUIRE(s != nullptr, "symm_create: out of host memory");
  s->rank = rank, s->world = world, s->device = device, s->tag = tag;
  s->timeout_s = timeout_s > 0 ? timeout_s : 120.0;
  const int rc = symm_create_impl(s, bytes);
  if (rc != B2D_OK) {
    b2d_symm_destroy(s);
    return rc;
  }
  *out = s;
  return B2D_OK;
}

static int symm_create_impl(b2d_symm* s, size_t bytes) {
  const int rank = s->rank, world = s->world, device = s->device;
  auto get_gran = DRV(cuMemGetAllocationGranularity);
  auto create = DRV(cuMemCreate);
  auto export_h = DRV(cuMemExportToShareableHandle);
  auto import_h = DRV(cuMemImportFromShareableHandle);
  if (!get_gran || !create || !export_h || !import_h) return B2D_ERR_UNSUPPORTED;

  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  B2D_CU_CHECK(get_gran(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  // Multicast objects have their own (coarser) granularity; honour it up front so the same
  // block can be bound later.
  int mc_attr = 0;
  if (auto get_attr = DRV(cuDeviceGetAttribute))
    get_attr(&mc_attr, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, device);
  s->mc_supported = mc_attr != 0;   // world == 1 (a one-device multicast object) is used by the single-GPU test
  if (s->mc_supported) {
    auto mc_gran = DRV(cuMulticastGetGranularity);
    CUmulticastObjectProp mp{};
    mp.numDevices = world, mp.size = bytes, mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t g2 = 0;
    if (mc_gran && mc_gran(&g2, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && g2 > gran)
      gran = g2;
  }
  s->bytes = (bytes + gran - 1) / gran * gran;
  B2D_CU_CHECK(create(&s->handle, s->bytes, &prop, 0));
  int rc = map_handle(s->handle, s->bytes, gran, device, &s->va[rank]);
  if (rc != B2D_OK) return rc;
  B2D_CUDA_CHECK(cudaMemset(reinterpret_cast<void*>(s->va[rank]), 0, s->bytes));
  B2D_CUDA_CHECK(cudaDeviceSynchronize());

  if (world > 1) {
    int my_fd = -1;
    B2D_CU_CHECK(export_h(&my_fd, s->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    int fds[kMaxRanks];
    for (int i = 0; i < kMaxRanks; ++i) fds[i] = -1;
    rc = exchange_fds(s, "mem", my_fd, -1, fds);
    close(my_fd);
    if (rc != B2D_OK) return rc;
    for (int p = 0; p < world && rc == B2D_OK; ++p) {
      if (p == rank) continue;
      const CUresult r = import_h(&s->peer_handle[p], reinterpret_cast<void*>(static_cast<uintptr_t>(fds[p])),
                                  CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
      if (r != CUDA_SUCCESS) {
        set_error("symm: importing the block of rank %d failed: CUresult %d", p, static_cast<int>(r));
        rc = B2D_ERR_CUDA;
        break;
      }
      rc = map_handle(s->peer_handle[p], s->bytes, gran, device, &s->va[p]);
    }
    for (int p = 0; p < world; ++p)
      if (fds[p] >= 0) close(fds[p]);   // the imported handles keep the allocations alive
    if (rc != B2D_OK) return rc;
  }
  return B2D_OK;
}

// NVLS phase 1 (collective): rank 0 creates the multicast object and ships it; everyone adds
// its device.  Callers must run a cross-rank barrier between phase 1 and phase 2.
extern "C" int b2d_symm_multicast_phase1(b2d_symm* s) {
  B2D_REQUIRE(s != nullptr, "symm_multicast: null handle");
  if (!s->mc_supported) {
    set_error("symm_multicast: device does not report multicast support");
    return B2D_ERR_UNSUPPORTED;
  }
  auto mc_create = DRV(cuMulticastCreate);
  auto mc_add = DRV(cuMulticastAddDevice);
  auto export_h = DRV(cuMemExportToShareableHandle);
  auto import_h = DRV(cuMemImportFromShareableHandle);
  if (!mc_create || !mc_add || !export_h || !import_h) return B2D_ERR_UNSUPPORTED;
  CUmulticastObjectProp mp{};
  mp.numDevices = s->world, mp.size = s->bytes, mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  int fds[kMaxRanks];
  for (int i = 0; i < kMaxRanks; ++i) fds[i] = -1;
  if (s->rank == 0) {
    B2D_CU_CHECK(mc_create(&s->mc_handle, &mp));
    int fd = -1;
    B2D_CU_CHECK(export_h(&fd, s->mc_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    const int rc = exchange_fds(s, "mc", fd, 0, fds);
    close(fd);
    if (rc != B2D_OK) return rc;
  } else {
    const int rc = exchange_fds(s, "mc", -1, 0, fds);
    if (rc != B2D_OK) return rc;
    B2D_CU_CHECK(import_h(&s->mc_handle, reinterpret_cast<void*>(static_cast<uintptr_t>(fds[0])),
                          CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    close(fds[0]);
  }
  B2D_CU_CHECK(mc_add(s->mc_handle, s->device));
  return B2D_OK;
}

// NVLS phase 2: bind this rank's block at offset 0 and map the multicast address.
extern "C" int b2d_symm_multicast_phase2(b2d_symm* s) {
  B2D_REQUIRE(s != nullptr && s->mc_handle != 0, "symm_multicast: phase 1 has not run");
  auto mc_bind = DRV(cuMulticastBindMem);
  auto mc_gran = DRV(cuMulticastGetGranularity);
  if (!mc_bind || !mc_gran) return B2D_ERR_UNSUPPORTED;
  B2D_CU_CHECK(mc_bind(s->mc_handle, 0, s->handle, 0, s->bytes, 0));
  CUmulticastObjectProp mp{};
  mp.numDevices = s->world, mp.size = s->bytes, mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  B2D_CU_CHECK(mc_gran(&gran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  const int rc = map_handle(s->mc_handle, s->bytes, gran, s->device, &s->mc_va);
  if (rc != B2D_OK) return rc;
  s->mc_bytes = s->bytes;
  s->mc_bound = true;
  return B2D_OK;
}

extern "C" void* b2d_symm_ptr(const b2d_symm* s, int rank) {
  return (s && rank >= 0 && rank < s->world) ? reinterpret_cast<void*>(s->va[rank]) : nullptr;
}
extern "C" void* b2d_symm_multicast_ptr(const b2d_symm* s) {
  return (s && s->mc_bound) ? reinterpret_cast<void*>(s->mc_va) : nullptr;
}
extern "C" size_t b2d_symm_bytes(const b2d_symm* s) { return s ? s->bytes : 0; }
extern "C" int b2d_symm_multicast_supported(const b2d_symm* s) { return s && s->mc_supported ? 1 : 0; }

extern "C" int b2d_symm_destroy(b2d_symm* s) {
  if (s == nullptr) return B2D_OK;
  auto unmap = DRV(cuMemUnmap);
  auto release = DRV(cuMemRelease);
  auto addr_free = DRV(cuMemAddressFree);
  cudaDeviceSynchronize();
  if (unmap && release && addr_free) {
    if (s->mc_va) {
      unmap(s->mc_va, s->mc_bytes);
      addr_free(s->mc_va, s->mc_bytes);
    }
    if (s->mc_handle) {
      if (auto unbind = DRV(cuMulticastUnbind))
        if (s->mc_bound) unbind(s->mc_handle, s->device, 0, s->bytes);
      release(s->mc_handle);
    }
    for (int p = 0; p < s->world; ++p) {
      if (s->va[p]) {
        unmap(s->va[p], s->bytes);
        addr_free(s->va[p], s->bytes);
      }
      if (p != s->rank && s->peer_handle[p]) release(s->peer_handle[p]);
    }
    if (s->handle) release(s->handle);
  }
  delete s;
  return B2D_OK;
}
