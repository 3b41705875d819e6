// comm_runtime.cu — the communication runtime under the peer-to-peer reducer (host code only).
//
// What the reference gets from torch.distributed + NCCL's bootstrap (hetseq/distributed_utils.py:11-41, and the
// buffers DDP's Reducer allocates, hetseq/controller.py:79-87) is here a small runtime of our own, shaped for one
// NVSwitch domain (<= 8 GPUs, one process per GPU):
//
//   * rendezvous and every later control message go over abstract-namespace Unix sockets (one listening socket
//     per rank, named by a job tag); file descriptors travel as SCM_RIGHTS ancillary data;
//   * a *symmetric buffer* is one physical allocation per rank (cuMemCreate, POSIX-FD shareable) that every rank
//     maps: b200_comm_alloc returns the local address, the address of every peer's copy in this process, and —
//     where the devices support it — an NVLS multicast address whose loads reduce over all copies inside the
//     switch (multimem.ld_reduce) and whose stores replicate to all copies (multimem.st);
//   * barriers are one round of empty messages.
//
// Nothing here launches a kernel; the kernels that use these mappings live in p2p_allreduce.cu.  All driver
// entry points are resolved at run time (cudaGetDriverEntryPoint), so the library still has no link-time
// dependency on libcuda and loads on machines without a driver.
#include <cuda.h>
#include "common.h"
#include "../include/b200seq.h"

#include <cerrno>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include <poll.h>
#include <sys/mman.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

namespace b200 {
namespace {

constexpr int kMaxRanks = 8;

// ---- driver API, resolved lazily ----------------------------------------------------------------------------
struct Driver {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
  CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*MemExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
  CUresult (*MemImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long) = nullptr;
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
  CUresult (*DeviceGet)(CUdevice*, int) = nullptr;
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
  CUresult (*GetErrorString)(CUresult, const char**) = nullptr;
  bool ok = false, multicast_api = false;
};

template <typename F>
bool resolve(const char* name, F& fn) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || p == nullptr) return false;
  fn = reinterpret_cast<F>(p);
  return true;
}

Driver& driver() {
  static Driver d;
  static bool tried = false;
  if (tried) return d;
  tried = true;
  bool ok = resolve("cuMemCreate", d.MemCreate) & resolve("cuMemRelease", d.MemRelease) &
            resolve("cuMemExportToShareableHandle", d.MemExport) & resolve("cuMemImportFromShareableHandle", d.MemImport) &
            resolve("cuMemAddressReserve", d.MemAddressReserve) & resolve("cuMemAddressFree", d.MemAddressFree) &
            resolve("cuMemMap", d.MemMap) & resolve("cuMemUnmap", d.MemUnmap) & resolve("cuMemSetAccess", d.MemSetAccess) &
            resolve("cuMemGetAllocationGranularity", d.MemGetAllocationGranularity) & resolve("cuDeviceGet", d.DeviceGet) &
            resolve("cuDeviceGetAttribute", d.DeviceGetAttribute) & resolve("cuGetErrorString", d.GetErrorString);
  d.ok = ok;
  d.multicast_api = resolve("cuMulticastCreate", d.MulticastCreate) & resolve("cuMulticastAddDevice", d.MulticastAddDevice) &
                    resolve("cuMulticastBindMem", d.MulticastBindMem) &
                    resolve("cuMulticastGetGranularity", d.MulticastGetGranularity) &
                    resolve("cuMulticastUnbind", d.MulticastUnbind);
  return d;
}

const char* cu_str(CUresult r) {
  const char* s = nullptr;
  if (driver().GetErrorString) driver().GetErrorString(r, &s);
  return s ? s : "unknown driver error";
}

#define B200_CU(expr)                                                                                  \
  do {                                                                                                 \
    CUresult _r = (expr);                                                                              \
    if (_r != CUDA_SUCCESS) return fail(B200_ERR_DRIVER, "%s: %s (%d)", #expr, cu_str(_r), (int)_r);    \
  } while (0)

// ---- control messages ---------------------------------------------------------------------------------------
struct Msg {
  uint32_t magic, seq;
  int32_t src, nfds;
  uint64_t word[4];
};
constexpr uint32_t kMagic = 0xB2005E91u;

struct Stashed {
  Msg m;
  int fds[2];
};

struct Buffer {
  size_t bytes = 0;                                  // mapped size (rounded up to the granularity)
  CUmemGenericAllocationHandle own = 0, mc = 0;
  CUmemGenericAllocationHandle peer[kMaxRanks] = {};
  CUdeviceptr va[kMaxRanks] = {};                    // va[rank] = local mapping
  CUdeviceptr mc_va = 0;
  bool mc_bound = false;
};

struct Comm {
  int rank = 0, world = 1, device = 0, listen_fd = -1, timeout_ms = 120000;
  uint32_t seq = 0;
  bool multicast = false;                            // every rank's device supports it and the API is present
  CUdevice cu_dev = 0;
  std::string tag;
  std::vector<Stashed> stash;
  std::vector<Buffer*> buffers;
};

socklen_t make_addr(sockaddr_un& a, const std::string& tag, int rank) {
  std::memset(&a, 0, sizeof(a));
  a.sun_family = AF_UNIX;
  const std::string name = "b200seq." + tag + "." + std::to_string(rank);
  const size_t n = name.size() < sizeof(a.sun_path) - 2 ? name.size() : sizeof(a.sun_path) - 2;
  a.sun_path[0] = '\0';                               // abstract namespace: vanishes with the process, no file to clean up
  std::memcpy(a.sun_path + 1, name.data(), n);
  return (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
}

long long now_ms() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (long long)ts.tv_sec * 1000 + ts.tv_nsec / 1000000;
}

int send_msg(Comm* c, int peer, const Msg& m, const int* fds) {
  sockaddr_un addr;
  const socklen_t len = make_addr(addr, c->tag, peer);
  const long long deadline = now_ms() + c->timeout_ms;
  int s = -1;
  for (;;) {
    s = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
    if (s < 0) return fail(B200_ERR_STATE, "comm: socket(): %s", strerror(errno));
    if (connect(s, reinterpret_cast<sockaddr*>(&addr), len) == 0) break;
    const int e = errno;
    close(s);
    if ((e != ECONNREFUSED && e != ENOENT && e != EAGAIN) || now_ms() > deadline)
      return fail(B200_ERR_STATE, "comm: rank %d cannot reach rank %d: %s", c->rank, peer, strerror(e));
    usleep(20000);                                    // the peer has not bound its socket yet
  }
  iovec iov{const_cast<Msg*>(&m), sizeof(Msg)};
  msghdr h{};
  h.msg_iov = &iov;
  h.msg_iovlen = 1;
  alignas(cmsghdr) char ctrl[CMSG_SPACE(2 * sizeof(int))];
  if (m.nfds > 0) {
    std::memset(ctrl, 0, sizeof(ctrl));
    h.msg_control = ctrl;
    h.msg_controllen = CMSG_SPACE(m.nfds * sizeof(int));
    cmsghdr* cm = CMSG_FIRSTHDR(&h);
    cm->cmsg_level = SOL_SOCKET;
    cm->cmsg_type = SCM_RIGHTS;
    cm->cmsg_len = CMSG_LEN(m.nfds * sizeof(int));
    std::memcpy(CMSG_DATA(cm), fds, m.nfds * sizeof(int));
  }
  const ssize_t n = sendmsg(s, &h, MSG_NOSIGNAL);
  const int e = errno;
  close(s);
  if (n != (ssize_t)sizeof(Msg)) return fail(B200_ERR_STATE, "comm: sendmsg to rank %d: %s", peer, strerror(e));
  return 0;
}

int recv_one(Comm* c, Stashed& out) {
  pollfd p{c->listen_fd, POLLIN, 0};
  const int pr = poll(&p, 1, c->timeout_ms);
  if (pr == 0) return fail(B200_ERR_STATE, "comm: rank %d waited %d ms for a peer message (a rank died or never started)", c->rank, c->timeout_ms);
  if (pr < 0) return fail(B200_ERR_STATE, "comm: poll(): %s", strerror(errno));
  const int s = accept4(c->listen_fd, nullptr, nullptr, SOCK_CLOEXEC);
  if (s < 0) return fail(B200_ERR_STATE, "comm: accept(): %s", strerror(errno));
  iovec iov{&out.m, sizeof(Msg)};
  msghdr h{};
  h.msg_iov = &iov;
  h.msg_iovlen = 1;
  alignas(cmsghdr) char ctrl[CMSG_SPACE(2 * sizeof(int))];
  h.msg_control = ctrl;
  h.msg_controllen = sizeof(ctrl);
  ssize_t n;
  do { n = recvmsg(s, &h, MSG_WAITALL | MSG_CMSG_CLOEXEC); } while (n < 0 && errno == EINTR);
  const int e = errno;
  close(s);
  if (n != (ssize_t)sizeof(Msg) || out.m.magic != kMagic) return fail(B200_ERR_STATE, "comm: short or foreign message (%zd bytes): %s", n, strerror(e));
  out.fds[0] = out.fds[1] = -1;
  int got = 0;
  for (cmsghdr* cm = CMSG_FIRSTHDR(&h); cm != nullptr; cm = CMSG_NXTHDR(&h, cm)) {
    if (cm->cmsg_level == SOL_SOCKET && cm->cmsg_type == SCM_RIGHTS) {
      const int k = (int)((cm->cmsg_len - CMSG_LEN(0)) / sizeof(int));
      for (int i = 0; i < k && got < 2; ++i) std::memcpy(&out.fds[got++], CMSG_DATA(cm) + i * sizeof(int), sizeof(int));
    }
  }
  if (got != out.m.nfds) return fail(B200_ERR_STATE, "comm: expected %d descriptors from rank %d, got %d", out.m.nfds, out.m.src, got);
  return 0;
}

// One all-to-all round: every rank sends `mine[p]` (+ fds) to every peer p and receives one message per peer.
// A fast peer may already be one round ahead; its message is stashed until that round is run here.
int exchange(Comm* c, const Msg* mine /*[world]*/, const int (*fds)[2] /*[world] or null*/, Stashed* in /*[world]*/) {
  const uint32_t seq = ++c->seq;
  for (int p = 0; p < c->world; ++p) {
    if (p == c->rank) continue;
    Msg m = mine[p];
    m.magic = kMagic; m.seq = seq; m.src = c->rank;
    const int rc = send_msg(c, p, m, fds ? fds[p] : nullptr);
    if (rc) return rc;
  }
  bool have[kMaxRanks] = {};
  int need = c->world - 1;
  for (size_t i = 0; i < c->stash.size();) {
    if (c->stash[i].m.seq == seq) {
      in[c->stash[i].m.src] = c->stash[i];
      have[c->stash[i].m.src] = true;
      --need;
      c->stash.erase(c->stash.begin() + i);
    } else {
      ++i;
    }
  }
  while (need > 0) {
    Stashed s;
    const int rc = recv_one(c, s);
    if (rc) return rc;
    if (s.m.src < 0 || s.m.src >= c->world || s.m.src == c->rank) return fail(B200_ERR_STATE, "comm: message from impossible rank %d", s.m.src);
    if (s.m.seq == seq && !have[s.m.src]) {
      in[s.m.src] = s;
      have[s.m.src] = true;
      --need;
    } else {
      c->stash.push_back(s);
    }
  }
  return 0;
}

int barrier(Comm* c, uint64_t word0, uint64_t* and_out) {
  Msg mine[kMaxRanks] = {};
  Stashed in[kMaxRanks];
  for (int p = 0; p < c->world; ++p) mine[p].word[0] = word0;
  const int rc = exchange(c, mine, nullptr, in);
  if (rc) return rc;
  uint64_t all = word0;
  for (int p = 0; p < c->world; ++p)
    if (p != c->rank) all &= in[p].m.word[0];
  if (and_out) *and_out = all;
  return 0;
}

int open_listener(Comm* c) {
  c->listen_fd = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
  if (c->listen_fd < 0) return fail(B200_ERR_STATE, "comm_init: socket(): %s", strerror(errno));
  sockaddr_un addr;
  const socklen_t len = make_addr(addr, c->tag, c->rank);
  if (bind(c->listen_fd, reinterpret_cast<sockaddr*>(&addr), len) != 0 || listen(c->listen_fd, 64) != 0) {
    const int e = errno;
    close(c->listen_fd);
    c->listen_fd = -1;
    return fail(B200_ERR_STATE, "comm_init: bind/listen: %s", strerror(e));
  }
  return 0;
}

size_t round_up(size_t n, size_t a) { return (n + a - 1) / a * a; }

void release_buffer(Comm* c, Buffer* b) {
  Driver& d = driver();
  if (b->mc_va) { d.MemUnmap(b->mc_va, b->bytes); d.MemAddressFree(b->mc_va, b->bytes); }
  if (b->mc_bound) d.MulticastUnbind(b->mc, c->cu_dev, 0, b->bytes);
  if (b->mc) d.MemRelease(b->mc);
  for (int p = 0; p < c->world; ++p) {
    if (b->va[p]) { d.MemUnmap(b->va[p], b->bytes); d.MemAddressFree(b->va[p], b->bytes); }
    if (p != c->rank && b->peer[p]) d.MemRelease(b->peer[p]);
  }
  if (b->own) d.MemRelease(b->own);
  delete b;
}

}  // namespace
}  // namespace b200

using namespace b200;

// Rendezvous: bind this rank's socket, then exchange device capabilities with every peer (which doubles as the
// "everyone is up" barrier).  `tag` must be the same string on every rank of the job and unique on the host.
extern "C" int b200_comm_init(int rank, int world, const char* tag, int timeout_ms, void** comm_out) {
  if (!comm_out || !tag) return fail(B200_ERR_ARG, "comm_init: null pointer");
  if (world < 1 || world > kMaxRanks || rank < 0 || rank >= world) return fail(B200_ERR_ARG, "comm_init: bad rank/world %d/%d", rank, world);
  Driver& d = driver();
  if (!d.ok) return fail(B200_ERR_DRIVER, "comm_init: CUDA virtual-memory driver API unavailable");
  B200_CUDA(cudaFree(nullptr));                      // make sure the primary context exists
  Comm* c = new Comm;
  c->rank = rank; c->world = world; c->tag = tag;
  if (timeout_ms > 0) c->timeout_ms = timeout_ms;
  B200_CUDA(cudaGetDevice(&c->device));
  B200_CU(d.DeviceGet(&c->cu_dev, c->device));
  int fd_ok = 0, mc_ok = 0;
  d.DeviceGetAttribute(&fd_ok, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, c->cu_dev);
  if (d.multicast_api) d.DeviceGetAttribute(&mc_ok, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, c->cu_dev);
  if (getenv("B200_NO_NVLS")) mc_ok = 0;
  if (const int lrc = open_listener(c)) { delete c; return lrc; }
  uint64_t all = 0;
  const uint64_t caps = (fd_ok ? 1u : 0u) | (mc_ok ? 2u : 0u);
  const int rc = world > 1 ? barrier(c, caps, &all) : 0;
  if (world == 1) all = caps;
  if (rc) { close(c->listen_fd); delete c; return rc; }
  if (!(all & 1u)) { close(c->listen_fd); delete c; return fail(B200_ERR_UNSUPPORTED, "comm_init: POSIX-FD shareable allocations unsupported on some rank"); }
  c->multicast = (all & 2u) != 0 && world > 1;
  *comm_out = c;
  return 0;
}

// Control-plane self-test, no GPU needed (tests/test_cpu_distributed.py runs it with three processes): rendezvous, an
// AND-agreement round, one descriptor per rank passed to every peer as SCM_RIGHTS (a memfd holding a known
// word, read back through the received descriptor), and two rounds with one rank lagging, so that a message of
// the next round can arrive before the current one is complete (the stash).
extern "C" int b200_comm_selftest(int rank, int world, const char* tag, int timeout_ms) {
  if (!tag || world < 1 || world > kMaxRanks || rank < 0 || rank >= world) return fail(B200_ERR_ARG, "comm_selftest: bad argument");
  Comm comm;
  Comm* c = &comm;
  c->rank = rank; c->world = world; c->tag = tag;
  if (timeout_ms > 0) c->timeout_ms = timeout_ms;
  if (const int rc = open_listener(c)) return rc;
  auto done = [&](int code) {
    for (auto& s : c->stash) for (int k = 0; k < 2; ++k) if (s.fds[k] >= 0) close(s.fds[k]);
    close(c->listen_fd);
    return code;
  };
  if (world == 1) return done(0);
  uint64_t all = 0;
  int rc = barrier(c, ~(1ull << rank), &all);
  if (rc) return done(rc);
  if (all != ~((1ull << world) - 1)) return done(fail(B200_ERR_STATE, "comm_selftest: agreement word %llx", (unsigned long long)all));
  const int mfd = memfd_create("b200seq.selftest", MFD_CLOEXEC);
  if (mfd < 0) return done(fail(B200_ERR_STATE, "comm_selftest: memfd_create: %s", strerror(errno)));
  const uint64_t token = 0xB200000000000000ull | (uint64_t)(rank * 1000 + 7);
  if (write(mfd, &token, sizeof(token)) != (ssize_t)sizeof(token)) { close(mfd); return done(fail(B200_ERR_STATE, "comm_selftest: write")); }
  Msg mine[kMaxRanks] = {};
  int fds[kMaxRanks][2];
  Stashed in[kMaxRanks];
  for (int p = 0; p < world; ++p) { mine[p].word[0] = (uint64_t)(rank * 16 + p); mine[p].nfds = 1; fds[p][0] = mfd; fds[p][1] = -1; }
  rc = exchange(c, mine, fds, in);
  close(mfd);
  if (rc) return done(rc);
  for (int p = 0; p < world; ++p) {
    if (p == rank) continue;
    uint64_t got = 0;
    const ssize_t n = pread(in[p].fds[0], &got, sizeof(got), 0);
    close(in[p].fds[0]);
    if (in[p].m.word[0] != (uint64_t)(p * 16 + rank)) return done(fail(B200_ERR_STATE, "comm_selftest: wrong payload from rank %d", p));
    if (n != (ssize_t)sizeof(got) || got != (0xB200000000000000ull | (uint64_t)(p * 1000 + 7)))
      return done(fail(B200_ERR_STATE, "comm_selftest: descriptor from rank %d does not read back", p));
  }
  // the last rank lags: a peer that completes a round first may deliver its next message early (stashed here)
  if (rank == world - 1) usleep(200000);
  for (int round = 0; round < 2; ++round) {
    rc = barrier(c, (uint64_t)(round + 1), &all);
    if (rc) return done(rc);
    if (all != (uint64_t)(round + 1)) return done(fail(B200_ERR_STATE, "comm_selftest: round %d word %llx", round, (unsigned long long)all));
  }
  return done(0);
}

extern "C" int b200_comm_multicast_supported(void* comm) { return comm && static_cast<Comm*>(comm)->multicast ? 1 : 0; }

extern "C" int b200_comm_barrier(void* comm) {
  Comm* c = static_cast<Comm*>(comm);
  if (!c) return fail(B200_ERR_ARG, "comm_barrier: null communicator");
  return c->world > 1 ? barrier(c, 1, nullptr) : 0;
}

// Collective: every rank calls it with the same `bytes`.  On return out->local is this rank's zero-filled buffer,
// out->peer[r] rank r's buffer as mapped here (peer[rank] == local), out->multicast the NVLS address or null.
extern "C" int b200_comm_alloc(void* comm, long long bytes, int want_multicast, b200_comm_buffer* out) {
  Comm* c = static_cast<Comm*>(comm);
  if (!c || !out || bytes <= 0) return fail(B200_ERR_ARG, "comm_alloc: bad argument");
  Driver& d = driver();
  const bool mc = want_multicast && c->multicast;
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = c->device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  B200_CU(d.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  CUmulticastObjectProp mprop{};
  mprop.numDevices = (unsigned)c->world;
  mprop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  if (mc) {
    size_t mg = 0;
    mprop.size = round_up((size_t)bytes, gran);
    B200_CU(d.MulticastGetGranularity(&mg, &mprop, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    if (mg > gran) gran = mg;
  }
  Buffer* b = new Buffer;
  b->bytes = round_up((size_t)bytes, gran);
  mprop.size = b->bytes;
  int rc = 0;
  int own_fd = -1, mc_fd = -1;
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = c->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  auto bail = [&](int code) { if (own_fd >= 0) close(own_fd); if (mc_fd >= 0) close(mc_fd); release_buffer(c, b); return code; };
#define TRY_CU(expr) do { CUresult _r = (expr); if (_r != CUDA_SUCCESS) return bail(fail(B200_ERR_DRIVER, "%s: %s (%d)", #expr, cu_str(_r), (int)_r)); } while (0)
  TRY_CU(d.MemCreate(&b->own, b->bytes, &prop, 0));
  TRY_CU(d.MemAddressReserve(&b->va[c->rank], b->bytes, gran, 0, 0));
  TRY_CU(d.MemMap(b->va[c->rank], b->bytes, 0, b->own, 0));
  TRY_CU(d.MemSetAccess(b->va[c->rank], b->bytes, &acc, 1));
  if (cudaMemset(reinterpret_cast<void*>(b->va[c->rank]), 0, b->bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess)
    return bail(fail(B200_ERR_DRIVER, "comm_alloc: clearing the new buffer failed"));
  if (c->world > 1) {
    TRY_CU(d.MemExport(&own_fd, b->own, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    bool mc_here = mc;
    if (mc && c->rank == 0) {
      if (d.MulticastCreate(&b->mc, &mprop) != CUDA_SUCCESS ||
          d.MemExport(&mc_fd, b->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) != CUDA_SUCCESS) {
        if (b->mc) { d.MemRelease(b->mc); b->mc = 0; }
        mc_fd = -1;
        mc_here = false;                             // the peers learn it from word[1] and skip the multicast steps too
      }
    }
    // round 1: memory handles (and rank 0's multicast handle)
    Msg mine[kMaxRanks] = {};
    int fds[kMaxRanks][2];
    Stashed in[kMaxRanks];
    for (int p = 0; p < c->world; ++p) {
      mine[p].word[0] = b->bytes;
      mine[p].word[1] = mc_here ? 1 : 0;
      fds[p][0] = own_fd; fds[p][1] = mc_fd;
      mine[p].nfds = (c->rank == 0 && mc_here) ? 2 : 1;
    }
    rc = exchange(c, mine, fds, in);
    close(own_fd); own_fd = -1;
    if (mc_fd >= 0) { close(mc_fd); mc_fd = -1; }
    if (rc) return bail(rc);
    if (c->rank != 0) mc_here = mc && in[0].m.word[1] == 1 && in[0].m.nfds == 2;
    int first_err = 0;
    for (int p = 0; p < c->world; ++p) {
      if (p == c->rank) continue;
      if (in[p].m.word[0] != b->bytes && !first_err) first_err = fail(B200_ERR_STATE, "comm_alloc: rank %d asked for a different size", p);
      if (!first_err) {
        CUresult r = d.MemImport(&b->peer[p], reinterpret_cast<void*>(static_cast<uintptr_t>(in[p].fds[0])), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
        if (r == CUDA_SUCCESS) r = d.MemAddressReserve(&b->va[p], b->bytes, gran, 0, 0);
        if (r == CUDA_SUCCESS) r = d.MemMap(b->va[p], b->bytes, 0, b->peer[p], 0);
        if (r == CUDA_SUCCESS) r = d.MemSetAccess(b->va[p], b->bytes, &acc, 1);
        if (r != CUDA_SUCCESS) first_err = fail(B200_ERR_DRIVER, "comm_alloc: mapping rank %d's buffer: %s (%d)", p, cu_str(r), (int)r);
      }
      if (p == 0 && mc_here && c->rank != 0 && !first_err) {
        if (d.MemImport(&b->mc, reinterpret_cast<void*>(static_cast<uintptr_t>(in[0].fds[1])), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) != CUDA_SUCCESS) {
          b->mc = 0;
        }
      }
      for (int k = 0; k < 2; ++k) if (in[p].fds[k] >= 0) close(in[p].fds[k]);
    }
    // round 2: did every rank map everything / obtain the multicast object?  (agreement before anyone binds)
    uint64_t all = 0;
    const uint64_t st = (first_err ? 0u : 1u) | ((mc_here && b->mc) ? 2u : 0u);
    rc = barrier(c, st, &all);
    if (rc) return bail(rc);
    if (!(all & 1u)) return bail(first_err ? first_err : fail(B200_ERR_STATE, "comm_alloc: a peer failed to map the buffers"));
    bool use_mc = (all & 2u) != 0;
    uint64_t mc_ok = 1;
    if (use_mc) {
      // add -> (driver blocks the bind until all devices are added) -> bind -> map the multicast range
      CUresult r = d.MulticastAddDevice(b->mc, c->cu_dev);
      if (r == CUDA_SUCCESS) {
        uint64_t added = 0;
        rc = barrier(c, 1, &added);
        if (rc) return bail(rc);
        if (added & 1u) {                             // every rank added its device (a bind would otherwise block for ever)
          r = d.MulticastBindMem(b->mc, 0, b->own, 0, b->bytes, 0);
          if (r == CUDA_SUCCESS) b->mc_bound = true;
        } else {
          r = CUDA_ERROR_NOT_READY;
        }
      } else {
        rc = barrier(c, 0, nullptr);                  // keep the round count equal on every rank
        if (rc) return bail(rc);
      }
      if (r == CUDA_SUCCESS) r = d.MemAddressReserve(&b->mc_va, b->bytes, gran, 0, 0);
      if (r == CUDA_SUCCESS) r = d.MemMap(b->mc_va, b->bytes, 0, b->mc, 0);
      if (r == CUDA_SUCCESS) r = d.MemSetAccess(b->mc_va, b->bytes, &acc, 1);
      if (r != CUDA_SUCCESS) {
        fail(B200_ERR_DRIVER, "comm_alloc: multicast set-up: %s (%d)", cu_str(r), (int)r);
        mc_ok = 0;
      }
    } else if (b->mc) {
      d.MemRelease(b->mc);
      b->mc = 0;
    }
    // round 3: binds are complete everywhere (and agreed to have worked) before the first multimem instruction
    uint64_t all_mc = 0;
    rc = barrier(c, mc_ok, &all_mc);
    if (rc) return bail(rc);
    if (use_mc && !(all_mc & 1u)) {                   // someone could not bind: nobody uses the multicast mapping
      if (b->mc_va) { d.MemUnmap(b->mc_va, b->bytes); d.MemAddressFree(b->mc_va, b->bytes); b->mc_va = 0; }
    }
  }
#undef TRY_CU
  c->buffers.push_back(b);
  std::memset(out, 0, sizeof(*out));
  out->handle = b;
  out->bytes = (long long)b->bytes;
  out->local = reinterpret_cast<void*>(b->va[c->rank]);
  for (int p = 0; p < c->world; ++p) out->peer[p] = reinterpret_cast<void*>(b->va[p]);
  out->multicast = reinterpret_cast<void*>(b->mc_va);
  return 0;
}

// Collective in effect (peers must not touch the buffer afterwards); callers barrier first.
extern "C" int b200_comm_free(void* comm, void* handle) {
  Comm* c = static_cast<Comm*>(comm);
  if (!c || !handle) return 0;
  for (size_t i = 0; i < c->buffers.size(); ++i) {
    if (c->buffers[i] == handle) {
      cudaDeviceSynchronize();
      release_buffer(c, c->buffers[i]);
      c->buffers.erase(c->buffers.begin() + i);
      return 0;
    }
  }
  return fail(B200_ERR_ARG, "comm_free: unknown buffer");
}

extern "C" int b200_comm_destroy(void* comm) {
  Comm* c = static_cast<Comm*>(comm);
  if (!c) return 0;
  cudaDeviceSynchronize();
  for (Buffer* b : c->buffers) release_buffer(c, b);
  for (auto& s : c->stash) for (int k = 0; k < 2; ++k) if (s.fds[k] >= 0) close(s.fds[k]);
  if (c->listen_fd >= 0) close(c->listen_fd);
  delete c;
  return 0;
}
