// See vmm_arena.h.  CUDA driver API through dlopen; fd passing through abstract unix datagram sockets.
#include "vmm_arena.h"

#include <cuda.h>
#include <dlfcn.h>
#include <errno.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/time.h>
#include <sys/un.h>
#include <unistd.h>

#include <mutex>
#include <sstream>

namespace b200ddp {
namespace {

struct Driver {
  void* lib = nullptr;
  std::string error;
#define B200_DRV(name) decltype(&::name) name = nullptr;
  B200_DRV(cuInit)
  B200_DRV(cuGetErrorString)
  B200_DRV(cuDeviceGet)
  B200_DRV(cuDeviceGetAttribute)
  B200_DRV(cuCtxGetCurrent)
  B200_DRV(cuCtxSetCurrent)
  B200_DRV(cuDevicePrimaryCtxRetain)
  B200_DRV(cuMemGetAllocationGranularity)
  B200_DRV(cuMemCreate)
  B200_DRV(cuMemRelease)
  B200_DRV(cuMemAddressReserve)
  B200_DRV(cuMemAddressFree)
  B200_DRV(cuMemMap)
  B200_DRV(cuMemUnmap)
  B200_DRV(cuMemSetAccess)
  B200_DRV(cuMemExportToShareableHandle)
  B200_DRV(cuMemImportFromShareableHandle)
  B200_DRV(cuMulticastCreate)
  B200_DRV(cuMulticastAddDevice)
  B200_DRV(cuMulticastBindMem)
  B200_DRV(cuMulticastGetGranularity)
  B200_DRV(cuMulticastUnbind)
#undef B200_DRV
};

// cuda.h renames a few entry points to versioned symbols (cuDevicePrimaryCtxRetain, cuMemAddressReserve ... are
// plain; cuCtx* are plain): stringify AFTER macro expansion so dlsym asks for the name the header means.
#define B200_STR2(x) #x
#define B200_STR(x) B200_STR2(x)

Driver& driver() {
  static Driver d;
  static std::once_flag once;
  std::call_once(once, [] {
    d.lib = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
    if (d.lib == nullptr) {
      d.error = std::string("libcuda.so.1 not loadable: ") + dlerror();
      return;
    }
#define B200_LOAD(name)                                                             \
  d.name = reinterpret_cast<decltype(d.name)>(dlsym(d.lib, B200_STR(name)));         \
  if (d.name == nullptr && d.error.empty()) d.error = "libcuda lacks " B200_STR(name);
    B200_LOAD(cuInit)
    B200_LOAD(cuGetErrorString)
    B200_LOAD(cuDeviceGet)
    B200_LOAD(cuDeviceGetAttribute)
    B200_LOAD(cuCtxGetCurrent)
    B200_LOAD(cuCtxSetCurrent)
    B200_LOAD(cuDevicePrimaryCtxRetain)
    B200_LOAD(cuMemGetAllocationGranularity)
    B200_LOAD(cuMemCreate)
    B200_LOAD(cuMemRelease)
    B200_LOAD(cuMemAddressReserve)
    B200_LOAD(cuMemAddressFree)
    B200_LOAD(cuMemMap)
    B200_LOAD(cuMemUnmap)
    B200_LOAD(cuMemSetAccess)
    B200_LOAD(cuMemExportToShareableHandle)
    B200_LOAD(cuMemImportFromShareableHandle)
    B200_LOAD(cuMulticastCreate)
    B200_LOAD(cuMulticastAddDevice)
    B200_LOAD(cuMulticastBindMem)
    B200_LOAD(cuMulticastGetGranularity)
    B200_LOAD(cuMulticastUnbind)
#undef B200_LOAD
  });
  return d;
}

std::string cu_err(const char* what, CUresult rc) {
  const char* s = nullptr;
  Driver& d = driver();
  if (d.cuGetErrorString) d.cuGetErrorString(rc, &s);
  std::ostringstream os;
  os << what << " failed: " << (s ? s : "unknown") << " (CUresult " << static_cast<int>(rc) << ")";
  return os.str();
}

#define B200_CU(call)                                         \
  do {                                                        \
    CUresult rc_ = d.call;                                    \
    if (rc_ != CUDA_SUCCESS) return cu_err(#call, rc_);       \
  } while (0)

std::string ensure_context(Driver& d, int device) {
  if (!d.error.empty()) return d.error;
  B200_CU(cuInit(0));
  CUcontext ctx = nullptr;
  B200_CU(cuCtxGetCurrent(&ctx));
  if (ctx == nullptr) {
    CUdevice dev;
    B200_CU(cuDeviceGet(&dev, device));
    B200_CU(cuDevicePrimaryCtxRetain(&ctx, dev));
    B200_CU(cuCtxSetCurrent(ctx));
  }
  return "";
}

CUmemAccessDesc rw_access(int device) {
  CUmemAccessDesc a{};
  a.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  a.location.id = device;
  a.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  return a;
}

struct Msg {
  int32_t kind, rank;
};

}  // namespace

VmmArena::VmmArena(int device, size_t nbytes, int rank, int world, std::string session)
    : device_(device), rank_(rank), world_(world), nbytes_(nbytes), session_(std::move(session)) {
  peer_fd_.assign(world, -1);
  peer_va_.assign(world, 0);
  peer_phys_.assign(world, 0);
}

VmmArena::~VmmArena() { close(); }

std::string VmmArena::sock_name(int r) const { return "b200ddp/" + session_ + "/" + std::to_string(r); }

std::string VmmArena::alloc_local() {
  Driver& d = driver();
  std::string e = ensure_context(d, device_);
  if (!e.empty()) return e;
  CUdevice dev;
  B200_CU(cuDeviceGet(&dev, device_));
  int vmm = 0, fd_ok = 0, mc = 0;
  B200_CU(cuDeviceGetAttribute(&vmm, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev));
  B200_CU(cuDeviceGetAttribute(&fd_ok, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev));
  if (!vmm || !fd_ok) return "device lacks the virtual-memory API or fd-exportable allocations";
  if (d.cuDeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) mc = 0;
  mc_supported_ = mc != 0 && world_ > 1;

  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device_;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  B200_CU(cuMemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (mc_supported_) {
    CUmulticastObjectProp mp{};
    mp.numDevices = static_cast<unsigned>(world_);
    mp.size = nbytes_;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (d.cuMulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > gran)
      gran = mg;
  }
  gran_ = gran ? gran : (2u << 20);
  nbytes_ = (nbytes_ + gran_ - 1) / gran_ * gran_;

  CUmemGenericAllocationHandle h = 0;
  B200_CU(cuMemCreate(&h, nbytes_, &prop, 0));
  phys_ = h;
  CUdeviceptr va = 0;
  B200_CU(cuMemAddressReserve(&va, nbytes_, gran_, 0, 0));
  base_ = va;
  B200_CU(cuMemMap(va, nbytes_, 0, h, 0));
  CUmemAccessDesc acc = rw_access(device_);
  B200_CU(cuMemSetAccess(va, nbytes_, &acc, 1));
  peer_va_[rank_] = base_;
  int fd = -1;
  B200_CU(cuMemExportToShareableHandle(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  my_fd_ = fd;

  // receiving socket (abstract namespace: no file to clean up, dies with the process)
  sock_ = ::socket(AF_UNIX, SOCK_DGRAM | SOCK_CLOEXEC, 0);
  if (sock_ < 0) return std::string("socket(): ") + strerror(errno);
  sockaddr_un addr{};
  addr.sun_family = AF_UNIX;
  const std::string name = sock_name(rank_);
  if (name.size() + 2 > sizeof(addr.sun_path)) return "session name too long for a unix socket";
  memcpy(addr.sun_path + 1, name.data(), name.size());
  const socklen_t len = static_cast<socklen_t>(offsetof(sockaddr_un, sun_path) + 1 + name.size());
  if (::bind(sock_, reinterpret_cast<sockaddr*>(&addr), len) != 0) return std::string("bind(): ") + strerror(errno);
  timeval tv{60, 0};
  setsockopt(sock_, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
  return "";
}

std::string VmmArena::send_fd(int to_rank, int kind, int fd) {
  sockaddr_un addr{};
  addr.sun_family = AF_UNIX;
  const std::string name = sock_name(to_rank);
  memcpy(addr.sun_path + 1, name.data(), name.size());
  const socklen_t len = static_cast<socklen_t>(offsetof(sockaddr_un, sun_path) + 1 + name.size());
  Msg m{kind, rank_};
  iovec iov{&m, sizeof(m)};
  alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))] = {};
  msghdr msg{};
  msg.msg_name = &addr;
  msg.msg_namelen = len;
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
    if (::sendmsg(sock_, &msg, 0) >= 0) return "";
    if (errno != ECONNREFUSED && errno != ENOENT && errno != EAGAIN) break;
    usleep(50 * 1000);       // peer's socket not bound yet
  }
  return std::string("sendmsg(fd) to rank ") + std::to_string(to_rank) + ": " + strerror(errno);
}

std::string VmmArena::recv_fd(int kind, int* from_rank, int* fd, int alt_kind, int* got_kind) {
  for (size_t i = 0; i < stash_.size(); ++i) {
    if (stash_[i].kind == kind || stash_[i].kind == alt_kind) {
      *from_rank = stash_[i].from;
      *fd = stash_[i].fd;
      if (got_kind) *got_kind = stash_[i].kind;
      stash_.erase(stash_.begin() + static_cast<long>(i));
      return "";
    }
  }
  for (;;) {
    Msg m{};
    iovec iov{&m, sizeof(m)};
    alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))] = {};
    msghdr msg{};
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof(ctrl);
    const ssize_t n = ::recvmsg(sock_, &msg, MSG_CMSG_CLOEXEC);
    if (n < 0) return std::string("recvmsg(fd): ") + strerror(errno);
    int got = -1;
    for (cmsghdr* c = CMSG_FIRSTHDR(&msg); c != nullptr; c = CMSG_NXTHDR(&msg, c))
      if (c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS) memcpy(&got, CMSG_DATA(c), sizeof(int));
    if (n != static_cast<ssize_t>(sizeof(m)) || got < 0) return "malformed fd message";
    if (m.kind == kind || m.kind == alt_kind) {
      *from_rank = m.rank;
      *fd = got;
      if (got_kind) *got_kind = m.kind;
      return "";
    }
    stash_.push_back(Stash{m.kind, m.rank, got});
  }
}

std::string VmmArena::exchange() {
  for (int r = 0; r < world_; ++r) {
    if (r == rank_) continue;
    std::string e = send_fd(r, 1, my_fd_);
    if (!e.empty()) return e;
  }
  for (int i = 0; i + 1 < world_; ++i) {
    int from = -1, fd = -1;
    std::string e = recv_fd(1, &from, &fd);
    if (!e.empty()) return e;
    if (from < 0 || from >= world_ || from == rank_ || peer_fd_[from] >= 0) return "unexpected sender in fd exchange";
    peer_fd_[from] = fd;
  }
  return "";
}

std::string VmmArena::map_peers() {
  Driver& d = driver();
  CUmemAccessDesc acc = rw_access(device_);
  for (int r = 0; r < world_; ++r) {
    if (r == rank_) continue;
    CUmemGenericAllocationHandle h = 0;
    B200_CU(cuMemImportFromShareableHandle(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(peer_fd_[r])),
                                           CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    peer_phys_[r] = h;
    ::close(peer_fd_[r]);
    peer_fd_[r] = -1;
    CUdeviceptr va = 0;
    B200_CU(cuMemAddressReserve(&va, nbytes_, gran_, 0, 0));
    peer_va_[r] = va;
    B200_CU(cuMemMap(va, nbytes_, 0, h, 0));
    B200_CU(cuMemSetAccess(va, nbytes_, &acc, 1));
  }
  return "";
}

std::string VmmArena::multicast_create() {
  if (!mc_supported_) return "unsupported";
  Driver& d = driver();
  CUmemGenericAllocationHandle mc = 0;
  if (rank_ == 0) {
    CUmulticastObjectProp mp{};
    mp.numDevices = static_cast<unsigned>(world_);
    mp.size = nbytes_;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CUresult rc = d.cuMulticastCreate(&mc, &mp);
    int fd = -1;
    if (rc == CUDA_SUCCESS) rc = d.cuMemExportToShareableHandle(&fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    // peers are waiting in recv_fd: always answer, with a dummy descriptor on failure, so nobody sits out the timeout
    const int send = rc == CUDA_SUCCESS ? fd : my_fd_;
    const int kind = rc == CUDA_SUCCESS ? 2 : 3;
    for (int r = 1; r < world_; ++r) send_fd(r, kind, send);
    if (fd >= 0) ::close(fd);
    if (rc != CUDA_SUCCESS) return cu_err("cuMulticastCreate/export", rc);
  } else {
    // kind 2 = the multicast object, kind 3 = rank 0 could not create one
    int from = -1, fd = -1, kind = 0;
    std::string e = recv_fd(2, &from, &fd, 3, &kind);
    if (!e.empty()) return e;
    if (kind == 3) {
      ::close(fd);
      return "rank 0 could not create the multicast object";
    }
    CUresult rc = d.cuMemImportFromShareableHandle(&mc, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                                   CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    ::close(fd);
    if (rc != CUDA_SUCCESS) return cu_err("cuMemImportFromShareableHandle(multicast)", rc);
  }
  mc_handle_ = mc;
  CUdevice dev;
  B200_CU(cuDeviceGet(&dev, device_));
  B200_CU(cuMulticastAddDevice(mc, dev));
  return "";
}

std::string VmmArena::multicast_bind() {
  if (!mc_supported_ || mc_handle_ == 0) return "unsupported";
  Driver& d = driver();
  B200_CU(cuMulticastBindMem(mc_handle_, 0, phys_, 0, nbytes_, 0));
  mc_bound_ = true;
  CUdeviceptr va = 0;
  B200_CU(cuMemAddressReserve(&va, nbytes_, gran_, 0, 0));
  B200_CU(cuMemMap(va, nbytes_, 0, mc_handle_, 0));
  CUmemAccessDesc acc = rw_access(device_);
  B200_CU(cuMemSetAccess(va, nbytes_, &acc, 1));
  mc_va_ = va;
  return "";
}

void VmmArena::close() {
  Driver& d = driver();
  if (d.error.empty() && d.cuMemUnmap != nullptr) {
    if (mc_va_) {
      d.cuMemUnmap(mc_va_, nbytes_);
      d.cuMemAddressFree(mc_va_, nbytes_);
      mc_va_ = 0;
    }
    if (mc_handle_) {
      if (mc_bound_) {
        CUdevice dev;
        if (d.cuDeviceGet(&dev, device_) == CUDA_SUCCESS) d.cuMulticastUnbind(mc_handle_, dev, 0, nbytes_);
        mc_bound_ = false;
      }
      d.cuMemRelease(mc_handle_);
      mc_handle_ = 0;
    }
    for (int r = 0; r < world_; ++r) {
      if (r == rank_) continue;
      if (peer_va_[r]) {
        d.cuMemUnmap(peer_va_[r], nbytes_);
        d.cuMemAddressFree(peer_va_[r], nbytes_);
        peer_va_[r] = 0;
      }
      if (peer_phys_[r]) {
        d.cuMemRelease(peer_phys_[r]);
        peer_phys_[r] = 0;
      }
    }
    if (base_) {
      d.cuMemUnmap(base_, nbytes_);
      d.cuMemAddressFree(base_, nbytes_);
      base_ = 0;
    }
    if (phys_) {
      d.cuMemRelease(phys_);
      phys_ = 0;
    }
  }
  for (int& fd : peer_fd_)
    if (fd >= 0) {
      ::close(fd);
      fd = -1;
    }
  for (auto& s : stash_) ::close(s.fd);
  stash_.clear();
  if (my_fd_ >= 0) {
    ::close(my_fd_);
    my_fd_ = -1;
  }
  if (sock_ >= 0) {
    ::close(sock_);
    sock_ = -1;
  }
}

}  // namespace b200ddp
