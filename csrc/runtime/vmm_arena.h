// Symmetric arena on the CUDA virtual-memory API, owned by this framework (host side, C++).
//
// Role in the reference stack: /root/reference/dpp.py:21 asks for the "nccl" backend, whose
// communicator sets up peer-mapped buffers and -- on NVSwitch systems -- an NVLink multicast
// (NVLS) object behind every collective.  Here that plumbing is
// explicit and ours: one physical allocation per rank (cuMemCreate), exported as a POSIX file
// descriptor, handed to every peer over a unix-domain socket (SCM_RIGHTS), imported and mapped
// by every peer (plain ld/st over NVLink), plus one multicast object that all ranks bind their
// allocation to, mapped a second time so that `multimem.*` instructions reduce / broadcast
// inside the switch.  No NCCL, no torch symmetric-memory allocator.
//
// The driver API is resolved with dlopen("libcuda.so.1") at first use, so the extension still
// imports on a machine without a GPU driver.
//
// Every stage returns "" on success or a message; nothing throws in the middle of the
// multi-rank sequence -- the caller gathers the per-rank outcomes and decides collectively
// (fall back to CUDA IPC, or to the c10d backend) so that no rank is left inside a collective.
#pragma once

#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace b200ddp {

class VmmArena {
 public:
  VmmArena(int device, size_t nbytes, int rank, int world, std::string session);
  ~VmmArena();

  // stage 1 (local): physical allocation + local mapping + exported fd + bound socket
  std::string alloc_local();
  // stage 2 (after every rank finished stage 1): send my fd to every peer, receive theirs
  std::string exchange();
  // stage 3 (local): import + map every peer's allocation
  std::string map_peers();
  // stage 4 (after stage 3 everywhere): rank 0 creates the multicast object and ships its fd; every rank
  // imports it and adds its device.  Returns "unsupported" when the device has no NVLink multicast.
  std::string multicast_create();
  // stage 5 (after stage 4 everywhere): bind the local allocation and map the multicast address
  std::string multicast_bind();
  void close();

  bool multicast_supported() const { return mc_supported_; }
  uint64_t base() const { return base_; }
  uint64_t peer_base(int r) const { return peer_va_.at(r); }
  uint64_t multicast_base() const { return mc_va_; }
  size_t nbytes() const { return nbytes_; }
  size_t granularity() const { return gran_; }

 private:
  std::string sock_name(int r) const;
  std::string send_fd(int to_rank, int kind, int fd);
  std::string recv_fd(int kind, int* from_rank, int* fd, int alt_kind = -1, int* got_kind = nullptr);

  int device_, rank_, world_;
  size_t nbytes_, gran_ = 0;
  std::string session_;
  bool mc_supported_ = false;
  int sock_ = -1;
  int my_fd_ = -1;
  std::vector<int> peer_fd_;
  uint64_t base_ = 0, mc_va_ = 0;
  std::vector<uint64_t> peer_va_;
  unsigned long long phys_ = 0, mc_handle_ = 0;
  std::vector<unsigned long long> peer_phys_;
  bool mc_bound_ = false;
  struct Stash { int kind, from, fd; };
  std::vector<Stash> stash_;
};

}  // namespace b200ddp
