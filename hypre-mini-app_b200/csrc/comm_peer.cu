// comm_peer.cu -- NVLink peer-memory arena for the halo exchange of the distributed SpMV ("one halo gather
// per SpMV ... overlapped with the diag SpMV over NVLink peer access").
//
// One process per GPU.  At communicator start every rank cudaMalloc's one arena, exports it with cudaIpcGetMemHandle,
// all-gathers the handles and maps every peer's arena (cudaIpcOpenMemHandle => loads/stores travel over NVLink /
// NVSwitch).  A ParCSR halo plan then owns a region of its rank's arena:
//     [ one 64-bit sequence flag per sending neighbour | halo buffer 0 | halo buffer 1 ]
// The sender's pack kernel stores x[send_idx] straight into the receiver's halo buffer (buffer = seq & 1) and the last
// block to finish publishes the sequence number in the receiver's flag with a system-scope release store; the receiver
// waits on its flags (system-scope acquire loads) right before the offd correction.  No NCCL call, no extra stream and
// no host work per exchange.  Two buffers suffice: a rank can be at most one exchange ahead of a neighbour, because it
// cannot finish exchange k of a plan before that neighbour has sent its part of k (parcsr.cu, halo_exchange_begin).
//
// Regions are handed out by a first-fit allocator on the host.  Reuse is safe without further synchronisation because
// regions are only allocated inside build_partition_and_halo() after a collective that every rank enters with all its
// earlier exchanges enqueued before it (NCCL is stream-ordered), and released when the owning plan dies.
#include "common.h"
#include <algorithm>
#include <cstring>
#include <map>

namespace b200 {

void allgather_dev(const void *in, void *out, size_t bytes);
void allreduce_max_host_i64(i64 *buf, int count);
bool shm_active();

namespace {

struct PeerArena {
  bool active = false;
  char *base = nullptr;
  size_t bytes = 0;
  std::vector<char *> peer;              // mapped base of every rank's arena (own rank: base)
  std::map<size_t, size_t> free_blocks;  // offset -> size
  std::map<size_t, size_t> used_blocks;
} g_arena;

constexpr size_t ALIGN = 256;

}  // namespace

bool peer_active() { return g_arena.active; }
char *peer_base(int rank) { return g_arena.active ? g_arena.peer[rank] : nullptr; }

// Collective over the NCCL communicator; called once by comm_init.  Leaves the arena inactive (NCCL halo path) when
// the ranks do not all own a distinct device with peer access to every other one, or when B200_HALO=nccl.
void peer_init() {
  Context &c = ctx();
  if (g_arena.active || c.nranks == 1 || shm_active()) return;
  const char *mode = getenv("B200_HALO");
  i64 ok = !(mode && strcmp(mode, "nccl") == 0);
  size_t mb = 256;
  if (const char *s = getenv("B200_PEER_ARENA_MB")) if (atoi(s) > 0) mb = (size_t)atoi(s);
  char *base = nullptr;
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  if (ok) {
    if (cudaMalloc(&base, mb << 20) != cudaSuccess) { cudaGetLastError(); ok = 0; base = nullptr; }
    else if (cudaIpcGetMemHandle(&mine, base) != cudaSuccess) { cudaGetLastError(); ok = 0; }
  }
  // gather {ok, device ordinal (PCI bus id hash is not needed on one box: the ordinal is global under torchrun), handle}
  struct Rec { i64 ok; i64 dev; cudaIpcMemHandle_t h; };
  Rec r;
  r.ok = ok;
  char bus[32] = {0};
  cudaDeviceGetPCIBusId(bus, sizeof(bus), c.device);
  i64 hsh = 1469598103934665603LL;
  for (char ch : bus) hsh = (hsh ^ (unsigned char)ch) * 1099511628211LL;
  r.dev = hsh;
  r.h = mine;
  DVec<char> din(sizeof(Rec)), dout(sizeof(Rec) * (size_t)c.nranks);
  B200_CUDA(cudaMemcpyAsync(din.p, &r, sizeof(Rec), cudaMemcpyHostToDevice, stream()));
  allgather_dev(din.p, dout.p, sizeof(Rec));
  std::vector<Rec> all((size_t)c.nranks);
  B200_CUDA(cudaMemcpyAsync(all.data(), dout.p, sizeof(Rec) * (size_t)c.nranks, cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
  for (int k = 0; k < c.nranks; ++k) {
    if (!all[k].ok) ok = 0;
    for (int j = 0; j < k; ++j) if (all[j].dev == all[k].dev) ok = 0;   // two ranks on one device: keep the NCCL path
  }
  std::vector<char *> peer((size_t)c.nranks, nullptr);
  if (ok) {
    for (int k = 0; k < c.nranks && ok; ++k) {
      if (k == c.rank) { peer[k] = base; continue; }
      void *p = nullptr;
      if (cudaIpcOpenMemHandle(&p, all[k].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = 0; }
      peer[k] = (char *)p;
    }
  }
  i64 bad = ok ? 0 : 1;
  allreduce_max_host_i64(&bad, 1);   // unanimous decision
  if (bad) {
    for (int k = 0; k < c.nranks; ++k) if (peer[k] && k != c.rank) cudaIpcCloseMemHandle(peer[k]);
    if (base) cudaFree(base);
    cudaGetLastError();
    if (c.rank == 0 && !(mode && strcmp(mode, "nccl") == 0)) fprintf(stderr, "[b200hypre] peer-memory halo exchange unavailable, using NCCL send/recv\n");
    return;
  }
  B200_CUDA(cudaMemset(base, 0, mb << 20));
  g_arena.base = base;
  g_arena.bytes = mb << 20;
  g_arena.peer = peer;
  g_arena.free_blocks.clear();
  g_arena.used_blocks.clear();
  g_arena.free_blocks[0] = g_arena.bytes;
  g_arena.active = true;
}

void peer_finalize() {
  if (!g_arena.active) return;
  Context &c = ctx();
  cudaDeviceSynchronize();
  for (int k = 0; k < (int)g_arena.peer.size(); ++k)
    if (k != c.rank && g_arena.peer[k]) cudaIpcCloseMemHandle(g_arena.peer[k]);
  cudaFree(g_arena.base);
  cudaGetLastError();
  g_arena = PeerArena();
}

// first fit; returns the offset of the region or -1 when the arena is exhausted (the caller then agrees with its peers on
// the NCCL path for that plan)
i64 peer_alloc(size_t bytes) {
  if (!g_arena.active) return -1;
  bytes = (std::max<size_t>(bytes, 1) + ALIGN - 1) / ALIGN * ALIGN;
  for (auto it = g_arena.free_blocks.begin(); it != g_arena.free_blocks.end(); ++it) {
    if (it->second < bytes) continue;
    size_t off = it->first, sz = it->second;
    g_arena.free_blocks.erase(it);
    if (sz > bytes) g_arena.free_blocks[off + bytes] = sz - bytes;
    g_arena.used_blocks[off] = bytes;
    return (i64)off;
  }
  return -1;
}

void peer_free(i64 off) {
  if (!g_arena.active || off < 0) return;
  auto u = g_arena.used_blocks.find((size_t)off);
  if (u == g_arena.used_blocks.end()) return;
  size_t o = u->first, sz = u->second;
  g_arena.used_blocks.erase(u);
  auto nx = g_arena.free_blocks.lower_bound(o);
  if (nx != g_arena.free_blocks.end() && o + sz == nx->first) { sz += nx->second; nx = g_arena.free_blocks.erase(nx); }
  if (nx != g_arena.free_blocks.begin()) {
    auto pv = std::prev(nx);
    if (pv->first + pv->second == o) { pv->second += sz; return; }
  }
  g_arena.free_blocks[o] = sz;
}

}  // namespace b200
