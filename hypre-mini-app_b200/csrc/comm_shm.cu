// comm_shm.cu -- host-staged inter-rank transport over a POSIX shared-memory segment.
//
// Purpose: lets N ranks (one process each) share ONE GPU, which NCCL refuses ("duplicate GPU"), so that the N-rank
// code paths of the library -- halo exchange, distributed AMG setup, Krylov reductions -- can be parity-tested against
// the N-virtual-rank oracle on a single-GPU box.  It is a transport, not a compute fallback: every kernel still runs on
// the device; only the bytes that NCCL would move over NVLink go device -> host segment -> device instead.  Production
// multi-GPU runs use NCCL (runtime.cu); this transport is selected explicitly with b200hypre_comm_init_shm().
//
// Layout of the segment: [Header | slot 0 | slot 1 | ... | slot N-1], one slot per rank.  A collective is:
// every rank writes its outgoing data + a small directory into its own slot, barrier, every rank reads what is
// addressed to it from the other slots, barrier.  Reductions are summed in rank order 0..N-1 on every rank, so the
// result is identical everywhere and independent of timing.
#include "common.h"
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <atomic>
#include <chrono>
#include <cstring>

namespace b200 {

namespace {

constexpr int MAX_RANKS = 64;
struct Header {
  std::atomic<int> arrived;
  std::atomic<int> generation;
  std::atomic<int> attached;
  int nranks;
  size_t slot_bytes;
};
struct Dir {  // at the start of every slot
  int nmsg;
  int dest[MAX_RANKS];
  size_t off[MAX_RANKS], bytes[MAX_RANKS];
};
constexpr size_t DIR_BYTES = (sizeof(Dir) + 255) & ~(size_t)255;

struct Shm {
  bool on = false;
  int rank = 0, nranks = 1;
  char *base = nullptr;
  size_t total = 0, slot_bytes = 0;
  std::string name;
  Header *hdr() { return (Header *)base; }
  char *slot(int r) { return base + 4096 + (size_t)r * slot_bytes; }
  Dir *dir(int r) { return (Dir *)slot(r); }
  char *payload(int r) { return slot(r) + DIR_BYTES; }
  size_t cap() const { return slot_bytes - DIR_BYTES; }
} g_shm;

void shm_barrier() {
  Header *h = g_shm.hdr();
  int gen = h->generation.load(std::memory_order_acquire);
  if (h->arrived.fetch_add(1, std::memory_order_acq_rel) + 1 == g_shm.nranks) {
    h->arrived.store(0, std::memory_order_relaxed);
    h->generation.store(gen + 1, std::memory_order_release);
    return;
  }
  auto t0 = std::chrono::steady_clock::now();
  long spins = 0;
  while (h->generation.load(std::memory_order_acquire) == gen) {
    if ((++spins & 1023) == 0) {
      usleep(50);
      double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
      static double limit = 0.0;   // B200_SHM_TIMEOUT_S: seconds a rank waits for its peers (default 300; sweeps use less)
      if (limit == 0.0) { const char *e = getenv("B200_SHM_TIMEOUT_S"); limit = (e && atof(e) > 0) ? atof(e) : 300.0; }
      if (dt > limit) fail(__FILE__, __LINE__, "shm transport: barrier timed out after " + std::to_string((int)limit) + " s (a peer rank died?)");
    }
  }
}

}  // namespace

bool shm_active() { return g_shm.on; }

void shm_init(int rank, int nranks, const char *name, size_t slot_bytes) {
  B200_REQUIRE(nranks >= 1 && nranks <= MAX_RANKS && rank >= 0 && rank < nranks, "shm transport: bad rank / size");
  if (slot_bytes < (1u << 20)) slot_bytes = 64u << 20;
  slot_bytes = (slot_bytes + 4095) & ~(size_t)4095;
  g_shm.rank = rank;
  g_shm.nranks = nranks;
  g_shm.slot_bytes = slot_bytes;
  g_shm.total = 4096 + slot_bytes * (size_t)nranks;
  g_shm.name = name;
  int fd = -1;
  if (rank == 0) {
    shm_unlink(name);
    fd = shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600);
    B200_REQUIRE(fd >= 0, std::string("shm transport: cannot create ") + name);
    B200_REQUIRE(ftruncate(fd, (off_t)g_shm.total) == 0, "shm transport: ftruncate failed");
  } else {
    for (int attempt = 0; attempt < 3000 && fd < 0; ++attempt) {
      fd = shm_open(name, O_RDWR, 0600);
      if (fd >= 0) {
        struct stat st;
        if (fstat(fd, &st) != 0 || (size_t)st.st_size < g_shm.total) { close(fd); fd = -1; }
      }
      if (fd < 0) usleep(20000);
    }
    B200_REQUIRE(fd >= 0, std::string("shm transport: cannot open ") + name);
  }
  void *p = mmap(nullptr, g_shm.total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  B200_REQUIRE(p != MAP_FAILED, "shm transport: mmap failed");
  g_shm.base = (char *)p;
  Header *h = g_shm.hdr();
  if (rank == 0) {  // the fresh segment is zero-filled: counters start at 0
    h->nranks = nranks;
    h->slot_bytes = slot_bytes;
  }
  h->attached.fetch_add(1, std::memory_order_acq_rel);
  auto t0 = std::chrono::steady_clock::now();
  while (h->attached.load(std::memory_order_acquire) < nranks) {
    usleep(200);
    double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    B200_REQUIRE(dt < 300.0, "shm transport: peers did not attach within 300 s");
  }
  g_shm.on = true;
  shm_barrier();
  if (rank == 0) shm_unlink(name);  // everybody has it mapped; the name can go
}

void shm_finalize() {
  if (!g_shm.on) return;
  munmap(g_shm.base, g_shm.total);
  g_shm = Shm();
}

// sum / max of `count` 8-byte values held on the host, in place.  kind: 0 = double sum, 1 = i64 sum, 2 = i64 max
void shm_allreduce_host(void *buf, int count, int kind) {
  size_t bytes = (size_t)count * 8;
  B200_REQUIRE(bytes <= g_shm.cap(), "shm transport: allreduce larger than the slot");
  memcpy(g_shm.payload(g_shm.rank), buf, bytes);
  shm_barrier();
  if (kind == 0) {
    double *out = (double *)buf;
    for (int k = 0; k < count; ++k) {
      double s = 0.0;
      for (int r = 0; r < g_shm.nranks; ++r) s += ((const double *)g_shm.payload(r))[k];
      out[k] = s;
    }
  } else {
    i64 *out = (i64 *)buf;
    for (int k = 0; k < count; ++k) {
      i64 s = ((const i64 *)g_shm.payload(0))[k];
      for (int r = 1; r < g_shm.nranks; ++r) {
        i64 v = ((const i64 *)g_shm.payload(r))[k];
        s = (kind == 1) ? s + v : (v > s ? v : s);
      }
      out[k] = s;
    }
  }
  shm_barrier();
}

void shm_allgather_host(const void *in, void *out, size_t bytes) {
  B200_REQUIRE(bytes <= g_shm.cap(), "shm transport: allgather larger than the slot");
  memcpy(g_shm.payload(g_shm.rank), in, bytes);
  shm_barrier();
  for (int r = 0; r < g_shm.nranks; ++r) memcpy((char *)out + (size_t)r * bytes, g_shm.payload(r), bytes);
  shm_barrier();
}

// device buffers; ordered on stream s (synchronous for the host)
void shm_allreduce_dev(double *buf, int count, cudaStream_t s) {
  std::vector<double> h((size_t)count);
  B200_CUDA(cudaMemcpyAsync(h.data(), buf, (size_t)count * 8, cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  shm_allreduce_host(h.data(), count, 0);
  B200_CUDA(cudaMemcpyAsync(buf, h.data(), (size_t)count * 8, cudaMemcpyHostToDevice, s));
  B200_CUDA(cudaStreamSynchronize(s));
}
void shm_allgather_dev(const void *in, void *out, size_t bytes, cudaStream_t s) {
  B200_REQUIRE(bytes <= g_shm.cap(), "shm transport: allgather larger than the slot");
  B200_CUDA(cudaMemcpyAsync(g_shm.payload(g_shm.rank), in, bytes, cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  shm_barrier();
  for (int r = 0; r < g_shm.nranks; ++r)
    B200_CUDA(cudaMemcpyAsync((char *)out + (size_t)r * bytes, g_shm.payload(r), bytes, cudaMemcpyHostToDevice, s));
  B200_CUDA(cudaStreamSynchronize(s));
  shm_barrier();
}

// point-to-point exchange: at most one message per (source, destination) pair and call
void shm_exchange_dev(int nsend, const int *send_rank, const void *const *send_ptr, const size_t *send_bytes, int nrecv,
                      const int *recv_rank, void *const *recv_ptr, const size_t *recv_bytes, cudaStream_t s) {
  Dir *d = g_shm.dir(g_shm.rank);
  char *pay = g_shm.payload(g_shm.rank);
  size_t off = 0;
  d->nmsg = 0;
  for (int k = 0; k < nsend; ++k) {
    if (!send_bytes[k]) continue;
    B200_REQUIRE(d->nmsg < MAX_RANKS, "shm transport: too many messages");
    for (int m = 0; m < d->nmsg; ++m) B200_REQUIRE(d->dest[m] != send_rank[k], "shm transport: two messages to one rank in one call");
    B200_REQUIRE(off + send_bytes[k] <= g_shm.cap(), "shm transport: messages exceed the slot (raise B200_SHM_SLOT_MB)");
    B200_CUDA(cudaMemcpyAsync(pay + off, send_ptr[k], send_bytes[k], cudaMemcpyDeviceToHost, s));
    d->dest[d->nmsg] = send_rank[k];
    d->off[d->nmsg] = off;
    d->bytes[d->nmsg] = send_bytes[k];
    ++d->nmsg;
    off += (send_bytes[k] + 15) & ~(size_t)15;
  }
  B200_CUDA(cudaStreamSynchronize(s));
  shm_barrier();
  for (int k = 0; k < nrecv; ++k) {
    if (!recv_bytes[k]) continue;
    int r = recv_rank[k];
    const Dir *o = g_shm.dir(r);
    int m = 0;
    for (; m < o->nmsg; ++m) if (o->dest[m] == g_shm.rank) break;
    B200_REQUIRE(m < o->nmsg, "shm transport: expected message is missing");
    B200_REQUIRE(o->bytes[m] == recv_bytes[k], "shm transport: message size mismatch");
    B200_CUDA(cudaMemcpyAsync(recv_ptr[k], g_shm.payload(r) + o->off[m], recv_bytes[k], cudaMemcpyHostToDevice, s));
  }
  B200_CUDA(cudaStreamSynchronize(s));
  shm_barrier();
}

}  // namespace b200
