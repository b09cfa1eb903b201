// runtime.cu -- process-global context of libb200hypre: device/stream setup, stream-ordered memory pool, error
// plumbing and the inter-rank communicator (NCCL, loaded lazily with dlopen so that a single-GPU run has no NCCL
// dependency).  Stands in for HYPRE_Init / HYPRE_Finalize / hypre's memory pools
// (/root/reference/src/main.cpp:82,97-115,218) and for the MPI calls hypre makes inside the solve path
// (SURVEY.md §2.5): Allreduce of inner products, barrier, halo send/recv.
//
// One process per GPU.  Rank / world size come from b200hypre_comm_init() (Python harness: the NCCL unique id is
// shipped with torch.distributed) or, for the stand-alone hypre_app, from the torchrun-style environment
// (RANK, WORLD_SIZE, LOCAL_RANK, MASTER_ADDR, MASTER_PORT) with a tiny TCP bootstrap of the NCCL unique id.
#include "common.h"
#include <arpa/inet.h>
#include <dlfcn.h>
#include <netinet/in.h>
#include <sys/socket.h>
#include <unistd.h>
#include <cstring>
#include <mutex>

namespace b200 {

[[noreturn]] void fail(const char *file, int line, const std::string &msg) {
  throw std::runtime_error(std::string(file) + ":" + std::to_string(line) + ": " + msg);
}

Context &ctx() {
  static Context c;
  return c;
}

void ensure_init() {
  Context &c = ctx();
  if (c.initialised) return;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    fail(__FILE__, __LINE__, "libb200hypre needs a CUDA device (sm_100a); there is no CPU fallback");
  }
  B200_CUDA(cudaGetDevice(&c.device));
  cudaDeviceProp prop;
  B200_CUDA(cudaGetDeviceProperties(&prop, c.device));
  c.sm_count = prop.multiProcessorCount;
  B200_CUDA(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
  B200_CUDA(cudaStreamCreateWithFlags(&c.comm_stream, cudaStreamNonBlocking));
  // stream-ordered pool that never gives memory back to the driver while the library lives: AMG setup allocates and
  // frees many temporaries and must not hit cudaMalloc in its loops (main.cpp:97-115 configures hypre's pools)
  cudaMemPool_t pool;
  B200_CUDA(cudaDeviceGetDefaultMemPool(&pool, c.device));
  unsigned long long thr = ~0ULL;
  B200_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
  c.initialised = true;
}

void *dev_alloc(size_t bytes) {
  ensure_init();
  void *p = nullptr;
  cudaError_t e = cudaMallocAsync(&p, bytes, stream());
  if (e != cudaSuccess) {
    cudaGetLastError();
    fail(__FILE__, __LINE__, "device allocation of " + std::to_string(bytes) + " bytes failed: " + cudaGetErrorString(e));
  }
  return p;
}
void dev_free(void *p) {
  if (p && ctx().initialised) cudaFreeAsync(p, stream());
}

// ---------------------------------------------------------------------------------------------------------------------
// kernel-class profiler
// ---------------------------------------------------------------------------------------------------------------------
namespace {
struct ProfRec { const char *cls; double bytes; cudaEvent_t e0, e1; };
struct Profiler {
  bool on = false;
  double min_bytes = 0;
  std::vector<ProfRec> recs;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pool;  // recycled event pairs
} g_prof;
}  // namespace

void prof_enable(bool on, double min_bytes) {
  for (ProfRec &r : g_prof.recs) g_prof.pool.emplace_back(r.e0, r.e1);
  g_prof.recs.clear();
  g_prof.on = on;
  g_prof.min_bytes = min_bytes;
}
ProfScope::ProfScope(const char *cls, double bytes) {
  if (!g_prof.on || bytes < g_prof.min_bytes || g_prof.recs.size() >= 65536) return;
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream(), &st) != cudaSuccess || st != cudaStreamCaptureStatusNone) { cudaGetLastError(); return; }
  ProfRec r;
  r.cls = cls;
  r.bytes = bytes;
  if (!g_prof.pool.empty()) { r.e0 = g_prof.pool.back().first; r.e1 = g_prof.pool.back().second; g_prof.pool.pop_back(); }
  else { B200_CUDA(cudaEventCreate(&r.e0)); B200_CUDA(cudaEventCreate(&r.e1)); }
  B200_CUDA(cudaEventRecord(r.e0, stream()));
  slot = (int)g_prof.recs.size();
  g_prof.recs.push_back(r);
}
ProfScope::~ProfScope() {
  if (slot >= 0) cudaEventRecord(g_prof.recs[slot].e1, stream());
}
int prof_summary(std::vector<std::string> &names, std::vector<double> &ms, std::vector<double> &calls, std::vector<double> &bytes) {
  names.clear(); ms.clear(); calls.clear(); bytes.clear();
  if (ctx().initialised) B200_CUDA(cudaStreamSynchronize(stream()));
  for (ProfRec &r : g_prof.recs) {
    float t = 0;
    if (cudaEventElapsedTime(&t, r.e0, r.e1) != cudaSuccess) { cudaGetLastError(); continue; }
    size_t k = 0;
    for (; k < names.size(); ++k) if (names[k] == r.cls) break;
    if (k == names.size()) { names.push_back(r.cls); ms.push_back(0); calls.push_back(0); bytes.push_back(0); }
    ms[k] += t; calls[k] += 1; bytes[k] += r.bytes;
  }
  return (int)names.size();
}

// ---------------------------------------------------------------------------------------------------------------------
// NCCL through dlopen
// ---------------------------------------------------------------------------------------------------------------------
struct Id128 { char b[128]; };  // ncclUniqueId (passed by value to ncclCommInitRank)
namespace {

struct NcclApi {
  void *h = nullptr;
  int (*GetUniqueId)(void *) = nullptr;
  int (*CommInitRank)(void **, int, Id128, int) = nullptr;
  int (*AllReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
  int (*AllGather)(const void *, void *, size_t, int, void *, cudaStream_t) = nullptr;
  int (*Send)(const void *, size_t, int, int, void *, cudaStream_t) = nullptr;
  int (*Recv)(void *, size_t, int, int, void *, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  int (*CommDestroy)(void *) = nullptr;
  const char *(*GetErrorString)(int) = nullptr;
};
NcclApi g_nccl;
// ncclDataType_t / ncclRedOp_t values (nccl.h): ncclInt8 0, ncclInt64 4, ncclFloat64 8; ncclSum 0, ncclMax 2
constexpr int NCCL_INT8 = 0, NCCL_INT64 = 4, NCCL_F64 = 8, NCCL_SUM = 0, NCCL_MAX = 2;

void nccl_load() {
  if (g_nccl.h) return;
  const char *names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char *n : names) {
    g_nccl.h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (g_nccl.h) break;
  }
  B200_REQUIRE(g_nccl.h, "multi-rank run requested but libnccl.so.2 could not be loaded");
#define LOADSYM(field, name)                                   \
  *(void **)(&g_nccl.field) = dlsym(g_nccl.h, name);           \
  B200_REQUIRE(g_nccl.field, std::string("NCCL symbol missing: ") + name)
  LOADSYM(GetUniqueId, "ncclGetUniqueId");
  LOADSYM(CommInitRank, "ncclCommInitRank");
  LOADSYM(AllReduce, "ncclAllReduce");
  LOADSYM(AllGather, "ncclAllGather");
  LOADSYM(Send, "ncclSend");
  LOADSYM(Recv, "ncclRecv");
  LOADSYM(GroupStart, "ncclGroupStart");
  LOADSYM(GroupEnd, "ncclGroupEnd");
  LOADSYM(CommDestroy, "ncclCommDestroy");
  LOADSYM(GetErrorString, "ncclGetErrorString");
#undef LOADSYM
}
#define B200_NCCL(call)                                                                                   \
  do {                                                                                                    \
    int r__ = (call);                                                                                     \
    if (r__ != 0) ::b200::fail(__FILE__, __LINE__, std::string("NCCL: ") + g_nccl.GetErrorString(r__));   \
  } while (0)

double *g_scratch = nullptr;  // 4 KiB device scratch for host-value reductions
double *scratch() {
  if (!g_scratch) B200_CUDA(cudaMalloc(&g_scratch, 4096));
  return g_scratch;
}
}  // namespace

// host-staged shared-memory transport for N ranks on one GPU (comm_shm.cu)
bool shm_active();
void shm_init(int rank, int nranks, const char *name, size_t slot_bytes);
void shm_finalize();
void shm_allreduce_host(void *buf, int count, int kind);
void shm_allgather_host(const void *in, void *out, size_t bytes);
void shm_allreduce_dev(double *buf, int count, cudaStream_t s);
void shm_allgather_dev(const void *in, void *out, size_t bytes, cudaStream_t s);
void shm_exchange_dev(int nsend, const int *send_rank, const void *const *send_ptr, const size_t *send_bytes, int nrecv,
                      const int *recv_rank, void *const *recv_ptr, const size_t *recv_bytes, cudaStream_t s);

void peer_init();       // comm_peer.cu
void peer_finalize();

void comm_init_shm(int rank, int nranks, const char *name, size_t slot_bytes) {
  ensure_init();
  Context &c = ctx();
  B200_REQUIRE(nranks >= 1 && rank >= 0 && rank < nranks, "comm_init_shm: bad rank / size");
  c.rank = rank;
  c.nranks = nranks;
  if (nranks == 1) return;
  const char *mb = getenv("B200_SHM_SLOT_MB");
  if (mb && atoi(mb) > 0) slot_bytes = (size_t)atoi(mb) << 20;
  shm_init(rank, nranks, name, slot_bytes);
}

void comm_get_unique_id(void *out128) {
  nccl_load();
  B200_NCCL(g_nccl.GetUniqueId(out128));
}

void comm_init(int rank, int nranks, const void *id128) {
  ensure_init();
  Context &c = ctx();
  B200_REQUIRE(nranks >= 1 && rank >= 0 && rank < nranks, "comm_init: bad rank / size");
  c.rank = rank;
  c.nranks = nranks;
  if (nranks == 1) return;
  nccl_load();
  Id128 id;
  memcpy(id.b, id128, 128);
  void *comm = nullptr;
  B200_NCCL(g_nccl.CommInitRank(&comm, nranks, id, rank));
  c.nccl = comm;
  peer_init();  // NVLink peer-memory arena for the halo exchange (falls back to NCCL send/recv when unavailable)
}

void comm_finalize() {
  Context &c = ctx();
  peer_finalize();
  if (c.nccl) { g_nccl.CommDestroy(c.nccl); c.nccl = nullptr; }
  shm_finalize();
  c.rank = 0;
  c.nranks = 1;
}

// torchrun-style environment bootstrap (stand-alone hypre_app): rank 0 serves the NCCL id over TCP on MASTER_PORT+23
void comm_init_from_env() {
  const char *ws = getenv("WORLD_SIZE"), *rk = getenv("RANK");
  int nranks = ws ? atoi(ws) : 1, rank = rk ? atoi(rk) : 0;
  if (nranks <= 1) { comm_init(0, 1, nullptr); return; }
  const char *lr = getenv("LOCAL_RANK");
  int ndev = 1;
  B200_CUDA(cudaGetDeviceCount(&ndev));
  B200_CUDA(cudaSetDevice((lr ? atoi(lr) : rank) % ndev));
  if (const char *shm = getenv("B200_COMM_SHM")) {  // N ranks on one GPU (tests): host-staged transport, comm_shm.cu
    comm_init_shm(rank, nranks, shm, 0);
    return;
  }
  const char *addr = getenv("MASTER_ADDR");
  const char *port = getenv("MASTER_PORT");
  int p = (port ? atoi(port) : 29500) + 23;
  Id128 id;
  if (rank == 0) {
    comm_get_unique_id(id.b);
    int srv = socket(AF_INET, SOCK_STREAM, 0);
    int one = 1;
    setsockopt(srv, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    sockaddr_in sa{};
    sa.sin_family = AF_INET;
    sa.sin_addr.s_addr = htonl(INADDR_ANY);
    sa.sin_port = htons((uint16_t)p);
    B200_REQUIRE(bind(srv, (sockaddr *)&sa, sizeof(sa)) == 0, "comm bootstrap: bind failed");
    B200_REQUIRE(listen(srv, nranks) == 0, "comm bootstrap: listen failed");
    for (int k = 1; k < nranks; ++k) {
      int fd = accept(srv, nullptr, nullptr);
      B200_REQUIRE(fd >= 0, "comm bootstrap: accept failed");
      B200_REQUIRE(write(fd, id.b, 128) == 128, "comm bootstrap: short write");
      close(fd);
    }
    close(srv);
  } else {
    sockaddr_in sa{};
    sa.sin_family = AF_INET;
    sa.sin_port = htons((uint16_t)p);
    inet_pton(AF_INET, (addr && strcmp(addr, "localhost") != 0) ? addr : "127.0.0.1", &sa.sin_addr);
    int fd = -1;
    for (int attempt = 0; attempt < 600; ++attempt) {
      fd = socket(AF_INET, SOCK_STREAM, 0);
      if (connect(fd, (sockaddr *)&sa, sizeof(sa)) == 0) break;
      close(fd);
      fd = -1;
      usleep(100000);
    }
    B200_REQUIRE(fd >= 0, "comm bootstrap: could not reach rank 0");
    size_t got = 0;
    while (got < 128) {
      ssize_t r = read(fd, id.b + got, 128 - got);
      B200_REQUIRE(r > 0, "comm bootstrap: short read");
      got += (size_t)r;
    }
    close(fd);
  }
  comm_init(rank, nranks, id.b);
}

void allreduce_sum_dev(double *buf, int count) {
  Context &c = ctx();
  if (c.nranks == 1 || count <= 0) return;
  if (shm_active()) { shm_allreduce_dev(buf, count, stream()); return; }
  B200_NCCL(g_nccl.AllReduce(buf, buf, (size_t)count, NCCL_F64, NCCL_SUM, c.nccl, stream()));
}
void allreduce_sum_host(double *buf, int count) {
  Context &c = ctx();
  if (c.nranks == 1 || count <= 0) return;
  B200_REQUIRE(count <= 512, "allreduce_sum_host: count > 512");
  if (shm_active()) { shm_allreduce_host(buf, count, 0); return; }
  double *d = scratch();
  B200_CUDA(cudaMemcpyAsync(d, buf, count * sizeof(double), cudaMemcpyHostToDevice, stream()));
  B200_NCCL(g_nccl.AllReduce(d, d, (size_t)count, NCCL_F64, NCCL_SUM, c.nccl, stream()));
  B200_CUDA(cudaMemcpyAsync(buf, d, count * sizeof(double), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
}
static void allreduce_host_i64(i64 *buf, int count, int op) {
  Context &c = ctx();
  if (c.nranks == 1 || count <= 0) return;
  B200_REQUIRE(count <= 512, "allreduce_host_i64: count > 512");
  if (shm_active()) { shm_allreduce_host(buf, count, op == NCCL_SUM ? 1 : 2); return; }
  i64 *d = (i64 *)scratch();
  B200_CUDA(cudaMemcpyAsync(d, buf, count * sizeof(i64), cudaMemcpyHostToDevice, stream()));
  B200_NCCL(g_nccl.AllReduce(d, d, (size_t)count, NCCL_INT64, op, c.nccl, stream()));
  B200_CUDA(cudaMemcpyAsync(buf, d, count * sizeof(i64), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
}
void allreduce_sum_host_i64(i64 *buf, int count) { allreduce_host_i64(buf, count, NCCL_SUM); }
void allreduce_max_host_i64(i64 *buf, int count) { allreduce_host_i64(buf, count, NCCL_MAX); }
void barrier() {
  B200_CUDA(cudaStreamSynchronize(stream()));
  i64 one = 1;
  allreduce_sum_host_i64(&one, 1);
}

// every rank contributes `bytes` bytes (device pointers); out holds nranks * bytes
void allgather_dev(const void *in, void *out, size_t bytes) {
  Context &c = ctx();
  if (c.nranks == 1) {
    if (in != out) B200_CUDA(cudaMemcpyAsync(out, in, bytes, cudaMemcpyDeviceToDevice, stream()));
    return;
  }
  if (shm_active()) { shm_allgather_dev(in, out, bytes, stream()); return; }
  B200_NCCL(g_nccl.AllGather(in, out, bytes, NCCL_INT8, c.nccl, stream()));
}
// host convenience: gathers `count` i64 per rank
void allgather_host_i64(const i64 *in, i64 *out, int count) {
  Context &c = ctx();
  if (c.nranks == 1) { memcpy(out, in, count * sizeof(i64)); return; }
  if (shm_active()) { shm_allgather_host(in, out, count * sizeof(i64)); return; }
  DVec<i64> di((size_t)count), dout((size_t)count * c.nranks);
  B200_CUDA(cudaMemcpyAsync(di.p, in, count * sizeof(i64), cudaMemcpyHostToDevice, stream()));
  allgather_dev(di.p, dout.p, count * sizeof(i64));
  B200_CUDA(cudaMemcpyAsync(out, dout.p, (size_t)count * c.nranks * sizeof(i64), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
}

// grouped point-to-point exchange of byte ranges (device pointers), on stream `s`
void exchange_dev(int nsend, const int *send_rank, const void *const *send_ptr, const size_t *send_bytes, int nrecv,
                  const int *recv_rank, void *const *recv_ptr, const size_t *recv_bytes, cudaStream_t s) {
  Context &c = ctx();
  if (c.nranks == 1) return;
  if (shm_active()) { shm_exchange_dev(nsend, send_rank, send_ptr, send_bytes, nrecv, recv_rank, recv_ptr, recv_bytes, s); return; }
  B200_NCCL(g_nccl.GroupStart());
  for (int k = 0; k < nsend; ++k)
    if (send_bytes[k]) B200_NCCL(g_nccl.Send(send_ptr[k], send_bytes[k], NCCL_INT8, send_rank[k], c.nccl, s));
  for (int k = 0; k < nrecv; ++k)
    if (recv_bytes[k]) B200_NCCL(g_nccl.Recv(recv_ptr[k], recv_bytes[k], NCCL_INT8, recv_rank[k], c.nccl, s));
  B200_NCCL(g_nccl.GroupEnd());
}

}  // namespace b200
