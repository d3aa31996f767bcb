// CPU collective backend ("gloo" role): full TCP mesh between the ranks of a group.
//
// The reference only ever names gloo in prose (README.md:133) but the plumbing config
// (ConvNet, mp.spawn, world_size 2, CPU) needs a CPU data plane.  Correctness first:
//  * allreduce = ring reduce-scatter + ring allgather, so every rank receives the SAME bytes
//    for every chunk (one owner reduces, everyone copies) -> replicas cannot drift.
//  * broadcast = root fans out; allgather = W-1 pairwise full-duplex exchanges.
// Endpoints are published through the rendezvous store under a per-group prefix so that
// sub-groups (new_group) get their own sockets and cannot interleave with the world group.
#include "net.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>

extern "C" {
int pdl_store_set(void* h, const char* key, const void* val, uint64_t n);
int pdl_store_get(void* h, const char* key, double timeout_s, void** out, uint64_t* out_n);
void pdl_free(void* p);
}

namespace pdl {

enum DType : int { F32 = 0, F64 = 1, I32 = 2, I64 = 3, U8 = 4 };
enum RedOp : int { SUM = 0, PROD = 1, MIN = 2, MAX = 3 };

static size_t dtype_size(int dt) {
  switch (dt) {
    case F32: case I32: return 4;
    case F64: case I64: return 8;
    default: return 1;
  }
}

template <typename T>
static void reduce_into(T* dst, const T* src, size_t n, int op) {
  switch (op) {
    case SUM: for (size_t i = 0; i < n; ++i) dst[i] += src[i]; break;
    case PROD: for (size_t i = 0; i < n; ++i) dst[i] *= src[i]; break;
    case MIN: for (size_t i = 0; i < n; ++i) dst[i] = std::min(dst[i], src[i]); break;
    case MAX: for (size_t i = 0; i < n; ++i) dst[i] = std::max(dst[i], src[i]); break;
  }
}

static void reduce_bytes(void* dst, const void* src, size_t n, int dt, int op) {
  switch (dt) {
    case F32: reduce_into(static_cast<float*>(dst), static_cast<const float*>(src), n, op); break;
    case F64: reduce_into(static_cast<double*>(dst), static_cast<const double*>(src), n, op); break;
    case I32: reduce_into(static_cast<int32_t*>(dst), static_cast<const int32_t*>(src), n, op); break;
    case I64: reduce_into(static_cast<int64_t*>(dst), static_cast<const int64_t*>(src), n, op); break;
    default: reduce_into(static_cast<uint8_t*>(dst), static_cast<const uint8_t*>(src), n, op); break;
  }
}

struct CpuComm {
  int rank = 0, world = 1;
  std::vector<int> peer_fd;  // peer_fd[r] = socket to rank r (-1 for self)
  std::vector<char> scratch;
  double timeout_s = 0;  // per collective; 0 = wait for ever
};

// Absolute deadline of a collective that starts now, and the error code of one that failed:
// -1 = a peer exited / the connection dropped, -2 = the deadline passed (a peer is alive but never joined).
static double deadline_of(const CpuComm* c) { return c->timeout_s > 0 ? now_s() + c->timeout_s : 0; }
static int failure() { return errno == ETIMEDOUT ? -2 : -1; }

}  // namespace pdl

using namespace pdl;

extern "C" {

// ranks are positions inside the group (0..world-1); `prefix` must be unique per group.
void* pdl_cpu_comm_create(void* store, const char* prefix, int rank, int world, const char* my_host, double timeout_s) {
  auto* c = new CpuComm();
  c->rank = rank;
  c->world = world;
  c->timeout_s = timeout_s;
  c->peer_fd.assign(world, -1);
  if (world == 1) return c;

  int port = 0;
  int lfd = listen_on(my_host, 0, &port);
  if (lfd < 0) { delete c; return nullptr; }
  char key[512], val[512];
  snprintf(key, sizeof(key), "%s/ep/%d", prefix, rank);
  snprintf(val, sizeof(val), "%s:%d", my_host, port);
  if (pdl_store_set(store, key, val, strlen(val)) != 0) { ::close(lfd); delete c; return nullptr; }

  // Dial every lower rank, accept every higher rank.
  for (int r = 0; r < rank; ++r) {
    snprintf(key, sizeof(key), "%s/ep/%d", prefix, r);
    void* out = nullptr;
    uint64_t n = 0;
    if (pdl_store_get(store, key, timeout_s, &out, &n) != 0) { ::close(lfd); delete c; return nullptr; }
    std::string ep(static_cast<char*>(out), n);
    pdl_free(out);
    size_t colon = ep.rfind(':');
    int fd = connect_to(ep.substr(0, colon).c_str(), atoi(ep.c_str() + colon + 1), timeout_s);
    int32_t me = rank;
    if (fd < 0 || !send_all(fd, &me, 4)) { ::close(lfd); delete c; return nullptr; }
    c->peer_fd[r] = fd;
  }
  double deadline = now_s() + timeout_s;
  for (int k = rank + 1; k < world; ++k) {
    int fd = -1;
    while (fd < 0) {
      pollfd p{lfd, POLLIN, 0};
      if (::poll(&p, 1, 200) > 0) fd = ::accept(lfd, nullptr, nullptr);
      if (fd < 0 && now_s() > deadline) { ::close(lfd); delete c; return nullptr; }
    }
    set_nodelay(fd);
    int32_t who = -1;
    if (!recv_all(fd, &who, 4) || who <= rank || who >= world) { ::close(fd); ::close(lfd); delete c; return nullptr; }
    c->peer_fd[who] = fd;
  }
  ::close(lfd);
  return c;
}

void pdl_cpu_comm_destroy(void* h) {
  auto* c = static_cast<CpuComm*>(h);
  if (!c) return;
  for (int fd : c->peer_fd)
    if (fd >= 0) ::close(fd);
  delete c;
}

int pdl_cpu_allreduce(void* h, void* buf, uint64_t count, int dtype, int op) {
  auto* c = static_cast<CpuComm*>(h);
  const int W = c->world, r = c->rank;
  if (W == 1 || count == 0) return 0;
  const size_t es = dtype_size(dtype);
  const uint64_t chunk = (count + W - 1) / W;
  auto lo = [&](int i) { return std::min<uint64_t>(count, chunk * i); };
  auto len = [&](int i) { return lo(i + 1) - lo(i); };
  char* base = static_cast<char*>(buf);
  c->scratch.resize(chunk * es);
  const int next = c->peer_fd[(r + 1) % W], prev = c->peer_fd[(r - 1 + W) % W];
  const double dl = deadline_of(c);
  // reduce-scatter: after W-1 steps rank r owns the fully reduced chunk (r+1)%W
  for (int s = 0; s < W - 1; ++s) {
    int send_i = ((r - s) % W + W) % W, recv_i = ((r - s - 1) % W + W) % W;
    if (!send_recv(next, base + lo(send_i) * es, len(send_i) * es, prev, c->scratch.data(), len(recv_i) * es, dl)) return failure();
    reduce_bytes(base + lo(recv_i) * es, c->scratch.data(), len(recv_i), dtype, op);
  }
  // allgather of the owned chunks
  for (int s = 0; s < W - 1; ++s) {
    int send_i = ((r + 1 - s) % W + W) % W, recv_i = ((r - s) % W + W) % W;
    if (!send_recv(next, base + lo(send_i) * es, len(send_i) * es, prev, base + lo(recv_i) * es, len(recv_i) * es, dl)) return failure();
  }
  return 0;
}

int pdl_cpu_broadcast(void* h, void* buf, uint64_t nbytes, int root) {
  auto* c = static_cast<CpuComm*>(h);
  if (c->world == 1 || nbytes == 0) return 0;
  const double dl = deadline_of(c);
  if (c->rank == root) {
    for (int p = 0; p < c->world; ++p)
      if (p != root && !send_all(c->peer_fd[p], buf, nbytes, dl)) return failure();
    return 0;
  }
  return recv_all(c->peer_fd[root], buf, nbytes, dl) ? 0 : failure();
}

// recv holds world blocks of nbytes; block r comes from rank r.
int pdl_cpu_allgather(void* h, const void* send, void* recv, uint64_t nbytes) {
  auto* c = static_cast<CpuComm*>(h);
  char* out = static_cast<char*>(recv);
  memcpy(out + static_cast<size_t>(c->rank) * nbytes, send, nbytes);
  const double dl = deadline_of(c);
  for (int s = 1; s < c->world; ++s) {
    int to = (c->rank + s) % c->world, from = (c->rank - s + c->world) % c->world;
    if (!send_recv(c->peer_fd[to], send, nbytes, c->peer_fd[from], out + static_cast<size_t>(from) * nbytes, nbytes, dl)) return failure();
  }
  return 0;
}

int pdl_cpu_barrier(void* h) {
  int32_t one = 1;
  return pdl_cpu_allreduce(h, &one, 1, I32, SUM);
}

int pdl_cpu_send(void* h, const void* buf, uint64_t nbytes, int dst) {
  auto* c = static_cast<CpuComm*>(h);
  return send_all(c->peer_fd[dst], buf, nbytes, deadline_of(c)) ? 0 : failure();
}

int pdl_cpu_recv(void* h, void* buf, uint64_t nbytes, int src) {
  auto* c = static_cast<CpuComm*>(h);
  return recv_all(c->peer_fd[src], buf, nbytes, deadline_of(c)) ? 0 : failure();
}

}  // extern "C"
