// TCP key-value store used for rendezvous (rank 0 hosts the server; every rank is a client).
//
// Plays the role torch's TCPStore plays for the reference's init_process_group calls
// (reference: mpspawn_dist.py:49-54 env://, example_mp.py:37-42 tcp://): set/get/add/wait
// with blocking reads, so ranks can publish socket endpoints and CUDA memory handles to
// each other.  One server thread per client connection; values live in a mutex-guarded map
// and blocked readers sleep on a condition variable.
#include "net.h"

#include <atomic>
#include <condition_variable>
#include <cstdlib>
#include <map>
#include <memory>
#include <mutex>

namespace pdl {

enum StoreOp : uint8_t { OP_SET = 1, OP_GET = 2, OP_ADD = 3, OP_WAIT = 4, OP_CHECK = 5, OP_DEL = 6, OP_NUMKEYS = 7, OP_CAS = 8 };

struct StoreServer {
  int listen_fd = -1;
  int port = 0;
  std::atomic<bool> stop{false};
  std::thread accept_thread;
  std::mutex mu;
  std::condition_variable cv;
  std::map<std::string, std::string> kv;
  std::mutex conn_mu;
  std::vector<int> conn_fds;
  std::vector<std::thread> conn_threads;

  static bool read_str(int fd, std::string* s) {
    uint32_t n;
    if (!recv_all(fd, &n, 4)) return false;
    s->resize(n);
    return n == 0 || recv_all(fd, &(*s)[0], n);
  }
  static bool read_blob(int fd, std::string* s) {
    uint64_t n;
    if (!recv_all(fd, &n, 8)) return false;
    s->resize(n);
    return n == 0 || recv_all(fd, &(*s)[0], n);
  }

  bool wait_keys(std::unique_lock<std::mutex>& lk, const std::vector<std::string>& keys, double timeout_s) {
    auto have_all = [&] {
      for (auto& k : keys)
        if (kv.find(k) == kv.end()) return false;
      return true;
    };
    if (timeout_s < 0) {
      cv.wait(lk, [&] { return have_all() || stop.load(); });
      return have_all();
    }
    return cv.wait_for(lk, std::chrono::duration<double>(timeout_s), [&] { return have_all() || stop.load(); }) && have_all();
  }

  void serve(int fd) {
    while (!stop.load()) {
      uint8_t op;
      if (!recv_all(fd, &op, 1)) break;
      if (op == OP_SET) {
        std::string k, v;
        if (!read_str(fd, &k) || !read_blob(fd, &v)) break;
        {
          std::lock_guard<std::mutex> lk(mu);
          kv[k] = std::move(v);
        }
        cv.notify_all();
        uint8_t ok = 1;
        if (!send_all(fd, &ok, 1)) break;
      } else if (op == OP_GET) {
        std::string k;
        double timeout;
        if (!read_str(fd, &k) || !recv_all(fd, &timeout, 8)) break;
        std::string v;
        uint8_t status = 0;
        {
          std::unique_lock<std::mutex> lk(mu);
          if (wait_keys(lk, {k}, timeout)) v = kv[k];
          else status = 1;
        }
        uint64_t n = v.size();
        if (!send_all(fd, &status, 1) || !send_all(fd, &n, 8) || (n && !send_all(fd, v.data(), n))) break;
      } else if (op == OP_ADD) {
        std::string k;
        int64_t delta;
        if (!read_str(fd, &k) || !recv_all(fd, &delta, 8)) break;
        int64_t nv;
        {
          std::lock_guard<std::mutex> lk(mu);
          auto it = kv.find(k);
          int64_t cur = (it == kv.end()) ? 0 : std::strtoll(it->second.c_str(), nullptr, 10);
          nv = cur + delta;
          kv[k] = std::to_string(nv);
        }
        cv.notify_all();
        if (!send_all(fd, &nv, 8)) break;
      } else if (op == OP_WAIT || op == OP_CHECK) {
        uint32_t nk;
        if (!recv_all(fd, &nk, 4)) break;
        std::vector<std::string> keys(nk);
        bool ok = true;
        for (auto& k : keys) ok = ok && read_str(fd, &k);
        double timeout = 0;
        if (!ok || (op == OP_WAIT && !recv_all(fd, &timeout, 8))) break;
        uint8_t status;
        {
          std::unique_lock<std::mutex> lk(mu);
          status = wait_keys(lk, keys, op == OP_WAIT ? timeout : 0.0) ? 0 : 1;
        }
        if (!send_all(fd, &status, 1)) break;
      } else if (op == OP_DEL) {
        std::string k;
        if (!read_str(fd, &k)) break;
        uint8_t erased;
        {
          std::lock_guard<std::mutex> lk(mu);
          erased = kv.erase(k) ? 1 : 0;
        }
        if (!send_all(fd, &erased, 1)) break;
      } else if (op == OP_NUMKEYS) {
        int64_t n;
        {
          std::lock_guard<std::mutex> lk(mu);
          n = static_cast<int64_t>(kv.size());
        }
        if (!send_all(fd, &n, 8)) break;
      } else if (op == OP_CAS) {
        std::string k, expect, desired;
        if (!read_str(fd, &k) || !read_blob(fd, &expect) || !read_blob(fd, &desired)) break;
        std::string out;
        {
          std::lock_guard<std::mutex> lk(mu);
          auto it = kv.find(k);
          if (it == kv.end()) {
            if (expect.empty()) { kv[k] = desired; out = desired; }
            else out = expect;
          } else if (it->second == expect) { it->second = desired; out = desired; }
          else out = it->second;
        }
        cv.notify_all();
        uint64_t n = out.size();
        if (!send_all(fd, &n, 8) || (n && !send_all(fd, out.data(), n))) break;
      } else {
        break;
      }
    }
    ::close(fd);
  }

  void accept_loop() {
    while (!stop.load()) {
      pollfd p{listen_fd, POLLIN, 0};
      int r = ::poll(&p, 1, 100);
      if (r <= 0) continue;
      int fd = ::accept(listen_fd, nullptr, nullptr);
      if (fd < 0) continue;
      set_nodelay(fd);
      std::lock_guard<std::mutex> lk(conn_mu);
      conn_fds.push_back(fd);
      conn_threads.emplace_back([this, fd] { serve(fd); });
    }
  }
};

struct StoreClient {
  int fd = -1;
  std::mutex mu;  // one request in flight per client
};

static bool write_str(int fd, const char* s) {
  uint32_t n = static_cast<uint32_t>(strlen(s));
  return send_all(fd, &n, 4) && (n == 0 || send_all(fd, s, n));
}

}  // namespace pdl

using namespace pdl;

extern "C" {

void* pdl_store_server_start(const char* host, int port, int* out_port) {
  auto* s = new StoreServer();
  s->listen_fd = listen_on(host, port, &s->port);
  if (s->listen_fd < 0) {
    delete s;
    return nullptr;
  }
  if (out_port) *out_port = s->port;
  s->accept_thread = std::thread([s] { s->accept_loop(); });
  return s;
}

void pdl_store_server_stop(void* h) {
  auto* s = static_cast<StoreServer*>(h);
  if (!s) return;
  s->stop.store(true);
  s->cv.notify_all();
  if (s->accept_thread.joinable()) s->accept_thread.join();
  {
    std::lock_guard<std::mutex> lk(s->conn_mu);
    for (int fd : s->conn_fds) ::shutdown(fd, SHUT_RDWR);
  }
  for (auto& t : s->conn_threads)
    if (t.joinable()) t.join();
  ::close(s->listen_fd);
  delete s;
}

void* pdl_store_client_connect(const char* host, int port, double timeout_s) {
  int fd = connect_to(host, port, timeout_s);
  if (fd < 0) return nullptr;
  auto* c = new StoreClient();
  c->fd = fd;
  return c;
}

void pdl_store_client_close(void* h) {
  auto* c = static_cast<StoreClient*>(h);
  if (!c) return;
  ::close(c->fd);
  delete c;
}

// 0 = ok, -1 = connection error, 1 = timeout
int pdl_store_set(void* h, const char* key, const void* val, uint64_t n) {
  auto* c = static_cast<StoreClient*>(h);
  std::lock_guard<std::mutex> lk(c->mu);
  uint8_t op = OP_SET, ok = 0;
  if (!send_all(c->fd, &op, 1) || !write_str(c->fd, key) || !send_all(c->fd, &n, 8) || (n && !send_all(c->fd, val, n)) ||
      !recv_all(c->fd, &ok, 1))
    return -1;
  return 0;
}

// On success *out is malloc'ed (free with pdl_free) and *out_n its length.
int pdl_store_get(void* h, const char* key, double timeout_s, void** out, uint64_t* out_n) {
  auto* c = static_cast<StoreClient*>(h);
  std::lock_guard<std::mutex> lk(c->mu);
  uint8_t op = OP_GET, status = 0;
  uint64_t n = 0;
  if (!send_all(c->fd, &op, 1) || !write_str(c->fd, key) || !send_all(c->fd, &timeout_s, 8) || !recv_all(c->fd, &status, 1) ||
      !recv_all(c->fd, &n, 8))
    return -1;
  void* buf = malloc(n ? n : 1);
  if (n && !recv_all(c->fd, buf, n)) {
    free(buf);
    return -1;
  }
  if (status != 0) {
    free(buf);
    return 1;
  }
  *out = buf;
  *out_n = n;
  return 0;
}

int pdl_store_add(void* h, const char* key, int64_t delta, int64_t* result) {
  auto* c = static_cast<StoreClient*>(h);
  std::lock_guard<std::mutex> lk(c->mu);
  uint8_t op = OP_ADD;
  if (!send_all(c->fd, &op, 1) || !write_str(c->fd, key) || !send_all(c->fd, &delta, 8) || !recv_all(c->fd, result, 8)) return -1;
  return 0;
}

static int keys_op(void* h, uint8_t op, const char** keys, int nkeys, double timeout_s) {
  auto* c = static_cast<StoreClient*>(h);
  std::lock_guard<std::mutex> lk(c->mu);
  uint32_t nk = static_cast<uint32_t>(nkeys);
  if (!send_all(c->fd, &op, 1) || !send_all(c->fd, &nk, 4)) return -1;
  for (int i = 0; i < nkeys; ++i)
    if (!write_str(c->fd, keys[i])) return -1;
  if (op == OP_WAIT && !send_all(c->fd, &timeout_s, 8)) return -1;
  uint8_t status;
  if (!recv_all(c->fd, &status, 1)) return -1;
  return status;
}

int pdl_store_wait(void* h, const char** keys, int nkeys, double timeout_s) { return keys_op(h, OP_WAIT, keys, nkeys, timeout_s); }
int pdl_store_check(void* h, const char** keys, int nkeys) { return keys_op(h, OP_CHECK, keys, nkeys, 0.0); }

int pdl_store_delete(void* h, const char* key) {
  auto* c = static_cast<StoreClient*>(h);
  std::lock_guard<std::mutex> lk(c->mu);
  uint8_t op = OP_DEL, erased = 0;
  if (!send_all(c->fd, &op, 1) || !write_str(c->fd, key) || !recv_all(c->fd, &erased, 1)) return -1;
  return erased;
}

int64_t pdl_store_num_keys(void* h) {
  auto* c = static_cast<StoreClient*>(h);
  std::lock_guard<std::mutex> lk(c->mu);
  uint8_t op = OP_NUMKEYS;
  int64_t n = -1;
  if (!send_all(c->fd, &op, 1) || !recv_all(c->fd, &n, 8)) return -1;
  return n;
}

int pdl_store_compare_set(void* h, const char* key, const void* expect, uint64_t en, const void* desired, uint64_t dn, void** out,
                          uint64_t* out_n) {
  auto* c = static_cast<StoreClient*>(h);
  std::lock_guard<std::mutex> lk(c->mu);
  uint8_t op = OP_CAS;
  uint64_t n = 0;
  if (!send_all(c->fd, &op, 1) || !write_str(c->fd, key) || !send_all(c->fd, &en, 8) || (en && !send_all(c->fd, expect, en)) ||
      !send_all(c->fd, &dn, 8) || (dn && !send_all(c->fd, desired, dn)) || !recv_all(c->fd, &n, 8))
    return -1;
  void* buf = malloc(n ? n : 1);
  if (n && !recv_all(c->fd, buf, n)) {
    free(buf);
    return -1;
  }
  *out = buf;
  *out_n = n;
  return 0;
}

void pdl_free(void* p) { free(p); }

}  // extern "C"
