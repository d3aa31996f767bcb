// Small blocking-socket helpers shared by the store and the CPU collective backend.
#pragma once
#include <arpa/inet.h>
#include <cerrno>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <string>
#include <sys/socket.h>
#include <sys/types.h>
#include <thread>
#include <unistd.h>
#include <vector>

namespace pdl {

inline double now_s() {
  using namespace std::chrono;
  return duration<double>(steady_clock::now().time_since_epoch()).count();
}

inline void set_nodelay(int fd) {
  int one = 1;
  setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
}

// Deadlines are absolute times on the now_s() clock; 0 means "wait for ever".  A missed deadline makes the call return
// false with errno = ETIMEDOUT (a peer that is alive but never shows up; a dead peer closes its socket and is seen at once).
inline int poll_ms(double deadline_s) {
  if (deadline_s <= 0) return 60000;
  const double left = deadline_s - now_s();
  if (left <= 0) return -1;
  return left > 60.0 ? 60000 : static_cast<int>(left * 1000.0) + 1;
}

// Wait until fd is ready for `events`.  false = deadline missed.
inline bool wait_ready(int fd, short events, double deadline_s) {
  for (;;) {
    const int ms = poll_ms(deadline_s);
    if (ms < 0) { errno = ETIMEDOUT; return false; }
    pollfd p{fd, events, 0};
    const int pr = ::poll(&p, 1, ms);
    if (pr > 0 || (pr < 0 && errno != EINTR)) return true;  // ready, or let the I/O call report the error
  }
}

// Write exactly n bytes. Returns false on a closed/broken socket (or a missed deadline).
inline bool send_all(int fd, const void* buf, size_t n, double deadline_s = 0) {
  const char* p = static_cast<const char*>(buf);
  while (n > 0) {
    if (deadline_s > 0 && !wait_ready(fd, POLLOUT, deadline_s)) return false;
    ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL | (deadline_s > 0 ? MSG_DONTWAIT : 0));
    if (k < 0) {
      if (errno == EINTR || errno == EAGAIN || errno == EWOULDBLOCK) continue;
      return false;
    }
    p += k;
    n -= static_cast<size_t>(k);
  }
  return true;
}

inline bool recv_all(int fd, void* buf, size_t n, double deadline_s = 0) {
  char* p = static_cast<char*>(buf);
  while (n > 0) {
    if (deadline_s > 0 && !wait_ready(fd, POLLIN, deadline_s)) return false;
    ssize_t k = ::recv(fd, p, n, deadline_s > 0 ? MSG_DONTWAIT : 0);
    if (k == 0) { errno = ECONNRESET; return false; }
    if (k < 0) {
      if (errno == EINTR || errno == EAGAIN || errno == EWOULDBLOCK) continue;
      return false;
    }
    p += k;
    n -= static_cast<size_t>(k);
  }
  return true;
}

// Full-duplex exchange on (possibly different) sockets without deadlocking on
// full kernel buffers: progress whichever direction is ready.
inline bool send_recv(int send_fd, const void* sbuf, size_t sn, int recv_fd, void* rbuf, size_t rn, double deadline_s = 0) {
  const char* sp = static_cast<const char*>(sbuf);
  char* rp = static_cast<char*>(rbuf);
  while (sn > 0 || rn > 0) {
    pollfd fds[2];
    int nf = 0, si = -1, ri = -1;
    if (sn > 0) { fds[nf] = {send_fd, POLLOUT, 0}; si = nf++; }
    if (rn > 0) { fds[nf] = {recv_fd, POLLIN, 0}; ri = nf++; }
    const int ms = poll_ms(deadline_s);
    if (ms < 0) { errno = ETIMEDOUT; return false; }
    int pr = ::poll(fds, nf, ms);
    if (pr < 0) { if (errno == EINTR) continue; return false; }
    if (pr == 0) continue;
    if (si >= 0 && (fds[si].revents & (POLLOUT | POLLERR | POLLHUP))) {
      ssize_t k = ::send(send_fd, sp, sn, MSG_NOSIGNAL | MSG_DONTWAIT);
      if (k < 0) { if (errno != EAGAIN && errno != EWOULDBLOCK && errno != EINTR) return false; }
      else { sp += k; sn -= static_cast<size_t>(k); }
    }
    if (ri >= 0 && (fds[ri].revents & (POLLIN | POLLERR | POLLHUP))) {
      ssize_t k = ::recv(recv_fd, rp, rn, MSG_DONTWAIT);
      if (k == 0) return false;
      if (k < 0) { if (errno != EAGAIN && errno != EWOULDBLOCK && errno != EINTR) return false; }
      else { rp += k; rn -= static_cast<size_t>(k); }
    }
  }
  return true;
}

// Listen on host:port (port 0 = ephemeral). Returns fd, fills bound port.
inline int listen_on(const char* host, int port, int* bound_port) {
  int fd = ::socket(AF_INET, SOCK_STREAM, 0);
  if (fd < 0) return -1;
  int one = 1;
  setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_port = htons(static_cast<uint16_t>(port));
  addr.sin_addr.s_addr = htonl(INADDR_ANY);
  (void)host;  // always bind every interface; peers pick the address they can reach
  if (::bind(fd, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) != 0 || ::listen(fd, 256) != 0) {
    ::close(fd);
    return -1;
  }
  socklen_t len = sizeof(addr);
  getsockname(fd, reinterpret_cast<sockaddr*>(&addr), &len);
  if (bound_port) *bound_port = ntohs(addr.sin_port);
  return fd;
}

// Connect with retry until the deadline (the server may not be up yet).
inline int connect_to(const char* host, int port, double timeout_s) {
  double deadline = now_s() + timeout_s;
  char portstr[16];
  snprintf(portstr, sizeof(portstr), "%d", port);
  while (true) {
    addrinfo hints{}, *res = nullptr;
    hints.ai_family = AF_INET;
    hints.ai_socktype = SOCK_STREAM;
    if (getaddrinfo(host, portstr, &hints, &res) == 0) {
      for (addrinfo* ai = res; ai; ai = ai->ai_next) {
        int fd = ::socket(ai->ai_family, ai->ai_socktype, ai->ai_protocol);
        if (fd < 0) continue;
        if (::connect(fd, ai->ai_addr, ai->ai_addrlen) == 0) {
          freeaddrinfo(res);
          set_nodelay(fd);
          return fd;
        }
        ::close(fd);
      }
      freeaddrinfo(res);
    }
    if (now_s() > deadline) return -1;
    std::this_thread::sleep_for(std::chrono::milliseconds(20));
  }
}

}  // namespace pdl
