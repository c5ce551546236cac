/*!
 * rabit-b200: POSIX TCP transport for the control plane and the host-memory engine.
 *
 * Covers the same ground as reference include/rabit/internal/socket.h (SockAddr 50-97,
 * Socket/TCPSocket 102-437, PollHelper 440-533) but is Linux-only and adds what the
 * device bootstrap needs: abstract-namespace unix sockets with SCM_RIGHTS so CUDA VMM
 * file-descriptor handles can travel between the per-GPU processes.
 *
 * Wire helpers keep the tracker protocol's encoding: native-endian int32, and strings as
 * int32 length + bytes (reference socket.h:414-436).
 */
#ifndef RABIT_INTERNAL_SOCKET_H_
#define RABIT_INTERNAL_SOCKET_H_

#include <arpa/inet.h>
#include <errno.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/types.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

#include "rabit/internal/utils.h"

namespace rabit {
namespace utils {

/*! \brief resolved IPv4 endpoint */
struct SockAddr {
  sockaddr_in addr;
  SockAddr() { std::memset(&addr, 0, sizeof(addr)); }
  SockAddr(const char *host, int port) { this->Set(host, port); }
  /*! \brief name of this machine as reported by the OS */
  static std::string GetHostName() {
    char buf[256];
    Check(gethostname(buf, sizeof(buf)) == 0, "gethostname failed: %s", strerror(errno));
    buf[sizeof(buf) - 1] = '\0';
    return std::string(buf);
  }
  /*! \brief resolve host (name or dotted quad) */
  void Set(const char *host, int port) {
    addrinfo hints; std::memset(&hints, 0, sizeof(hints));
    hints.ai_family = AF_INET;
    hints.ai_socktype = SOCK_STREAM;
    addrinfo *res = nullptr;
    int rc = getaddrinfo(host, nullptr, &hints, &res);
    Check(rc == 0 && res != nullptr, "cannot resolve host \"%s\": %s", host, gai_strerror(rc));
    std::memcpy(&addr, res->ai_addr, sizeof(addr));
    addr.sin_port = htons(static_cast<uint16_t>(port));
    freeaddrinfo(res);
  }
  int Port() const { return ntohs(addr.sin_port); }
  std::string AddrStr() const {
    char buf[INET_ADDRSTRLEN];
    const char *s = inet_ntop(AF_INET, &addr.sin_addr, buf, sizeof(buf));
    Assert(s != nullptr, "inet_ntop failed");
    return std::string(s);
  }
};

/*! \brief monotonic seconds (deadlines inside the socket helpers) */
inline double NowSeconds() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return static_cast<double>(ts.tv_sec) + static_cast<double>(ts.tv_nsec) * 1e-9;
}

/*! \brief thin value-type handle over a TCP socket fd; Close() is explicit */
class TCPSocket {
 public:
  int fd;
  TCPSocket() : fd(-1) {}
  explicit TCPSocket(int fd_) : fd(fd_) {}
  bool IsClosed() const { return fd < 0; }
  bool BadSocket() const {
    if (fd < 0) return true;
    int err = 0; socklen_t len = sizeof(err);
    if (getsockopt(fd, SOL_SOCKET, SO_ERROR, &err, &len) != 0) return true;
    return err == EBADF || err == EINTR;
  }
  void Create() {
    fd = socket(AF_INET, SOCK_STREAM, 0);
    Check(fd >= 0, "socket(): %s", strerror(errno));
  }
  void Close() {
    if (fd >= 0) { ::close(fd); fd = -1; }
  }
  void SetNonBlock(bool on) {
    int fl = fcntl(fd, F_GETFL, 0);
    Check(fl != -1, "fcntl(F_GETFL): %s", strerror(errno));
    fl = on ? (fl | O_NONBLOCK) : (fl & ~O_NONBLOCK);
    Check(fcntl(fd, F_SETFL, fl) != -1, "fcntl(F_SETFL): %s", strerror(errno));
  }
  void SetKeepAlive(bool on) {
    int v = on ? 1 : 0;
    setsockopt(fd, SOL_SOCKET, SO_KEEPALIVE, &v, sizeof(v));
  }
  void SetNoDelay(bool on) {
    int v = on ? 1 : 0;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &v, sizeof(v));
  }
  void SetReuseAddr(bool on) {
    int v = on ? 1 : 0;
    setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &v, sizeof(v));
  }
  void SetLinger(int seconds) {
    linger lg; lg.l_onoff = 1; lg.l_linger = seconds;
    setsockopt(fd, SOL_SOCKET, SO_LINGER, &lg, sizeof(lg));
  }
  void SetBufSize(int bytes) {
    setsockopt(fd, SOL_SOCKET, SO_SNDBUF, &bytes, sizeof(bytes));
    setsockopt(fd, SOL_SOCKET, SO_RCVBUF, &bytes, sizeof(bytes));
  }
  /*! \brief bind to the first free port in [start, end); returns it or -1 */
  int TryBindHost(int start, int end) {
    for (int port = start; port < end; ++port) {
      sockaddr_in a; std::memset(&a, 0, sizeof(a));
      a.sin_family = AF_INET; a.sin_addr.s_addr = htonl(INADDR_ANY);
      a.sin_port = htons(static_cast<uint16_t>(port));
      if (bind(fd, reinterpret_cast<sockaddr *>(&a), sizeof(a)) == 0) return port;
      if (errno != EADDRINUSE && errno != EACCES) {
        Error("bind(): %s", strerror(errno));
      }
    }
    return -1;
  }
  /*! \brief bind to an OS-chosen port, return it */
  int BindAny() {
    sockaddr_in a; std::memset(&a, 0, sizeof(a));
    a.sin_family = AF_INET; a.sin_addr.s_addr = htonl(INADDR_ANY); a.sin_port = 0;
    Check(bind(fd, reinterpret_cast<sockaddr *>(&a), sizeof(a)) == 0, "bind(): %s", strerror(errno));
    socklen_t len = sizeof(a);
    Check(getsockname(fd, reinterpret_cast<sockaddr *>(&a), &len) == 0, "getsockname failed");
    return ntohs(a.sin_port);
  }
  void Listen(int backlog = 128) {
    Check(listen(fd, backlog) == 0, "listen(): %s", strerror(errno));
  }
  TCPSocket Accept() {
    int c;
    do { c = accept(fd, nullptr, nullptr); } while (c < 0 && errno == EINTR);
    Check(c >= 0, "accept(): %s", strerror(errno));
    return TCPSocket(c);
  }
  bool Connect(const SockAddr &a) {
    int rc;
    do {
      rc = connect(fd, reinterpret_cast<const sockaddr *>(&a.addr), sizeof(a.addr));
    } while (rc != 0 && errno == EINTR);
    return rc == 0;
  }
  /*! \brief one send(2); returns bytes or -1 (errno set) */
  ssize_t Send(const void *buf, size_t len) {
    return send(fd, buf, len, MSG_NOSIGNAL);
  }
  ssize_t Recv(void *buf, size_t len) { return recv(fd, buf, len, 0); }
  /*! \brief blocking loop until everything is sent; returns bytes sent (< len on error) */
  size_t SendAll(const void *buf_, size_t len) {
    const char *buf = static_cast<const char *>(buf_);
    size_t done = 0;
    while (done < len) {
      ssize_t n = send(fd, buf + done, len - done, MSG_NOSIGNAL);
      if (n < 0) {
        if (errno == EINTR) continue;
        if (errno == EAGAIN || errno == EWOULDBLOCK) { WaitFor(POLLOUT); continue; }
        return done;
      }
      done += static_cast<size_t>(n);
    }
    return done;
  }
  size_t RecvAll(void *buf_, size_t len) {
    char *buf = static_cast<char *>(buf_);
    size_t done = 0;
    while (done < len) {
      ssize_t n = recv(fd, buf + done, len - done, 0);
      if (n < 0) {
        if (errno == EINTR) continue;
        if (errno == EAGAIN || errno == EWOULDBLOCK) { WaitFor(POLLIN); continue; }
        return done;
      }
      if (n == 0) return done;
      done += static_cast<size_t>(n);
    }
    return done;
  }
  /*!
   * \brief RecvAll with a deadline, for handshakes between two workers: both ends are busy with exactly this
   *  connection, so silence means the other end is gone or wedged. Returns the bytes received (< len on
   *  timeout, close or error). Works in blocking and non-blocking mode.
   */
  size_t RecvAllTimed(void *buf_, size_t len, int timeout_ms) {
    char *buf = static_cast<char *>(buf_);
    size_t done = 0;
    const double t_end = NowSeconds() + timeout_ms * 1e-3;
    while (done < len) {
      const double left = t_end - NowSeconds();
      if (left <= 0) return done;
      pollfd p; p.fd = fd; p.events = POLLIN; p.revents = 0;
      const int rc = poll(&p, 1, static_cast<int>(left * 1e3) + 1);
      if (rc < 0 && errno == EINTR) continue;
      if (rc <= 0) return done;
      ssize_t n = recv(fd, buf + done, len - done, MSG_DONTWAIT);
      if (n < 0) {
        if (errno == EINTR || errno == EAGAIN || errno == EWOULDBLOCK) continue;
        return done;
      }
      if (n == 0) return done;
      done += static_cast<size_t>(n);
    }
    return done;
  }
  bool RecvIntTimed(int *v, int timeout_ms) { return RecvAllTimed(v, sizeof(*v), timeout_ms) == sizeof(*v); }
  // ---- tracker wire helpers ----
  bool SendInt(int v) { return SendAll(&v, sizeof(v)) == sizeof(v); }
  bool RecvInt(int *v) { return RecvAll(v, sizeof(*v)) == sizeof(*v); }
  bool SendStr(const std::string &s) {
    int n = static_cast<int>(s.length());
    if (!SendInt(n)) return false;
    return n == 0 || SendAll(s.data(), s.length()) == s.length();
  }
  bool RecvStr(std::string *s) {
    int n = 0;
    if (!RecvInt(&n) || n < 0) return false;
    s->resize(n);
    return n == 0 || RecvAll(&(*s)[0], s->length()) == s->length();
  }

 private:
  void WaitFor(short ev) {  // NOLINT(*)
    pollfd p; p.fd = fd; p.events = ev; p.revents = 0;
    poll(&p, 1, 1000);
  }
};

/*! \brief poll(2) set keyed by fd */
class PollHelper {
 public:
  void WatchRead(int fd) { Slot(fd).events |= POLLIN; }
  void WatchWrite(int fd) { Slot(fd).events |= POLLOUT; }
  void WatchException(int fd) { Slot(fd).events |= POLLPRI; }
  /*! \brief wake up when the peer closes its end (no effect on what CheckRead reports for a socket not watched for reading) */
  void WatchHangup(int fd) { Slot(fd).events |= POLLRDHUP; }
  bool CheckHangup(int fd) const { return Revents(fd) & (POLLRDHUP | POLLHUP | POLLERR); }
  bool CheckRead(int fd) const { return Revents(fd) & (POLLIN | POLLHUP | POLLERR); }
  bool CheckWrite(int fd) const { return Revents(fd) & POLLOUT; }
  bool CheckExcept(int fd) const { return Revents(fd) & POLLPRI; }
  void Clear() { fds_.clear(); index_.clear(); }
  /*! \brief wait until something is ready; timeout_ms < 0 blocks */
  int Poll(int timeout_ms = -1) {
    int rc;
    do { rc = poll(fds_.data(), fds_.size(), timeout_ms); } while (rc < 0 && errno == EINTR);
    Check(rc >= 0, "poll(): %s", strerror(errno));
    return rc;
  }

 private:
  pollfd &Slot(int fd) {
    auto it = index_.find(fd);
    if (it == index_.end()) {
      pollfd p; p.fd = fd; p.events = 0; p.revents = 0;
      index_[fd] = fds_.size();
      fds_.push_back(p);
      return fds_.back();
    }
    return fds_[it->second];
  }
  short Revents(int fd) const {  // NOLINT(*)
    auto it = index_.find(fd);
    return it == index_.end() ? 0 : fds_[it->second].revents;
  }
  std::vector<pollfd> fds_;
  std::unordered_map<int, size_t> index_;
};

/*!
 * \brief unix-domain datagram endpoint in the abstract namespace, used to hand CUDA VMM
 *  file descriptors (SCM_RIGHTS) to the other per-GPU processes of the same host.
 *  No counterpart in the reference (it has no device memory to share).
 */
class FdChannel {
 public:
  FdChannel() : fd_(-1) {}
  ~FdChannel() { Close(); }
  void Close() { if (fd_ >= 0) { ::close(fd_); fd_ = -1; } }
  /*! \brief bind to "\0<name>" */
  bool Open(const std::string &name) {
    fd_ = socket(AF_UNIX, SOCK_DGRAM, 0);
    if (fd_ < 0) return false;
    sockaddr_un a; socklen_t len = MakeAddr(name, &a);
    return bind(fd_, reinterpret_cast<sockaddr *>(&a), len) == 0;
  }
  /*! \brief send one fd plus a 64-bit tag to the endpoint called name; retries while the peer is not bound yet */
  bool SendFd(const std::string &name, int fd, uint64_t tag, int timeout_ms = 30000) {
    sockaddr_un a; socklen_t alen = MakeAddr(name, &a);
    msghdr msg; std::memset(&msg, 0, sizeof(msg));
    iovec io; io.iov_base = &tag; io.iov_len = sizeof(tag);
    char ctl[CMSG_SPACE(sizeof(int))]; std::memset(ctl, 0, sizeof(ctl));
    msg.msg_name = &a; msg.msg_namelen = alen;
    msg.msg_iov = &io; msg.msg_iovlen = 1;
    msg.msg_control = ctl; msg.msg_controllen = sizeof(ctl);
    cmsghdr *c = CMSG_FIRSTHDR(&msg);
    c->cmsg_level = SOL_SOCKET; c->cmsg_type = SCM_RIGHTS; c->cmsg_len = CMSG_LEN(sizeof(int));
    std::memcpy(CMSG_DATA(c), &fd, sizeof(int));
    for (int waited = 0; waited <= timeout_ms; waited += 5) {
      if (sendmsg(fd_, &msg, 0) >= 0) return true;
      if (errno != ECONNREFUSED && errno != ENOENT && errno != EAGAIN && errno != EINTR) return false;
      usleep(5000);
    }
    return false;
  }
  /*! \brief receive one fd and its tag; returns -1 on timeout/error */
  int RecvFd(uint64_t *tag, int timeout_ms = 30000) {
    pollfd p; p.fd = fd_; p.events = POLLIN; p.revents = 0;
    int rc;
    do { rc = poll(&p, 1, timeout_ms); } while (rc < 0 && errno == EINTR);
    if (rc <= 0) return -1;
    msghdr msg; std::memset(&msg, 0, sizeof(msg));
    iovec io; io.iov_base = tag; io.iov_len = sizeof(*tag);
    char ctl[CMSG_SPACE(sizeof(int))]; std::memset(ctl, 0, sizeof(ctl));
    msg.msg_iov = &io; msg.msg_iovlen = 1;
    msg.msg_control = ctl; msg.msg_controllen = sizeof(ctl);
    if (recvmsg(fd_, &msg, 0) < 0) return -1;
    cmsghdr *c = CMSG_FIRSTHDR(&msg);
    if (c == nullptr || c->cmsg_type != SCM_RIGHTS) return -1;
    int fd; std::memcpy(&fd, CMSG_DATA(c), sizeof(int));
    return fd;
  }

 private:
  static socklen_t MakeAddr(const std::string &name, sockaddr_un *a) {
    std::memset(a, 0, sizeof(*a));
    a->sun_family = AF_UNIX;
    size_t n = std::min(name.length(), sizeof(a->sun_path) - 2);
    a->sun_path[0] = '\0';
    std::memcpy(a->sun_path + 1, name.data(), n);
    return static_cast<socklen_t>(offsetof(sockaddr_un, sun_path) + 1 + n);
  }
  int fd_;
};

}  // namespace utils
}  // namespace rabit
#endif  // RABIT_INTERNAL_SOCKET_H_
