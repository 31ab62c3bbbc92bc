// Minimal TCP key/value store used for rendezvous: rank 0 hosts it on
// MASTER_ADDR:MASTER_PORT, every rank connects, and the framework exchanges through it whatever the
// reference exchanges through torch's TCPStore (ref:pytorch/resnet/main.py:148 -> init_process_group):
// here that is the CUDA IPC handles of the symmetric arenas, the bucket-plan hash, and barriers.
//
// Wire format (little endian):  u8 op | u32 keylen | key | u64 arg | payload(arg bytes, SET only)
//   SET  -> u8 status
//   GET  -> waits server-side until the key exists or `arg` ms elapsed -> i64 len (-1 timeout) | payload
//   ADD  -> i64 new value (arg is the signed delta)
//   CHECK-> u8 exists
// One detached thread per connection; world sizes here are <= a few hundred.
#include <arpa/inet.h>
#include <errno.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <stdint.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/types.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

enum Op : uint8_t { OP_SET = 1, OP_GET = 2, OP_ADD = 3, OP_CHECK = 4 };

bool read_all(int fd, void* buf, size_t n) {
  char* p = static_cast<char*>(buf);
  while (n > 0) {
    ssize_t r = ::recv(fd, p, n, 0);
    if (r == 0) return false;
    if (r < 0) {
      if (errno == EINTR) continue;
      return false;
    }
    p += r;
    n -= static_cast<size_t>(r);
  }
  return true;
}

bool write_all(int fd, const void* buf, size_t n) {
  const char* p = static_cast<const char*>(buf);
  while (n > 0) {
    ssize_t r = ::send(fd, p, n, MSG_NOSIGNAL);
    if (r < 0) {
      if (errno == EINTR) continue;
      return false;
    }
    p += r;
    n -= static_cast<size_t>(r);
  }
  return true;
}

struct Server {
  int listen_fd = -1;
  int port = 0;
  std::atomic<bool> stop{false};
  std::thread accept_thread;
  std::mutex mu;
  std::condition_variable cv;
  std::map<std::string, std::string> kv;
  std::vector<int> conns;
  std::atomic<int> live{0};

  void serve(int fd) {
    live++;
    for (;;) {
      uint8_t op;
      uint32_t klen;
      uint64_t arg;
      if (!read_all(fd, &op, 1) || !read_all(fd, &klen, 4) || klen > (1u << 20)) break;
      std::string key(klen, '\0');
      if (klen && !read_all(fd, &key[0], klen)) break;
      if (!read_all(fd, &arg, 8)) break;
      if (op == OP_SET) {
        std::string val(arg, '\0');
        if (arg && !read_all(fd, &val[0], arg)) break;
        {
          std::lock_guard<std::mutex> lk(mu);
          kv[key] = std::move(val);
        }
        cv.notify_all();
        uint8_t ok = 0;
        if (!write_all(fd, &ok, 1)) break;
      } else if (op == OP_GET) {
        std::unique_lock<std::mutex> lk(mu);
        const bool found = cv.wait_for(lk, std::chrono::milliseconds(arg),
                                       [&] { return stop.load() || kv.find(key) != kv.end(); });
        if (!found || kv.find(key) == kv.end()) {
          lk.unlock();
          int64_t len = -1;
          if (!write_all(fd, &len, 8)) break;
        } else {
          std::string val = kv[key];
          lk.unlock();
          int64_t len = static_cast<int64_t>(val.size());
          if (!write_all(fd, &len, 8) || (len && !write_all(fd, val.data(), val.size()))) break;
        }
      } else if (op == OP_ADD) {
        int64_t nv;
        {
          std::lock_guard<std::mutex> lk(mu);
          int64_t cur = 0;
          auto it = kv.find(key);
          if (it != kv.end() && it->second.size() == 8) memcpy(&cur, it->second.data(), 8);
          nv = cur + static_cast<int64_t>(arg);
          kv[key] = std::string(reinterpret_cast<const char*>(&nv), 8);
        }
        cv.notify_all();
        if (!write_all(fd, &nv, 8)) break;
      } else if (op == OP_CHECK) {
        uint8_t ex;
        {
          std::lock_guard<std::mutex> lk(mu);
          ex = kv.find(key) != kv.end() ? 1 : 0;
        }
        if (!write_all(fd, &ex, 1)) break;
      } else {
        break;
      }
    }
    ::close(fd);
    live--;
  }

  void accept_loop() {
    while (!stop.load()) {
      sockaddr_in peer;
      socklen_t len = sizeof(peer);
      int fd = ::accept(listen_fd, reinterpret_cast<sockaddr*>(&peer), &len);
      if (fd < 0) {
        if (stop.load()) break;
        if (errno == EINTR || errno == EAGAIN) continue;
        break;
      }
      int one = 1;
      setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
      {
        std::lock_guard<std::mutex> lk(mu);
        conns.push_back(fd);
      }
      std::thread(&Server::serve, this, fd).detach();
    }
  }
};

struct Client {
  int fd = -1;
  std::mutex mu;
};

bool send_header(int fd, uint8_t op, const char* key, uint64_t arg) {
  uint32_t klen = static_cast<uint32_t>(strlen(key));
  return write_all(fd, &op, 1) && write_all(fd, &klen, 4) && write_all(fd, key, klen) && write_all(fd, &arg, 8);
}

}  // namespace

extern "C" void* dmpi_store_server_start(int port, int* out_port) {
  int fd = ::socket(AF_INET, SOCK_STREAM, 0);
  if (fd < 0) return nullptr;
  int one = 1;
  setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  sockaddr_in addr;
  memset(&addr, 0, sizeof(addr));
  addr.sin_family = AF_INET;
  addr.sin_addr.s_addr = htonl(INADDR_ANY);
  addr.sin_port = htons(static_cast<uint16_t>(port));
  if (::bind(fd, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) != 0 || ::listen(fd, 512) != 0) {
    ::close(fd);
    return nullptr;
  }
  socklen_t len = sizeof(addr);
  getsockname(fd, reinterpret_cast<sockaddr*>(&addr), &len);
  Server* s = new Server();
  s->listen_fd = fd;
  s->port = ntohs(addr.sin_port);
  if (out_port) *out_port = s->port;
  s->accept_thread = std::thread(&Server::accept_loop, s);
  return s;
}

extern "C" void dmpi_store_server_stop(void* handle) {
  Server* s = static_cast<Server*>(handle);
  if (!s) return;
  s->stop.store(true);
  ::shutdown(s->listen_fd, SHUT_RDWR);
  ::close(s->listen_fd);
  s->cv.notify_all();
  if (s->accept_thread.joinable()) s->accept_thread.join();
  {
    std::lock_guard<std::mutex> lk(s->mu);
    for (int fd : s->conns) ::shutdown(fd, SHUT_RDWR);
  }
  for (int i = 0; i < 200 && s->live.load() > 0; ++i) std::this_thread::sleep_for(std::chrono::milliseconds(5));
  if (s->live.load() == 0) delete s;  // otherwise leak: a detached thread may still touch it
}

extern "C" void* dmpi_store_connect(const char* host, int port, int timeout_ms) {
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
  char portstr[16];
  snprintf(portstr, sizeof(portstr), "%d", port);
  for (;;) {
    addrinfo hints;
    memset(&hints, 0, sizeof(hints));
    hints.ai_family = AF_INET;
    hints.ai_socktype = SOCK_STREAM;
    addrinfo* res = nullptr;
    if (getaddrinfo(host, portstr, &hints, &res) == 0 && res) {
      int fd = ::socket(res->ai_family, res->ai_socktype, res->ai_protocol);
      if (fd >= 0) {
        if (::connect(fd, res->ai_addr, res->ai_addrlen) == 0) {
          freeaddrinfo(res);
          int one = 1;
          setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
          Client* c = new Client();
          c->fd = fd;
          return c;
        }
        ::close(fd);
      }
      freeaddrinfo(res);
    }
    if (std::chrono::steady_clock::now() >= deadline) return nullptr;
    std::this_thread::sleep_for(std::chrono::milliseconds(50));
  }
}

extern "C" void dmpi_store_close(void* handle) {
  Client* c = static_cast<Client*>(handle);
  if (!c) return;
  ::shutdown(c->fd, SHUT_RDWR);
  ::close(c->fd);
  delete c;
}

extern "C" int dmpi_store_set(void* handle, const char* key, const void* data, long long len) {
  Client* c = static_cast<Client*>(handle);
  std::lock_guard<std::mutex> lk(c->mu);
  uint8_t st;
  if (!send_header(c->fd, OP_SET, key, static_cast<uint64_t>(len))) return -1;
  if (len && !write_all(c->fd, data, static_cast<size_t>(len))) return -1;
  return read_all(c->fd, &st, 1) ? 0 : -1;
}

// returns the value length, -1 on timeout / error, -2 when the buffer is too small (value is drained)
extern "C" long long dmpi_store_get(void* handle, const char* key, void* buf, long long buflen, int timeout_ms) {
  Client* c = static_cast<Client*>(handle);
  std::lock_guard<std::mutex> lk(c->mu);
  if (!send_header(c->fd, OP_GET, key, static_cast<uint64_t>(timeout_ms))) return -1;
  int64_t len;
  if (!read_all(c->fd, &len, 8)) return -1;
  if (len < 0) return -1;
  if (len > buflen) {
    std::vector<char> sink(static_cast<size_t>(len));
    read_all(c->fd, sink.data(), sink.size());
    return -2;
  }
  if (len && !read_all(c->fd, buf, static_cast<size_t>(len))) return -1;
  return len;
}

extern "C" long long dmpi_store_add(void* handle, const char* key, long long delta) {
  Client* c = static_cast<Client*>(handle);
  std::lock_guard<std::mutex> lk(c->mu);
  if (!send_header(c->fd, OP_ADD, key, static_cast<uint64_t>(delta))) return INT64_MIN;
  int64_t nv;
  return read_all(c->fd, &nv, 8) ? nv : INT64_MIN;
}

extern "C" int dmpi_store_check(void* handle, const char* key) {
  Client* c = static_cast<Client*>(handle);
  std::lock_guard<std::mutex> lk(c->mu);
  if (!send_header(c->fd, OP_CHECK, key, 0)) return -1;
  uint8_t ex;
  return read_all(c->fd, &ex, 1) ? ex : -1;
}
