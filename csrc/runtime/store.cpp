// Control-plane key-value store over TCP: what c10d::TCPStore is to the reference (rendezvous.py:198-287,
// multigpu.py:30-32) - rank 0 hosts it, every rank connects, and the bootstrap exchanges small blobs through it
// (CUDA-IPC handles of the symmetric buffers, bucket orders, barriers).  No torch, no Python in the data path.
//
//   server : one thread, poll() over the listening socket and all clients, whole requests handled one at a time
//   client : blocking request / response; get() polls until the key exists or the timeout expires
//   wire   : request  = u8 op | u32 key_len | key | u64 val_len | val
//            response = i64 status (>= 0: payload length or counter value, -1: no such key) | payload
#include <arpa/inet.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/time.h>
#include <unistd.h>

#include <atomic>
#include <cerrno>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

enum : uint8_t { OP_SET = 1, OP_GET = 2, OP_ADD = 3, OP_DEL = 4, OP_NKEYS = 5 };

bool read_all(int fd, void* buf, size_t n) {
  auto* p = static_cast<uint8_t*>(buf);
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
  auto* p = static_cast<const uint8_t*>(buf);
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
  std::thread thread;
  std::atomic<bool> stop{false};
  std::map<std::string, std::vector<uint8_t>> kv;

  bool handle(int fd) {
    uint8_t op;
    uint32_t klen;
    uint64_t vlen;
    if (!read_all(fd, &op, 1) || !read_all(fd, &klen, 4)) return false;
    if (klen > (1u << 20)) return false;
    std::string key(klen, '\0');
    if (klen && !read_all(fd, &key[0], klen)) return false;
    if (!read_all(fd, &vlen, 8)) return false;
    if (vlen > (1ull << 30)) return false;
    std::vector<uint8_t> val(vlen);
    if (vlen && !read_all(fd, val.data(), vlen)) return false;
    int64_t status = 0;
    const std::vector<uint8_t>* payload = nullptr;
    switch (op) {
      case OP_SET:
        kv[key] = std::move(val);
        break;
      case OP_GET: {
        auto it = kv.find(key);
        if (it == kv.end()) status = -1;
        else { status = static_cast<int64_t>(it->second.size()); payload = &it->second; }
        break;
      }
      case OP_ADD: {
        int64_t delta = 0;
        if (val.size() == 8) memcpy(&delta, val.data(), 8);
        int64_t cur = 0;
        auto it = kv.find(key);
        if (it != kv.end() && it->second.size() == 8) memcpy(&cur, it->second.data(), 8);
        cur += delta;
        std::vector<uint8_t> v(8);
        memcpy(v.data(), &cur, 8);
        kv[key] = std::move(v);
        status = cur;
        break;
      }
      case OP_DEL:
        status = static_cast<int64_t>(kv.erase(key));
        break;
      case OP_NKEYS:
        status = static_cast<int64_t>(kv.size());
        break;
      default:
        return false;
    }
    if (!write_all(fd, &status, 8)) return false;
    if (payload && !payload->empty() && !write_all(fd, payload->data(), payload->size())) return false;
    return true;
  }

  void run() {
    std::vector<pollfd> fds;
    fds.push_back({listen_fd, POLLIN, 0});
    while (!stop.load()) {
      int n = ::poll(fds.data(), fds.size(), 50);
      if (n <= 0) continue;
      if (fds[0].revents & POLLIN) {
        int c = ::accept(listen_fd, nullptr, nullptr);
        if (c >= 0) {
          int one = 1;
          setsockopt(c, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
          // a client that stalls in the middle of a request must not freeze the server (one thread serves every rank):
          // its blocking reads give up after two seconds and the connection is dropped
          timeval tv{2, 0};
          setsockopt(c, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
          fds.push_back({c, POLLIN, 0});
        }
      }
      for (size_t i = 1; i < fds.size();) {
        bool drop = false;
        if (fds[i].revents & (POLLHUP | POLLERR | POLLNVAL)) drop = !(fds[i].revents & POLLIN);
        if (!drop && (fds[i].revents & POLLIN)) drop = !handle(fds[i].fd);
        if (drop) {
          ::close(fds[i].fd);
          fds.erase(fds.begin() + static_cast<long>(i));
        } else {
          ++i;
        }
      }
    }
    for (size_t i = 1; i < fds.size(); ++i) ::close(fds[i].fd);
  }
};

struct Client {
  int fd = -1;
  std::mutex mu;

  // returns status; fills out (if not null) with the payload
  int64_t request(uint8_t op, const char* key, const void* val, uint64_t vlen, std::vector<uint8_t>* out) {
    std::lock_guard<std::mutex> lock(mu);
    const uint32_t klen = static_cast<uint32_t>(strlen(key));
    std::vector<uint8_t> msg(1 + 4 + klen + 8 + vlen);
    size_t o = 0;
    msg[o++] = op;
    memcpy(&msg[o], &klen, 4); o += 4;
    memcpy(&msg[o], key, klen); o += klen;
    memcpy(&msg[o], &vlen, 8); o += 8;
    if (vlen) memcpy(&msg[o], val, vlen);
    if (!write_all(fd, msg.data(), msg.size())) return -2;
    int64_t status;
    if (!read_all(fd, &status, 8)) return -2;
    if (op == OP_GET && status > 0) {
      std::vector<uint8_t> buf(static_cast<size_t>(status));
      if (!read_all(fd, buf.data(), buf.size())) return -2;
      if (out) *out = std::move(buf);
    }
    return status;
  }
};

}  // namespace

// ------------------------------------------------------------------------------------------------ C ABI
// Errors: -1 = no such key / timeout, -2 = connection problem, -3 = buffer too small (needed size is returned negated - 16)
// bind_host: dotted IPv4 address to listen on (the job's MASTER_ADDR; "127.0.0.1" for the single-host jobs this framework
// runs).  NULL / "" / "0.0.0.0" listens on every interface.  The store is unauthenticated and its Python client unpickles
// what it reads, so it should never be reachable from outside the job: the default of the Python side is the loopback.
extern "C" void* ddpx_store_server_start_on(const char* bind_host, int port, int* bound_port) {
  int fd = ::socket(AF_INET, SOCK_STREAM, 0);
  if (fd < 0) return nullptr;
  int one = 1;
  setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_addr.s_addr = htonl(INADDR_ANY);
  if (bind_host != nullptr && bind_host[0] != 0) {
    in_addr a{};
    if (inet_pton(AF_INET, bind_host, &a) == 1) addr.sin_addr = a;
    else { ::close(fd); return nullptr; }
  }
  addr.sin_port = htons(static_cast<uint16_t>(port));
  if (::bind(fd, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) != 0 || ::listen(fd, 128) != 0) {
    ::close(fd);
    return nullptr;
  }
  socklen_t len = sizeof(addr);
  getsockname(fd, reinterpret_cast<sockaddr*>(&addr), &len);
  auto* s = new Server();
  s->listen_fd = fd;
  s->port = ntohs(addr.sin_port);
  if (bound_port) *bound_port = s->port;
  s->thread = std::thread([s] { s->run(); });
  return s;
}

extern "C" void* ddpx_store_server_start(int port, int* bound_port) {
  return ddpx_store_server_start_on(nullptr, port, bound_port);
}

extern "C" void ddpx_store_server_stop(void* h) {
  auto* s = static_cast<Server*>(h);
  if (!s) return;
  s->stop.store(true);
  if (s->thread.joinable()) s->thread.join();
  ::close(s->listen_fd);
  delete s;
}

extern "C" void* ddpx_store_connect(const char* host, int port, int timeout_ms) {
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
  addrinfo hints{};
  hints.ai_family = AF_INET;
  hints.ai_socktype = SOCK_STREAM;
  const std::string ports = std::to_string(port);
  do {
    addrinfo* res = nullptr;
    if (getaddrinfo(host, ports.c_str(), &hints, &res) == 0 && res) {
      int fd = ::socket(res->ai_family, res->ai_socktype, res->ai_protocol);
      if (fd >= 0) {
        if (::connect(fd, res->ai_addr, res->ai_addrlen) == 0) {
          int one = 1;
          setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
          freeaddrinfo(res);
          auto* c = new Client();
          c->fd = fd;
          return c;
        }
        ::close(fd);
      }
      freeaddrinfo(res);
    }
    std::this_thread::sleep_for(std::chrono::milliseconds(20));     // the server may not be listening yet
  } while (std::chrono::steady_clock::now() < deadline);
  return nullptr;
}

extern "C" void ddpx_store_close(void* h) {
  auto* c = static_cast<Client*>(h);
  if (!c) return;
  ::close(c->fd);
  delete c;
}

extern "C" long long ddpx_store_set(void* h, const char* key, const void* data, long long len) {
  return static_cast<Client*>(h)->request(OP_SET, key, data, static_cast<uint64_t>(len), nullptr);
}

// Blocks (polling) until the key exists or timeout_ms passed.  Returns the value length, copies min(len, cap) bytes.
extern "C" long long ddpx_store_get(void* h, const char* key, void* buf, long long cap, int timeout_ms) {
  auto* c = static_cast<Client*>(h);
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
  int backoff_us = 200;
  for (;;) {
    std::vector<uint8_t> out;
    const int64_t st = c->request(OP_GET, key, nullptr, 0, &out);
    if (st >= 0) {
      if (st > cap) return -16 - st;
      if (st > 0) memcpy(buf, out.data(), static_cast<size_t>(st));
      return st;
    }
    if (st == -2) return -2;
    if (std::chrono::steady_clock::now() >= deadline) return -1;
    std::this_thread::sleep_for(std::chrono::microseconds(backoff_us));
    if (backoff_us < 5000) backoff_us *= 2;
  }
}

extern "C" long long ddpx_store_add(void* h, const char* key, long long delta) {
  int64_t d = delta;
  return static_cast<Client*>(h)->request(OP_ADD, key, &d, 8, nullptr);
}

extern "C" long long ddpx_store_delete(void* h, const char* key) {
  return static_cast<Client*>(h)->request(OP_DEL, key, nullptr, 0, nullptr);
}

extern "C" long long ddpx_store_num_keys(void* h) {
  return static_cast<Client*>(h)->request(OP_NKEYS, "", nullptr, 0, nullptr);
}
