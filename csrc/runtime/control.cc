// TCP control plane + ring all-reduce for CPU workers.
//
// Stands in for what TensorFlow starts behind MultiWorkerMirroredStrategy
// (/root/reference/README.md:21-23,65-68): a server per task on the task's TF_CONFIG address,
// peers that wait for each other before training starts, the `RING` collective (ring
// all-reduce over the same TCP ports), and shutdown of the service when training ends.
// Design differences from TF's gRPC runtime: one listener per task and three kinds of plain TCP
// connections, each introduced by a 16-byte hello {magic, rank, channel, world}:
//   ring      rank r -> rank (r+1)%W       bulk gradient traffic (reduce-scatter + all-gather)
//   control   rank r -> leader (rank 0)    barrier / all-gather / broadcast of small blobs
//                                          (rendezvous, CUDA IPC handle exchange, metric sums)
//   heartbeat rank r -> leader             liveness; the leader fans a failure notice out
// Every blocking wait is a poll() loop bounded by the group timeout and by the failure flag,
// so a dead peer turns into an error that names the rank instead of a hang.
#include <arpa/inet.h>
#include <errno.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <signal.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <thread>

#include "rt_common.h"
#include "../../include/tfdl_runtime.h"

using namespace tfdl_rt;
using Clock = std::chrono::steady_clock;

namespace {

constexpr uint32_t kMagic = 0x7F0D1B20u;
enum Channel : uint32_t { CH_RING = 1, CH_CTRL = 2, CH_HB = 3 };

struct Hello {
  uint32_t magic, rank, channel, world;
};

int64_t now_ms() {
  return std::chrono::duration_cast<std::chrono::milliseconds>(Clock::now().time_since_epoch()).count();
}

void set_nodelay(int fd) {
  int one = 1;
  setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
  int buf = 4 << 20;
  setsockopt(fd, SOL_SOCKET, SO_SNDBUF, &buf, sizeof(buf));
  setsockopt(fd, SOL_SOCKET, SO_RCVBUF, &buf, sizeof(buf));
}
void set_nonblock(int fd, bool on) {
  int fl = fcntl(fd, F_GETFL, 0);
  fcntl(fd, F_SETFL, on ? (fl | O_NONBLOCK) : (fl & ~O_NONBLOCK));
}

bool split_addr(const std::string& a, std::string* host, int* port) {
  size_t c = a.rfind(':');
  if (c == std::string::npos) return false;
  *host = a.substr(0, c);
  *port = atoi(a.c_str() + c + 1);
  return *port > 0 && *port < 65536;
}

struct Group {
  int rank = 0, world = 1;
  int timeout_ms = 60000;
  int hb_interval_ms = 200, hb_timeout_ms = 5000;
  std::vector<std::string> addrs;
  int listen_fd = -1;
  int ring_send = -1, ring_recv = -1;
  std::vector<int> ctrl;  // leader: per rank; worker: [0] = leader
  std::vector<int> hb;    // same indexing
  std::thread hb_thread;
  std::atomic<int> failed_rank{-1};
  uint64_t ring_seq = 0;  // collectives issued on the ring so far (header check in ring_allreduce)
  // host-visible word that device kernels poll inside their spin loops (pinned, device-mapped memory owned by the
  // caller): the heartbeat thread stores failed_rank + 1 there the moment a peer is declared dead, which releases every
  // kernel waiting for that peer's gradients instead of letting it spin into its trap limit
  std::atomic<uint32_t*> abort_word{nullptr};
  std::atomic<bool> stop{false};
  std::mutex ctrl_mu;
  std::vector<char> scratch;
  uint64_t bytes_sent = 0, bytes_recv = 0;

  // Send half of a large ring step on a second thread: at these sizes a step is two memory copies into / out of the
  // socket buffers plus the accumulation, and one thread doing all three is what bounds the bandwidth. The helper is
  // created on the first large collective and parked on a condition variable between steps.
  struct Sender {
    std::thread th;
    std::mutex mu;
    std::condition_variable cv;
    const char* ptr = nullptr;
    size_t len = 0;
    int peer = 0;
    bool has_job = false, done = true, quit = false;
    int rc = 0;
    uint64_t sent = 0;
    char err[512] = "";
    std::atomic<bool> abort{false};
  } sender;

  ~Group() { shutdown_all(); }

  void shutdown_all() {
    stop = true;
    if (sender.th.joinable()) {
      {
        std::lock_guard<std::mutex> lk(sender.mu);
        sender.quit = true;
        sender.abort = true;
      }
      sender.cv.notify_all();
      sender.th.join();
    }
    if (hb_thread.joinable()) hb_thread.join();
    auto cl = [](int& fd) {
      if (fd >= 0) {
        ::shutdown(fd, SHUT_RDWR);
        ::close(fd);
        fd = -1;
      }
    };
    cl(ring_send);
    cl(ring_recv);
    for (auto& f : ctrl) cl(f);
    for (auto& f : hb) cl(f);
    cl(listen_fd);
  }

  bool check_failed(const char* what) {
    int f = failed_rank.load();
    if (f >= 0) {
      set_error("%s aborted: rank %d (%s) is unreachable or has exited", what, f,
                f < (int)addrs.size() ? addrs[f].c_str() : "?");
      return true;
    }
    return false;
  }

  // Blocking exact-size IO with timeout + failure awareness. peer is only for the message.
  int io_all(int fd, void* buf, size_t n, bool write, int peer, const char* what) {
    char* p = static_cast<char*>(buf);
    int64_t deadline = now_ms() + timeout_ms;
    while (n) {
      if (check_failed(what)) return -2;
      pollfd pf{fd, (short)(write ? POLLOUT : POLLIN), 0};
      int pr = poll(&pf, 1, 100);
      if (pr < 0 && errno == EINTR) continue;
      if (pr == 0) {
        if (now_ms() > deadline) {
          set_error("%s timed out after %d ms waiting for rank %d (%s)", what, timeout_ms, peer,
                    peer >= 0 && peer < (int)addrs.size() ? addrs[peer].c_str() : "?");
          return -3;
        }
        continue;
      }
      ssize_t r = write ? ::send(fd, p, n, MSG_NOSIGNAL) : ::recv(fd, p, n, 0);
      if (r > 0) {
        p += r;
        n -= r;
        (write ? bytes_sent : bytes_recv) += r;
        deadline = now_ms() + timeout_ms;
      } else if (r == 0 || (errno != EAGAIN && errno != EWOULDBLOCK && errno != EINTR)) {
        if (failed_rank.load() < 0) failed_rank = peer;
        set_error("%s: connection to rank %d (%s) lost (%s)", what, peer,
                  peer >= 0 && peer < (int)addrs.size() ? addrs[peer].c_str() : "?",
                  r == 0 ? "closed by peer" : strerror(errno));
        return -2;
      }
    }
    return 0;
  }
};

int connect_retry(Group* g, int peer, Channel ch) {
  std::string host;
  int port;
  if (!split_addr(g->addrs[peer], &host, &port)) {
    set_error("bad address %s", g->addrs[peer].c_str());
    return -1;
  }
  addrinfo hints{}, *res = nullptr;
  hints.ai_family = AF_INET;
  hints.ai_socktype = SOCK_STREAM;
  char ps[16];
  snprintf(ps, sizeof(ps), "%d", port);
  int64_t deadline = now_ms() + g->timeout_ms;
  while (true) {
    if (getaddrinfo(host.c_str(), ps, &hints, &res) == 0 && res) {
      int fd = socket(AF_INET, SOCK_STREAM, 0);
      if (fd >= 0) {
        // non-blocking connect bounded by the rendezvous deadline: a blocking connect() to an unroutable peer can
        // sit in the kernel for minutes, far past timeout_ms
        set_nonblock(fd, true);
        int cr = connect(fd, res->ai_addr, res->ai_addrlen);
        if (cr != 0 && errno == EINPROGRESS) {
          int64_t left = deadline - now_ms();
          pollfd pf{fd, POLLOUT, 0};
          int pr = poll(&pf, 1, (int)std::max<int64_t>(1, std::min<int64_t>(left, 1000)));
          int soerr = 0;
          socklen_t sl = sizeof(soerr);
          if (pr > 0 && getsockopt(fd, SOL_SOCKET, SO_ERROR, &soerr, &sl) == 0 && soerr == 0) cr = 0;
        }
        if (cr == 0) {
          set_nonblock(fd, false);
          freeaddrinfo(res);
          set_nodelay(fd);
          Hello h{kMagic, (uint32_t)g->rank, (uint32_t)ch, (uint32_t)g->world};
          if (::send(fd, &h, sizeof(h), MSG_NOSIGNAL) == (ssize_t)sizeof(h)) return fd;
          ::close(fd);
        } else {
          ::close(fd);
        }
      }
      freeaddrinfo(res);
      res = nullptr;
    }
    if (now_ms() > deadline) {
      set_error("rendezvous: rank %d could not reach rank %d at %s within %d ms (is that task running?)", g->rank,
                peer, g->addrs[peer].c_str(), g->timeout_ms);
      return -1;
    }
    std::this_thread::sleep_for(std::chrono::milliseconds(50));
  }
}

void heartbeat_loop(Group* g) {
  const bool leader = g->rank == 0;
  std::vector<int64_t> last(g->world, now_ms());
  std::vector<bool> quit(g->world, false);
  std::vector<bool> want_rank(g->world, false);  // an 'F' was the last byte of a recv: its rank byte comes next
  int64_t last_ping = 0;
  int announced = -1;
  while (!g->stop.load()) {
    int64_t t = now_ms();
    if (int f = g->failed_rank.load(); f >= 0) {
      if (uint32_t* w = g->abort_word.load()) *reinterpret_cast<volatile uint32_t*>(w) = (uint32_t)f + 1u;
    }
    if (!leader && t - last_ping >= g->hb_interval_ms) {
      char c = 'P';
      if (::send(g->hb[0], &c, 1, MSG_NOSIGNAL | MSG_DONTWAIT) < 0 && errno != EAGAIN && errno != EWOULDBLOCK) {
        if (g->failed_rank.load() < 0 && !g->stop.load()) g->failed_rank = 0;
      }
      last_ping = t;
    }
    std::vector<pollfd> pfs;
    std::vector<int> who;
    if (leader) {
      for (int r = 1; r < g->world; ++r)
        if (g->hb[r] >= 0 && !quit[r]) {
          pfs.push_back({g->hb[r], POLLIN, 0});
          who.push_back(r);
        }
    } else {
      pfs.push_back({g->hb[0], POLLIN, 0});
      who.push_back(0);
    }
    if (pfs.empty()) {
      std::this_thread::sleep_for(std::chrono::milliseconds(50));
      continue;
    }
    int pr = poll(pfs.data(), pfs.size(), 50);
    if (pr > 0) {
      for (size_t i = 0; i < pfs.size(); ++i) {
        if (!(pfs[i].revents & (POLLIN | POLLHUP | POLLERR))) continue;
        char buf[64];
        ssize_t r = ::recv(pfs[i].fd, buf, sizeof(buf), MSG_DONTWAIT);
        int peer = who[i];
        if (r > 0) {
          last[peer] = now_ms();
          for (ssize_t k = 0; k < r; ++k) {
            if (want_rank[peer]) {  // rank byte of a failure notice whose 'F' ended the previous recv
              want_rank[peer] = false;
              if (g->failed_rank.load() < 0) g->failed_rank = (unsigned char)buf[k];
              continue;
            }
            if (buf[k] == 'Q') quit[peer] = true;  // graceful goodbye
            if (buf[k] == 'F') {                   // failure notice from the leader: next byte = rank (groups are <= 255 ranks)
              if (k + 1 < r) {
                if (g->failed_rank.load() < 0) g->failed_rank = (unsigned char)buf[k + 1];
                ++k;
              } else {
                want_rank[peer] = true;
              }
            }
          }
        } else if (r == 0 || (errno != EAGAIN && errno != EWOULDBLOCK && errno != EINTR)) {
          if (!quit[peer] && !g->stop.load() && g->failed_rank.load() < 0) g->failed_rank = peer;
          quit[peer] = true;
        }
      }
    }
    if (leader) {
      t = now_ms();
      for (int r = 1; r < g->world; ++r)
        if (!quit[r] && t - last[r] > g->hb_timeout_ms && g->failed_rank.load() < 0) g->failed_rank = r;
      int f = g->failed_rank.load();
      if (f >= 0 && announced != f) {
        announced = f;
        char msg[2] = {'F', (char)f};
        for (int r = 1; r < g->world; ++r)
          if (r != f && g->hb[r] >= 0) ::send(g->hb[r], msg, 2, MSG_NOSIGNAL | MSG_DONTWAIT);
      }
    }
  }
  // graceful goodbye so the other side does not flag us as failed
  char q = 'Q';
  if (leader) {
    for (int r = 1; r < g->world; ++r)
      if (g->hb[r] >= 0) ::send(g->hb[r], &q, 1, MSG_NOSIGNAL | MSG_DONTWAIT);
  } else if (g->hb[0] >= 0) {
    ::send(g->hb[0], &q, 1, MSG_NOSIGNAL | MSG_DONTWAIT);
  }
}

}  // namespace

extern "C" void* tfdl_rt_group_create(int rank, int world, const char* const* addrs, int timeout_ms,
                                      int hb_interval_ms, int hb_timeout_ms) {
  signal(SIGPIPE, SIG_IGN);
  if (world > 255) {  // the heartbeat's failure notice carries the rank in one byte
    set_error("a control group holds at most 255 tasks, got %d", world);
    return nullptr;
  }
  if (world < 1 || rank < 0 || rank >= world) {
    set_error("group_create: rank %d out of range for world %d", rank, world);
    return nullptr;
  }
  auto* g = new Group();
  g->rank = rank;
  g->world = world;
  if (timeout_ms > 0) g->timeout_ms = timeout_ms;
  if (hb_interval_ms > 0) g->hb_interval_ms = hb_interval_ms;
  if (hb_timeout_ms > 0) g->hb_timeout_ms = hb_timeout_ms;
  for (int i = 0; i < world; ++i) g->addrs.emplace_back(addrs[i]);
  g->ctrl.assign(world, -1);
  g->hb.assign(world, -1);
  if (world == 1) return g;

  std::string host;
  int port;
  if (!split_addr(g->addrs[rank], &host, &port)) {
    set_error("group_create: bad address %s", g->addrs[rank].c_str());
    delete g;
    return nullptr;
  }
  g->listen_fd = socket(AF_INET, SOCK_STREAM, 0);
  int one = 1;
  setsockopt(g->listen_fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  sockaddr_in sa{};
  sa.sin_family = AF_INET;
  sa.sin_addr.s_addr = htonl(INADDR_ANY);
  sa.sin_port = htons(port);
  if (bind(g->listen_fd, (sockaddr*)&sa, sizeof(sa)) != 0 || listen(g->listen_fd, 4 * world + 8) != 0) {
    set_error("group_create: cannot listen on port %d of %s: %s", port, g->addrs[rank].c_str(), strerror(errno));
    delete g;
    return nullptr;
  }
  // outgoing
  g->ring_send = connect_retry(g, (rank + 1) % world, CH_RING);
  if (g->ring_send < 0) { delete g; return nullptr; }
  if (rank != 0) {
    g->ctrl[0] = connect_retry(g, 0, CH_CTRL);
    g->hb[0] = g->ctrl[0] >= 0 ? connect_retry(g, 0, CH_HB) : -1;
    if (g->ctrl[0] < 0 || g->hb[0] < 0) { delete g; return nullptr; }
  }
  // incoming
  int expect = 1 + (rank == 0 ? 2 * (world - 1) : 0);
  int64_t deadline = now_ms() + g->timeout_ms;
  while (expect > 0) {
    pollfd pf{g->listen_fd, POLLIN, 0};
    int pr = poll(&pf, 1, 100);
    if (pr <= 0) {
      if (now_ms() > deadline) {
        set_error("rendezvous: rank %d timed out after %d ms waiting for %d more peer connection(s)", rank,
                  g->timeout_ms, expect);
        delete g;
        return nullptr;
      }
      continue;
    }
    int fd = accept(g->listen_fd, nullptr, nullptr);
    if (fd < 0) continue;
    set_nodelay(fd);
    Hello h{};
    timeval tv{5, 0};
    setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
    if (::recv(fd, &h, sizeof(h), MSG_WAITALL) != (ssize_t)sizeof(h) || h.magic != kMagic) {
      ::close(fd);
      continue;
    }
    timeval tv0{0, 0};
    setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv0, sizeof(tv0));
    if ((int)h.world != world || (int)h.rank >= world) {
      set_error("rendezvous: peer rank %u reports a cluster of %u tasks but this task has %d: "
                "TF_CONFIG 'cluster' must be identical on every node", h.rank, h.world, world);
      ::close(fd);
      delete g;
      return nullptr;
    }
    if (h.channel == CH_RING && g->ring_recv < 0 && (int)h.rank == (rank + world - 1) % world) {
      g->ring_recv = fd;
    } else if (h.channel == CH_CTRL && rank == 0 && g->ctrl[h.rank] < 0) {
      g->ctrl[h.rank] = fd;
    } else if (h.channel == CH_HB && rank == 0 && g->hb[h.rank] < 0) {
      g->hb[h.rank] = fd;
    } else {
      ::close(fd);
      continue;
    }
    --expect;
  }
  g->hb_thread = std::thread(heartbeat_loop, g);
  return g;
}

extern "C" void tfdl_rt_group_destroy(void* h) { delete static_cast<Group*>(h); }
extern "C" int tfdl_rt_group_rank(void* h) { return static_cast<Group*>(h)->rank; }
extern "C" int tfdl_rt_group_world(void* h) { return static_cast<Group*>(h)->world; }
extern "C" int tfdl_rt_group_failed_rank(void* h) { return static_cast<Group*>(h)->failed_rank.load(); }
extern "C" void tfdl_rt_group_set_abort_word(void* h, uint32_t* word) {
  auto* g = static_cast<Group*>(h);
  g->abort_word.store(word);
  if (int f = g->failed_rank.load(); f >= 0 && word) *reinterpret_cast<volatile uint32_t*>(word) = (uint32_t)f + 1u;
}
extern "C" void tfdl_rt_group_set_timeout(void* h, int ms) { static_cast<Group*>(h)->timeout_ms = ms; }
extern "C" uint64_t tfdl_rt_group_bytes_sent(void* h) { return static_cast<Group*>(h)->bytes_sent; }

// All-gather of equal-sized blobs through the leader. out must hold world * nbytes.
extern "C" int tfdl_rt_allgather(void* h, const void* in, uint64_t nbytes, void* out) {
  auto* g = static_cast<Group*>(h);
  char* o = static_cast<char*>(out);
  if (g->world == 1) {
    if (nbytes) memcpy(o, in, nbytes);
    return 0;
  }
  std::lock_guard<std::mutex> lk(g->ctrl_mu);
  uint64_t hdr = nbytes;
  if (g->rank == 0) {
    if (nbytes) memcpy(o, in, nbytes);
    for (int r = 1; r < g->world; ++r) {
      uint64_t got = 0;
      if (int rc = g->io_all(g->ctrl[r], &got, 8, false, r, "allgather")) return rc;
      if (got != nbytes) {
        set_error("allgather: rank %d sent %llu bytes, expected %llu", r, (unsigned long long)got,
                  (unsigned long long)nbytes);
        return -1;
      }
      if (nbytes)
        if (int rc = g->io_all(g->ctrl[r], o + r * nbytes, nbytes, false, r, "allgather")) return rc;
    }
    for (int r = 1; r < g->world; ++r) {
      if (int rc = g->io_all(g->ctrl[r], &hdr, 8, true, r, "allgather")) return rc;
      if (nbytes)
        if (int rc = g->io_all(g->ctrl[r], o, nbytes * g->world, true, r, "allgather")) return rc;
    }
  } else {
    if (int rc = g->io_all(g->ctrl[0], &hdr, 8, true, 0, "allgather")) return rc;
    if (nbytes)
      if (int rc = g->io_all(g->ctrl[0], const_cast<void*>(in), nbytes, true, 0, "allgather")) return rc;
    uint64_t back = 0;
    if (int rc = g->io_all(g->ctrl[0], &back, 8, false, 0, "allgather")) return rc;
    if (nbytes)
      if (int rc = g->io_all(g->ctrl[0], o, nbytes * g->world, false, 0, "allgather")) return rc;
  }
  return 0;
}

extern "C" int tfdl_rt_barrier(void* h) { return tfdl_rt_allgather(h, nullptr, 0, nullptr); }

// Broadcast from rank 0 (the chief: initial variable values, README.md:51 / tf_dist_example.py:56-57).
extern "C" int tfdl_rt_broadcast0(void* h, void* buf, uint64_t nbytes) {
  auto* g = static_cast<Group*>(h);
  if (g->world == 1 || nbytes == 0) return 0;
  std::lock_guard<std::mutex> lk(g->ctrl_mu);
  if (g->rank == 0) {
    for (int r = 1; r < g->world; ++r)
      if (int rc = g->io_all(g->ctrl[r], buf, nbytes, true, r, "broadcast")) return rc;
  } else {
    if (int rc = g->io_all(g->ctrl[0], buf, nbytes, false, 0, "broadcast")) return rc;
  }
  return 0;
}

// ---- helper thread for the send half of large ring steps ---------------------------------------------------
static void sender_loop(Group* g) {
  Group::Sender& s = g->sender;
  for (;;) {
    const char* p;
    size_t left;
    int peer;
    {
      std::unique_lock<std::mutex> lk(s.mu);
      s.cv.wait(lk, [&] { return s.has_job || s.quit; });
      if (s.quit) return;
      p = s.ptr;
      left = s.len;
      peer = s.peer;
    }
    int rc = 0;
    uint64_t sent = 0;
    char err[512] = "";
    int64_t deadline = now_ms() + g->timeout_ms;
    while (left && rc == 0) {
      if (s.abort.load()) { rc = -5; break; }  // the receiving side gave up: its error is the one that is reported
      if (int f = g->failed_rank.load(); f >= 0) {
        snprintf(err, sizeof err, "ring all-reduce aborted: rank %d (%s) is unreachable or has exited", f,
                 f < (int)g->addrs.size() ? g->addrs[f].c_str() : "?");
        rc = -2;
        break;
      }
      pollfd pf{g->ring_send, POLLOUT, 0};
      int pr = poll(&pf, 1, 100);
      if (pr < 0 && errno == EINTR) continue;
      if (pr == 0) {
        if (now_ms() > deadline) {
          snprintf(err, sizeof err, "ring all-reduce timed out after %d ms (sending to rank %d)", g->timeout_ms, peer);
          rc = -3;
        }
        continue;
      }
      ssize_t w = ::send(g->ring_send, p, left, MSG_NOSIGNAL);
      if (w > 0) {
        p += w;
        left -= (size_t)w;
        sent += (uint64_t)w;
        deadline = now_ms() + g->timeout_ms;
      } else if (w < 0 && errno != EAGAIN && errno != EWOULDBLOCK && errno != EINTR) {
        if (g->failed_rank.load() < 0) g->failed_rank = peer;
        snprintf(err, sizeof err, "ring all-reduce: connection to rank %d (%s) lost: %s", peer, g->addrs[peer].c_str(),
                 strerror(errno));
        rc = -2;
      }
    }
    {
      std::lock_guard<std::mutex> lk(s.mu);
      s.rc = rc;
      s.sent = sent;
      memcpy(s.err, err, sizeof err);
      s.has_job = false;
      s.done = true;
    }
    s.cv.notify_all();
  }
}
static void sender_start(Group* g, const char* p, size_t len, int peer) {
  Group::Sender& s = g->sender;
  if (!s.th.joinable()) s.th = std::thread(sender_loop, g);
  {
    std::lock_guard<std::mutex> lk(s.mu);
    s.ptr = p;
    s.len = len;
    s.peer = peer;
    s.abort = false;
    s.done = false;
    s.has_job = true;
  }
  s.cv.notify_all();
}
// Waits for the helper's step; gives up = true makes it stop early. Returns its rc (error text already set).
static int sender_join(Group* g, bool give_up) {
  Group::Sender& s = g->sender;
  if (give_up) s.abort = true;
  std::unique_lock<std::mutex> lk(s.mu);
  s.cv.wait(lk, [&] { return s.done; });
  g->bytes_sent += s.sent;
  s.sent = 0;
  if (s.rc != 0 && s.rc != -5 && !give_up) set_error("%s", s.err);
  return s.rc;
}
// Below this size a step is latency, not copy bandwidth (TFDL_RING_THREAD_BYTES overrides; 0 = never use the helper).
static size_t threaded_step_bytes() {
  static const size_t v = [] {
    if (const char* e = getenv("TFDL_RING_THREAD_BYTES")) {
      const long long b = atoll(e);
      return b <= 0 ? (size_t)-1 : (size_t)b;
    }
    return (size_t)(1 << 20);
  }();
  return v;
}

// Ring all-reduce (sum): W-1 reduce-scatter steps then W-1 all-gather steps; each step sends one
// segment to the next rank while receiving one from the previous rank, both non-blocking under a
// single poll() loop, and the incoming bytes are accumulated as they arrive. From 1 MB per step the
// send half moves to a helper thread (sender_loop above).
template <class T>
static int ring_allreduce(Group* g, T* buf, int64_t n) {
  const int W = g->world, R = g->rank;
  if (W == 1 || n == 0) return 0;
  std::vector<int64_t> off(W + 1);
  for (int i = 0; i <= W; ++i) off[i] = n * i / W;
  size_t maxseg = 0;
  for (int i = 0; i < W; ++i) maxseg = std::max<size_t>(maxseg, (off[i + 1] - off[i]) * sizeof(T));
  if (g->scratch.size() < maxseg) g->scratch.resize(maxseg);
  const int prev = (R + W - 1) % W, next = (R + 1) % W;
  // Every ring collective starts with a header {magic, element size, count, sequence number} to the next rank, checked
  // against the previous rank's: workers that fell out of step (one left its epoch early and started the float64 metric
  // reduction while its peers are still in the float32 gradient reduction) fail here with a message instead of consuming
  // each other's bytes as data.
  {
    struct Hdr { uint32_t magic, elem; int64_t n; uint64_t seq; } mine{0x52444654u, (uint32_t)sizeof(T), n, g->ring_seq++}, theirs{};
    if (int hrc = g->io_all(g->ring_send, &mine, sizeof(mine), true, next, "ring all-reduce header")) return hrc;
    if (int hrc = g->io_all(g->ring_recv, &theirs, sizeof(theirs), false, prev, "ring all-reduce header")) return hrc;
    if (theirs.magic != mine.magic || theirs.elem != mine.elem || theirs.n != mine.n || theirs.seq != mine.seq) {
      set_error("ring all-reduce out of step: this rank (%d) is in collective #%llu (%lld elements of %u bytes) but rank %d "
                "sent the header of collective #%llu (%lld elements of %u bytes) - the workers ran different numbers of steps",
                R, (unsigned long long)mine.seq, (long long)mine.n, mine.elem, prev, (unsigned long long)theirs.seq,
                (long long)theirs.n, theirs.elem);
      return -4;
    }
  }
  set_nonblock(g->ring_send, true);
  set_nonblock(g->ring_recv, true);
  int rc = 0;
  for (int phase = 0; phase < 2 && rc == 0; ++phase) {
    for (int s = 0; s < W - 1 && rc == 0; ++s) {
      int send_seg = phase == 0 ? (R - s + W) % W : (R + 1 - s + W) % W;
      int recv_seg = phase == 0 ? (R - s - 1 + 2 * W) % W : (R - s + W) % W;
      const char* sp = reinterpret_cast<const char*>(buf + off[send_seg]);
      size_t sleft = (off[send_seg + 1] - off[send_seg]) * sizeof(T);
      T* dst = buf + off[recv_seg];
      size_t rtotal = (off[recv_seg + 1] - off[recv_seg]) * sizeof(T), rgot = 0, rdone_elems = 0;
      char* rp = phase == 0 ? g->scratch.data() : reinterpret_cast<char*>(dst);
      int64_t deadline = now_ms() + g->timeout_ms;
      const bool threaded = sleft >= threaded_step_bytes();
      if (threaded) {  // the helper sends this step's segment; this thread receives and accumulates
        sender_start(g, sp, sleft, next);
        sleft = 0;
      }
      while ((sleft || rgot < rtotal) && rc == 0) {
        if (g->check_failed("ring all-reduce")) { rc = -2; break; }
        pollfd pf[2] = {{g->ring_send, (short)(sleft ? POLLOUT : 0), 0}, {g->ring_recv, (short)(rgot < rtotal ? POLLIN : 0), 0}};
        int pr = poll(pf, 2, 100);
        if (pr < 0 && errno == EINTR) continue;
        if (pr == 0) {
          if (now_ms() > deadline) {
            set_error("ring all-reduce timed out after %d ms (waiting on rank %d / rank %d)", g->timeout_ms, prev, next);
            rc = -3;
          }
          continue;
        }
        if (sleft && (pf[0].revents & (POLLOUT | POLLERR | POLLHUP))) {
          ssize_t w = ::send(g->ring_send, sp, sleft, MSG_NOSIGNAL);
          if (w > 0) { sp += w; sleft -= w; g->bytes_sent += w; deadline = now_ms() + g->timeout_ms; }
          else if (w < 0 && errno != EAGAIN && errno != EWOULDBLOCK && errno != EINTR) {
            if (g->failed_rank.load() < 0) g->failed_rank = next;
            set_error("ring all-reduce: connection to rank %d (%s) lost: %s", next, g->addrs[next].c_str(), strerror(errno));
            rc = -2;
          }
        }
        if (rgot < rtotal && (pf[1].revents & (POLLIN | POLLERR | POLLHUP))) {
          ssize_t r = ::recv(g->ring_recv, rp + rgot, rtotal - rgot, 0);
          if (r > 0) {
            rgot += r; g->bytes_recv += r; deadline = now_ms() + g->timeout_ms;
            if (phase == 0) {
              size_t elems = rgot / sizeof(T);
              const T* src = reinterpret_cast<const T*>(g->scratch.data());
              for (size_t i = rdone_elems; i < elems; ++i) dst[i] += src[i];
              rdone_elems = elems;
            }
          } else if (r == 0 || (errno != EAGAIN && errno != EWOULDBLOCK && errno != EINTR)) {
            if (g->failed_rank.load() < 0) g->failed_rank = prev;
            set_error("ring all-reduce: connection from rank %d (%s) lost", prev, g->addrs[prev].c_str());
            rc = -2;
          }
        }
      }
      if (threaded) {
        const int src = sender_join(g, rc != 0);
        if (rc == 0 && src != 0) rc = src;
      }
    }
  }
  set_nonblock(g->ring_send, false);
  set_nonblock(g->ring_recv, false);
  return rc;
}

extern "C" int tfdl_rt_ring_allreduce_f32(void* h, float* buf, int64_t n) {
  return ring_allreduce(static_cast<Group*>(h), buf, n);
}
extern "C" int tfdl_rt_ring_allreduce_f64(void* h, double* buf, int64_t n) {
  return ring_allreduce(static_cast<Group*>(h), buf, n);
}
