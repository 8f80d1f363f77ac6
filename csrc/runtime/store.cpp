// N1: rendezvous key-value store over TCP (rank 0 hosts a server thread).
//
// What the job needs at start-up -- exchange of IPC handles and device pointers, and a host
// barrier -- is what Open MPI gets from the PMIx "modex" + fence (ompi:ompi/runtime/
// ompi_mpi_init.c:657-724) over ORTE's out-of-band TCP tree (orte/mca/oob/tcp).  This is the
// minimal equivalent: set / blocking get / barrier, length-prefixed binary protocol.
#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/time.h>
#include <unistd.h>

#include <atomic>
#include <cerrno>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/ewdml.h"

void ewdml_set_error(const char* fmt, ...);

namespace {

// Largest value the store carries (optimizer state of a VGG-size model under Adam is ~80 MB; the length prefix is 32 bits).
constexpr uint32_t kMaxBlob = 1u << 30;

bool send_all(int fd, const void* buf, size_t n) {
    const char* p = (const char*)buf;
    while (n) {
        ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
        if (k <= 0) return false;
        p += k; n -= (size_t)k;
    }
    return true;
}
bool recv_all(int fd, void* buf, size_t n) {
    char* p = (char*)buf;
    while (n) {
        ssize_t k = ::recv(fd, p, n, 0);
        if (k <= 0) return false;
        p += k; n -= (size_t)k;
    }
    return true;
}
bool send_blob(int fd, const std::string& s) {
    uint32_t n = (uint32_t)s.size();
    return send_all(fd, &n, 4) && (n == 0 || send_all(fd, s.data(), n));
}
bool recv_blob(int fd, std::string* s) {
    uint32_t n = 0;
    if (!recv_all(fd, &n, 4)) return false;
    if (n > kMaxBlob) return false;
    s->resize(n);
    return n == 0 || recv_all(fd, &(*s)[0], n);
}

enum : uint8_t { OP_SET = 1, OP_GET = 2, OP_ADD = 3 };

// Has the other end closed?  (Requests are strictly one at a time per connection, so while a reply is owed nothing but a
// hang-up can arrive.)
bool peer_gone(int fd) {
    pollfd p{fd, (short)(POLLIN | POLLRDHUP), 0};
    if (::poll(&p, 1, 0) <= 0) return false;
    if (p.revents & (POLLRDHUP | POLLHUP | POLLERR | POLLNVAL)) return true;
    if (p.revents & POLLIN) {
        char c;
        return ::recv(fd, &c, 1, MSG_PEEK | MSG_DONTWAIT) == 0;
    }
    return false;
}

struct Server {
    int listen_fd = -1;
    std::thread accept_thread;
    std::vector<std::thread> workers;
    std::mutex mu;
    std::condition_variable cv;
    std::map<std::string, std::string> kv;
    std::map<std::string, int64_t> counters;
    std::atomic<bool> stop{false};
    std::atomic<int> active{0};        // connections being served (the hosting rank's own included)
    std::vector<int> fds;

    void serve(int fd) {
        for (;;) {
            uint8_t op;
            std::string key, val;
            if (!recv_all(fd, &op, 1) || !recv_blob(fd, &key)) break;
            if (op == OP_SET) {
                if (!recv_blob(fd, &val)) break;
                { std::lock_guard<std::mutex> lk(mu); kv[key] = val; }
                cv.notify_all();
                uint8_t ok = 1;
                if (!send_all(fd, &ok, 1)) break;
            } else if (op == OP_GET) {
                std::unique_lock<std::mutex> lk(mu);
                // Wait for the key, but look at the connection now and then: a client that died while blocked here would
                // otherwise keep this thread (and the server's shutdown, which waits for its connections) until the end.
                bool gone = false;
                while (!stop.load() && !kv.count(key) && !gone) {
                    cv.wait_for(lk, std::chrono::milliseconds(250));
                    if (stop.load() || kv.count(key)) break;
                    lk.unlock();
                    gone = peer_gone(fd);
                    lk.lock();
                }
                if (!kv.count(key)) break;
                std::string v = kv[key];
                lk.unlock();
                if (!send_blob(fd, v)) break;
            } else if (op == OP_ADD) {
                int64_t delta = 0, now = 0;
                if (!recv_all(fd, &delta, 8)) break;
                { std::lock_guard<std::mutex> lk(mu); now = (counters[key] += delta); kv["__ctr__" + key + "#" + std::to_string(now)] = "1"; }
                cv.notify_all();
                if (!send_all(fd, &now, 8)) break;
            } else {
                break;
            }
        }
        // Forget the descriptor before closing it: the number may be handed to an unrelated socket of this process right
        // away, and destroy() shuts down everything still listed.
        {
            std::lock_guard<std::mutex> lk(mu);
            for (size_t i = 0; i < fds.size(); ++i)
                if (fds[i] == fd) { fds[i] = fds.back(); fds.pop_back(); break; }
            ::close(fd);
        }
        active.fetch_sub(1);
    }

    void accept_loop() {
        while (!stop.load()) {
            pollfd p{listen_fd, POLLIN, 0};
            if (::poll(&p, 1, 100) <= 0) continue;
            int fd = ::accept(listen_fd, nullptr, nullptr);
            if (fd < 0) continue;
            int one = 1;
            setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
            std::lock_guard<std::mutex> lk(mu);
            fds.push_back(fd);
            active.fetch_add(1);
            workers.emplace_back([this, fd] { serve(fd); });
        }
    }
};

}  // namespace

struct ewdml_store {
    int rank = 0, world = 1;
    int fd = -1;
    Server* server = nullptr;
    std::mutex mu;
    std::map<std::string, int> barrier_epoch;
};

extern "C" int ewdml_store_create(ewdml_store** out, const char* addr, int port, int rank, int world, double timeout_s) {
    if (!out || !addr || port <= 0 || rank < 0 || world < 1 || rank >= world) { ewdml_set_error("store_create: bad args"); return EWDML_ERR_ARG; }
    ewdml_store* s = new ewdml_store();
    s->rank = rank; s->world = world;
    if (rank == 0) {
        Server* sv = new Server();
        sv->listen_fd = ::socket(AF_INET, SOCK_STREAM, 0);
        int one = 1;
        setsockopt(sv->listen_fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
        sockaddr_in sa{};
        sa.sin_family = AF_INET; sa.sin_port = htons((uint16_t)port);
        // Listen only where the job says the master lives (MASTER_ADDR; loopback for single-node jobs) instead of on every
        // interface: the store is unauthenticated, so it should not be reachable from networks the job does not use.
        // EWDML_STORE_BIND_ANY=1 restores the wildcard (multi-homed masters whose MASTER_ADDR is not a local address).
        const char* any = getenv("EWDML_STORE_BIND_ANY");
        bool bound = false;
        if (!(any && atoi(any)) && inet_pton(AF_INET, addr, &sa.sin_addr) == 1)
            bound = ::bind(sv->listen_fd, (sockaddr*)&sa, sizeof(sa)) == 0;
        if (!bound) {
            sa.sin_addr.s_addr = htonl(INADDR_ANY);
            bound = ::bind(sv->listen_fd, (sockaddr*)&sa, sizeof(sa)) == 0;
        }
        if (!bound || ::listen(sv->listen_fd, 128) != 0) {
            ewdml_set_error("store: cannot listen on port %d: %s", port, strerror(errno));
            ::close(sv->listen_fd); delete sv; delete s;
            return EWDML_ERR_IO;
        }
        sv->accept_thread = std::thread([sv] { sv->accept_loop(); });
        s->server = sv;
    }
    auto deadline = std::chrono::steady_clock::now() + std::chrono::duration<double>(timeout_s > 0 ? timeout_s : 60.0);
    for (;;) {
        s->fd = ::socket(AF_INET, SOCK_STREAM, 0);
        sockaddr_in sa{};
        sa.sin_family = AF_INET; sa.sin_port = htons((uint16_t)port);
        if (inet_pton(AF_INET, addr, &sa.sin_addr) != 1) {
            ewdml_set_error("store: address must be an IPv4 literal, got %s", addr);
            ::close(s->fd); ewdml_store_destroy(s);
            return EWDML_ERR_ARG;
        }
        if (::connect(s->fd, (sockaddr*)&sa, sizeof(sa)) == 0) break;
        ::close(s->fd); s->fd = -1;
        if (std::chrono::steady_clock::now() > deadline) {
            ewdml_set_error("store: rank %d could not reach %s:%d", rank, addr, port);
            ewdml_store_destroy(s);
            return EWDML_ERR_TIMEOUT;
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(50));
    }
    int one = 1;
    setsockopt(s->fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
    // A blocking get / barrier waits for a peer.  A peer that died never arrives, so the wait has a deadline: after
    // EWDML_STORE_TIMEOUT_S (default 1800 s: longer than any legitimate skew between ranks, e.g. a server waiting for its
    // workers' last step) the call fails with EWDML_ERR_TIMEOUT instead of hanging the job.
    const char* te = getenv("EWDML_STORE_TIMEOUT_S");
    double wait_s = te ? atof(te) : 1800.0;
    if (wait_s > 0) {
        timeval tv;
        tv.tv_sec = (time_t)wait_s;
        tv.tv_usec = (suseconds_t)((wait_s - (double)tv.tv_sec) * 1e6);
        setsockopt(s->fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
    }
    *out = s;
    return 0;
}

static int io_error(const char* what, const char* key) {
    if (errno == EAGAIN || errno == EWOULDBLOCK) {
        ewdml_set_error("%s(%s): timed out waiting for a peer (EWDML_STORE_TIMEOUT_S); the store connection is now unusable", what, key);
        return EWDML_ERR_TIMEOUT;
    }
    ewdml_set_error("%s(%s): connection lost", what, key);
    return EWDML_ERR_IO;
}

extern "C" int ewdml_store_set(ewdml_store* s, const char* key, const void* val, size_t len) {
    if (!s || !key) return EWDML_ERR_ARG;
    if (len > kMaxBlob) {
        ewdml_set_error("store_set(%s): value of %zu bytes exceeds the store limit of %u bytes", key, len, kMaxBlob);
        return EWDML_ERR_OVERFLOW;
    }
    std::lock_guard<std::mutex> lk(s->mu);
    uint8_t op = OP_SET, ok = 0;
    errno = 0;
    if (!send_all(s->fd, &op, 1) || !send_blob(s->fd, key) || !send_blob(s->fd, std::string((const char*)val, len)) ||
        !recv_all(s->fd, &ok, 1)) return io_error("store_set", key);
    return 0;
}

extern "C" int ewdml_store_get(ewdml_store* s, const char* key, void* val, size_t cap, size_t* len) {
    if (!s || !key) return EWDML_ERR_ARG;
    std::lock_guard<std::mutex> lk(s->mu);
    uint8_t op = OP_GET;
    std::string v;
    errno = 0;
    if (!send_all(s->fd, &op, 1) || !send_blob(s->fd, key) || !recv_blob(s->fd, &v)) return io_error("store_get", key);
    if (len) *len = v.size();
    if (v.size() > cap) { ewdml_set_error("store_get(%s): value of %zu bytes exceeds buffer %zu", key, v.size(), cap); return EWDML_ERR_OVERFLOW; }
    if (val && !v.empty()) memcpy(val, v.data(), v.size());
    return 0;
}

extern "C" int ewdml_store_barrier(ewdml_store* s, const char* name) {
    if (!s || !name) return EWDML_ERR_ARG;
    int epoch;
    { std::lock_guard<std::mutex> lk(s->mu); epoch = ++s->barrier_epoch[name]; }
    std::string key = std::string("__bar__") + name;
    {
        std::lock_guard<std::mutex> lk(s->mu);
        uint8_t op = OP_ADD;
        int64_t one = 1, now = 0;
        errno = 0;
        if (!send_all(s->fd, &op, 1) || !send_blob(s->fd, key) || !send_all(s->fd, &one, 8) || !recv_all(s->fd, &now, 8))
            return io_error("store_barrier", name);
    }
    std::string wait_key = "__ctr__" + key + "#" + std::to_string((int64_t)epoch * s->world);
    return ewdml_store_get(s, wait_key.c_str(), nullptr, 16, nullptr);
}

extern "C" int ewdml_store_destroy(ewdml_store* s) {
    if (!s) return 0;
    if (s->fd >= 0) { ::shutdown(s->fd, SHUT_RDWR); ::close(s->fd); }
    if (s->server) {
        Server* sv = s->server;
        // The hosting rank usually leaves a final barrier first (its reply travels over loopback) and closes right away.
        // Stopping the server then would cut the replies still owed to the other ranks ("connection lost" in THEIR barrier,
        // seen once in a 4-GPU test run), so the server lingers until every client has disconnected -- or 20 s, for clients
        // that died without closing.
        for (int waited_ms = 0; sv->active.load() > 0 && waited_ms < 20000; waited_ms += 5)
            std::this_thread::sleep_for(std::chrono::milliseconds(5));
        // stop is set under the mutex the waiters check their predicate under, so no waiter can miss the wake-up.
        { std::lock_guard<std::mutex> lk(sv->mu); sv->stop.store(true); for (int fd : sv->fds) ::shutdown(fd, SHUT_RDWR); }
        sv->cv.notify_all();
        if (sv->accept_thread.joinable()) sv->accept_thread.join();
        for (auto& t : sv->workers) if (t.joinable()) t.join();
        ::close(sv->listen_fd);
        delete sv;
    }
    delete s;
    return 0;
}
