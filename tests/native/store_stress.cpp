// Host-side stress of the rendezvous store, meant to be built with -fsanitize=thread or -fsanitize=address,undefined
// (tests/test_runtime_native.py::test_store_under_sanitizers does both).  The Open MPI tree the reference vendors has
// no sanitizer or race-detection target of its own (only the Valgrind memchecker hooks,
// ompi:opal/mca/memchecker/valgrind/memchecker_valgrind_module.c:36-42); this is the host half of that sub-system here,
// the device half being scripts/sanitize.sh.
//
// WORLD "ranks" run as threads of one process, each with its own client connection; rank 0 also hosts the server.
// Every rank: publishes values of growing size, reads every peer's values back and checks them byte for byte, crosses
// named barriers, and finally leaves in a staggered order with the HOST FIRST, which is the case the server's
// linger-until-clients-leave logic exists for.
//
//   usage: store_stress PORT [WORLD=4] [ROUNDS=40]
#include <arpa/inet.h>
#include <netinet/in.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/ewdml.h"

// csrc/runtime/core.cu owns the real one (it needs the CUDA runtime); the store only formats messages through it.
static thread_local char g_err[512];
void ewdml_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static std::atomic<int> g_fail{0};
#define CHECK(cond, ...)                                                   \
    do {                                                                   \
        if (!(cond)) {                                                     \
            fprintf(stderr, "rank %d line %d: ", rank, __LINE__);          \
            fprintf(stderr, __VA_ARGS__);                                  \
            fprintf(stderr, " (last error: %s)\n", g_err);                 \
            g_fail.fetch_add(1);                                           \
            return;                                                        \
        }                                                                  \
    } while (0)

static std::string value_of(int rank, int round) {
    // sizes sweep 0 B .. ~96 KiB so that both the first-guess buffer and the resize path are taken
    size_t n = (size_t)((round * 2477 + rank * 131) % 98304);
    if (round % 7 == 0) n = 0;
    std::string v(n, '\0');
    for (size_t i = 0; i < n; ++i) v[i] = (char)((i * 31 + (size_t)rank * 17 + (size_t)round) & 0xff);
    return v;
}

// A client that asks for a key nobody will ever set and then dies: the server thread serving it has to notice the hang-up,
// or the host's destroy() would wait out its whole linger period for a connection that is no longer there.
static void dead_waiter(int port) {
    int fd = socket(AF_INET, SOCK_STREAM, 0);
    sockaddr_in sa{};
    sa.sin_family = AF_INET; sa.sin_port = htons((uint16_t)port);
    inet_pton(AF_INET, "127.0.0.1", &sa.sin_addr);
    if (connect(fd, (sockaddr*)&sa, sizeof(sa)) == 0) {
        const char key[] = "never-set";
        uint8_t op = 2;                              // OP_GET
        uint32_t n = sizeof(key) - 1;
        (void)!send(fd, &op, 1, MSG_NOSIGNAL); (void)!send(fd, &n, 4, MSG_NOSIGNAL); (void)!send(fd, key, n, MSG_NOSIGNAL);
        std::this_thread::sleep_for(std::chrono::milliseconds(50));
    }
    close(fd);
}

static void run_rank(int rank, int world, int port, int rounds) {
    ewdml_store* st = nullptr;
    int rc = ewdml_store_create(&st, "127.0.0.1", port, rank, world, 30.0);
    CHECK(rc == 0 && st, "create rc=%d", rc);
    std::vector<char> buf(1024);
    if (rank == 1 || world == 1) dead_waiter(port);
    for (int r = 0; r < rounds; ++r) {
        std::string mine = value_of(rank, r);
        std::string key = "v/" + std::to_string(r) + "/" + std::to_string(rank);
        rc = ewdml_store_set(st, key.c_str(), mine.data(), mine.size());
        CHECK(rc == 0, "set rc=%d", rc);
        for (int p = 0; p < world; ++p) {
            int src = (rank + p) % world;
            std::string k2 = "v/" + std::to_string(r) + "/" + std::to_string(src);
            size_t len = 0;
            rc = ewdml_store_get(st, k2.c_str(), buf.data(), buf.size(), &len);   // blocks until src has published
            if (rc == EWDML_ERR_OVERFLOW) {                                        // same retry the Python binding does
                buf.resize(len);
                rc = ewdml_store_get(st, k2.c_str(), buf.data(), buf.size(), &len);
            }
            CHECK(rc == 0, "get rc=%d", rc);
            std::string want = value_of(src, r);
            CHECK(len == want.size() && (len == 0 || memcmp(buf.data(), want.data(), len) == 0), "value of rank %d round %d differs", src, r);
        }
        if (r % 4 == 0) {
            rc = ewdml_store_barrier(st, r % 8 == 0 ? "even" : "odd");
            CHECK(rc == 0, "barrier rc=%d", rc);
        }
    }
    rc = ewdml_store_barrier(st, "final");
    CHECK(rc == 0, "final barrier rc=%d", rc);
    // host leaves first; the others a little later, so its destroy() has live clients to wait for
    if (rank != 0) std::this_thread::sleep_for(std::chrono::milliseconds(20 * rank));
    auto t0 = std::chrono::steady_clock::now();
    rc = ewdml_store_destroy(st);
    double took = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    CHECK(rc == 0, "destroy rc=%d", rc);
    CHECK(took < 10.0, "destroy took %.1f s: the server waited for a connection that was gone", took);
}

int main(int argc, char** argv) {
    if (argc < 2) { fprintf(stderr, "usage: %s PORT [WORLD] [ROUNDS]\n", argv[0]); return 2; }
    int port = atoi(argv[1]);
    int world = argc > 2 ? atoi(argv[2]) : 4;
    int rounds = argc > 3 ? atoi(argv[3]) : 40;
    std::vector<std::thread> th;
    for (int r = 0; r < world; ++r) th.emplace_back(run_rank, r, world, port, rounds);
    for (auto& t : th) t.join();
    if (g_fail.load()) { fprintf(stderr, "FAILED: %d ranks\n", g_fail.load()); return 1; }
    printf("store_stress ok: world %d, %d rounds\n", world, rounds);
    return 0;
}
