// N4: local launcher + watchdog.
//
// Starts `nproc` ranks of a command on this node with the env contract the trainer reads
// (RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT, distributed_nn.py:75-78) and watches
// them: when any rank exits non-zero or dies on a signal, the rest of the job is torn down
// (SIGTERM, then SIGKILL after a grace period) instead of dead-locking on the next collective,
// which is what happens in the reference when a worker disappears.  This is the
// role of mpirun = orterun + plm/odls (fork/exec of local ranks) + errmgr/default_hnp (abort the
// job on ABORTED / COMM_FAILED, ompi:orte/mca/errmgr/default_hnp/errmgr_default_hnp.c:352-703),
// reduced to the single-node case this framework targets.  Like mpirun it also passes SIGINT / SIGTERM / SIGHUP on:
// every rank lives in its own process group (so a rank's own children die with it), which also means the terminal's
// Ctrl-C no longer reaches the ranks by itself -- the launcher catches the signal and tears the job down.
#include <errno.h>
#include <signal.h>
#include <sys/wait.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/ewdml.h"

void ewdml_set_error(const char* fmt, ...);

namespace {

volatile sig_atomic_t g_signal = 0;
void on_signal(int sig) { g_signal = sig; }

// Signal a rank's whole process group; the group may not exist yet (the child has not reached setpgid) or the leader
// may have changed it, so fall back to the process itself.
void signal_rank(pid_t pid, int sig) {
    if (pid <= 0) return;
    if (kill(-pid, sig) != 0) kill(pid, sig);
}

}  // namespace

extern "C" int ewdml_launch(int nproc, const char* master_addr, int master_port, char* const argv[], double grace_s,
                            int* failed_rank) {
    if (nproc < 1 || !argv || !argv[0]) { ewdml_set_error("launch: bad args"); return EWDML_ERR_ARG; }
    if (failed_rank) *failed_rank = -1;
    std::vector<pid_t> pids(nproc, -1);
    // multi-node: NODE_RANK / NNODES in the launcher's environment (set by `ewdml.cli.launch --node-rank/--nnodes` or
    // tools/cluster.py) place this node's ranks at node_rank * nproc + local_rank in a world of nnodes * nproc
    const char* nr = getenv("NODE_RANK");
    const char* nn = getenv("NNODES");
    const int node_rank = nr ? atoi(nr) : 0, nnodes = (nn && atoi(nn) > 0) ? atoi(nn) : 1;
    if (node_rank < 0 || node_rank >= nnodes) { ewdml_set_error("launch: NODE_RANK %d outside NNODES %d", node_rank, nnodes); return EWDML_ERR_ARG; }

    // our handlers for the duration of the call; whatever the host program (Python) had installed is restored on return
    const int forwarded[3] = {SIGINT, SIGTERM, SIGHUP};
    struct sigaction old_sa[3], sa;
    memset(&sa, 0, sizeof(sa));
    sa.sa_handler = on_signal;
    sigemptyset(&sa.sa_mask);
    g_signal = 0;
    for (int i = 0; i < 3; ++i) sigaction(forwarded[i], &sa, &old_sa[i]);
    auto restore = [&] { for (int i = 0; i < 3; ++i) sigaction(forwarded[i], &old_sa[i], nullptr); };

    const auto grace = std::chrono::duration_cast<std::chrono::steady_clock::duration>(
        std::chrono::duration<double>(grace_s > 0 ? grace_s : 5.0));
    int started = 0;
    bool fork_failed = false;
    for (int r = 0; r < nproc; ++r) {
        pid_t pid = fork();
        if (pid < 0) { ewdml_set_error("launch: fork failed for rank %d: %s", r, strerror(errno)); fork_failed = true; break; }
        if (pid == 0) {
            setpgid(0, 0);
            for (int i = 0; i < 3; ++i) signal(forwarded[i], SIG_DFL);
            setenv("RANK", std::to_string(node_rank * nproc + r).c_str(), 1);
            setenv("LOCAL_RANK", std::to_string(r).c_str(), 1);
            setenv("WORLD_SIZE", std::to_string(nnodes * nproc).c_str(), 1);
            setenv("GROUP_RANK", std::to_string(node_rank).c_str(), 1);
            setenv("LOCAL_WORLD_SIZE", std::to_string(nproc).c_str(), 1);
            setenv("MASTER_ADDR", master_addr ? master_addr : "127.0.0.1", 1);
            setenv("MASTER_PORT", std::to_string(master_port).c_str(), 1);
            execvp(argv[0], argv);
            fprintf(stderr, "[ewdml-launch] exec %s failed: %s\n", argv[0], strerror(errno));
            _exit(127);
        }
        setpgid(pid, pid);      // also from this side: the group exists before anybody can want to signal it
        pids[r] = pid;
        ++started;
    }

    int alive = started, job_status = fork_failed ? EWDML_ERR_STATE : 0;
    bool tearing_down = false;
    auto kill_deadline = std::chrono::steady_clock::time_point::max();
    auto tear_down = [&] {
        tearing_down = true;
        for (int r = 0; r < nproc; ++r) signal_rank(pids[r], SIGTERM);
        kill_deadline = std::chrono::steady_clock::now() + grace;
    };
    if (fork_failed && alive > 0) tear_down();      // a partial job would only block in its first collective
    while (alive > 0) {
        if (g_signal && !tearing_down) {
            fprintf(stderr, "[ewdml-launch] caught signal %d; tearing the job down\n", (int)g_signal);
            if (job_status == 0) job_status = 128 + (int)g_signal;
            tear_down();
        }
        // Only this job's ranks are waited for (never waitpid(-1)): the caller may have children of its own.
        bool reaped = false;
        for (int rank = 0; rank < nproc; ++rank) {
            if (pids[rank] <= 0) continue;
            int st = 0;
            pid_t got = waitpid(pids[rank], &st, WNOHANG);
            if (got == 0) continue;
            if (got < 0 && errno == EINTR) continue;
            // got < 0 otherwise: somebody else reaped it (ECHILD); count it as gone with an unknown status
            int code = got < 0 ? 0 : (WIFEXITED(st) ? WEXITSTATUS(st) : 128 + (WIFSIGNALED(st) ? WTERMSIG(st) : 0));
            pid_t pid = pids[rank];
            pids[rank] = -1;
            --alive;
            reaped = true;
            if (tearing_down) kill(-pid, SIGKILL);      // what the rank itself started goes with it
            if (code != 0 && !tearing_down) {
                job_status = code;
                if (failed_rank) *failed_rank = rank;
                fprintf(stderr, "[ewdml-launch] rank %d exited with status %d; tearing the job down\n", rank, code);
                tear_down();
            }
        }
        if (reaped) continue;
        if (tearing_down && std::chrono::steady_clock::now() > kill_deadline) {
            for (int r = 0; r < nproc; ++r) signal_rank(pids[r], SIGKILL);
            kill_deadline = std::chrono::steady_clock::time_point::max();
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(20));
    }
    restore();
    return job_status;
}
