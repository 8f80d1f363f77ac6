// N4: local launcher + watchdog.
//
// Starts `nproc` ranks of a command on this node with the env contract the trainer reads
// (RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT, distributed_nn.py:75-78) and watches
// them: when any rank exits non-zero or dies on a signal, the rest of the job is torn down
// (SIGTERM, then SIGKILL after a grace period) instead of dead-locking on the next collective,
// which is what happens in the reference when a worker disappears.  This is the
// role of mpirun = orterun + plm/odls (fork/exec of local ranks) + errmgr/default_hnp (abort the
// job on ABORTED / COMM_FAILED, ompi:orte/mca/errmgr/default_hnp/errmgr_default_hnp.c:352-703),
// reduced to the single-node case this framework targets.
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
    for (int r = 0; r < nproc; ++r) {
        pid_t pid = fork();
        if (pid < 0) { ewdml_set_error("launch: fork failed: %s", strerror(errno)); return EWDML_ERR_STATE; }
        if (pid == 0) {
            setpgid(0, 0);
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
        pids[r] = pid;
    }
    int alive = nproc, job_status = 0;
    bool tearing_down = false;
    auto kill_deadline = std::chrono::steady_clock::time_point::max();
    while (alive > 0) {
        int st = 0;
        pid_t pid = waitpid(-1, &st, WNOHANG);
        if (pid == 0) {
            if (tearing_down && std::chrono::steady_clock::now() > kill_deadline) {
                for (int r = 0; r < nproc; ++r) if (pids[r] > 0) kill(-pids[r], SIGKILL);
                kill_deadline = std::chrono::steady_clock::time_point::max();
            }
            std::this_thread::sleep_for(std::chrono::milliseconds(20));
            continue;
        }
        if (pid < 0) break;
        int rank = -1;
        for (int r = 0; r < nproc; ++r) if (pids[r] == pid) rank = r;
        if (rank < 0) continue;
        pids[rank] = -1;
        --alive;
        int code = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + (WIFSIGNALED(st) ? WTERMSIG(st) : 0);
        if (code != 0 && !tearing_down) {
            job_status = code;
            if (failed_rank) *failed_rank = rank;
            fprintf(stderr, "[ewdml-launch] rank %d exited with status %d; tearing the job down\n", rank, code);
            tearing_down = true;
            for (int r = 0; r < nproc; ++r) if (pids[r] > 0) kill(-pids[r], SIGTERM);
            kill_deadline = std::chrono::steady_clock::now() +
                            std::chrono::duration_cast<std::chrono::steady_clock::duration>(
                                std::chrono::duration<double>(grace_s > 0 ? grace_s : 5.0));
        }
    }
    return job_status;
}
