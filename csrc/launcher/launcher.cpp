// detb200-launch: native single-box process launcher + worker watchdog + rank log prefixer.
//
// One binary replaces the reference's launch chain
//   pid_server -> torch.distributed.run -> pid_client -> wrap_rank -> script
// (harness/determined/launch/torch_distributed.py:15-106, ipc.py:272-562 PIDServer/PIDClient,
//  launch/wrap_rank.py:103-187): it spawns N ranks with the torchrun environment contract
// (RANK, LOCAL_RANK, WORLD_SIZE, LOCAL_WORLD_SIZE, GROUP_RANK, GROUP_WORLD_SIZE, MASTER_ADDR,
// MASTER_PORT), prefixes every output line with "[rank=N] ", and watches the workers through
// pidfds in the same epoll loop: the first worker that dies abnormally takes the whole group down
// (SIGTERM, then SIGKILL after a grace period) and its exit status becomes ours — there is no
// in-job elasticity, recovery is a restart from the last checkpoint (max_restarts 0 semantics).
//
// usage: detb200-launch --nproc N [--master-addr A] [--master-port P] [--grace SECONDS]
//                       [--no-prefix] [--node-rank R --nnodes M] -- cmd [args...]
#include <errno.h>
#include <fcntl.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/epoll.h>
#include <sys/syscall.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

#include <string>
#include <vector>

#ifndef SYS_pidfd_open
#define SYS_pidfd_open 434
#endif

struct Worker {
  pid_t pid = -1;
  int out_fd = -1;   // read end of the worker's stdout+stderr pipe
  int pidfd = -1;
  bool exited = false;
  int status = 0;
  std::string partial;
};

static std::vector<Worker> g_workers;
static volatile sig_atomic_t g_signal = 0;

static void on_signal(int sig) { g_signal = sig; }

static double now_s() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec + ts.tv_nsec * 1e-9;
}

static void write_all(int fd, const char* p, size_t n) {
  while (n > 0) {
    ssize_t w = write(fd, p, n);
    if (w < 0) {
      if (errno == EINTR) continue;
      return;
    }
    p += w;
    n -= (size_t)w;
  }
}

static void flush_lines(Worker& w, int rank, bool prefix, bool final_flush) {
  size_t start = 0;
  for (;;) {
    size_t nl = w.partial.find('\n', start);
    if (nl == std::string::npos) break;
    std::string line;
    if (prefix) line = "[rank=" + std::to_string(rank) + "] ";
    line.append(w.partial, start, nl - start + 1);
    write_all(STDOUT_FILENO, line.data(), line.size());
    start = nl + 1;
  }
  w.partial.erase(0, start);
  if (final_flush && !w.partial.empty()) {
    std::string line;
    if (prefix) line = "[rank=" + std::to_string(rank) + "] ";
    line += w.partial + "\n";
    write_all(STDOUT_FILENO, line.data(), line.size());
    w.partial.clear();
  }
}

static void signal_all(int sig) {
  for (auto& w : g_workers)
    if (w.pid > 0 && !w.exited) kill(w.pid, sig);
}

int main(int argc, char** argv) {
  int nproc = 1, nnodes = 1, node_rank = 0;
  std::string master_addr = "127.0.0.1", master_port = "29400";
  double grace = 10.0;
  bool prefix = true;
  int i = 1;
  for (; i < argc; ++i) {
    std::string a = argv[i];
    auto need = [&](const char* name) -> const char* {
      if (i + 1 >= argc) { fprintf(stderr, "detb200-launch: %s needs a value\n", name); exit(2); }
      return argv[++i];
    };
    if (a == "--") { ++i; break; }
    else if (a == "--nproc" || a == "--nproc-per-node" || a == "--nproc_per_node") nproc = atoi(need("--nproc"));
    else if (a == "--nnodes") nnodes = atoi(need("--nnodes"));
    else if (a == "--node-rank" || a == "--node_rank") node_rank = atoi(need("--node-rank"));
    else if (a == "--master-addr" || a == "--master_addr") master_addr = need("--master-addr");
    else if (a == "--master-port" || a == "--master_port") master_port = need("--master-port");
    else if (a == "--grace") grace = atof(need("--grace"));
    else if (a == "--no-prefix") prefix = false;
    else { fprintf(stderr, "detb200-launch: unknown option %s\n", a.c_str()); return 2; }
  }
  if (i >= argc || nproc < 1) {
    fprintf(stderr, "usage: detb200-launch --nproc N [--master-addr A] [--master-port P] [--grace S] [--no-prefix] -- cmd [args...]\n");
    return 2;
  }
  char** cmd = argv + i;

  struct sigaction sa;
  memset(&sa, 0, sizeof(sa));
  sa.sa_handler = on_signal;
  sigaction(SIGTERM, &sa, nullptr);
  sigaction(SIGINT, &sa, nullptr);
  signal(SIGPIPE, SIG_IGN);

  int ep = epoll_create1(EPOLL_CLOEXEC);
  g_workers.resize(nproc);
  const int world = nproc * nnodes;
  for (int r = 0; r < nproc; ++r) {
    int pfd[2];
    if (pipe2(pfd, O_CLOEXEC) != 0) { perror("pipe"); return 1; }
    pid_t pid = fork();
    if (pid < 0) { perror("fork"); signal_all(SIGKILL); return 1; }
    if (pid == 0) {
      dup2(pfd[1], STDOUT_FILENO);
      dup2(pfd[1], STDERR_FILENO);
      setenv("RANK", std::to_string(node_rank * nproc + r).c_str(), 1);
      setenv("LOCAL_RANK", std::to_string(r).c_str(), 1);
      setenv("WORLD_SIZE", std::to_string(world).c_str(), 1);
      setenv("LOCAL_WORLD_SIZE", std::to_string(nproc).c_str(), 1);
      setenv("LOCAL_SIZE", std::to_string(nproc).c_str(), 1);
      setenv("GROUP_RANK", std::to_string(node_rank).c_str(), 1);
      setenv("CROSS_RANK", std::to_string(node_rank).c_str(), 1);
      setenv("GROUP_WORLD_SIZE", std::to_string(nnodes).c_str(), 1);
      setenv("CROSS_SIZE", std::to_string(nnodes).c_str(), 1);
      setenv("MASTER_ADDR", master_addr.c_str(), 1);
      setenv("MASTER_PORT", master_port.c_str(), 1);
      setenv("USE_TORCH_DISTRIBUTED", "1", 1);
      setenv("DET_CHIEF_IP", master_addr.c_str(), 0);
      setenv("PYTHONUNBUFFERED", "1", 0);
      execvp(cmd[0], cmd);
      fprintf(stderr, "detb200-launch: exec %s failed: %s\n", cmd[0], strerror(errno));
      _exit(127);
    }
    close(pfd[1]);
    Worker& w = g_workers[r];
    w.pid = pid;
    w.out_fd = pfd[0];
    fcntl(w.out_fd, F_SETFL, fcntl(w.out_fd, F_GETFL) | O_NONBLOCK);
    w.pidfd = (int)syscall(SYS_pidfd_open, pid, 0);
    epoll_event ev;
    ev.events = EPOLLIN;
    ev.data.u64 = ((uint64_t)r << 1);          // even: output pipe
    epoll_ctl(ep, EPOLL_CTL_ADD, w.out_fd, &ev);
    if (w.pidfd >= 0) {
      ev.data.u64 = ((uint64_t)r << 1) | 1;    // odd: pidfd
      epoll_ctl(ep, EPOLL_CTL_ADD, w.pidfd, &ev);
    }
  }

  int first_bad = 0;
  bool terminating = false;
  double kill_deadline = 0;
  int alive = nproc, open_pipes = nproc;
  char buf[65536];
  while (alive > 0 || open_pipes > 0) {
    epoll_event evs[64];
    int n = epoll_wait(ep, evs, 64, 200);
    if (n < 0 && errno != EINTR) break;
    if (g_signal && !terminating) {
      // forwarded: let the ranks shut down (the harness treats SIGTERM as "preempt now")
      signal_all(g_signal);
      terminating = true;
      kill_deadline = now_s() + grace;
      if (first_bad == 0) first_bad = 128 + g_signal;
    }
    for (int k = 0; k < n; ++k) {
      int r = (int)(evs[k].data.u64 >> 1);
      Worker& w = g_workers[r];
      if ((evs[k].data.u64 & 1) == 0) {
        for (;;) {
          ssize_t got = read(w.out_fd, buf, sizeof(buf));
          if (got > 0) {
            w.partial.append(buf, (size_t)got);
            flush_lines(w, r, prefix, false);
            continue;
          }
          if (got == 0) {
            flush_lines(w, r, prefix, true);
            epoll_ctl(ep, EPOLL_CTL_DEL, w.out_fd, nullptr);
            close(w.out_fd);
            w.out_fd = -1;
            --open_pipes;
          }
          break;
        }
      }
    }
    // reap: pidfd readiness or plain polling fallback
    for (int r = 0; r < nproc; ++r) {
      Worker& w = g_workers[r];
      if (w.exited) continue;
      int st = 0;
      pid_t got = waitpid(w.pid, &st, WNOHANG);
      if (got == w.pid) {
        w.exited = true;
        w.status = st;
        --alive;
        if (w.pidfd >= 0) { epoll_ctl(ep, EPOLL_CTL_DEL, w.pidfd, nullptr); close(w.pidfd); w.pidfd = -1; }
        int code = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + WTERMSIG(st);
        if (code != 0 && first_bad == 0) {
          first_bad = code;
          fprintf(stderr, "detb200-launch: rank %d (pid %d) exited with status %d; stopping the other %d workers\n", r,
                  (int)w.pid, code, alive);
        }
        if (code != 0 && !terminating && alive > 0) {
          signal_all(SIGTERM);
          terminating = true;
          kill_deadline = now_s() + grace;
        }
      }
    }
    if (terminating && alive > 0 && now_s() > kill_deadline) {
      fprintf(stderr, "detb200-launch: workers ignored SIGTERM for %.0fs; sending SIGKILL\n", grace);
      signal_all(SIGKILL);
      kill_deadline = now_s() + 3600;
    }
    if (alive == 0 && open_pipes > 0) {
      // children of the workers may hold the pipe open; drain what is there and stop
      for (int r = 0; r < nproc; ++r) {
        Worker& w = g_workers[r];
        if (w.out_fd < 0) continue;
        for (;;) {
          ssize_t got = read(w.out_fd, buf, sizeof(buf));
          if (got > 0) { w.partial.append(buf, (size_t)got); continue; }
          break;
        }
        flush_lines(w, r, prefix, true);
        close(w.out_fd);
        w.out_fd = -1;
      }
      open_pipes = 0;
    }
  }
  return first_bad;
}
