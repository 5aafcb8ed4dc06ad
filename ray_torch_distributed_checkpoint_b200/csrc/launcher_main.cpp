// rtdc-launch: native worker-group launcher (one rank per GPU / CPU worker on this node).
//
//   rtdc-launch --nproc N [--master-addr A] --master-port P [--bind-cores] [--grace SEC] -- prog args...
//
// Forks N children, each with RANK / LOCAL_RANK / WORLD_SIZE / LOCAL_WORLD_SIZE / NODE_RANK /
// MASTER_ADDR / MASTER_PORT set, optionally pinned to an equal slice of the CPU cores (keeps each
// rank's host threads next to each other on a two-socket B200 host). The group lives and dies
// together: the first non-zero exit (or a signal to the launcher) terminates every other rank, first
// with SIGTERM, after the grace period with SIGKILL; the launcher's exit status is that of the first
// failing rank. Children get PR_SET_PDEATHSIG so they cannot outlive the launcher.
//
// The reference gets this behaviour from Ray actors + @metaflow_ray gang scheduling
// (train_flow.py:39,42; my_ray_module.py:244-250); there is no launcher source in the reference.
#include <errno.h>
#include <sched.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/prctl.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

#include <string>
#include <vector>

static volatile sig_atomic_t g_signal = 0;
static void on_signal(int s) { g_signal = s; }

static double now_s() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec + ts.tv_nsec * 1e-9;
}

int main(int argc, char** argv) {
  int nproc = 1, port = 29500, bind = 0;
  double grace = 5.0;
  std::string addr = "127.0.0.1";
  int i = 1;
  for (; i < argc; ++i) {
    std::string a = argv[i];
    if (a == "--") { ++i; break; }
    if (a == "--nproc" && i + 1 < argc) nproc = atoi(argv[++i]);
    else if (a == "--master-port" && i + 1 < argc) port = atoi(argv[++i]);
    else if (a == "--master-addr" && i + 1 < argc) addr = argv[++i];
    else if (a == "--grace" && i + 1 < argc) grace = atof(argv[++i]);
    else if (a == "--bind-cores") bind = 1;
    else { fprintf(stderr, "rtdc-launch: unknown argument %s\n", a.c_str()); return 2; }
  }
  if (i >= argc || nproc < 1) {
    fprintf(stderr, "usage: rtdc-launch --nproc N --master-port P [--master-addr A] [--bind-cores] -- prog args...\n");
    return 2;
  }
  struct sigaction sa;
  memset(&sa, 0, sizeof(sa));
  sa.sa_handler = on_signal;
  sigaction(SIGTERM, &sa, nullptr);
  sigaction(SIGINT, &sa, nullptr);

  cpu_set_t avail;
  CPU_ZERO(&avail);
  std::vector<int> cores;
  if (bind && sched_getaffinity(0, sizeof(avail), &avail) == 0) {
    for (int c = 0; c < CPU_SETSIZE; ++c)
      if (CPU_ISSET(c, &avail)) cores.push_back(c);
  }

  std::vector<pid_t> pids(nproc, -1);
  const pid_t parent = getpid();
  for (int r = 0; r < nproc; ++r) {
    pid_t pid = fork();
    if (pid < 0) { perror("rtdc-launch: fork"); g_signal = SIGTERM; break; }
    if (pid == 0) {
      prctl(PR_SET_PDEATHSIG, SIGKILL);
      if (getppid() != parent) _exit(1);
      char buf[32];
      snprintf(buf, sizeof(buf), "%d", r); setenv("RANK", buf, 1); setenv("LOCAL_RANK", buf, 1);
      snprintf(buf, sizeof(buf), "%d", nproc); setenv("WORLD_SIZE", buf, 1); setenv("LOCAL_WORLD_SIZE", buf, 1);
      setenv("NODE_RANK", "0", 1);
      setenv("MASTER_ADDR", addr.c_str(), 1);
      snprintf(buf, sizeof(buf), "%d", port); setenv("MASTER_PORT", buf, 1);
      if (bind && (int)cores.size() >= nproc) {
        const int per = (int)cores.size() / nproc;
        cpu_set_t mine;
        CPU_ZERO(&mine);
        for (int c = r * per; c < (r + 1) * per; ++c) CPU_SET(cores[c], &mine);
        sched_setaffinity(0, sizeof(mine), &mine);
        // leave OMP/MKL thread counts consistent with the slice unless the user already chose
        snprintf(buf, sizeof(buf), "%d", per);
        setenv("OMP_NUM_THREADS", buf, 0);
      }
      execvp(argv[i], argv + i);
      fprintf(stderr, "rtdc-launch: exec %s failed: %s\n", argv[i], strerror(errno));
      _exit(127);
    }
    pids[r] = pid;
  }

  int alive = 0;
  for (pid_t p : pids) alive += p > 0;
  int first_fail = 0;
  bool terminating = false;
  double kill_at = 0;
  while (alive > 0) {
    if (g_signal && !terminating) {
      terminating = true;
      if (!first_fail) first_fail = 128 + g_signal;
      for (pid_t p : pids) if (p > 0) kill(p, SIGTERM);
      kill_at = now_s() + grace;
    }
    int status = 0;
    pid_t done = waitpid(-1, &status, WNOHANG);
    if (done == 0) {
      if (terminating && now_s() > kill_at) {
        for (pid_t p : pids) if (p > 0) kill(p, SIGKILL);
        kill_at = now_s() + 1e9;
      }
      usleep(2000);
      continue;
    }
    if (done < 0) {
      if (errno == EINTR) continue;
      break;
    }
    int rank = -1;
    for (int r = 0; r < nproc; ++r) if (pids[r] == done) { rank = r; pids[r] = -1; }
    if (rank < 0) continue;
    --alive;
    int code = WIFEXITED(status) ? WEXITSTATUS(status) : 128 + WTERMSIG(status);
    if (code != 0 && !terminating) {
      fprintf(stderr, "rtdc-launch: rank %d (pid %d) exited with status %d; stopping the other %d rank(s)\n", rank,
              (int)done, code, alive);
      first_fail = code;
      terminating = true;
      for (pid_t p : pids) if (p > 0) kill(p, SIGTERM);
      kill_at = now_s() + grace;
    }
  }
  return first_fail;
}
