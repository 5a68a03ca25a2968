// knnrun — process launcher + rendezvous for the cudaMpiKNN_* executables (replaces mpirun; there is
// no MPI on the target image):   knnrun -n <ranks> <exe> [args...]
// Forks <ranks> copies of <exe> with KNNRUN_RANK / KNNRUN_SIZE / KNNRUN_SOCK set and serves their two
// collectives (allgather of equal-sized blobs, barrier = allgather of zero bytes) over one unix
// stream socket per rank. Bulk data never passes through here: GPUs exchange it over NVLink.
// Exit code = first non-zero child exit code.
// Role of `mpirun -n <ranks>` in the reference's command lines (/root/reference/README.md:31,39) and of
// MPI_Init / MPI_Comm_rank / MPI_Comm_size / MPI_Finalize (/root/reference/unorderedDataVariant.cu:34-35,107,238).
#include <errno.h>
#include <poll.h>
#include <signal.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <sys/wait.h>
#include <unistd.h>
#include <string>
#include <vector>

struct Wire { uint32_t op; uint32_t rank; uint64_t bytes; };

static bool io_all(int fd, void* buf, size_t n, bool writing) {
  size_t done = 0;
  while (done < n) {
    const ssize_t r = writing ? write(fd, (char*)buf + done, n - done) : read(fd, (char*)buf + done, n - done);
    if (r <= 0) { if (r < 0 && errno == EINTR) continue; return false; }
    done += (size_t)r;
  }
  return true;
}

int main(int ac, char** av) {
  int ranks = 0, first = 1;
  if (ac >= 4 && std::string(av[1]) == "-n") { ranks = atoi(av[2]); first = 3; }
  if (ranks < 1 || first >= ac) { fprintf(stderr, "usage: knnrun -n <ranks> <exe> [args...]\n"); return 1; }
  char sock_path[108];
  snprintf(sock_path, sizeof sock_path, "/tmp/knnrun_%d.sock", (int)getpid());
  unlink(sock_path);
  const int srv = socket(AF_UNIX, SOCK_STREAM, 0);
  sockaddr_un addr;
  memset(&addr, 0, sizeof addr);
  addr.sun_family = AF_UNIX;
  strncpy(addr.sun_path, sock_path, sizeof(addr.sun_path) - 1);
  if (bind(srv, (sockaddr*)&addr, sizeof addr) != 0 || listen(srv, ranks) != 0) { perror("knnrun: bind/listen"); return 1; }
  signal(SIGPIPE, SIG_IGN);

  std::vector<pid_t> pids;
  for (int r = 0; r < ranks; ++r) {
    const pid_t pid = fork();
    if (pid == 0) {
      setenv("KNNRUN_RANK", std::to_string(r).c_str(), 1);
      setenv("KNNRUN_SIZE", std::to_string(ranks).c_str(), 1);
      setenv("KNNRUN_SOCK", sock_path, 1);
      close(srv);
      execvp(av[first], av + first);
      perror("knnrun: exec");
      _exit(127);
    }
    pids.push_back(pid);
  }

  int rc = 0;
  std::vector<int> fds(ranks, -1);
  int live = ranks;
  if (ranks > 1) {
    // accept one connection per rank; the first message of each carries its rank
    std::vector<int> pending;
    std::vector<Wire> head(ranks);
    std::vector<bool> have_head(ranks, false);
    for (int i = 0; i < ranks && rc == 0; ++i) {
      // wait for the next rank to connect, but notice ranks that die first (bad arguments etc.)
      for (;;) {
        pollfd pf{srv, POLLIN, 0};
        if (poll(&pf, 1, 100) > 0) break;
        for (size_t c = 0; c < pids.size() && rc == 0; ++c) {
          int st = 0;
          if (pids[c] > 0 && waitpid(pids[c], &st, WNOHANG) == pids[c]) {
            pids[c] = -1;
            rc = WIFEXITED(st) ? WEXITSTATUS(st) : 128;
            if (rc == 0) rc = 1;   // a rank that left without joining breaks the job
          }
        }
        if (rc) break;
      }
      if (rc) break;
      const int fd = accept(srv, nullptr, nullptr);
      if (fd < 0) { perror("knnrun: accept"); rc = 1; break; }
      Wire w;
      if (!io_all(fd, &w, sizeof w, false) || w.rank >= (uint32_t)ranks) { rc = 1; break; }
      fds[w.rank] = fd;
      head[w.rank] = w;
      have_head[w.rank] = true;
    }
    std::vector<char> blob;
    while (rc == 0 && live == ranks) {
      // one collective: headers (already read for the first round), payloads, reply to everyone
      uint64_t bytes = 0;
      bool bye = false;
      for (int r = 0; r < ranks && rc == 0; ++r) {
        if (!have_head[r] && !io_all(fds[r], &head[r], sizeof(Wire), false)) { rc = 1; break; }
        have_head[r] = false;
        if (head[r].op == 2) bye = true;
        if (r == 0) bytes = head[r].bytes;
        else if (head[r].op != 2 && head[r].bytes != bytes) { fprintf(stderr, "knnrun: ranks disagree on allgather size\n"); rc = 1; }
      }
      if (rc || bye) break;
      blob.resize((size_t)bytes * ranks + 1);
      for (int r = 0; r < ranks && rc == 0; ++r)
        if (bytes && !io_all(fds[r], blob.data() + (size_t)r * bytes, (size_t)bytes, false)) rc = 1;
      for (int r = 0; r < ranks && rc == 0; ++r)
        if (!io_all(fds[r], blob.data(), bytes ? (size_t)bytes * ranks : 1, true)) rc = 1;
    }
  }
  // Hang up first: a rank still blocked in a collective (because a peer left early) sees EOF and exits.
  for (int& fd : fds) if (fd >= 0) { close(fd); fd = -1; }
  for (pid_t pid : pids) {
    if (pid <= 0) continue;
    int st = 0;
    if (rc != 0) kill(pid, SIGTERM);
    waitpid(pid, &st, 0);
    const int code = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + (WIFSIGNALED(st) ? WTERMSIG(st) : 0);
    if (code != 0 && rc == 0) rc = code;
  }
  for (int fd : fds) if (fd >= 0) close(fd);
  close(srv);
  unlink(sock_path);
  return rc;
}
