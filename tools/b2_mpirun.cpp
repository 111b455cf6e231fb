// tools/b2_mpirun.cpp -- launcher for programs built against tools/mpi_shim/mpi.h:  b2_mpirun -n N prog [args...]
// Creates the shared segment, forks N ranks (B2_MPI_SHM / B2_MPI_RANK / B2_MPI_SIZE in their environment) and waits.
// When rank 0 (the reference's learner) exits, the remaining ranks get a grace period and are then terminated:
// upstream's env ranks block in MPI_Recv for an action that never comes once the learner has left.
#include <fcntl.h>
#include <signal.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "mpi_shim/mpi.h"

int main(int argc, char** argv) {
  int n = 0, quiet = 0, i = 1;
  for (; i < argc; i++) {
    if (!strcmp(argv[i], "-n") && i + 1 < argc) n = atoi(argv[++i]);
    else if (!strcmp(argv[i], "-q")) quiet = 1;   // ranks' stdout -> /dev/null (the reference prints every tensor)
    else break;
  }
  if (n < 1 || n > B2MPI_MAX_RANKS || i >= argc) {
    fprintf(stderr, "usage: b2_mpirun -n N [-q] prog [args...]   (1 <= N <= %d)\n", B2MPI_MAX_RANKS);
    return 2;
  }
  const std::string name = "/b2mpi_" + std::to_string((long)getpid());
  const int fd = shm_open(name.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
  if (fd < 0 || ftruncate(fd, sizeof(b2mpi_shm)) != 0) { perror("b2_mpirun: shm"); return 1; }
  void* p = mmap(nullptr, sizeof(b2mpi_shm), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (p == MAP_FAILED) { perror("b2_mpirun: mmap"); shm_unlink(name.c_str()); return 1; }
  b2mpi_shm* s = new (p) b2mpi_shm();   // zero-initialised rings and barriers
  s->size = n;
  std::vector<pid_t> pids((size_t)n);
  for (int r = 0; r < n; r++) {
    const pid_t pid = fork();
    if (pid == 0) {
      setenv("B2_MPI_SHM", name.c_str(), 1);
      setenv("B2_MPI_RANK", std::to_string(r).c_str(), 1);
      setenv("B2_MPI_SIZE", std::to_string(n).c_str(), 1);
      if (quiet) {
        const int dn = open("/dev/null", O_WRONLY);
        if (dn >= 0) { dup2(dn, 1); close(dn); }
      }
      execvp(argv[i], argv + i);
      perror("b2_mpirun: exec");
      _exit(127);
    }
    pids[(size_t)r] = pid;
  }
  int rc = 0, left = n;
  bool rank0_done = false;
  double grace_s = 5.0;
  while (left > 0) {
    int st = 0;
    const pid_t w = waitpid(-1, &st, rank0_done ? WNOHANG : 0);
    if (w > 0) {
      left--;
      const int code = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + WTERMSIG(st);
      for (int r = 0; r < n; r++)
        if (pids[(size_t)r] == w) {
          pids[(size_t)r] = -1;
          if (r == 0) { rank0_done = true; rc = code; }
          else if (code != 0 && !rank0_done && rc == 0) rc = code;
        }
      continue;
    }
    if (rank0_done) {
      if (grace_s <= 0) {
        for (int r = 0; r < n; r++)
          if (pids[(size_t)r] > 0) kill(pids[(size_t)r], SIGTERM);   // exact PIDs this launcher forked
        rank0_done = false;   // back to blocking waits for the terminated ranks
        continue;
      }
      usleep(100000);
      grace_s -= 0.1;
    }
  }
  shm_unlink(name.c_str());
  return rc;
}
