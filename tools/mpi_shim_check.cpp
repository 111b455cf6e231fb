// tools/mpi_shim_check.cpp -- CPU-only check of tools/mpi_shim/mpi.h with the reference's message pattern:
// rank 0 polls with MPI_Irecv/MPI_Test and answers with MPI_Isend, ranks >= 1 send float[8] / receive float[1] and
// meet on a split communicator's barrier.  Run as: b2_mpirun -n 3 mpi_shim_check
#include <cstdio>
#include <vector>

#include "mpi_shim/mpi.h"

int main(int argc, char** argv) {
  MPI_Init(&argc, &argv);
  int n = 0, me = 0;
  MPI_Comm_size(MPI_COMM_WORLD, &n);
  MPI_Comm_rank(MPI_COMM_WORLD, &me);
  MPI_Comm sub;
  MPI_Comm_split(MPI_COMM_WORLD, me == 0, me, &sub);
  int sn = 0, sr = 0;
  MPI_Comm_size(sub, &sn);
  MPI_Comm_rank(sub, &sr);
  const int rounds = 200;
  int bad = 0;
  if (me == 0) {
    if (sn != 1 || sr != 0) bad++;
    std::vector<std::vector<float>> in((size_t)n, std::vector<float>(8));
    std::vector<MPI_Request> req((size_t)n, MPI_REQUEST_NULL);
    std::vector<int> got((size_t)n, 0);
    int done = 0;
    while (done < (n - 1) * rounds) {
      for (int i = 1; i < n; i++) {
        if (got[(size_t)i] == rounds) continue;
        if (req[(size_t)i] == MPI_REQUEST_NULL) MPI_Irecv(in[(size_t)i].data(), 8, MPI_FLOAT, i, i, MPI_COMM_WORLD, &req[(size_t)i]);
        int flag = 0;
        MPI_Status st;
        MPI_Test(&req[(size_t)i], &flag, &st);
        if (!flag) continue;
        if (in[(size_t)i][0] != (float)i || in[(size_t)i][7] != (float)got[(size_t)i] || st.MPI_SOURCE != i) bad++;
        float reply = in[(size_t)i][7] * 2.0f + (float)i;
        MPI_Request r;
        MPI_Isend(&reply, 1, MPI_FLOAT, i, i + 2 * n, MPI_COMM_WORLD, &r);
        got[(size_t)i]++;
        done++;
      }
    }
  } else {
    if (sn != n - 1 || sr != me - 1) bad++;
    for (int k = 0; k < rounds; k++) {
      float msg[8] = {(float)me, 0, 0, 0, 0, 0, 0, (float)k};
      MPI_Send(msg, 8, MPI_FLOAT, 0, me, MPI_COMM_WORLD);
      float reply = -1.f;
      MPI_Status st;
      MPI_Recv(&reply, 1, MPI_FLOAT, 0, me + 2 * n, MPI_COMM_WORLD, &st);
      if (reply != (float)k * 2.0f + (float)me) bad++;
      if (k % 50 == 0) MPI_Barrier(sub);
    }
  }
  MPI_Finalize();
  std::printf("rank %d of %d: %s\n", me, n, bad ? "FAIL" : "ok");
  return bad ? 1 : 0;
}
