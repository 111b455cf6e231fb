// tools/mpi_shim/mpi.h -- just enough MPI, over POSIX shared memory, to run the reference's UNMODIFIED driver
// (src/main.cpp + include/cart_MPI.h) on a box without an MPI installation.
//
// Implements exactly the calls the reference makes: MPI_Init, MPI_Comm_size, MPI_Comm_rank,
// MPI_Comm_split, MPI_Send, MPI_Recv, MPI_Isend, MPI_Irecv, MPI_Test, MPI_Barrier, MPI_Finalize, for MPI_FLOAT
// payloads of at most 64 bytes (the reference sends float[8] up and float[1] down).  Every (source, destination)
// pair is a single-producer single-consumer ring in a shared segment created by tools/b2_mpirun, which forks the
// ranks and passes B2_MPI_SHM / B2_MPI_RANK / B2_MPI_SIZE.  Sends are buffered (MPI_Isend completes immediately);
// message order per pair is FIFO, which is all the reference relies on (one tag per direction per pair).
// Header-only, single translation unit use (the reference has one).  Not a general MPI.
#ifndef B2_MPI_SHIM_H
#define B2_MPI_SHIM_H
#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <unistd.h>

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#define B2MPI_MAX_RANKS 16
#define B2MPI_QCAP 64
#define B2MPI_MSG_BYTES 64

struct b2mpi_msg {
  int tag, bytes;
  char data[B2MPI_MSG_BYTES];
};
struct b2mpi_queue {
  std::atomic<unsigned> head, tail;  // consumer advances head, producer advances tail
  b2mpi_msg slot[B2MPI_QCAP];
};
struct b2mpi_barrier {
  std::atomic<int> count, gen;
};
struct b2mpi_shm {
  int size;
  std::atomic<int> color[B2MPI_MAX_RANKS];
  b2mpi_barrier bar[4];                                   // 0 = world, 1 + color for split communicators
  b2mpi_queue q[B2MPI_MAX_RANKS][B2MPI_MAX_RANKS];        // q[src][dst]
};

struct b2mpi_comm {
  int id, size, rank;
};
typedef b2mpi_comm* MPI_Comm;
struct MPI_Status {
  int MPI_SOURCE, MPI_TAG, MPI_ERROR;
};
struct b2mpi_req {
  void* buf;
  int bytes, src, tag;
};
typedef b2mpi_req* MPI_Request;
typedef int MPI_Datatype;
#define MPI_FLOAT 4
#define MPI_DOUBLE 8
#define MPI_INT 104
#define MPI_SUCCESS 0
#define MPI_REQUEST_NULL ((MPI_Request) nullptr)
#define MPI_STATUS_IGNORE ((MPI_Status*)nullptr)

namespace b2mpi {
inline b2mpi_shm*& shm() { static b2mpi_shm* p = nullptr; return p; }
inline int& my_rank() { static int r = 0; return r; }
inline b2mpi_comm& world() { static b2mpi_comm c{0, 1, 0}; return c; }
inline int type_bytes(MPI_Datatype t) { return t == MPI_INT ? 4 : t; }
inline void relax() { sched_yield(); }
inline void die(const char* what) { std::fprintf(stderr, "mpi shim: %s\n", what); std::_Exit(3); }

inline void barrier(b2mpi_barrier& b, int n) {
  const int gen = b.gen.load(std::memory_order_acquire);
  if (b.count.fetch_add(1, std::memory_order_acq_rel) + 1 == n) {
    b.count.store(0, std::memory_order_relaxed);
    b.gen.fetch_add(1, std::memory_order_release);
  } else {
    while (b.gen.load(std::memory_order_acquire) == gen) relax();
  }
}
inline void push(int dst, int tag, const void* buf, int bytes) {
  if (bytes > B2MPI_MSG_BYTES) die("message larger than 64 bytes");
  b2mpi_queue& q = shm()->q[my_rank()][dst];
  const unsigned t = q.tail.load(std::memory_order_relaxed);
  while (t - q.head.load(std::memory_order_acquire) >= B2MPI_QCAP) relax();
  b2mpi_msg& m = q.slot[t % B2MPI_QCAP];
  m.tag = tag; m.bytes = bytes;
  std::memcpy(m.data, buf, (size_t)bytes);
  q.tail.store(t + 1, std::memory_order_release);
}
inline bool try_pop(int src, int tag, void* buf, int bytes, MPI_Status* st) {
  b2mpi_queue& q = shm()->q[src][my_rank()];
  const unsigned h = q.head.load(std::memory_order_relaxed);
  if (q.tail.load(std::memory_order_acquire) == h) return false;
  const b2mpi_msg& m = q.slot[h % B2MPI_QCAP];
  if (m.tag != tag) die("tag mismatch (the shim is FIFO per pair)");
  std::memcpy(buf, m.data, (size_t)(m.bytes < bytes ? m.bytes : bytes));
  if (st) { st->MPI_SOURCE = src; st->MPI_TAG = m.tag; st->MPI_ERROR = MPI_SUCCESS; }
  q.head.store(h + 1, std::memory_order_release);
  return true;
}
}  // namespace b2mpi

#define MPI_COMM_WORLD (&b2mpi::world())

inline int MPI_Init(int*, char***) {
  const char* name = std::getenv("B2_MPI_SHM");
  const char* rank = std::getenv("B2_MPI_RANK");
  const char* size = std::getenv("B2_MPI_SIZE");
  if (!name || !rank || !size) b2mpi::die("not started by tools/b2_mpirun (B2_MPI_SHM/RANK/SIZE unset)");
  const int fd = shm_open(name, O_RDWR, 0600);
  if (fd < 0) b2mpi::die("shm_open failed");
  void* p = mmap(nullptr, sizeof(b2mpi_shm), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (p == MAP_FAILED) b2mpi::die("mmap failed");
  b2mpi::shm() = static_cast<b2mpi_shm*>(p);
  b2mpi::my_rank() = std::atoi(rank);
  b2mpi::world() = b2mpi_comm{0, std::atoi(size), std::atoi(rank)};
  if (b2mpi::world().size > B2MPI_MAX_RANKS) b2mpi::die("too many ranks");
  return MPI_SUCCESS;
}
inline int MPI_Finalize() { return MPI_SUCCESS; }
inline int MPI_Comm_size(MPI_Comm c, int* n) { *n = c->size; return MPI_SUCCESS; }
inline int MPI_Comm_rank(MPI_Comm c, int* r) { *r = c->rank; return MPI_SUCCESS; }
inline int MPI_Barrier(MPI_Comm c) { b2mpi::barrier(b2mpi::shm()->bar[c->id], c->size); return MPI_SUCCESS; }
// colors 0/1 only (the reference splits learner from env ranks, src/main.cpp:424); rank order inside = world order
inline int MPI_Comm_split(MPI_Comm c, int color, int /*key*/, MPI_Comm* out) {
  if (c->id != 0 || color < 0 || color > 2) b2mpi::die("MPI_Comm_split: only a 0/1/2 colouring of MPI_COMM_WORLD is supported");
  b2mpi_shm* s = b2mpi::shm();
  s->color[c->rank].store(color, std::memory_order_release);
  MPI_Barrier(c);
  int n = 0, r = 0;
  for (int i = 0; i < c->size; i++)
    if (s->color[i].load(std::memory_order_acquire) == color) { if (i < c->rank) r++; n++; }
  MPI_Barrier(c);
  *out = new b2mpi_comm{1 + color, n, r};
  return MPI_SUCCESS;
}
inline int MPI_Send(const void* buf, int count, MPI_Datatype t, int dst, int tag, MPI_Comm) {
  b2mpi::push(dst, tag, buf, count * b2mpi::type_bytes(t));
  return MPI_SUCCESS;
}
inline int MPI_Isend(const void* buf, int count, MPI_Datatype t, int dst, int tag, MPI_Comm c, MPI_Request* req) {
  MPI_Send(buf, count, t, dst, tag, c);   // buffered: complete on return
  if (req) *req = MPI_REQUEST_NULL;
  return MPI_SUCCESS;
}
inline int MPI_Recv(void* buf, int count, MPI_Datatype t, int src, int tag, MPI_Comm, MPI_Status* st) {
  while (!b2mpi::try_pop(src, tag, buf, count * b2mpi::type_bytes(t), st)) b2mpi::relax();
  return MPI_SUCCESS;
}
inline int MPI_Irecv(void* buf, int count, MPI_Datatype t, int src, int tag, MPI_Comm, MPI_Request* req) {
  *req = new b2mpi_req{buf, count * b2mpi::type_bytes(t), src, tag};
  return MPI_SUCCESS;
}
inline int MPI_Test(MPI_Request* req, int* flag, MPI_Status* st) {
  if (*req == MPI_REQUEST_NULL) { *flag = 1; return MPI_SUCCESS; }
  b2mpi_req* r = *req;
  if (b2mpi::try_pop(r->src, r->tag, r->buf, r->bytes, st)) {
    delete r;
    *req = MPI_REQUEST_NULL;
    *flag = 1;
  } else {
    *flag = 0;
  }
  return MPI_SUCCESS;
}
#endif  // B2_MPI_SHIM_H
