// Communication of the mesh-partitioned path: ghost-value exchange before mat-vecs and sum reductions of
// Krylov scalars.  One process per GPU; NCCL (over NVLink / NVSwitch) does the transport.  libnccl is opened
// at run time (dlopen) so that the single-GPU library has no link-time dependency on it.
//
// Both operations are enqueued on the caller's stream: no host synchronisation is added.  Messages are small
// (a ghost layer is O((N/P)^(2/3)) values, a reduction 1..30 doubles), so the exchange is latency bound; the
// design keeps the NUMBER of collectives per Krylov iteration low rather than optimising bandwidth.
#pragma once
#include "common.cuh"

namespace s35 {

// ghost exchange plan of one DoF kind on this rank (device copy of s35_halo)
struct HaloDev
{
  int32_t n_owned = 0, n_local = 0;
  std::vector<int32_t> neighbor_rank, send_ptr, recv_ptr;
  DevBuf<int32_t> send_idx;
  DevBuf<double> send_buf;
  void create(const s35_halo &h, int32_t n_owned, int32_t n_local, cudaStream_t s);
  bool empty() const { return neighbor_rank.empty(); }
};

struct Comm
{
  int rank = 0, n_ranks = 1;
  virtual ~Comm() {}
  // buf[0..count) <- sum over ranks, in place, device memory
  virtual void allreduce_sum(double *buf, int count, cudaStream_t s) = 0;
  // x[n_owned ..) <- owned values of the neighbours; x[send_idx] -> neighbours
  virtual void exchange(HaloDev &halo, double *x, cudaStream_t s) = 0;
  // An independent communicator over the same ranks (i = 0, 1, 2), for work that runs concurrently with the
  // work on this one -- the velocity components, each on its own stream and host thread.  nullptr if the
  // transport has none; the callers then take turns on the main communicator.
  virtual Comm *sub(int i) { (void)i; return nullptr; }
  // The exchange in two parts for callers that bundle several operations (combiner.h): the transfers alone,
  // after the owned values have been packed (pack_for_send), between group_begin() and group_end().
  virtual void exchange_transfers(HaloDev &halo, double *x, cudaStream_t s) = 0;
  virtual void group_begin() {}
  virtual void group_end() {}
  // Only meaningful where the sub-communicators share this one through the collective combiner (combiner.h):
  // enter_all() on the parent before the concurrent solves start, leave() on a sub-communicator when its
  // solve is over (also when it failed).  No-ops everywhere else.
  virtual void enter_all() {}
  virtual void leave() {}
};

// x[send_idx] -> halo.send_buf, on stream s
void pack_for_send(HaloDev &halo, const double *x, cudaStream_t s);

// the communicator s35_comm_init created for this process, or nullptr
Comm *process_comm();
void comm_unique_id(unsigned char id[128]);
void comm_init(int rank, int n_ranks, const unsigned char id[128]);
void comm_finalize();
// ranks as threads of one process sharing one GPU (single-GPU tests of the partitioned path)
void comm_thread_group(int n_ranks);
void comm_thread_bind(int rank);

// sqrt in place of a device scalar (after an all-reduced sum of squares)
void scalar_sqrt(double *x, cudaStream_t s);
// x *= factor
void scalar_scale(double *x, double factor, cudaStream_t s);

} // namespace s35
