#include "comm.cuh"

#include "combiner.h"

#include <dlfcn.h>
#include <nccl.h> // types only; the library is opened with dlopen

#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>

namespace s35 {

namespace {

__global__ void k_pack(double *__restrict__ out, const double *__restrict__ x, const int32_t *__restrict__ idx, int32_t n)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    out[i] = x[idx[i]];
}

__global__ void k_scalar_sqrt(double *x) { *x = sqrt(*x); }
__global__ void k_scalar_scale(double *x, double f) { *x *= f; }

// ------------------------------------------------------------------------------------------------ NCCL
struct NcclApi
{
  void *handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*CommSplit)(ncclComm_t, int, int, ncclComm_t *, ncclConfig_t *) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;

  void load()
  {
    if (handle)
      return;
    // a process that has already loaded NCCL (e.g. through torch) gets that copy
    handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!handle)
      handle = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!handle)
      throw Error(S35_ERR_CUDA, std::string("cannot load libnccl: ") + dlerror());
    auto sym = [&](const char *name) {
      void *p = dlsym(handle, name);
      if (!p)
        throw Error(S35_ERR_CUDA, std::string("libnccl lacks ") + name);
      return p;
    };
    GetUniqueId = (decltype(GetUniqueId))sym("ncclGetUniqueId");
    CommInitRank = (decltype(CommInitRank))sym("ncclCommInitRank");
    CommDestroy = (decltype(CommDestroy))sym("ncclCommDestroy");
    CommSplit = (decltype(CommSplit))dlsym(handle, "ncclCommSplit"); // NCCL >= 2.18; optional
    AllReduce = (decltype(AllReduce))sym("ncclAllReduce");
    Send = (decltype(Send))sym("ncclSend");
    Recv = (decltype(Recv))sym("ncclRecv");
    GroupStart = (decltype(GroupStart))sym("ncclGroupStart");
    GroupEnd = (decltype(GroupEnd))sym("ncclGroupEnd");
    GetErrorString = (decltype(GetErrorString))sym("ncclGetErrorString");
  }
  void check(ncclResult_t r, const char *what)
  {
    if (r != ncclSuccess)
      throw Error(S35_ERR_CUDA, std::string("NCCL error in ") + what + ": " + (GetErrorString ? GetErrorString(r) : "?"));
  }
};

NcclApi g_nccl;

// What the collective combiner (combiner.h) needs from CUDA and from the transport: every round goes out on one
// communication stream of the one communicator, inside one group.
struct RoundsBackend
{
  using Stream = cudaStream_t;
  using Event = cudaEvent_t;
  struct Op
  {
    HaloDev *halo = nullptr; // a ghost exchange of x if set, else a sum of count doubles at x
    double *x = nullptr;
    int count = 0;
  };
  Comm *transport = nullptr;
  cudaStream_t stream = nullptr;
  cudaEvent_t party[3] = {nullptr, nullptr, nullptr}, round[2] = {nullptr, nullptr};
  explicit RoundsBackend(Comm *t) : transport(t)
  {
    S35_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    for (cudaEvent_t &e : party)
      S35_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (cudaEvent_t &e : round)
      S35_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  RoundsBackend(const RoundsBackend &) = delete;
  RoundsBackend &operator=(const RoundsBackend &) = delete;
  ~RoundsBackend()
  {
    for (cudaEvent_t e : party)
      if (e)
        cudaEventDestroy(e);
    for (cudaEvent_t e : round)
      if (e)
        cudaEventDestroy(e);
    if (stream)
      cudaStreamDestroy(stream);
  }
  Stream comm_stream() { return stream; }
  Event party_event(int p) { return party[p]; }
  Event round_event(int parity) { return round[parity]; }
  void record(Event e, Stream s) { S35_CUDA(cudaEventRecord(e, s)); }
  void wait(Stream s, Event e) { S35_CUDA(cudaStreamWaitEvent(s, e, 0)); }
  void group_begin() { transport->group_begin(); }
  void group_end() { transport->group_end(); }
  void issue(const Op &op, Stream s)
  {
    if (op.halo)
      transport->exchange_transfers(*op.halo, op.x, s);
    else
      transport->allreduce_sum(op.x, op.count, s);
  }
};

// one velocity component's view of the shared communicator
struct CombinedComm : Comm
{
  Combiner<RoundsBackend> *rounds = nullptr;
  int party = 0;
  void allreduce_sum(double *buf, int count, cudaStream_t s) override { rounds->submit(party, RoundsBackend::Op{nullptr, buf, count}, s); }
  void exchange(HaloDev &h, double *x, cudaStream_t s) override
  {
    // handed in even when this rank has nothing to exchange: all ranks have to form the same rounds
    pack_for_send(h, x, s);
    rounds->submit(party, RoundsBackend::Op{&h, x, 0}, s);
  }
  void exchange_transfers(HaloDev &, double *, cudaStream_t) override
  {
    throw Error(S35_ERR_INVALID, "a combined communicator is not a transport");
  }
  void leave() override { rounds->leave(party); }
};

// three views of one transport that share it through the collective combiner
struct SharedRounds
{
  std::unique_ptr<RoundsBackend> backend;
  std::unique_ptr<Combiner<RoundsBackend>> rounds;
  std::vector<std::unique_ptr<Comm>> views;
  void create(Comm *transport)
  {
    backend.reset(new RoundsBackend(transport));
    rounds.reset(new Combiner<RoundsBackend>(*backend, 3));
    for (int i = 0; i < 3; ++i)
      {
        std::unique_ptr<CombinedComm> c(new CombinedComm);
        c->rank = transport->rank;
        c->n_ranks = transport->n_ranks;
        c->rounds = rounds.get();
        c->party = i;
        views.push_back(std::move(c));
      }
  }
  void clear()
  {
    views.clear();
    rounds.reset();
    backend.reset();
  }
};

// How the velocity components of a partitioned run share the communicator (S35_DIST_CONCURRENT):
//   2 (default)  the component solves run concurrently and hand their collectives to the collective combiner, which
//                issues them in common rounds on ONE communicator (measured on 2 B200s, r1: 237 ms per step against
//                335 ms with the components in turn; bit-identical results);
//   0            the components take turns on the communicator;
//   1            a communicator of its own per component (NCCL only; dead-locked on 8 GPUs in round 1: kept for the
//                comparison only).
int concurrent_mode()
{
  const char *conc = std::getenv("S35_DIST_CONCURRENT");
  return conc && *conc ? std::atoi(conc) : 2;
}

struct NcclComm : Comm
{
  ncclComm_t comm = nullptr;
  std::vector<std::unique_ptr<NcclComm>> split;
  SharedRounds shared;
  ~NcclComm() override
  {
    split.clear();
    shared.clear();
    if (comm)
      g_nccl.CommDestroy(comm);
  }
  Comm *sub(int i) override
  {
    if (i >= 0 && i < (int)shared.views.size())
      return shared.views[i].get();
    return i >= 0 && i < (int)split.size() ? split[i].get() : nullptr;
  }
  void enter_all() override
  {
    if (shared.rounds)
      shared.rounds->enter_all();
  }
  void allreduce_sum(double *buf, int count, cudaStream_t s) override
  {
    g_nccl.check(g_nccl.AllReduce(buf, buf, (size_t)count, ncclDouble, ncclSum, comm, s), "ncclAllReduce");
  }
  void group_begin() override { g_nccl.check(g_nccl.GroupStart(), "ncclGroupStart"); }
  void group_end() override { g_nccl.check(g_nccl.GroupEnd(), "ncclGroupEnd"); }
  void exchange_transfers(HaloDev &h, double *x, cudaStream_t s) override
  {
    for (size_t q = 0; q < h.neighbor_rank.size(); ++q)
      {
        const int32_t ns = h.send_ptr[q + 1] - h.send_ptr[q], nr = h.recv_ptr[q + 1] - h.recv_ptr[q];
        if (ns)
          g_nccl.check(g_nccl.Send(h.send_buf.ptr + h.send_ptr[q], (size_t)ns, ncclDouble, h.neighbor_rank[q], comm, s), "ncclSend");
        if (nr)
          g_nccl.check(g_nccl.Recv(x + h.n_owned + h.recv_ptr[q], (size_t)nr, ncclDouble, h.neighbor_rank[q], comm, s), "ncclRecv");
      }
  }
  void exchange(HaloDev &h, double *x, cudaStream_t s) override
  {
    if (h.empty())
      return;
    pack_for_send(h, x, s);
    group_begin();
    exchange_transfers(h, x, s);
    group_end();
  }
};

std::unique_ptr<Comm> g_process_comm;

// ------------------------------------------------------------------------ ranks as threads of one process
// N ranks as N host threads sharing ONE GPU: the same partitioned code path without NCCL (which refuses two
// ranks on one device).  Used by the single-GPU tests of the multi-rank path.  The exchange is a gather from
// the peer's vector, the reduction a sum over the peers' buffers in rank order (identical on all ranks).
struct ThreadGroup
{
  int n = 0;
  std::mutex m;
  std::condition_variable cv;
  int waiting = 0;
  long generation = 0;
  std::vector<double *> ptr;
  std::vector<HaloDev *> halo;
  void barrier()
  {
    std::unique_lock<std::mutex> lock(m);
    const long gen = generation;
    if (++waiting == n)
      {
        waiting = 0;
        ++generation;
        cv.notify_all();
      }
    else
      cv.wait(lock, [&] { return gen != generation; });
  }
};

struct PeerPtrs
{
  const double *p[16];
};

__global__ void k_sum_peers(double *out, PeerPtrs peers, int n_ranks, int count)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count)
    {
      double s = 0.0;
      for (int r = 0; r < n_ranks; ++r)
        s += peers.p[r][i];
      out[i] = s;
    }
}

__global__ void k_gather_from_peer(double *__restrict__ dst, const double *__restrict__ peer_x,
                                   const int32_t *__restrict__ peer_send_idx, int32_t n)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    dst[i] = peer_x[peer_send_idx[i]];
}

struct ThreadComm : Comm
{
  std::shared_ptr<ThreadGroup> g;
  DevBuf<double> tmp;
  void allreduce_sum(double *buf, int count, cudaStream_t s) override
  {
    S35_REQUIRE(count <= 64, "thread communicator: reduction too long");
    if (tmp.n < 64)
      tmp.alloc(64);
    S35_CUDA(cudaStreamSynchronize(s));
    g->ptr[rank] = buf;
    g->barrier();
    PeerPtrs peers;
    for (int r = 0; r < n_ranks; ++r)
      peers.p[r] = g->ptr[r];
    S35_LAUNCH(k_sum_peers, 1, 64, 0, s, tmp.ptr, peers, n_ranks, count);
    S35_CUDA(cudaStreamSynchronize(s));
    g->barrier(); // every rank has read every buffer
    S35_CUDA(cudaMemcpyAsync(buf, tmp.ptr, count * sizeof(double), cudaMemcpyDeviceToDevice, s));
  }
  SharedRounds shared;
  Comm *sub(int i) override { return i >= 0 && i < (int)shared.views.size() ? shared.views[i].get() : nullptr; }
  void enter_all() override
  {
    if (shared.rounds)
      shared.rounds->enter_all();
  }
  void exchange(HaloDev &h, double *x, cudaStream_t s) override { exchange_transfers(h, x, s); }
  void exchange_transfers(HaloDev &h, double *x, cudaStream_t s) override
  {
    S35_CUDA(cudaStreamSynchronize(s));
    g->ptr[rank] = x;
    g->halo[rank] = &h;
    g->barrier();
    for (size_t q = 0; q < h.neighbor_rank.size(); ++q)
      {
        const int peer = h.neighbor_rank[q];
        const int32_t nr = h.recv_ptr[q + 1] - h.recv_ptr[q];
        if (!nr)
          continue;
        HaloDev &ph = *g->halo[peer];
        size_t me = 0;
        while (me < ph.neighbor_rank.size() && ph.neighbor_rank[me] != rank)
          ++me;
        S35_REQUIRE(me < ph.neighbor_rank.size() && ph.send_ptr[me + 1] - ph.send_ptr[me] == nr,
                    "thread communicator: halo plans of two ranks do not match");
        S35_LAUNCH(k_gather_from_peer, (nr + 255) / 256, 256, 0, s, x + h.n_owned + h.recv_ptr[q], g->ptr[peer],
                   ph.send_idx.ptr + ph.send_ptr[me], nr);
      }
    S35_CUDA(cudaStreamSynchronize(s));
    g->barrier(); // nobody overwrites owned values before every neighbour has read them
  }
};

std::shared_ptr<ThreadGroup> g_thread_group;
thread_local std::unique_ptr<ThreadComm> t_thread_comm;

} // namespace

void HaloDev::create(const s35_halo &h, int32_t n_owned_, int32_t n_local_, cudaStream_t s)
{
  n_owned = n_owned_;
  n_local = n_local_;
  neighbor_rank.assign(h.neighbor_rank, h.neighbor_rank + h.n_neighbors);
  send_ptr.assign(h.send_ptr, h.send_ptr + h.n_neighbors + 1);
  recv_ptr.assign(h.recv_ptr, h.recv_ptr + h.n_neighbors + 1);
  S35_REQUIRE(recv_ptr.back() == n_local - n_owned, "halo: the receive segments do not cover the ghost DoFs");
  for (int32_t i = 0; i < send_ptr.back(); ++i)
    S35_REQUIRE(h.send_idx[i] >= 0 && h.send_idx[i] < n_owned, "halo: send index is not an owned DoF");
  send_idx.upload(h.send_idx, (size_t)send_ptr.back(), s);
  send_buf.alloc((size_t)std::max<int32_t>(send_ptr.back(), 1));
}

void pack_for_send(HaloDev &h, const double *x, cudaStream_t s)
{
  const int32_t n_send = h.empty() ? 0 : h.send_ptr.back();
  if (n_send)
    S35_LAUNCH(k_pack, (n_send + 255) / 256, 256, 0, s, h.send_buf.ptr, x, h.send_idx.ptr, n_send);
}

Comm *process_comm()
{
  if (t_thread_comm)
    return t_thread_comm.get();
  return g_process_comm.get();
}

void comm_unique_id(unsigned char id[128])
{
  g_nccl.load();
  ncclUniqueId uid;
  g_nccl.check(g_nccl.GetUniqueId(&uid), "ncclGetUniqueId");
  static_assert(sizeof(uid) == 128, "ncclUniqueId is expected to be 128 bytes");
  std::memcpy(id, &uid, 128);
}

void comm_init(int rank, int n_ranks, const unsigned char id[128])
{
  S35_REQUIRE(n_ranks >= 1 && rank >= 0 && rank < n_ranks, "comm_init: bad rank / n_ranks");
  g_nccl.load();
  std::unique_ptr<NcclComm> c(new NcclComm);
  c->rank = rank;
  c->n_ranks = n_ranks;
  ncclUniqueId uid;
  std::memcpy(&uid, id, 128);
  g_nccl.check(g_nccl.CommInitRank(&c->comm, n_ranks, uid, rank), "ncclCommInitRank");
  const int mode = concurrent_mode();
  if (mode == 2)
    c->shared.create(c.get());
  if (g_nccl.CommSplit && mode == 1)
    for (int i = 0; i < 3; ++i)
      {
        std::unique_ptr<NcclComm> sc(new NcclComm);
        sc->rank = rank;
        sc->n_ranks = n_ranks;
        g_nccl.check(g_nccl.CommSplit(c->comm, 0, rank, &sc->comm, nullptr), "ncclCommSplit");
        c->split.push_back(std::move(sc));
      }
  g_process_comm = std::move(c);
}

void comm_finalize()
{
  g_process_comm.reset();
  t_thread_comm.reset();
}

void comm_thread_group(int n_ranks)
{
  S35_REQUIRE(n_ranks >= 1 && n_ranks <= 16, "thread communicator: 1..16 ranks");
  std::shared_ptr<ThreadGroup> g(new ThreadGroup);
  g->n = n_ranks;
  g->ptr.assign(n_ranks, nullptr);
  g->halo.assign(n_ranks, nullptr);
  g_thread_group = g;
}

void comm_thread_bind(int rank)
{
  S35_REQUIRE(g_thread_group && rank >= 0 && rank < g_thread_group->n, "thread communicator: no group or bad rank");
  t_thread_comm.reset(new ThreadComm);
  t_thread_comm->g = g_thread_group;
  t_thread_comm->rank = rank;
  t_thread_comm->n_ranks = g_thread_group->n;
  if (concurrent_mode() == 2)
    t_thread_comm->shared.create(t_thread_comm.get());
}

void scalar_sqrt(double *x, cudaStream_t s) { S35_LAUNCH(k_scalar_sqrt, 1, 1, 0, s, x); }
void scalar_scale(double *x, double f, cudaStream_t s) { S35_LAUNCH(k_scalar_scale, 1, 1, 0, s, x, f); }

} // namespace s35
