// Micro-benchmarks behind the design of the substitution sweeps and the assembly roofline (run on the GPU box):
//   fp64 FMA peak, dependent-DFMA latency, what ONE SM can stream (bulk copies and plain loads), and what a hand-off
//   of a value costs between warps of a block (named barrier, mbarrier), between blocks of a cluster (remote store +
//   mbarrier transaction, cluster barrier) and between blocks through L2.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o latency latency.cu && ./latency
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

namespace cg = cooperative_groups;
#define CK(x)                                                                                    \
  do                                                                                             \
    {                                                                                            \
      cudaError_t e = (x);                                                                       \
      if (e != cudaSuccess)                                                                      \
        {                                                                                        \
          std::printf("CUDA error %s at line %d: %s\n", cudaGetErrorName(e), __LINE__, #x);      \
          std::exit(1);                                                                          \
        }                                                                                        \
    }                                                                                            \
  while (0)

__device__ __forceinline__ uint32_t saddr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, int c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(saddr(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint64_t *b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(saddr(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t *b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(saddr(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity)
{
  asm volatile("{\n.reg .pred p;\nLAB_WAIT:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra LAB_DONE;\nbra LAB_WAIT;\nLAB_DONE:\n}" ::"r"(saddr(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(saddr(dst)), "l"(src), "r"(bytes), "r"(saddr(bar)) : "memory");
}

// ---------------------------------------------------------------- 1. FP64 FMA peak
__global__ void k_fp64_peak(double *out, int iters)
{
  double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  const double m = 1.0000001, c = 1e-7;
  for (int i = 0; i < iters; ++i)
    {
      a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
      a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

// ---------------------------------------------------------------- 2. dependent DFMA latency (one warp)
__global__ void k_dfma_latency(double *out, long long *cycles, int iters)
{
  double a = threadIdx.x * 1e-9;
  const double m = 1.0000001, c = 1e-7;
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i)
    {
      a = fma(a, m, c); a = fma(a, m, c); a = fma(a, m, c); a = fma(a, m, c);
      a = fma(a, m, c); a = fma(a, m, c); a = fma(a, m, c); a = fma(a, m, c);
    }
  const long long t1 = clock64();
  out[threadIdx.x] = a;
  if (threadIdx.x == 0)
    cycles[0] = t1 - t0;
}

// ---------------------------------------------------------------- 3. streaming into one SM
// bulk copies: lane 0 of warp 0 keeps `stages` chunks in flight; warp 1 consumes (touches one word) and frees
__global__ void __launch_bounds__(64, 1) k_stream_bulk(const char *src, size_t bytes_per_block, int chunk, int stages, long long *cycles, double *sink)
{
  extern __shared__ __align__(128) unsigned char smem[];
  uint64_t *full = reinterpret_cast<uint64_t *>(smem), *empty = full + 16;
  unsigned char *buf = smem + 256;
  const char *mine = src + (size_t)blockIdx.x * bytes_per_block;
  const int n_chunks = (int)(bytes_per_block / chunk);
  if (threadIdx.x == 0)
    {
      for (int s = 0; s < stages; ++s)
        {
          mbar_init(full + s, 1);
          mbar_init(empty + s, 1);
        }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
  __syncthreads();
  const long long t0 = clock64();
  if (threadIdx.x == 0)
    {
      for (int c = 0; c < n_chunks; ++c)
        {
          const int s = c % stages;
          if (c >= stages)
            mbar_wait(empty + s, (uint32_t)((c / stages - 1) & 1));
          mbar_expect(full + s, (uint32_t)chunk);
          bulk_g2s(buf + (size_t)s * chunk, mine + (size_t)c * chunk, (uint32_t)chunk, full + s);
        }
    }
  else if (threadIdx.x == 32)
    {
      double acc = 0;
      for (int c = 0; c < n_chunks; ++c)
        {
          const int s = c % stages;
          mbar_wait(full + s, (uint32_t)((c / stages) & 1));
          acc += *reinterpret_cast<double *>(buf + (size_t)s * chunk);
          mbar_arrive(empty + s);
        }
      sink[blockIdx.x] = acc;
      cycles[blockIdx.x] = clock64() - t0;
    }
}
// plain loads: every thread of the block streams 16-byte words, 8 in flight
__global__ void __launch_bounds__(1024, 1) k_stream_ldg(const double2 *src, size_t words_per_block, long long *cycles, double *sink)
{
  const double2 *mine = src + (size_t)blockIdx.x * words_per_block;
  double acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (size_t i = threadIdx.x; i + 7 * blockDim.x < words_per_block; i += 8 * blockDim.x)
    {
      double2 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u)
        v[u] = __ldcs(mine + i + (size_t)u * blockDim.x);
#pragma unroll
      for (int u = 0; u < 8; ++u)
        acc += v[u].x + v[u].y;
    }
  __syncthreads();
  if (threadIdx.x == 0)
    cycles[blockIdx.x] = clock64() - t0;
  if (acc == 123.456)
    sink[0] = acc;
}

// ---------------------------------------------------------------- 4. hand-off between two warp groups of a block
// group A (threads 0..W) and group B (W..2W) alternate: store to shared, signal, the other waits, loads, stores, ...
template <int MODE> // 0: named barriers (arrive / sync), 1: mbarrier (one arrive per warp)
__global__ void __launch_bounds__(512, 1) k_handoff_block(int W, int rounds, long long *cycles, double *sink)
{
  __shared__ double slot[512];
  __shared__ uint64_t bar[2];
  const int tid = threadIdx.x, g = tid / W, i = tid % W;
  if (tid == 0)
    {
      mbar_init(bar + 0, W / 32);
      mbar_init(bar + 1, W / 32);
    }
  slot[tid] = 1.0;
  __syncthreads();
  double v = 1.0;
  const long long t0 = clock64();
  for (int r = 0; r < rounds; ++r)
    {
      if (g == 0)
        {
          // wait for B's previous result (round r-1), compute, publish
          if (r > 0)
            {
              if (MODE == 0)
                asm volatile("bar.sync 2, %0;" ::"r"(2 * W) : "memory");
              else
                mbar_wait(bar + 1, (uint32_t)((r - 1) & 1));
            }
          v = fma(slot[W + ((i + 1) % W)], 0.999, 0.001);
          slot[i] = v;
          if (MODE == 0)
            asm volatile("bar.arrive 1, %0;" ::"r"(2 * W) : "memory");
          else
            {
              __syncwarp();
              if ((tid & 31) == 0)
                mbar_arrive(bar + 0);
            }
        }
      else
        {
          if (MODE == 0)
            asm volatile("bar.sync 1, %0;" ::"r"(2 * W) : "memory");
          else
            mbar_wait(bar + 0, (uint32_t)(r & 1));
          v = fma(slot[(i + 1) % W], 0.999, 0.001);
          slot[W + i] = v;
          if (MODE == 0)
            asm volatile("bar.arrive 2, %0;" ::"r"(2 * W) : "memory");
          else
            {
              __syncwarp();
              if ((tid & 31) == 0)
                mbar_arrive(bar + 1);
            }
        }
    }
  const long long t1 = clock64();
  if (MODE == 0 && g == 0) // drain the last arrive of B
    asm volatile("bar.sync 2, %0;" ::"r"(2 * W) : "memory");
  if (tid == 0)
    cycles[0] = t1 - t0;
  sink[tid] = v;
}
// every thread of ONE group: load, fma, store, __syncthreads-like named barrier (the plain level loop)
__global__ void __launch_bounds__(1024, 1) k_level_loop(int rounds, long long *cycles, double *sink)
{
  __shared__ double slot[1024];
  const int tid = threadIdx.x, n = blockDim.x;
  slot[tid] = 1.0;
  __syncthreads();
  double v = 1.0;
  const long long t0 = clock64();
  for (int r = 0; r < rounds; ++r)
    {
      v = fma(slot[(tid + 1) % n], 0.999, 0.001);
      asm volatile("bar.sync 1, %0;" ::"r"(n) : "memory");
      slot[tid] = v;
      asm volatile("bar.sync 1, %0;" ::"r"(n) : "memory");
    }
  const long long t1 = clock64();
  if (tid == 0)
    cycles[0] = t1 - t0;
  sink[tid] = v;
}

// ---------------------------------------------------------------- 5. hand-off between blocks of a cluster
// MODE 0: remote 8-byte store that completes a transaction on the receiver's mbarrier (st.async); ring of all CTAs
// MODE 1: plain remote store + barrier.cluster arrive/wait per round (all CTAs exchange)
template <int MODE>
__global__ void __launch_bounds__(128, 1) k_handoff_cluster(int rounds, long long *cycles, double *sink)
{
  __shared__ __align__(16) double inbox[8];
  __shared__ uint64_t bar;
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank(), size = cluster.num_blocks();
  const unsigned next = (rank + 1) % size;
  if (threadIdx.x == 0)
    {
      mbar_init(&bar, 1);
      inbox[0] = 1.0;
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
  cluster.sync();
  double v = 1.0;
  long long t0 = clock64();
  if (MODE == 0)
    {
      if (threadIdx.x == 0)
        {
          uint32_t remote_inbox, remote_bar;
          asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote_inbox) : "r"(saddr(inbox)), "r"(next));
          asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote_bar) : "r"(saddr(&bar)), "r"(next));
          // a token travels round the ring: rank 0 starts it
          for (int r = 0; r < rounds; ++r)
            {
              const bool my_turn_first = rank == 0;
              if (!(my_turn_first && r == 0))
                {
                  mbar_expect(&bar, 8); // armed before the sender can possibly send: see below
                }
              if (my_turn_first)
                {
                  if (r > 0)
                    {
                      mbar_wait(&bar, (uint32_t)((r - 1) & 1));
                      v = fma(inbox[0], 0.999, 0.001);
                    }
                  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(remote_inbox), "l"(__double_as_longlong(v)), "r"(remote_bar) : "memory");
                }
              else
                {
                  mbar_wait(&bar, (uint32_t)(r & 1));
                  v = fma(inbox[0], 0.999, 0.001);
                  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(remote_inbox), "l"(__double_as_longlong(v)), "r"(remote_bar) : "memory");
                }
            }
          if (rank == 0)
            {
              mbar_expect(&bar, 8);
              mbar_wait(&bar, (uint32_t)((rounds - 1) & 1));
            }
        }
    }
  else
    {
      for (int r = 0; r < rounds; ++r)
        {
          if (threadIdx.x == 0)
            {
              double *remote = cluster.map_shared_rank(inbox, next);
              remote[1 + (r & 1)] = v;
            }
          asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
          v = fma(inbox[1 + (r & 1)], 0.999, 0.001);
        }
    }
  const long long t1 = clock64();
  if (threadIdx.x == 0 && rank == 0)
    cycles[0] = t1 - t0;
  cluster.sync();
  sink[blockIdx.x * blockDim.x + threadIdx.x] = v;
}

// ---------------------------------------------------------------- 6. hand-off between two blocks through L2
__global__ void k_handoff_l2(volatile double *box, int rounds, long long *cycles)
{
  // box[0]: written by block 0, box[16]: written by block 1; values count up
  if (threadIdx.x != 0)
    return;
  const int me = blockIdx.x;
  const long long t0 = clock64();
  for (int r = 1; r <= rounds; ++r)
    {
      if (me == 0)
        {
          box[0] = (double)r;
          while (box[16] != (double)r)
            ;
        }
      else
        {
          while (box[0] != (double)r)
            ;
          box[16] = (double)r;
        }
    }
  if (me == 0)
    cycles[0] = clock64() - t0;
}

int main()
{
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  int clock_khz = 0;
  CK(cudaDeviceGetAttribute(&clock_khz, cudaDevAttrClockRate, 0));
  std::printf("{\"device\": \"%s\", \"sms\": %d, \"clock_mhz\": %.0f", prop.name, prop.multiProcessorCount, clock_khz / 1000.0);
  long long *cycles;
  double *sink;
  CK(cudaMallocManaged(&cycles, 1024 * sizeof(long long)));
  CK(cudaMalloc(&sink, (1 << 22) * sizeof(double)));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  float ms;

  { // 1
    const int iters = 1 << 16, blocks = prop.multiProcessorCount * 8, threads = 256;
    k_fp64_peak<<<blocks, threads>>>(sink, 1024);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    k_fp64_peak<<<blocks, threads>>>(sink, iters);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    CK(cudaEventElapsedTime(&ms, e0, e1));
    const double flop = 2.0 * 8 * iters * (double)blocks * threads;
    std::printf(", \"fp64_fma_tflops\": %.2f", flop / (ms * 1e-3) / 1e12);
  }
  { // 2
    const int iters = 4096;
    k_dfma_latency<<<1, 32>>>(sink, cycles, iters);
    CK(cudaDeviceSynchronize());
    std::printf(", \"dfma_dependent_latency_cycles\": %.2f", (double)cycles[0] / (8.0 * iters));
  }
  { // 3
    const size_t total = (size_t)1 << 30;
    char *src;
    CK(cudaMalloc(&src, total));
    CK(cudaMemset(src, 1, total));
    std::printf(", \"one_sm_stream\": {");
    bool first = true;
    for (int blocks : {1, 2, 4, 8})
      for (int chunk : {4096, 16384, 32768})
        for (int stages : {2, 4, 6})
          {
            if ((size_t)chunk * stages > 200 * 1024)
              continue;
            const size_t per_block = (size_t)64 << 20; // 64 MB per block from HBM
            const size_t smem = 256 + (size_t)chunk * stages;
            CK(cudaFuncSetAttribute(k_stream_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            k_stream_bulk<<<blocks, 64, smem>>>(src, per_block, chunk, stages, cycles, sink);
            CK(cudaDeviceSynchronize());
            long long worst = 0;
            for (int b = 0; b < blocks; ++b)
              worst = std::max(worst, cycles[b]);
            std::printf("%s\"bulk_b%d_c%d_s%d_B_per_clk\": %.1f", first ? "" : ", ", blocks, chunk, stages, (double)per_block / worst);
            first = false;
          }
    // the same from L2 (a 32 MB region read twice: second pass timed)
    for (int chunk : {16384})
      {
        const size_t per_block = (size_t)32 << 20, smem = 256 + (size_t)chunk * 6;
        CK(cudaFuncSetAttribute(k_stream_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_stream_bulk<<<1, 64, smem>>>(src, per_block, chunk, 6, cycles, sink);
        CK(cudaDeviceSynchronize());
        k_stream_bulk<<<1, 64, smem>>>(src, per_block, chunk, 6, cycles, sink);
        CK(cudaDeviceSynchronize());
        std::printf(", \"bulk_from_L2_b1_c%d_s6_B_per_clk\": %.1f", chunk, (double)per_block / cycles[0]);
      }
    for (int blocks : {1, 2, 4})
      {
        const size_t per_block = (size_t)64 << 20;
        k_stream_ldg<<<blocks, 1024>>>(reinterpret_cast<const double2 *>(src), per_block / 16, cycles, sink);
        CK(cudaDeviceSynchronize());
        long long worst = 0;
        for (int b = 0; b < blocks; ++b)
          worst = std::max(worst, cycles[b]);
        std::printf(", \"ldg128_b%d_B_per_clk\": %.1f", blocks, (double)per_block / worst);
      }
    std::printf("}");
    CK(cudaFree(src));
  }
  { // 4
    const int rounds = 20000;
    std::printf(", \"block_handoff_cycles\": {");
    for (int W : {32, 128, 256})
      {
        k_handoff_block<0><<<1, 2 * W>>>(W, rounds, cycles, sink);
        CK(cudaDeviceSynchronize());
        std::printf("\"named_barrier_W%d\": %.1f, ", W, (double)cycles[0] / (2.0 * rounds));
        k_handoff_block<1><<<1, 2 * W>>>(W, rounds, cycles, sink);
        CK(cudaDeviceSynchronize());
        std::printf("\"mbarrier_W%d\": %.1f, ", W, (double)cycles[0] / (2.0 * rounds));
      }
    for (int n : {128, 256, 512, 1024})
      {
        k_level_loop<<<1, n>>>(rounds, cycles, sink);
        CK(cudaDeviceSynchronize());
        std::printf("\"two_barriers_per_level_T%d\": %.1f%s", n, (double)cycles[0] / rounds, n == 1024 ? "" : ", ");
      }
    std::printf("}");
  }
  { // 5
    const int rounds = 20000;
    std::printf(", \"cluster_handoff_cycles\": {");
    bool first = true;
    for (int size : {2, 4, 8})
      {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(size);
        cfg.blockDim = dim3(128);
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = size;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        CK(cudaLaunchKernelEx(&cfg, k_handoff_cluster<0>, rounds, cycles, sink));
        CK(cudaDeviceSynchronize());
        std::printf("%s\"st_async_per_hop_size%d\": %.1f", first ? "" : ", ", size, (double)cycles[0] / ((double)rounds * size));
        first = false;
        CK(cudaLaunchKernelEx(&cfg, k_handoff_cluster<1>, rounds, cycles, sink));
        CK(cudaDeviceSynchronize());
        std::printf(", \"store_plus_cluster_barrier_size%d\": %.1f", size, (double)cycles[0] / rounds);
      }
    std::printf("}");
  }
  { // 6
    double *box;
    CK(cudaMalloc(&box, 64 * sizeof(double)));
    CK(cudaMemset(box, 0, 64 * sizeof(double)));
    k_handoff_l2<<<2, 32>>>(box, 20000, cycles);
    CK(cudaDeviceSynchronize());
    std::printf(", \"l2_handoff_cycles_per_hop\": %.1f", (double)cycles[0] / (2.0 * 20000));
  }
  std::printf("}\n");
  return 0;
}
