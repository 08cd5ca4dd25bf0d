// How long does a shared-memory load take on a busy SM, and does it matter on which sub-partition (SMSP) the busy
// warps sit?  Background for the wavefront sweep (csrc/ilu.cu): its hand-off chain contains one gather from shared
// memory, issued while up to 20 other warps stream their own loads.
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o lds_queue lds_queue.cu && ./lds_queue
//
// One block of 25 warps.  Warp 0 is the probe: every ~600 cycles it times four independent 8-byte gathers and the
// sum that consumes them.  The other warps are "hammers": bursts of `burst` independent 8-byte gathers plus 16-byte
// streaming loads, as the sweep's look-ahead phase issues them.  mode: 0 no hammer; 1 hammers on all sub-partitions;
// 2 hammers only on sub-partitions 1..3 (warp index % 4 != 0), the probe's sub-partition is otherwise empty;
// 3 hammers only on the probe's sub-partition.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CHECK(x)                                                                                  \
  do                                                                                              \
    {                                                                                             \
      cudaError_t e = (x);                                                                        \
      if (e != cudaSuccess)                                                                       \
        {                                                                                         \
          std::printf("CUDA error %s at line %d: %s\n", cudaGetErrorName(e), __LINE__, #x);      \
          std::exit(1);                                                                           \
        }                                                                                         \
    }                                                                                             \
  while (0)

constexpr int kWords = 8192; // doubles in the gathered window

template <int BURST>
__device__ __forceinline__ double hammer_burst(const double *win, const double2 *stream, unsigned &state, int lane)
{
  double x[BURST];
#pragma unroll
  for (int k = 0; k < BURST; ++k)
    {
      state = state * 1664525u + 1013904223u;
      x[k] = win[(state >> 8) & (kWords - 1)];
    }
  double2 v[BURST / 2];
#pragma unroll
  for (int k = 0; k < BURST / 2; ++k)
    v[k] = stream[((state >> 4) + 32 * k) & 2047 & ~31 | lane];
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < BURST; ++k)
    s = fma(x[k], k & 1 ? v[k / 2].y : v[k / 2].x, s);
  return s;
}

template <int BURST>
__global__ void __launch_bounds__(800, 1) k_probe(int mode, int rounds, long long *cycles, double *sink)
{
  extern __shared__ __align__(16) unsigned char smem[];
  double *win = reinterpret_cast<double *>(smem);
  double2 *stream = reinterpret_cast<double2 *>(smem + kWords * 8);
  volatile int *stop = reinterpret_cast<volatile int *>(smem + kWords * 8 + 2048 * 16);
  for (int k = threadIdx.x; k < kWords; k += blockDim.x)
    win[k] = 1.0 / (1 + k);
  for (int k = threadIdx.x; k < 2048; k += blockDim.x)
    stream[k] = make_double2(1e-3 * k, 2e-3 * k);
  if (threadIdx.x == 0)
    *stop = 0;
  __syncthreads();
  const int warp = threadIdx.x / 32, lane = threadIdx.x & 31;
  unsigned state = 12345u + 977u * threadIdx.x;
  if (warp == 0)
    {
      long long total = 0;
      double s = 0.0;
      for (int r = 0; r < rounds; ++r)
        {
          unsigned a[4];
#pragma unroll
          for (int k = 0; k < 4; ++k)
            {
              state = state * 1664525u + 1013904223u;
              a[k] = (state >> 8) & (kWords - 1);
            }
          const long long t0 = clock64();
          const double x0 = win[a[0]], x1 = win[a[1]], x2 = win[a[2]], x3 = win[a[3]];
          s = fma(x0, 1.0000001, s);
          s = fma(x1, 1.0000001, s);
          s = fma(x2, 1.0000001, s);
          s = fma(x3, 1.0000001, s);
          // the clock must not be read before the sum exists
          const long long t1 = clock64() + (long long)(__double_as_longlong(s) & 0);
          total += t1 - t0;
          __nanosleep(300);
        }
      if (lane == 0)
        {
          cycles[0] = total;
          *stop = 1;
        }
      sink[threadIdx.x] = s;
      return;
    }
  const bool hammer = mode == 1 || (mode == 2 && warp % 4 != 0) || (mode == 3 && warp % 4 == 0);
  if (!hammer)
    return;
  double s = 0.0;
  while (!*stop)
    s += hammer_burst<BURST>(win, stream, state, lane);
  sink[threadIdx.x] = s;
}

template <int BURST>
void run(long long *cycles, double *sink)
{
  const int rounds = 2000;
  const size_t smem = kWords * 8 + 2048 * 16 + 16;
  CHECK(cudaFuncSetAttribute(k_probe<BURST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  std::printf("\"burst%d\": {", BURST);
  for (int mode = 0; mode < 4; ++mode)
    {
      k_probe<BURST><<<1, 800, smem>>>(mode, rounds, cycles, sink);
      CHECK(cudaDeviceSynchronize());
      static const char *names[] = {"idle", "hammers_everywhere", "hammers_on_other_subpartitions", "hammers_on_probe_subpartition"};
      std::printf("%s\"%s\": %.1f", mode ? ", " : "", names[mode], (double)cycles[0] / rounds);
    }
  std::printf("}");
}

int main()
{
  long long *cycles;
  double *sink;
  CHECK(cudaMallocManaged(&cycles, 64));
  CHECK(cudaMalloc(&sink, 1024 * 8));
  std::printf("{\"what\": \"cycles for four independent 8-byte shared-memory gathers and their sum, one probe warp\", ");
  run<2>(cycles, sink);
  std::printf(", ");
  run<8>(cycles, sink);
  std::printf(", ");
  run<16>(cycles, sink);
  std::printf("}\n");
  return 0;
}
