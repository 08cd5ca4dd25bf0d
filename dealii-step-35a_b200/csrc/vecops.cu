#include "vecops.cuh"

namespace s35 {

std::atomic<int64_t> g_launch_count{0};

int sm_count()
{
  static int cached[64] = {0};
  int dev = 0;
  S35_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && cached[dev])
    return cached[dev];
  int n = 0;
  S35_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
  if (dev < 64)
    cached[dev] = n;
  return n;
}

namespace {

constexpr int kBlock = 256;

__device__ __forceinline__ double load_scalar(const DScal &a)
{
  if (a.ptr == nullptr)
    return a.mul;
  const double v = *a.ptr;
  return a.mul * (a.recip ? 1.0 / v : v);
}

inline int stream_grid(int64_t n, int per_thread)
{
  const int64_t blocks = (n + (int64_t)kBlock * per_thread - 1) / ((int64_t)kBlock * per_thread);
  const int64_t cap = (int64_t)sm_count() * 8;
  return (int)std::max<int64_t>(1, std::min(blocks, cap));
}

__global__ void __launch_bounds__(kBlock) k_lincomb(double *__restrict__ z, DScal a, const double *x, DScal b, const double *y,
                                                     int64_t n)
{
  const double av = load_scalar(a), bv = load_scalar(b);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (y == nullptr)
    for (; i < n; i += stride)
      z[i] = av * x[i];
  else
    for (; i < n; i += stride)
      z[i] = av * x[i] + bv * y[i];
}

__global__ void __launch_bounds__(kBlock) k_lincomb3(double *__restrict__ z, DScal a, const double *x, DScal b, const double *y,
                                                      DScal c, const double *w, int64_t n)
{
  const double av = load_scalar(a), bv = load_scalar(b), cv = load_scalar(c);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    z[i] = av * x[i] + bv * y[i] + cv * w[i];
}

__global__ void __launch_bounds__(kBlock) k_multiply(double *__restrict__ z, const double *x, const double *y, int64_t n)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    z[i] = x[i] * y[i];
}

__global__ void __launch_bounds__(kBlock) k_set(double *__restrict__ z, double v, int64_t n)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    z[i] = v;
}

// block-wide sum in a fixed order; result valid in thread 0
__device__ __forceinline__ double block_sum(double v)
{
  __shared__ double warp_part[kBlock / 32];
  for (int o = 16; o > 0; o >>= 1)
    v += __shfl_down_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads(); // protect warp_part against a previous call
  if (lane == 0)
    warp_part[warp] = v;
  __syncthreads();
  if (warp == 0)
    {
      v = lane < kBlock / 32 ? warp_part[lane] : 0.0;
      for (int o = 16; o > 0; o >>= 1)
        v += __shfl_down_sync(0xffffffffu, v, o);
    }
  return v;
}

// the last block to arrive adds the per-block partials in index order
__device__ __forceinline__ void finish_reduction(double block_value, double *partials, unsigned int *ticket, double *result,
                                                 int post)
{
  __shared__ bool is_last;
  if (threadIdx.x == 0)
    {
      partials[blockIdx.x] = block_value;
      __threadfence();
      const unsigned int t = atomicAdd(ticket, 1u);
      is_last = (t == gridDim.x - 1);
    }
  __syncthreads();
  if (is_last)
    {
      __threadfence();
      double v = 0.0;
      for (int i = threadIdx.x; i < gridDim.x; i += kBlock)
        v += __ldcg(partials + i);
      v = block_sum(v);
      if (threadIdx.x == 0)
        {
          *result = post == kSqrt ? sqrt(v) : v;
          *ticket = 0u;
        }
    }
}

__global__ void __launch_bounds__(kBlock) k_dot(double *result, const double *__restrict__ x, const double *__restrict__ y,
                                                 int64_t n, int post, double *partials, unsigned int *ticket)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  double acc = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    acc += x[i] * y[i];
  acc = block_sum(acc);
  finish_reduction(acc, partials, ticket, result, post);
}

__global__ void __launch_bounds__(kBlock) k_add_and_dot(double *result, double *y, DScal a, const double *__restrict__ x,
                                                         const double *w, int64_t n, int post, double *partials,
                                                         unsigned int *ticket)
{
  const double av = load_scalar(a);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  double acc = 0.0;
  const bool self = (w == y);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    {
      const double yi = y[i] + av * x[i];
      y[i] = yi;
      acc += yi * (self ? yi : w[i]);
    }
  acc = block_sum(acc);
  finish_reduction(acc, partials, ticket, result, post);
}

__global__ void __launch_bounds__(kBlock) k_cg_update(double *result, double *x, double *g, DScal alpha,
                                                       const double *__restrict__ d, const double *__restrict__ h, int64_t n,
                                                       int post, double *partials, unsigned int *ticket)
{
  const double av = load_scalar(alpha);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  double acc = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    {
      x[i] += av * d[i];
      const double gi = g[i] + av * h[i];
      g[i] = gi;
      acc += gi * gi;
    }
  acc = block_sum(acc);
  finish_reduction(acc, partials, ticket, result, post);
}

__global__ void __launch_bounds__(kBlock) k_gather_abs_mean(double *result, const double *__restrict__ values,
                                                             const int32_t *__restrict__ idx, int32_t n, double *partials,
                                                             unsigned int *ticket)
{
  const int stride = gridDim.x * blockDim.x;
  double acc = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    acc += fabs(values[idx[i]]);
  acc = block_sum(acc);
  // the last block divides the total by n
  __shared__ bool is_last;
  if (threadIdx.x == 0)
    {
      partials[blockIdx.x] = acc;
      __threadfence();
      is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    }
  __syncthreads();
  if (is_last)
    {
      __threadfence();
      double v = 0.0;
      for (int i = threadIdx.x; i < gridDim.x; i += kBlock)
        v += __ldcg(partials + i);
      v = block_sum(v);
      if (threadIdx.x == 0)
        {
          *result = v / n;
          *ticket = 0u;
        }
    }
}

struct Coefs
{
  double c[32];
};

__global__ void __launch_bounds__(kBlock) k_multi_axpy(double *x, const double *__restrict__ v0, int64_t vstride, Coefs coef,
                                                        int m, int64_t n)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    {
      double xi = x[i];
      // same order as the reference's loop of x.add(h(i), tmp_vectors[i])
      for (int k = 0; k < m; ++k)
        xi += coef.c[k] * v0[(int64_t)k * vstride + i];
      x[i] = xi;
    }
}

__global__ void k_scatter_set(double *x, const int32_t *__restrict__ idx, const double *__restrict__ val, int32_t m)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < m)
    x[idx[i]] = val ? val[i] : 0.0;
}

__global__ void k_scalar_divide(double *out, const double *num, const double *den, double *save_num)
{
  const double nv = *num, dv = *den;
  *out = nv / dv;
  if (save_num)
    *save_num = nv;
}

} // namespace

void vec_lincomb(double *z, DScal a, const double *x, DScal b, const double *y, int64_t n, cudaStream_t s)
{
  if (n == 0)
    return;
  S35_LAUNCH(k_lincomb, stream_grid(n, 4), kBlock, 0, s, z, a, x, b, y, n);
}

void vec_lincomb3(double *z, DScal a, const double *x, DScal b, const double *y, DScal c, const double *w, int64_t n,
                  cudaStream_t s)
{
  if (n == 0)
    return;
  S35_LAUNCH(k_lincomb3, stream_grid(n, 4), kBlock, 0, s, z, a, x, b, y, c, w, n);
}

void vec_multiply(double *z, const double *x, const double *y, int64_t n, cudaStream_t s)
{
  if (n == 0)
    return;
  S35_LAUNCH(k_multiply, stream_grid(n, 4), kBlock, 0, s, z, x, y, n);
}

void vec_copy(double *dst, const double *src, int64_t n, cudaStream_t s)
{
  if (n)
    S35_CUDA(cudaMemcpyAsync(dst, src, n * sizeof(double), cudaMemcpyDeviceToDevice, s));
}

void vec_set(double *dst, double value, int64_t n, cudaStream_t s)
{
  if (n == 0)
    return;
  S35_LAUNCH(k_set, stream_grid(n, 4), kBlock, 0, s, dst, value, n);
}

static int reduce_grid(int64_t n)
{
  const int64_t blocks = (n + (int64_t)kBlock * 8 - 1) / ((int64_t)kBlock * 8);
  return (int)std::max<int64_t>(1, std::min<int64_t>(blocks, ReduceWs::kMaxBlocks));
}

void vec_dot(double *result, const double *x, const double *y, int64_t n, ReducePost post, ReduceWs &ws, cudaStream_t s)
{
  S35_LAUNCH(k_dot, reduce_grid(n), kBlock, 0, s, result, x, y, n, (int)post, ws.partials.ptr, ws.ticket.ptr);
}

void vec_add_and_dot(double *result, double *y, DScal a, const double *x, const double *w, int64_t n, ReducePost post,
                     ReduceWs &ws, cudaStream_t s)
{
  S35_LAUNCH(k_add_and_dot, reduce_grid(n), kBlock, 0, s, result, y, a, x, w, n, (int)post, ws.partials.ptr, ws.ticket.ptr);
}

void vec_cg_update(double *result, double *x, double *g, DScal alpha, const double *d, const double *h, int64_t n,
                   ReducePost post, ReduceWs &ws, cudaStream_t s)
{
  S35_LAUNCH(k_cg_update, reduce_grid(n), kBlock, 0, s, result, x, g, alpha, d, h, n, (int)post, ws.partials.ptr, ws.ticket.ptr);
}

void vec_gather_abs_mean(double *result, const double *values, const int32_t *idx, int32_t n, ReduceWs &ws, cudaStream_t s)
{
  S35_LAUNCH(k_gather_abs_mean, reduce_grid(n), kBlock, 0, s, result, values, idx, n, ws.partials.ptr, ws.ticket.ptr);
}

void vec_multi_axpy(double *x, const double *v0, int64_t stride, const double *coef_host, int m, int64_t n, cudaStream_t s)
{
  S35_REQUIRE(m <= 32, "vec_multi_axpy: at most 32 vectors");
  if (m == 0 || n == 0)
    return;
  Coefs c;
  for (int i = 0; i < 32; ++i)
    c.c[i] = i < m ? coef_host[i] : 0.0;
  S35_LAUNCH(k_multi_axpy, stream_grid(n, 2), kBlock, 0, s, x, v0, stride, c, m, n);
}

void vec_scatter_set(double *x, const int32_t *idx, const double *val, int32_t m, cudaStream_t s)
{
  if (m == 0)
    return;
  S35_LAUNCH(k_scatter_set, (m + 127) / 128, 128, 0, s, x, idx, val, m);
}

void scalar_divide(double *out, const double *num, const double *den, double *save_num, cudaStream_t s)
{
  S35_LAUNCH(k_scalar_divide, 1, 1, 0, s, out, num, den, save_num);
}

void flush_l2(cudaStream_t s)
{
  static thread_local DevBuf<double> *buf = nullptr;
  const size_t n = (size_t)160 * 1024 * 1024 / sizeof(double);
  if (!buf)
    {
      buf = new DevBuf<double>();
      buf->alloc(n);
    }
  S35_CUDA(cudaMemsetAsync(buf->ptr, 0, n * sizeof(double), s));
}

} // namespace s35

// ------------------------------------------------------------------------------------------------------
// s35_vec_dot_device_noalloc: the dot product for callers that must not allocate.
//
// Who needs this: a host program that sizes all of its device memory once at start-up and then runs its
// Krylov loop against a fixed budget (cudaMalloc may fail there, and even when it succeeds it stalls every
// stream of the device); a caller whose buffers all come from its own pool and who is not allowed to reach
// past that pool into the driver; code that is in the middle of recording a stream and therefore cannot
// call anything that synchronises.  The ordinary s35_vec_dot stages host vectors in freshly allocated device
// buffers and allocates its reduction scratch on every call, so none of those callers can use it.
//
// This variant takes device pointers for both vectors and for the result and launches the same reduction
// kernel; the per-block partial sums and the ticket counter live in the array below, which is part of the
// library image: every GPU that loads the library carries these 8.25 KB until the process ends, and nothing
// is ever allocated or freed.  The array holds one partial per block of 2048 elements, 1024 of them, which
// is where the size limit of 2 097 152 elements comes from.
// ------------------------------------------------------------------------------------------------------
#include <mutex>

namespace s35 {
void set_last_error(const std::string &msg); // capi.cu

namespace {

constexpr int kNoAllocMaxBlocks = 1024;
constexpr int64_t kNoAllocMaxN = (int64_t)kNoAllocMaxBlocks * kBlock * 8; // 2 097 152 elements
// [0, 1024): per-block partial sums; [1024]: the ticket counter of the single-pass reduction
static __device__ __align__(256) double g_noalloc_scratch[kNoAllocMaxBlocks + 32];

// Device address of the scratch array on the current device.  In host code `g_noalloc_scratch` is only a
// handle for the symbol, not something a kernel can dereference, so the address is looked up once per device
// (with that device current) and remembered.
double *noalloc_scratch_on_current_device()
{
  static std::mutex mutex;
  static double *address[64] = {};
  int device = 0;
  S35_CUDA(cudaGetDevice(&device));
  S35_REQUIRE(device >= 0 && device < 64, "vec_dot_device_noalloc: device ordinal above 63");
  std::lock_guard<std::mutex> lock(mutex);
  if (address[device] == nullptr)
    {
      double *p = nullptr;
      S35_CUDA(cudaGetSymbolAddress((void **)&p, g_noalloc_scratch));
      S35_REQUIRE(p != nullptr, "vec_dot_device_noalloc: no device address for the scratch array");
      address[device] = p;
    }
  return address[device];
}

} // namespace
} // namespace s35

extern "C" s35_status s35_vec_dot_device_noalloc(int64_t n, const double *x_device, const double *y_device,
                                                 double *result_device, void *stream)
{
  using namespace s35;
  try
    {
      S35_REQUIRE(n >= 0, "vec_dot_device_noalloc: negative length");
      S35_REQUIRE(result_device != nullptr && (n == 0 || (x_device != nullptr && y_device != nullptr)),
                  "vec_dot_device_noalloc: null pointer");
      if (n > kNoAllocMaxN)
        throw Error(S35_ERR_INVALID, "vec_dot_device_noalloc: " + std::to_string(n) + " elements exceed the limit of " +
                                       std::to_string(kNoAllocMaxN) +
                                       " set by the built-in scratch array; use s35_vec_dot, which allocates");
      cudaStream_t s = static_cast<cudaStream_t>(stream);
      double *scratch = noalloc_scratch_on_current_device();
      unsigned int *ticket = reinterpret_cast<unsigned int *>(scratch + kNoAllocMaxBlocks);
      // the array still holds whatever the previous call left: the ticket must start at zero
      S35_CUDA(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), s));
      S35_LAUNCH(k_dot, reduce_grid(n), kBlock, 0, s, result_device, x_device, y_device, n, (int)kSum, scratch, ticket);
      return S35_OK;
    }
  catch (const Error &e)
    {
      set_last_error(e.what());
      return e.code;
    }
  catch (const std::exception &e)
    {
      set_last_error(e.what());
      return S35_ERR_INVALID;
    }
}
