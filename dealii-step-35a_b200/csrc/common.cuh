// Shared plumbing for the sm_100a kernels: error handling, RAII device
// buffers, launch accounting.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/step35a_b200.h"

namespace s35 {

struct Error : std::runtime_error
{
  s35_status code;
  Error(s35_status c, const std::string &msg) : std::runtime_error(msg), code(c) {}
};

inline void cuda_check(cudaError_t e, const char *what, const char *file, int line)
{
  if (e != cudaSuccess)
    {
      char buf[512];
      std::snprintf(buf, sizeof(buf), "CUDA error %s at %s:%d: %s", cudaGetErrorName(e), file, line, what);
      throw Error(S35_ERR_CUDA, std::string(buf) + " (" + cudaGetErrorString(e) + ")");
    }
}
#define S35_CUDA(call) ::s35::cuda_check((call), #call, __FILE__, __LINE__)
#define S35_REQUIRE(cond, msg)                                    \
  do                                                              \
    {                                                             \
      if (!(cond))                                                \
        throw ::s35::Error(S35_ERR_INVALID, std::string(msg));    \
    }                                                             \
  while (0)

extern std::atomic<int64_t> g_launch_count;
// every kernel launch goes through this so that bench.py can report gpu_launches
#define S35_LAUNCH(kernel, grid, block, smem, stream, ...)                  \
  do                                                                        \
    {                                                                       \
      kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);           \
      ::s35::g_launch_count.fetch_add(1, std::memory_order_relaxed);        \
      S35_CUDA(cudaGetLastError());                                         \
    }                                                                       \
  while (0)

int sm_count();

template <typename T>
struct DevBuf
{
  T *ptr = nullptr;
  size_t n = 0;
  DevBuf() = default;
  explicit DevBuf(size_t count) { alloc(count); }
  DevBuf(const DevBuf &) = delete;
  DevBuf &operator=(const DevBuf &) = delete;
  DevBuf(DevBuf &&o) noexcept : ptr(o.ptr), n(o.n) { o.ptr = nullptr; o.n = 0; }
  DevBuf &operator=(DevBuf &&o) noexcept
  {
    if (this != &o)
      {
        release();
        ptr = o.ptr;
        n = o.n;
        o.ptr = nullptr;
        o.n = 0;
      }
    return *this;
  }
  ~DevBuf() { release(); }
  void release()
  {
    if (ptr)
      cudaFree(ptr);
    ptr = nullptr;
    n = 0;
  }
  void alloc(size_t count)
  {
    release();
    n = count;
    if (count)
      S35_CUDA(cudaMalloc((void **)&ptr, count * sizeof(T)));
  }
  void upload(const T *host, size_t count, cudaStream_t s = 0)
  {
    if (count != n)
      alloc(count);
    if (count)
      {
        S35_CUDA(cudaMemcpyAsync(ptr, host, count * sizeof(T), cudaMemcpyHostToDevice, s));
        S35_CUDA(cudaStreamSynchronize(s));
      }
  }
  void upload(const std::vector<T> &host, cudaStream_t s = 0) { upload(host.data(), host.size(), s); }
  void download(T *host, cudaStream_t s = 0) const
  {
    if (n)
      {
        S35_CUDA(cudaMemcpyAsync(host, ptr, n * sizeof(T), cudaMemcpyDeviceToHost, s));
        S35_CUDA(cudaStreamSynchronize(s));
      }
  }
  void zero(cudaStream_t s = 0)
  {
    if (n)
      S35_CUDA(cudaMemsetAsync(ptr, 0, n * sizeof(T), s));
  }
  operator T *() const { return ptr; }
};

// pinned host scalar block for reading solver scalars back without a pageable copy
template <typename T>
struct PinnedBuf
{
  T *ptr = nullptr;
  size_t n = 0;
  PinnedBuf() = default;
  explicit PinnedBuf(size_t count) : n(count) { S35_CUDA(cudaMallocHost((void **)&ptr, count * sizeof(T))); }
  PinnedBuf(const PinnedBuf &) = delete;
  PinnedBuf &operator=(const PinnedBuf &) = delete;
  ~PinnedBuf()
  {
    if (ptr)
      cudaFreeHost(ptr);
  }
  void alloc(size_t count)
  {
    if (ptr)
      cudaFreeHost(ptr);
    n = count;
    S35_CUDA(cudaMallocHost((void **)&ptr, count * sizeof(T)));
  }
};

} // namespace s35
