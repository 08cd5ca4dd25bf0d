#include "ilu.cuh"

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace s35 {

namespace {

constexpr int kBlock = 128;

__device__ __forceinline__ double ld_poll(const double *p)
{
  double v;
  asm volatile("ld.volatile.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ int ld_poll_int(const int *p)
{
  int v;
  asm volatile("ld.volatile.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ bool is_sentinel(double v) { return __double_as_longlong(v) == -1LL; }
__device__ __forceinline__ double sentinel() { return __longlong_as_double(-1LL); }
// a NaN result must never look like the sentinel
__device__ __forceinline__ double canonical(double v) { return v != v ? __longlong_as_double(0x7ff8000000000000LL) : v; }

template <int G>
__device__ __forceinline__ unsigned group_mask()
{
  if (G == 32)
    return 0xffffffffu;
  const int first = ((threadIdx.x & 31) / G) * G;
  return ((1u << G) - 1u) << first;
}

// Forward substitution with the unit lower factor: y_r = src_r - sum_{c<r} L_rc y_c.
// Also arms dst (the output of the backward sweep) with the sentinel.
template <int G>
__global__ void __launch_bounds__(kBlock)
k_ilu_forward(int32_t n, const int32_t *__restrict__ order, const int32_t *__restrict__ rowptr,
              const int32_t *__restrict__ diag, const int32_t *__restrict__ col, const double *__restrict__ lu,
              const double *__restrict__ src, double *y, double *dst, int32_t *tickets)
{
  constexpr int ROWS = kBlock / G;
  __shared__ int s_base;
  if (threadIdx.x == 0)
    {
      const int t = atomicAdd(&tickets[0], 1);
      if (t == 0)
        tickets[1] = 0; // the backward sweep is not running: reset its ticket
      s_base = t * ROWS;
    }
  __syncthreads();
  const int lane = threadIdx.x % G;
  const int p = s_base + threadIdx.x / G;
  const unsigned gmask = group_mask<G>();
  int r = -1, k = 0, end = 0;
  if (p < n)
    {
      r = order[p];
      k = rowptr[r] + lane;
      end = diag[r];
    }
  double acc = 0.0, a = 0.0;
  int c = 0;
  bool loaded = false;
  for (;;)
    {
      if (k < end)
        {
          if (!loaded)
            {
              c = col[k];
              a = lu[k];
              loaded = true;
            }
          const double v = ld_poll(y + c);
          if (!is_sentinel(v))
            {
              acc += a * v;
              k += G;
              loaded = false;
            }
        }
      if (!__any_sync(gmask, k < end))
        break;
    }
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1)
    acc += __shfl_down_sync(gmask, acc, o, G);
  if (lane == 0 && r >= 0)
    {
      dst[r] = sentinel();
      y[r] = canonical(src[r] - acc);
    }
}

// Backward substitution: x_r = (y_r - sum_{c>r} U_rc x_c) / u_rr, inverse pivot stored on the diagonal.
// Re-arms y for the next application.
template <int G>
__global__ void __launch_bounds__(kBlock)
k_ilu_backward(int32_t n, const int32_t *__restrict__ order, const int32_t *__restrict__ rowptr,
               const int32_t *__restrict__ diag, const int32_t *__restrict__ col, const double *__restrict__ lu, double *y,
               double *x, int32_t *tickets)
{
  constexpr int ROWS = kBlock / G;
  __shared__ int s_base;
  if (threadIdx.x == 0)
    {
      const int t = atomicAdd(&tickets[1], 1);
      if (t == 0)
        tickets[0] = 0;
      s_base = t * ROWS;
    }
  __syncthreads();
  const int lane = threadIdx.x % G;
  const int p = s_base + threadIdx.x / G;
  const unsigned gmask = group_mask<G>();
  int r = -1, k = 0, end = 0;
  if (p < n)
    {
      r = order[p];
      k = diag[r] + 1 + lane;
      end = rowptr[r + 1];
    }
  double acc = 0.0, a = 0.0;
  int c = 0;
  bool loaded = false;
  for (;;)
    {
      if (k < end)
        {
          if (!loaded)
            {
              c = col[k];
              a = lu[k];
              loaded = true;
            }
          const double v = ld_poll(x + c);
          if (!is_sentinel(v))
            {
              acc += a * v;
              k += G;
              loaded = false;
            }
        }
      if (!__any_sync(gmask, k < end))
        break;
    }
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1)
    acc += __shfl_down_sync(gmask, acc, o, G);
  if (lane == 0 && r >= 0)
    {
      const double yr = y[r];
      y[r] = sentinel();
      x[r] = canonical((yr - acc) * lu[diag[r]]);
    }
}

// lu = a, optionally with deal.II's diagonal strengthening:
// a_rr += strengthen * sum_{c != r} |a_rc|  (SparseLUDecomposition::strengthen_diagonal_impl)
__global__ void k_ilu_prepare(int32_t n, const int32_t *__restrict__ rowptr, const int32_t *__restrict__ diag,
                              const double *__restrict__ a, double *__restrict__ lu, double strengthen)
{
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n)
    return;
  const int beg = rowptr[r], end = rowptr[r + 1], d = diag[r];
  double rowsum = 0.0;
  for (int k = beg; k < end; ++k)
    {
      const double v = a[k];
      lu[k] = v;
      if (k != d)
        rowsum += fabs(v);
    }
  if (strengthen > 0.0)
    lu[d] = a[d] + strengthen * rowsum;
}

// IKJ ILU(0), one warp per row k:
//   for stored j < k ascending:  t = a_kj / u_jj;  a_kj = t;  a_kc -= t * u_jc for stored (j,c), c > j, (k,c) stored
//   a_kk <- 1 / a_kk
// Row j must be finished first; `ready[j]` says so.  Values of other rows are read through L2 (ld.cg)
// because an L1 line may still hold what was there before that row was finished.
__global__ void __launch_bounds__(kBlock)
k_ilu_factor(int32_t n, const int32_t *__restrict__ order, const int32_t *__restrict__ rowptr,
             const int32_t *__restrict__ diag, const int32_t *__restrict__ col, double *lu, int32_t *ready, int32_t *tickets)
{
  constexpr int ROWS = kBlock / 32;
  __shared__ int s_base;
  if (threadIdx.x == 0)
    s_base = atomicAdd(&tickets[2], 1) * ROWS;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int p = s_base + (threadIdx.x >> 5);
  if (p >= n)
    return;
  const int row = order[p];
  const int beg = rowptr[row], dpos = diag[row], end = rowptr[row + 1];
  for (int jpos = beg; jpos < dpos; ++jpos)
    {
      const int j = col[jpos];
      while (ld_poll_int(ready + j) == 0)
        ;
      __threadfence();
      const int jd = diag[j], jend = rowptr[j + 1];
      const double t = __ldcg(lu + jpos) * __ldcg(lu + jd);
      __syncwarp();
      if (lane == 0)
        lu[jpos] = t;
      for (int jj = jd + 1 + lane; jj < jend; jj += 32)
        {
          const int c = col[jj];
          // binary search for column c right of jpos in row `row`
          int lo = jpos + 1, hi = end;
          while (lo < hi)
            {
              const int mid = (lo + hi) >> 1;
              if (col[mid] < c)
                lo = mid + 1;
              else
                hi = mid;
            }
          if (lo < end && col[lo] == c)
            lu[lo] = __ldcg(lu + lo) - t * __ldcg(lu + jj);
        }
      __syncwarp();
    }
  if (lane == 0)
    {
      const double pivot = __ldcg(lu + dpos);
      if (pivot == 0.0)
        tickets[3] = 1;
      lu[dpos] = 1.0 / pivot;
      __threadfence();
      atomicExch(ready + row, 1);
    }
}

template <int G>
void launch_apply(DevIlu &p, double *dst, const double *src, cudaStream_t s)
{
  const DevMatrix &A = *p.pattern;
  constexpr int ROWS = kBlock / G;
  const int grid = (p.n + ROWS - 1) / ROWS;
  S35_LAUNCH((k_ilu_forward<G>), grid, kBlock, 0, s, p.n, p.order_fwd.ptr, A.rowptr.ptr, A.diag.ptr, A.col.ptr, p.lu.ptr, src,
             p.y.ptr, dst, p.tickets.ptr);
  S35_LAUNCH((k_ilu_backward<G>), grid, kBlock, 0, s, p.n, p.order_bwd.ptr, A.rowptr.ptr, A.diag.ptr, A.col.ptr, p.lu.ptr,
             p.y.ptr, dst, p.tickets.ptr);
}

} // namespace

void DevIlu::create(const DevMatrix &A, cudaStream_t s)
{
  S35_REQUIRE(A.n_rows == A.n_cols, "ILU: matrix must be square");
  pattern = &A;
  n = A.n_rows;
  const std::vector<int32_t> &rp = A.h_rowptr, &cl = A.h_col;
  std::vector<int32_t> level_f(n, 0), level_b(n, 0);
  levels_fwd = levels_bwd = n ? 1 : 0;
  int64_t lower = 0;
  for (int32_t r = 0; r < n; ++r)
    {
      bool has_diag = false;
      int32_t lv = 0;
      for (int32_t k = rp[r]; k < rp[r + 1]; ++k)
        {
          if (cl[k] < r)
            {
              lv = std::max(lv, level_f[cl[k]] + 1);
              ++lower;
            }
          has_diag = has_diag || cl[k] == r;
        }
      S35_REQUIRE(has_diag, "ILU: every row must store its diagonal entry");
      level_f[r] = lv;
      levels_fwd = std::max(levels_fwd, lv + 1);
    }
  for (int32_t r = n - 1; r >= 0; --r)
    {
      int32_t lv = 0;
      for (int32_t k = rp[r]; k < rp[r + 1]; ++k)
        if (cl[k] > r)
          lv = std::max(lv, level_b[cl[k]] + 1);
      level_b[r] = lv;
      levels_bwd = std::max(levels_bwd, lv + 1);
    }
  std::vector<int32_t> of(n), ob(n);
  const char *order_env = std::getenv("S35_ILU_ORDER");
  if (order_env && std::strcmp(order_env, "natural") == 0)
    for (int32_t i = 0; i < n; ++i)
      {
        of[i] = i;
        ob[i] = n - 1 - i;
      }
  else
    {
      // counting sort by level: rows of one level are adjacent, ascending row index inside a level
      auto sort_by_level = [&](const std::vector<int32_t> &level, int32_t n_levels, std::vector<int32_t> &out, bool descending) {
        std::vector<int32_t> start(n_levels + 1, 0);
        for (int32_t r = 0; r < n; ++r)
          start[level[r] + 1]++;
        for (int32_t l = 0; l < n_levels; ++l)
          start[l + 1] += start[l];
        if (!descending)
          for (int32_t r = 0; r < n; ++r)
            out[start[level[r]]++] = r;
        else
          for (int32_t r = n - 1; r >= 0; --r)
            out[start[level[r]]++] = r;
      };
      sort_by_level(level_f, levels_fwd, of, false);
      sort_by_level(level_b, levels_bwd, ob, true);
    }
  order_fwd.upload(of, s);
  order_bwd.upload(ob, s);
  lu.alloc(A.nnz);
  y.alloc(n);
  if (n)
    S35_CUDA(cudaMemsetAsync(y.ptr, 0xFF, (size_t)n * sizeof(double), s));
  ready.alloc(n);
  tickets.alloc(4);
  tickets.zero(s);
  const double mean_lower = n ? (double)lower / n : 1.0;
  group = 4;
  while (group < 32 && group * 4 <= mean_lower)
    group *= 2;
  if (const char *g = std::getenv("S35_ILU_GROUP"))
    {
      const int v = std::atoi(g);
      if (v == 4 || v == 8 || v == 16 || v == 32)
        group = v;
    }
  factored = false;
}

void DevIlu::factor(const double *values, double strengthen_diagonal, cudaStream_t s)
{
  S35_REQUIRE(pattern != nullptr, "ILU: not created");
  if (n == 0)
    return;
  const DevMatrix &A = *pattern;
  S35_LAUNCH(k_ilu_prepare, (n + 127) / 128, 128, 0, s, n, A.rowptr.ptr, A.diag.ptr, values, lu.ptr, strengthen_diagonal);
  ready.zero(s);
  S35_CUDA(cudaMemsetAsync(tickets.ptr + 2, 0, 2 * sizeof(int32_t), s));
  constexpr int ROWS = kBlock / 32;
  S35_LAUNCH(k_ilu_factor, (n + ROWS - 1) / ROWS, kBlock, 0, s, n, order_fwd.ptr, A.rowptr.ptr, A.diag.ptr, A.col.ptr, lu.ptr,
             ready.ptr, tickets.ptr);
  factored = true;
}

void DevIlu::check_pivots(cudaStream_t s)
{
  int32_t flag = 0;
  S35_CUDA(cudaMemcpyAsync(&flag, tickets.ptr + 3, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  S35_CUDA(cudaStreamSynchronize(s));
  if (flag)
    throw Error(S35_ERR_ZERO_PIVOT, "SparseILU: zero pivot met during the factorisation");
}

void DevIlu::apply(double *dst, const double *src, cudaStream_t s)
{
  S35_REQUIRE(factored, "ILU: apply before factor");
  S35_REQUIRE(dst != src, "ILU: dst must not alias src");
  if (n == 0)
    return;
  switch (group)
    {
    case 4: launch_apply<4>(*this, dst, src, s); break;
    case 8: launch_apply<8>(*this, dst, src, s); break;
    case 16: launch_apply<16>(*this, dst, src, s); break;
    default: launch_apply<32>(*this, dst, src, s); break;
    }
}

} // namespace s35
