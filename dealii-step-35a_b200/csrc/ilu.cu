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

// One substitution sweep, forward (unit lower factor: y_r = src_r - sum_{c<r} L_rc y_c) or backward
// (x_r = (y_r - sum_{c>r} U_rc x_c) / u_rr, inverse pivot stored on the diagonal).
//
// One warp per row.  A lane first loads its (column, value) pairs of the row into registers -- P per lane
// and chunk, all loads in flight together -- and only then polls the solution entries it needs, all pending
// ones together and only the pending ones, so that rows far ahead of the wave front put one request per
// unresolved dependency and polling round on L2 (an optional growing pause between rounds exists for
// experiments; it only added latency on B200 and is off).  The wait for the
// dependency that is produced last costs one round trip through L2 plus the warp reduction.  Every lane
// adds its products in index order and the warp reduction is a fixed tree: the result does not depend on
// timing.
// Blocks are persistent and take groups of rows from a ticket counter in topological order (dependency
// level, then row), which makes the spin-waiting deadlock free: a row only waits for rows of earlier
// tickets, and those are resident or done.  Rows of one warp never wait for each other (one row per warp).
// The forward sweep also arms x (the backward sweep's output) with the sentinel; the backward sweep
// re-arms y for the next application.
template <int P, bool FORWARD>
__global__ void __launch_bounds__(kBlock)
k_ilu_sweep(int32_t n, const int32_t *__restrict__ order, const int32_t *__restrict__ rowptr,
            const int32_t *__restrict__ diag, const int32_t *__restrict__ col, const double *__restrict__ lu,
            const double *__restrict__ src, double *y, double *x, int32_t *tickets, unsigned backoff_cap_ns)
{
  constexpr int ROWS = kBlock / 32;
  __shared__ int s_base;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int my_ticket = FORWARD ? 0 : 1;
  const double *wait_on = FORWARD ? y : x;
  for (;;)
    {
      __syncthreads();
      if (threadIdx.x == 0)
        {
          const int t = atomicAdd(&tickets[my_ticket], 1);
          if (t == 0)
            tickets[1 - my_ticket] = 0; // the other sweep is not running: reset its ticket
          s_base = t * ROWS;
        }
      __syncthreads();
      const int base = s_base;
      if (base >= n)
        break;
      const int p = base + warp;
      if (p >= n)
        continue;
      const int r = order[p];
      const int beg = FORWARD ? rowptr[r] : diag[r] + 1;
      const int end = FORWARD ? diag[r] : rowptr[r + 1];
      // everything the last step of the row needs is fetched before the waiting starts
      const double rhs = FORWARD ? src[r] : y[r];
      const double inv_pivot = FORWARD ? 1.0 : lu[diag[r]];
      double acc = 0.0;
      for (int chunk = beg; chunk < end; chunk += 32 * P)
        {
          int c[P];
          double a[P], v[P];
          unsigned pending = 0;
#pragma unroll
          for (int i = 0; i < P; ++i)
            {
              const int k = chunk + i * 32 + lane;
              const bool valid = k < end;
              c[i] = valid ? col[k] : 0;
              a[i] = valid ? lu[k] : 0.0;
              v[i] = 0.0;
              pending |= valid ? (1u << i) : 0u;
            }
          unsigned pause = 0;
          for (;;)
            {
#pragma unroll
              for (int i = 0; i < P; ++i)
                if ((pending >> i) & 1u)
                  v[i] = ld_poll(wait_on + c[i]);
#pragma unroll
              for (int i = 0; i < P; ++i)
                if (((pending >> i) & 1u) && !is_sentinel(v[i]))
                  pending &= ~(1u << i);
              if (!__any_sync(0xffffffffu, pending != 0u))
                break;
              if (pause)
                __nanosleep(pause);
              pause = min(pause + 16u, backoff_cap_ns);
            }
#pragma unroll
          for (int i = 0; i < P; ++i)
            acc += a[i] * v[i];
        }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1)
        acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (lane == 0)
        {
          if (FORWARD)
            {
              x[r] = sentinel();
              y[r] = canonical(rhs - acc);
            }
          else
            {
              y[r] = sentinel();
              x[r] = canonical((rhs - acc) * inv_pivot);
            }
        }
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

template <int P>
void launch_apply(DevIlu &p, double *dst, const double *src, cudaStream_t s)
{
  const DevMatrix &A = *p.pattern;
  constexpr int ROWS = kBlock / 32;
  const int chunks = (p.n + ROWS - 1) / ROWS;
  const int grid = std::min(chunks, sm_count() * p.blocks_per_sm);
  S35_LAUNCH((k_ilu_sweep<P, true>), grid, kBlock, 0, s, p.n, p.order_fwd.ptr, A.rowptr.ptr, A.diag.ptr, A.col.ptr, p.lu.ptr,
             src, p.y.ptr, dst, p.tickets.ptr, (unsigned)p.backoff_ns);
  S35_LAUNCH((k_ilu_sweep<P, false>), grid, kBlock, 0, s, p.n, p.order_bwd.ptr, A.rowptr.ptr, A.diag.ptr, A.col.ptr, p.lu.ptr,
             src, p.y.ptr, dst, p.tickets.ptr, (unsigned)p.backoff_ns);
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
  for (int32_t r = 0; r < n; ++r)
    {
      bool has_diag = false;
      int32_t lv = 0;
      for (int32_t k = rp[r]; k < rp[r + 1]; ++k)
        {
          if (cl[k] < r)
            lv = std::max(lv, level_f[cl[k]] + 1);
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
  // register slots per lane: one chunk of 32 * per_lane entries should hold a typical row half
  int32_t max_half = 0;
  for (int32_t r = 0; r < n; ++r)
    {
      int32_t lo = 0;
      for (int32_t k = rp[r]; k < rp[r + 1]; ++k)
        lo += cl[k] < r;
      max_half = std::max(max_half, std::max(lo, rp[r + 1] - rp[r] - lo - 1));
    }
  per_lane = max_half <= 128 ? 2 : 4; // measured on B200: 2 beats 1 and 4 for Q1 (<= 13) and Q2 (<= ~100) rows
  if (const char *g = std::getenv("S35_ILU_PER_LANE"))
    {
      const int v = std::atoi(g);
      if (v == 1 || v == 2 || v == 4)
        per_lane = v;
    }
  blocks_per_sm = 4;
  if (const char *g = std::getenv("S35_ILU_BLOCKS_PER_SM"))
    blocks_per_sm = std::max(1, std::min(16, std::atoi(g)));
  backoff_ns = 0; // measured on B200: pausing between polling rounds only adds latency
  if (const char *g = std::getenv("S35_ILU_BACKOFF_NS"))
    backoff_ns = std::max(0, std::min(10000, std::atoi(g)));
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
  switch (per_lane)
    {
    case 1: launch_apply<1>(*this, dst, src, s); break;
    case 2: launch_apply<2>(*this, dst, src, s); break;
    default: launch_apply<4>(*this, dst, src, s); break;
    }
}

} // namespace s35
