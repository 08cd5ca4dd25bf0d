#include "sparse.cuh"

#include <algorithm>
#include <cstdlib>

namespace s35 {

namespace {

constexpr int kBlock = 256;

// LANES threads share one row; a warp therefore reads 32/LANES rows at a time,
// each as a contiguous run of (col, val) pairs.  x is gathered through the
// read-only path.
template <int LANES, int NV>
__global__ void __launch_bounds__(kBlock)
k_spmv(int32_t n_rows, const int32_t *__restrict__ rowptr, const int32_t *__restrict__ col, const double *__restrict__ v0,
       const double *__restrict__ v1, const double *__restrict__ v2, const double *__restrict__ x0,
       const double *__restrict__ x1, const double *__restrict__ x2, double *__restrict__ y, int add)
{
  const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t row = gtid / LANES;
  const int lane = (int)(gtid % LANES);
  double sum = 0.0;
  if (row < n_rows)
    {
      const int32_t beg = rowptr[row], end = rowptr[row + 1];
      for (int32_t k = beg + lane; k < end; k += LANES)
        {
          const int32_t c = col[k];
          sum += v0[k] * __ldg(x0 + c);
          if (NV > 1)
            sum += v1[k] * __ldg(x1 + c);
          if (NV > 2)
            sum += v2[k] * __ldg(x2 + c);
        }
    }
#pragma unroll
  for (int o = LANES / 2; o > 0; o >>= 1)
    sum += __shfl_down_sync(0xffffffffu, sum, o, LANES);
  if (row < n_rows && lane == 0)
    y[row] = add ? y[row] + sum : sum;
}

// Warp-per-row variant with R consecutive rows of a warp in flight together: the R + 1 row boundaries come from
// one load, then every trip issues the (col, val) loads of all R rows before the first gather of x is needed.
// The row-per-warp kernel above is bound by its chain of dependent loads (row pointer -> col/val -> x), not by
// bandwidth; this keeps R chains going per warp.  Per row the arithmetic is the same as in k_spmv<32, 1> -- lane
// j adds its entries beg + j, beg + j + 32, ... in that order, then the same shuffle tree -- so the results are
// bit-identical to it.
template <int R>
__global__ void __launch_bounds__(kBlock)
k_spmv_rows(int32_t n_rows, const int32_t *__restrict__ rowptr, const int32_t *__restrict__ col, const double *__restrict__ v0,
            const double *__restrict__ x0, double *__restrict__ y, int add)
{
  const int lane = threadIdx.x & 31;
  const int64_t row0 = ((((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5)) * R;
  if (row0 >= n_rows)
    return;
  const int32_t bound = rowptr[min(row0 + lane, (int64_t)n_rows)]; // lanes 0..R are used; rows past the end come out empty
  int32_t k[R], end[R];
  double sum[R];
  int32_t longest = 0;
#pragma unroll
  for (int i = 0; i < R; ++i)
    {
      const int32_t beg = __shfl_sync(0xffffffffu, bound, i);
      end[i] = __shfl_sync(0xffffffffu, bound, i + 1);
      longest = max(longest, end[i] - beg);
      k[i] = beg + lane;
      sum[i] = 0.0;
    }
  for (int32_t done = 0; done < longest; done += 32)
    {
      int32_t c[R];
      double a[R];
#pragma unroll
      for (int i = 0; i < R; ++i)
        {
          const bool in = k[i] < end[i];
          c[i] = in ? col[k[i]] : -1;
          a[i] = in ? v0[k[i]] : 0.0;
        }
#pragma unroll
      for (int i = 0; i < R; ++i)
        {
          if (c[i] >= 0)
            sum[i] += a[i] * __ldg(x0 + c[i]);
          k[i] += 32;
        }
    }
  double mine = 0.0;
#pragma unroll
  for (int i = 0; i < R; ++i)
    {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1)
        sum[i] += __shfl_down_sync(0xffffffffu, sum[i], o, 32);
      const double total = __shfl_sync(0xffffffffu, sum[i], 0);
      if (lane == i)
        mine = total;
    }
  if (lane < R && row0 + lane < n_rows)
    y[row0 + lane] = add ? y[row0 + lane] + mine : mine;
}

// rows in flight per warp for matrices whose rows take a whole warp (0: the plain row-per-warp kernel; S35_SPMV_ROWS
// overrides, tools/tune_spmv.py compares)
int rows_in_flight()
{
  if (const char *t = std::getenv("S35_SPMV_ROWS"))
    return std::atoi(t);
  // measured on B200 (velocity matrix, r1, L2 flushed): 0.350 ms plain, 0.303 / 0.283 / 0.311 ms with 2 / 4 / 8 rows in
  // flight.  A CSR-stream kernel (products staged in shared memory, block nonzero ranges cut on the host) was tried in
  // round 2 and reached 0.302 ms on this matrix and lost on the others: not kept.
  return 4;
}

template <int R>
void launch_spmv_rows(int32_t n_rows, const int32_t *rowptr, const int32_t *col, const double *v, const double *x, double *y,
                      bool add, cudaStream_t s)
{
  const int64_t warps = ((int64_t)n_rows + R - 1) / R;
  const int grid = (int)((warps * 32 + kBlock - 1) / kBlock);
  S35_LAUNCH((k_spmv_rows<R>), grid, kBlock, 0, s, n_rows, rowptr, col, v, x, y, (int)add);
}

template <int NV>
void launch_spmv(int lanes, int32_t n_rows, const int32_t *rowptr, const int32_t *col, const double *const *v,
                 const double *const *x, double *y, bool add, cudaStream_t s)
{
  if (n_rows == 0)
    return;
  if (NV == 1 && lanes == 32)
    switch (rows_in_flight())
      {
      case 2: launch_spmv_rows<2>(n_rows, rowptr, col, v[0], x[0], y, add, s); return;
      case 4: launch_spmv_rows<4>(n_rows, rowptr, col, v[0], x[0], y, add, s); return;
      case 8: launch_spmv_rows<8>(n_rows, rowptr, col, v[0], x[0], y, add, s); return;
      default: break;
      }
  const int64_t threads = (int64_t)n_rows * lanes;
  const int grid = (int)((threads + kBlock - 1) / kBlock);
  const double *v1 = NV > 1 ? v[1] : nullptr, *v2 = NV > 2 ? v[2] : nullptr;
  const double *x1 = NV > 1 ? x[1] : nullptr, *x2 = NV > 2 ? x[2] : nullptr;
  switch (lanes)
    {
    case 2: S35_LAUNCH((k_spmv<2, NV>), grid, kBlock, 0, s, n_rows, rowptr, col, v[0], v1, v2, x[0], x1, x2, y, (int)add); break;
    case 4: S35_LAUNCH((k_spmv<4, NV>), grid, kBlock, 0, s, n_rows, rowptr, col, v[0], v1, v2, x[0], x1, x2, y, (int)add); break;
    case 8: S35_LAUNCH((k_spmv<8, NV>), grid, kBlock, 0, s, n_rows, rowptr, col, v[0], v1, v2, x[0], x1, x2, y, (int)add); break;
    case 16: S35_LAUNCH((k_spmv<16, NV>), grid, kBlock, 0, s, n_rows, rowptr, col, v[0], v1, v2, x[0], x1, x2, y, (int)add); break;
    default: S35_LAUNCH((k_spmv<32, NV>), grid, kBlock, 0, s, n_rows, rowptr, col, v[0], v1, v2, x[0], x1, x2, y, (int)add); break;
    }
}

__global__ void k_gather(double *__restrict__ dst, const double *__restrict__ src, const int32_t *__restrict__ perm, int64_t n)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    {
      const int32_t from = perm[i]; // < 0: an entry the source pattern does not have (widened ILU pattern)
      dst[i] = from >= 0 ? src[from] : 0.0;
    }
}

int pick_lanes(int64_t nnz, int32_t n_rows)
{
  const double mean = n_rows ? (double)nnz / n_rows : 1.0;
  int lanes = 2;
  while (lanes < 32 && lanes * 2 <= mean)
    lanes *= 2;
  return lanes;
}

} // namespace

void DevMatrix::create(int32_t nr, int32_t nc, const int32_t *h_rp, const int32_t *h_c, const double *h_val, cudaStream_t s)
{
  S35_REQUIRE(nr >= 0 && nc >= 0 && h_rp != nullptr, "matrix: bad sizes or null rowptr");
  S35_REQUIRE(h_rp[0] == 0, "matrix: rowptr[0] must be 0");
  n_rows = nr;
  n_cols = nc;
  nnz = h_rp[nr];
  S35_REQUIRE(nnz == 0 || h_c != nullptr, "matrix: null column array");
  h_rowptr.assign(h_rp, h_rp + nr + 1);
  h_col.assign(h_c, h_c + nnz);
  std::vector<int32_t> h_diag;
  if (nr <= nc) // square, or owned rows x (owned + ghost) columns
    h_diag.assign(nr, -1);
  for (int32_t r = 0; r < nr; ++r)
    {
      S35_REQUIRE(h_rp[r + 1] >= h_rp[r], "matrix: rowptr not monotone");
      for (int32_t k = h_rp[r]; k < h_rp[r + 1]; ++k)
        {
          S35_REQUIRE(h_c[k] >= 0 && h_c[k] < nc, "matrix: column index out of range");
          S35_REQUIRE(k == h_rp[r] || h_c[k] > h_c[k - 1], "matrix: columns must be strictly ascending within a row");
          if (nr <= nc && h_c[k] == r)
            h_diag[r] = k;
        }
    }
  rowptr.upload(h_rowptr, s);
  col.upload(h_col, s);
  val.alloc(nnz);
  if (h_val)
    val.upload(h_val, nnz, s);
  else
    val.zero(s);
  if (nr <= nc)
    diag.upload(h_diag, s);
  lanes = pick_lanes(nnz, nr);
  transposed.reset();
}

void DevMatrix::set_values(const double *host_val, cudaStream_t s)
{
  val.upload(host_val, nnz, s);
  values_changed();
}

const DevMatrix &DevMatrix::transpose(cudaStream_t s) const
{
  if (!transposed)
    {
      transposed.reset(new Transposed);
      std::vector<int32_t> t_rowptr(n_cols + 1, 0), t_col(nnz), perm(nnz);
      for (int64_t k = 0; k < nnz; ++k)
        t_rowptr[h_col[k] + 1]++;
      for (int32_t c = 0; c < n_cols; ++c)
        t_rowptr[c + 1] += t_rowptr[c];
      std::vector<int32_t> fill(t_rowptr.begin(), t_rowptr.end() - 1);
      for (int32_t r = 0; r < n_rows; ++r)
        for (int32_t k = h_rowptr[r]; k < h_rowptr[r + 1]; ++k)
          {
            const int32_t p = fill[h_col[k]]++;
            t_col[p] = r; // rows visited in ascending order: columns of T come out sorted
            perm[p] = k;
          }
      transposed->T.reset(new DevMatrix);
      transposed->T->create(n_cols, n_rows, t_rowptr.data(), t_col.data(), nullptr, s);
      transposed->perm.upload(perm, s);
    }
  if (!transposed->values_current)
    {
      gather_values(transposed->T->val, val, transposed->perm, nnz, s);
      transposed->values_current = true;
    }
  return *transposed->T;
}

void spmv_values(const DevMatrix &P, const double *val, const double *x, double *y, bool add, cudaStream_t s)
{
  const double *v[1] = {val}, *xs[1] = {x};
  launch_spmv<1>(P.lanes, P.n_rows, P.rowptr, P.col, v, xs, y, add, s);
}

void spmv(const DevMatrix &A, const double *x, double *y, bool add, cudaStream_t s) { spmv_values(A, A.val, x, y, add, s); }

void spmv_multi(const DevMatrix &P, int m, const double *const *val, const double *const *x, double *y, bool add, cudaStream_t s)
{
  S35_REQUIRE(m >= 1 && m <= 3, "spmv_multi: 1..3 operands");
  const int lanes = std::min(32, P.lanes * (m > 1 ? 2 : 1));
  if (m == 1)
    launch_spmv<1>(lanes, P.n_rows, P.rowptr, P.col, val, x, y, add, s);
  else if (m == 2)
    launch_spmv<2>(lanes, P.n_rows, P.rowptr, P.col, val, x, y, add, s);
  else
    launch_spmv<3>(lanes, P.n_rows, P.rowptr, P.col, val, x, y, add, s);
}

void gather_values(double *dst, const double *src, const int32_t *perm, int64_t n, cudaStream_t s)
{
  if (n == 0)
    return;
  const int grid = (int)std::min<int64_t>((n + kBlock * 4 - 1) / (kBlock * 4), (int64_t)sm_count() * 8);
  S35_LAUNCH(k_gather, std::max(grid, 1), kBlock, 0, s, dst, src, perm, n);
}

} // namespace s35
