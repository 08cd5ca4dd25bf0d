#include "ilu.cuh"

#include "wavefront.h"

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

// ---------------------------------------------------------------------------------------------- wavefront mode
// One block per sweep: 256 consumer threads (thread i owns row i of the current tile) and one producer warp that
// keeps kWaveStages tiles of the packed factor (wavefront.h) in flight with bulk copies.  Per tile a consumer reads
// its row's entries and their columns' solution values from shared memory (or, for the few "far" ones, from the
// position-ordered solution in global memory), writes its result into the window and to global memory, and meets
// the others at a named barrier: a level costs shared-memory latency and a barrier instead of a trip through L2.
constexpr int kWaveConsumers = 256; // = WavefrontOptions::tile_rows
constexpr int kWaveThreads = kWaveConsumers + 32;
constexpr int kWaveStages = 4;

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
  asm volatile("{\n"
               ".reg .pred p;\n"
               "WAVE_WAIT:\n"
               "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
               "@p bra WAVE_DONE;\n"
               "bra WAVE_WAIT;\n"
               "WAVE_DONE:\n"
               "}" ::"r"(smem_addr(bar)),
               "r"(parity)
               : "memory");
}
// one attempt; its latency overlaps with whatever the thread does until it looks at the answer
__device__ __forceinline__ uint32_t mbar_try(uint64_t *bar, uint32_t parity)
{
  uint32_t done;
  asm volatile("{\n"
               ".reg .pred p;\n"
               "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
               "selp.u32 %0, 1, 0, p;\n"
               "}"
               : "=r"(done)
               : "r"(smem_addr(bar)), "r"(parity)
               : "memory");
  return done;
}
// global -> shared bulk copy (16-byte aligned, a multiple of 16 bytes) that reports to an mbarrier
__device__ __forceinline__ void bulk_copy_to_shared(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(dst)), "l"(src),
               "r"(bytes), "r"(smem_addr(bar))
               : "memory");
}
__device__ __forceinline__ double ld_l2(const double *p)
{
  double v;
  asm volatile("ld.global.cg.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}

struct WaveArgs
{
  const int32_t *meta, *directory;
  const double *values;
  int32_t n_tiles, window, stage_meta, stage_vals;
  const double *rhs;   // right-hand sides, indexed through the plan
  double *by_position; // the sweep's result in position order (window overflow and the next sweep's input)
  double *natural;     // the sweep's result by row index, or nullptr
};

template <bool FORWARD>
__global__ void __launch_bounds__(kWaveThreads, 1) k_wave_sweep(const WaveArgs a)
{
  extern __shared__ __align__(16) unsigned char wave_smem[];
  double *const window = reinterpret_cast<double *>(wave_smem);
  double *const stage_vals = window + a.window;
  int32_t *const stage_meta = reinterpret_cast<int32_t *>(stage_vals + (size_t)kWaveStages * a.stage_vals);
  uint64_t *const full = reinterpret_cast<uint64_t *>(stage_meta + (size_t)kWaveStages * a.stage_meta);
  uint64_t *const empty = full + kWaveStages;
  const int tid = threadIdx.x;
  if (tid == 0)
    {
      for (int s = 0; s < kWaveStages; ++s)
        {
          mbar_init(full + s, 1);
          mbar_init(empty + s, 1);
        }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
  __syncthreads();

  if (tid >= kWaveConsumers)
    {
      // producer warp: lane 0 keeps the stages filled.  The tile directory is read 32 tiles at a time by the whole
      // warp, the next batch while the current one is being issued, so that the producer never waits for it.
      const int lane = tid - kWaveConsumers;
      const int4 *const directory = reinterpret_cast<const int4 *>(a.directory);
      int4 next = lane < a.n_tiles ? directory[lane] : make_int4(0, 0, 0, 0);
      for (int base = 0; base < a.n_tiles; base += 32)
        {
          const int4 batch = next;
          const int ahead = base + 32 + lane;
          next = ahead < a.n_tiles ? directory[ahead] : make_int4(0, 0, 0, 0);
          const int count = min(32, a.n_tiles - base);
          for (int u = 0; u < count; ++u)
            {
              const int32_t meta_off = __shfl_sync(0xffffffffu, batch.x, u), meta_len = __shfl_sync(0xffffffffu, batch.y, u);
              const int32_t val_off = __shfl_sync(0xffffffffu, batch.z, u), val_len = __shfl_sync(0xffffffffu, batch.w, u);
              if (lane == 0)
                {
                  const int t = base + u, s = t % kWaveStages;
                  if (t >= kWaveStages)
                    mbar_wait(empty + s, (uint32_t)((t / kWaveStages - 1) & 1));
                  mbar_expect_tx(full + s, (uint32_t)meta_len * 4u + (uint32_t)val_len * 8u);
                  bulk_copy_to_shared(stage_meta + (size_t)s * a.stage_meta, a.meta + meta_off, (uint32_t)meta_len * 4u, full + s);
                  if (val_len)
                    bulk_copy_to_shared(stage_vals + (size_t)s * a.stage_vals, a.values + val_off, (uint32_t)val_len * 8u, full + s);
                }
              __syncwarp();
            }
        }
      return;
    }

  // consumers
  double rhs_next = 0.0;
  if (a.n_tiles > 0)
    {
      mbar_wait(full + 0, 0u);
      const WaveTile first = wave_tile(stage_meta, stage_vals);
      if (tid < first.rows)
        rhs_next = ld_l2(a.rhs + first.rhs[tid]);
    }
  for (int t = 0; t < a.n_tiles; ++t)
    {
      const int s = t % kWaveStages;
      // the stage's arrival was awaited when its right-hand sides were fetched
      const WaveTile tile = wave_tile(stage_meta + (size_t)s * a.stage_meta, stage_vals + (size_t)s * a.stage_vals);
      const double rhs = rhs_next;
      // the next tile's stage has normally been complete for a long time: ask now, look at the answer after the row
      const int s1 = (t + 1) % kWaveStages;
      const uint32_t parity1 = (uint32_t)(((t + 1) / kWaveStages) & 1);
      const uint32_t next_landed = t + 1 < a.n_tiles ? mbar_try(full + s1, parity1) : 1u;
      if (tid < tile.rows)
        {
          const double result =
            wave_row(tile, tid, rhs, window, WaveFarLoad{a.by_position}, FORWARD);
          const int32_t pos = tile.first_pos + tid;
          window[pos & (a.window - 1)] = result;
          a.by_position[pos] = result;
          if (a.natural)
            a.natural[tile.row[tid]] = result;
        }
      if (t + 1 < a.n_tiles)
        {
          // the next tile's right-hand sides travel while this tile finishes
          if (!next_landed)
            mbar_wait(full + s1, parity1);
          const WaveTile next = wave_tile(stage_meta + (size_t)s1 * a.stage_meta, stage_vals + (size_t)s1 * a.stage_vals);
          if (tid < next.rows)
            rhs_next = ld_l2(a.rhs + next.rhs[tid]);
        }
      asm volatile("bar.sync 1, %0;" ::"n"(kWaveConsumers) : "memory"); // window complete for the next tile; stage s read
      if (tid == 0)
        mbar_arrive(empty + s);
    }
}

__global__ void k_wave_gather_values(int64_t n, const int32_t *__restrict__ source, const double *__restrict__ lu,
                                     double *__restrict__ values)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    values[i] = source[i] >= 0 ? lu[source[i]] : 0.0;
}

void wave_launch(DevIlu &p, double *dst, const double *src, cudaStream_t s)
{
  DevIlu::Wave &w = p.wave;
  WaveArgs fwd{w.meta[0].ptr, w.directory[0].ptr, w.values[0].ptr, w.n_tiles[0], w.window, w.stage_meta, w.stage_vals,
               src,           w.by_position[0].ptr, nullptr};
  WaveArgs bwd{w.meta[1].ptr, w.directory[1].ptr, w.values[1].ptr, w.n_tiles[1], w.window, w.stage_meta, w.stage_vals,
               w.by_position[0].ptr, w.by_position[1].ptr, dst};
  S35_LAUNCH((k_wave_sweep<true>), 1, kWaveThreads, w.smem_bytes, s, fwd);
  S35_LAUNCH((k_wave_sweep<false>), 1, kWaveThreads, w.smem_bytes, s, bwd);
}

// packs both sweeps of the pattern; returns false if the mode does not suit this factor
bool wave_create(DevIlu &p, const DevMatrix &A, cudaStream_t s)
{
  const int32_t n = A.n_rows;
  if (n == 0 || (double)A.nnz / n > 40.0)
    return false; // long rows: the wide sweep streams them with many SMs, one SM cannot
  std::vector<int32_t> diag(n);
  for (int32_t r = 0; r < n; ++r)
    for (int32_t k = A.h_rowptr[r]; k < A.h_rowptr[r + 1]; ++k)
      if (A.h_col[k] == r)
        diag[r] = k;
  WavefrontOptions opt;
  opt.tile_rows = kWaveConsumers;
  opt.window = 8192;
  if (const char *t = std::getenv("S35_WAVE_WINDOW"))
    opt.window = std::atoi(t);
  // Tiles as large as the shared memory allows: on the r1 pressure factor 3072 entries per tile give one tile per
  // level (1 758 tiles for 1 755 levels), 1536 entries split a quarter of the levels (2 402 tiles).
  std::vector<int> candidates = {3072, 1536, 768};
  if (const char *t = std::getenv("S35_WAVE_TILE_ENTRIES"))
    candidates.assign(1, std::atoi(t));
  DevIlu::Wave &w = p.wave;
  WavefrontPlan F, B;
  bool fits = false;
  for (int tile_entries : candidates)
    {
      opt.tile_entries = tile_entries;
      F = wavefront_pack(n, A.h_rowptr.data(), A.h_col.data(), diag.data(), true, nullptr, opt);
      // the backward sweep takes its right-hand sides from the forward sweep's position-ordered result
      B = wavefront_pack(n, A.h_rowptr.data(), A.h_col.data(), diag.data(), false, F.position.data(), opt);
      w.window = opt.window;
      w.stages = kWaveStages;
      w.stage_meta = round_up(std::max(F.max_meta, B.max_meta), 4);
      w.stage_vals = round_up(std::max(std::max(F.max_values, B.max_values), 2), 2);
      w.smem_bytes = (size_t)w.window * 8 + (size_t)kWaveStages * ((size_t)w.stage_vals * 8 + (size_t)w.stage_meta * 4) + 2 * kWaveStages * 8;
      if (w.smem_bytes <= 227 * 1024)
        {
          fits = true;
          break;
        }
    }
  if (!fits)
    return false;
  const WavefrontPlan *plans[2] = {&F, &B};
  for (int d = 0; d < 2; ++d)
    {
      const WavefrontPlan &P = *plans[d];
      w.n_tiles[d] = P.n_tiles;
      w.n_values[d] = P.n_values;
      w.far_fraction[d] = P.entries ? (double)P.far_entries / (double)P.entries : 0.0;
      w.meta[d].upload(P.meta, s);
      w.directory[d].upload(P.directory, s);
      std::vector<int32_t> source(P.value_source.begin(), P.value_source.end()); // factor positions fit 32 bits (nnz is int32-indexed)
      w.value_source[d].upload(source, s);
      w.values[d].alloc((size_t)std::max<int64_t>(P.n_values, 2));
      w.by_position[d].alloc(n);
    }
  S35_CUDA(cudaFuncSetAttribute(k_wave_sweep<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)w.smem_bytes));
  S35_CUDA(cudaFuncSetAttribute(k_wave_sweep<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)w.smem_bytes));
  w.on = true;
  return true;
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
  wave = Wave();
  if (const char *mode = std::getenv("S35_ILU_MODE"))
    if (std::strcmp(mode, "wavefront") == 0)
      wave_create(*this, A, s);
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
  if (wave.on)
    for (int d = 0; d < 2; ++d)
      if (wave.n_values[d])
        {
          const int grid = (int)std::min<int64_t>((wave.n_values[d] + 255) / 256, (int64_t)sm_count() * 8);
          S35_LAUNCH(k_wave_gather_values, grid, 256, 0, s, wave.n_values[d], wave.value_source[d].ptr, lu.ptr, wave.values[d].ptr);
        }
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
  if (wave.on)
    {
      wave_launch(*this, dst, src, s);
      return;
    }
  switch (per_lane)
    {
    case 1: launch_apply<1>(*this, dst, src, s); break;
    case 2: launch_apply<2>(*this, dst, src, s); break;
    default: launch_apply<4>(*this, dst, src, s); break;
    }
}

} // namespace s35
