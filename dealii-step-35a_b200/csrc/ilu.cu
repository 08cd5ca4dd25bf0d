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
// One block per sweep (wavefront.h has the schedule and the stream format).  `groups` warp groups of `group_rows`
// threads take the tiles in turn; one more warp keeps the shared-memory ring filled with bulk copies.  For tile t,
// thread i of group t % groups
//   prepares: waits for the tile's bytes, sums the row's OLD entries (values of tiles <= t - groups - 1, complete
//             since the group's previous tile), fetches the right-hand side and the first fresh entries;
//   finishes: waits on the named barrier of tile t - 1, adds the FRESH entries from the window, stores the result
//             into the window, arrives on the barrier of tile t (without waiting), then writes global memory.
// The hand-off of a level is therefore: barrier release -> shared-memory load -> a few dependent FMAs -> store ->
// arrive; everything else overlaps with the other groups' tiles.
constexpr int kWaveBarriers = 8;    // named barriers 1..8, tile t uses 1 + t % 8; a group may wait on tile t + groups - 1's
                                    // barrier while tile t's is still open, so groups < kWaveBarriers
constexpr int kWaveMaxThreads = 800;  // 80 registers per thread
#ifdef S35_WAVE_TRACING
constexpr bool kWaveTracing = true; // clocks of every tile into WaveArgs::trace (a build for tools/, see Makefile)
#else
constexpr bool kWaveTracing = false;
#endif
constexpr int kWaveSlots = 32;      // = WavefrontOptions::slots: tiles in flight, one pair of mbarriers each

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
__device__ __forceinline__ void group_wait(int tile, int threads)
{
  asm volatile("bar.sync %0, %1;" ::"r"(1 + tile % kWaveBarriers), "r"(threads) : "memory");
}
__device__ __forceinline__ void group_signal(int tile, int threads)
{
  asm volatile("bar.arrive %0, %1;" ::"r"(1 + tile % kWaveBarriers), "r"(threads) : "memory");
}

struct WaveArgs
{
  const unsigned char *blob;
  const int4 *directory;  // per tile: blob offset / 16, ring offset / 16, bytes, tile that must be finished before the load
  const int2 *rhs_dir;    // per tile: offset (doubles) and size (bytes) of its right-hand sides in rhs layout
  int32_t n_tiles, groups, group_rows, window, ring_bytes;
  const double *rhs;      // right-hand sides in this sweep's rhs layout
  double *by_position;    // the sweep's result in position order (read again for values that left the window)
  double *next;           // forward sweep: the backward sweep's right-hand sides (its rhs layout); backward: the result by row
  long long *trace;       // S35_WAVE_TRACING builds: per tile, clocks of its warps (else nullptr)
};

template <bool FORWARD>
__global__ void __launch_bounds__(kWaveMaxThreads, 1) k_wave_sweep(const WaveArgs a)
{
  extern __shared__ __align__(128) unsigned char wave_smem[];
  double *const window = reinterpret_cast<double *>(wave_smem);
  unsigned char *const ring = wave_smem + (size_t)a.window * 8;
  uint64_t *const full = reinterpret_cast<uint64_t *>(ring + a.ring_bytes);
  uint64_t *const empty = full + kWaveSlots;
  int32_t *const ring_at = reinterpret_cast<int32_t *>(empty + kWaveSlots); // where in the ring the tile of a slot lies
  const int tid = threadIdx.x;
  const int n_consumers = a.groups * a.group_rows;
  if (tid == 0)
    {
      for (int s = 0; s < kWaveSlots; ++s)
        {
          mbar_init(full + s, 1);
          mbar_init(empty + s, a.group_rows / 32);
        }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
  for (int k = tid; k < a.window; k += blockDim.x)
    window[k] = 0.0; // a padding entry multiplies 0.0 with slot 0: it must hold a finite number from the start
  __syncthreads();

  if (tid >= n_consumers)
    {
      // producer: ONE thread issues the copies (no warp-wide instruction may sit between two copies: measured on
      // B200, a shuffle in this loop waited > 1000 cycles behind the consumers' barrier waits).  The directory is
      // read four tiles ahead.
      if (tid != n_consumers)
        return;
      constexpr int kAhead = 4;
      int4 dir[kAhead];
      int2 rdir[kAhead];
#pragma unroll
      for (int u = 0; u < kAhead; ++u)
        {
          dir[u] = u < a.n_tiles ? __ldg(a.directory + u) : make_int4(0, 0, 0, 0);
          rdir[u] = u < a.n_tiles ? __ldg(a.rhs_dir + u) : make_int2(0, 0);
        }
      int freed = 0; // tiles known to be finished
      for (int base = 0; base < a.n_tiles; base += kAhead)
        {
#pragma unroll
          for (int u = 0; u < kAhead; ++u)
            {
              const int t = base + u;
              if (t >= a.n_tiles)
                break;
              const int4 d = dir[u];
              const int2 r = rdir[u];
              if (t + kAhead < a.n_tiles)
                {
                  dir[u] = __ldg(a.directory + t + kAhead);
                  rdir[u] = __ldg(a.rhs_dir + t + kAhead);
                }
              const int s = t % kWaveSlots;
              if (kWaveTracing && a.trace)
                a.trace[24 * (size_t)t + 5] = clock64();
              while (freed <= d.w)
                {
                  mbar_wait(empty + freed % kWaveSlots, (uint32_t)((freed / kWaveSlots) & 1));
                  ++freed;
                }
              if (kWaveTracing && a.trace)
                a.trace[24 * (size_t)t + 4] = clock64();
              // tile and right-hand sides land side by side; one barrier counts both
              ring_at[s] = d.y * 16;
              mbar_expect_tx(full + s, (uint32_t)(d.z + r.y));
              bulk_copy_to_shared(ring + (size_t)d.y * 16, a.blob + (size_t)d.x * 16, (uint32_t)d.z, full + s);
              bulk_copy_to_shared(ring + (size_t)d.y * 16 + d.z, a.rhs + r.x, (uint32_t)r.y, full + s);
            }
        }
      return;
    }

  const int group = tid / a.group_rows, i = tid - group * a.group_rows, lane = tid & 31;
  const int warp_first_row = i - lane;
  const int pair = 2 * a.group_rows; // threads on a tile's barrier: its own group arrives, the next tile's group waits
  const uint32_t mask = (uint32_t)a.window - 1u;
  // The consumers read nothing from global memory (a pending global load holds a warp up at its next shared-memory
  // result, however early it was issued): tile, right-hand sides and the tile's place in the ring all come through
  // shared memory, behind the tile's `full` barrier.
  for (int t = group; t < a.n_tiles; t += a.groups)
    {
      const int s = t % kWaveSlots;
      if (kWaveTracing && a.trace && i == 0)
        a.trace[24 * (size_t)t + 0] = clock64();
      mbar_wait(full + s, (uint32_t)((t / kWaveSlots) & 1));
      if (kWaveTracing && a.trace && i == 0)
        a.trace[24 * (size_t)t + 1] = clock64();
      const unsigned char *const tile = ring + ring_at[s];
      const int4 head = *reinterpret_cast<const int4 *>(tile); // first position, rows, warps, n_far
      const int rows = head.y, desc_first_pos = head.x;
      if (warp_first_row >= rows)
        {
          // no row for this warp: it only keeps the barriers' counts right
          if (t > 0)
            group_wait(t - 1, pair);
          if (t + 1 < a.n_tiles)
            group_signal(t, pair);
          if (lane == 0)
            mbar_arrive(empty + s);
          continue;
        }
      const bool tracing = kWaveTracing && a.trace != nullptr && i == 0;
      const uint32_t word = reinterpret_cast<const uint32_t *>(tile)[4 + (i >> 5)];
      const unsigned char *const block = tile + (size_t)(word >> 16) * 128;
      const int n_pre = (int)((word >> 8) & 0xffu), n_chunks = (int)(word & 0xffu);
      const bool active = i < rows;
      constexpr int kHead = FORWARD ? kWaveForwardHead : kWaveBackwardHead;
      const double rhs = reinterpret_cast<const double *>(tile + reinterpret_cast<const int32_t *>(tile)[12])[active ? i : 0];
      // this lane's first chunk: slots at +0, values 0-1 at +256, values 2-3 at +768; the next chunk 1280 bytes on
      const unsigned char *const lane_slots = block + kHead + 8 * lane;
      const unsigned char *const lane_vals = block + kHead + 256 + 16 * lane;
      double acc = 0.0, piv = 1.0, result = 0.0;
      int32_t scatter_to; // where the result goes in a.next
      if (FORWARD)
        scatter_to = reinterpret_cast<const int32_t *>(block)[lane];
      else
        {
          piv = reinterpret_cast<const double *>(block)[lane];
          scatter_to = reinterpret_cast<const int32_t *>(block + 256)[lane];
        }
      // one chunk: four products, the solution values straight from the window (the slots are byte offsets)
      struct Chunk
      {
        uint2 s4;
        double2 lo, hi;
      };
      auto load_chunk = [&](int c) {
        Chunk k;
        k.s4 = *reinterpret_cast<const uint2 *>(lane_slots + (size_t)c * kWaveChunkBytes);
        k.lo = *reinterpret_cast<const double2 *>(lane_vals + (size_t)c * kWaveChunkBytes);
        k.hi = *reinterpret_cast<const double2 *>(lane_vals + (size_t)c * kWaveChunkBytes + 512);
        return k;
      };
      auto sum_chunk = [&](const Chunk &k, double sum) {
        const unsigned char *const w = reinterpret_cast<const unsigned char *>(window);
        const double x0 = *reinterpret_cast<const double *>(w + (k.s4.x & 0xffffu)), x1 = *reinterpret_cast<const double *>(w + (k.s4.x >> 16));
        const double x2 = *reinterpret_cast<const double *>(w + (k.s4.y & 0xffffu)), x3 = *reinterpret_cast<const double *>(w + (k.s4.y >> 16));
        sum = fma(k.lo.x, x0, sum);
        sum = fma(k.lo.y, x1, sum);
        sum = fma(k.hi.x, x2, sum);
        return fma(k.hi.y, x3, sum);
      };
      // one chunk's loads in flight at a time: a deeper look-ahead only lengthens the queue in front of the load unit,
      // in which the gather of the group that is on turn has to wait (unroll 2: 1.138 ms per application, 1: 1.120)
#pragma unroll 1
      for (int c = 0; c < n_pre; ++c)
        acc = sum_chunk(load_chunk(c), acc);
      if (head.w > 0)
        {
          // values that have left the window: a short list per tile, each thread picks its own
          const WaveFar *const far = reinterpret_cast<const WaveFar *>(tile + 64);
          for (int f = 0; f < head.w; ++f)
            if (far[f].row == i)
              acc = fma(far[f].value, ld_l2(a.by_position + far[f].position), acc);
        }
      // the fresh chunks' factor entries are fetched before the wait (the first two; more are rare)
      const int n_post = n_chunks - n_pre;
      Chunk post0 = {make_uint2(0u, 0u), make_double2(0., 0.), make_double2(0., 0.)}, post1 = post0;
      if (n_post > 0)
        post0 = load_chunk(n_pre);
      if (n_post > 1)
        post1 = load_chunk(n_pre + 1);
      __syncwarp();
      if (tracing)
        a.trace[24 * (size_t)t + 2] = clock64();
      if (t > 0)
        group_wait(t - 1, pair);
      if (tracing)
        a.trace[24 * (size_t)t + 3] = clock64();
      if (kWaveTracing && a.trace && lane == 0)
        a.trace[24 * (size_t)t + 8 + (i >> 5)] = clock64();
      if (n_post > 0)
        acc = sum_chunk(post0, acc);
      if (n_post > 1)
        acc = sum_chunk(post1, acc);
      for (int c = n_pre + 2; c < n_pre + n_post; ++c)
        acc = sum_chunk(load_chunk(c), acc);
      if (active)
        {
          result = rhs - acc;
          if (!FORWARD)
            result *= piv;
          window[(uint32_t)(desc_first_pos + i) & mask] = result;
        }
      const int32_t first_pos = desc_first_pos;
      __syncwarp();
      if (t + 1 < a.n_tiles)
        group_signal(t, pair);
      if (tracing)
        a.trace[24 * (size_t)t + 6] = clock64();
      if (kWaveTracing && a.trace && lane == 0)
        a.trace[24 * (size_t)t + 16 + (i >> 5)] = clock64();
      if (lane == 0)
        mbar_arrive(empty + s); // the tile's bytes are no longer needed by this warp (before the global stores: the
                                // arrive is a release and would wait for them)
      if (active)
        {
          a.by_position[first_pos + i] = result;
          a.next[scatter_to] = result;
        }
    }
}
#undef S35_WAVE_GATHER

// stream[dst[i]] = lu[src[i]]: the factor's values into the packed stream after a factorisation
__global__ void k_wave_fill_values(int64_t n, const int32_t *__restrict__ dst, const int32_t *__restrict__ src,
                                   const double *__restrict__ lu, double *__restrict__ stream)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    stream[dst[i]] = lu[src[i]];
}

// dst[slot[r]] = src[r]: the sweep's input into rhs layout
__global__ void k_wave_to_layout(int32_t n, const int32_t *__restrict__ slot, const double *__restrict__ src, double *__restrict__ dst)
{
  const int32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n)
    dst[slot[r]] = src[r];
}

void wave_launch(DevIlu &p, double *dst, const double *src, cudaStream_t s)
{
  DevIlu::Wave &w = p.wave;
  S35_LAUNCH(k_wave_to_layout, (p.n + 255) / 256, 256, 0, s, p.n, w.rhs_slot.ptr, src, w.rhs[0].ptr);
  WaveArgs fwd{w.blob[0].ptr, reinterpret_cast<const int4 *>(w.directory[0].ptr), reinterpret_cast<const int2 *>(w.rhs_dir[0].ptr),
               w.n_tiles[0], w.groups, w.group_rows, w.window, w.ring_bytes, w.rhs[0].ptr, w.by_position[0].ptr, w.rhs[1].ptr,
               w.trace.ptr};
  WaveArgs bwd{w.blob[1].ptr, reinterpret_cast<const int4 *>(w.directory[1].ptr), reinterpret_cast<const int2 *>(w.rhs_dir[1].ptr),
               w.n_tiles[1], w.groups, w.group_rows, w.window, w.ring_bytes, w.rhs[1].ptr, w.by_position[1].ptr, dst, nullptr};
  const int threads = w.groups * w.group_rows + 32;
  S35_LAUNCH((k_wave_sweep<true>), 1, threads, w.smem_bytes, s, fwd);
  S35_LAUNCH((k_wave_sweep<false>), 1, threads, w.smem_bytes, s, bwd);
  if (kWaveTracing && w.trace.ptr && ++w.trace_calls == 4)
    {
      // S35_WAVE_TRACE=<file>: the forward sweep's clocks of the fourth application, one line per tile
      std::vector<long long> h(w.trace.n);
      w.trace.download(h.data(), s);
      if (FILE *f = std::fopen(std::getenv("S35_WAVE_TRACE"), "w"))
        {
          for (int t = 0; t < w.n_tiles[0]; ++t)
            {
              std::fprintf(f, "%d", t);
              for (int k = 0; k < 24; ++k)
                std::fprintf(f, " %lld", h[24 * (size_t)t + k]);
              std::fprintf(f, "\n");
            }
          std::fclose(f);
        }
    }
}

int env_int(const char *name, int fallback)
{
  const char *t = std::getenv(name);
  return t && *t ? std::atoi(t) : fallback;
}

// packs both sweeps of the pattern; returns false if the mode does not suit this factor
bool wave_create(DevIlu &p, const DevMatrix &A, cudaStream_t s, bool force = false)
{
  const int32_t n = A.n_rows;
  if (n == 0)
    return false;
  std::vector<int32_t> diag(n);
  for (int32_t r = 0; r < n; ++r)
    for (int32_t k = A.h_rowptr[r]; k < A.h_rowptr[r + 1]; ++k)
      if (A.h_col[k] == r)
        diag[r] = k;
  WavefrontOptions opt;
  opt.groups = env_int("S35_WAVE_GROUPS", opt.groups);
  opt.group_rows = env_int("S35_WAVE_GROUP_ROWS", opt.group_rows);
  opt.window = env_int("S35_WAVE_WINDOW", opt.window);
  opt.ring_bytes = env_int("S35_WAVE_RING_KB", opt.ring_bytes / 1024) * 1024;
  opt.max_tile_bytes = env_int("S35_WAVE_TILE_KB", opt.max_tile_bytes / 1024) * 1024;
  DevIlu::Wave &w = p.wave;
  w = DevIlu::Wave();
  opt.slots = kWaveSlots;
  w.smem_bytes = (size_t)opt.window * 8 + (size_t)opt.ring_bytes + (size_t)kWaveSlots * 20 + 16;
  if (opt.groups >= kWaveBarriers || opt.groups * opt.group_rows + 32 > kWaveMaxThreads || w.smem_bytes > 227 * 1024)
    return false;
  WavefrontPlan F, B;
  try
    {
      // the backward sweep first: the forward sweep writes its results straight into the backward sweep's rhs layout
      B = wavefront_pack(n, A.h_rowptr.data(), A.h_col.data(), diag.data(), false, nullptr, opt);
      F = wavefront_pack(n, A.h_rowptr.data(), A.h_col.data(), diag.data(), true, B.rhs_slot.data(), opt);
    }
  catch (const std::invalid_argument &)
    {
      return false; // e.g. a row that does not fit a tile: the wide sweep takes this factor
    }
  if (!force && (double)(F.far_entries + B.far_entries) > 0.02 * (double)(F.entries + B.entries + 1))
    return false; // too many values outside the window: every one is a trip to L2 in the middle of a tile
  // Levels narrower than a warp on average (coarse meshes, BASELINE configs 2 and 3: ~16 rows per level): five groups of
  // ONE warp hand the levels round (measured, tools/tune_ilu.py narrow: 8-17 % faster sweeps than the default shape, of which
  // most would idle and still take part in every barrier).
  const double rows_per_level = (double)n / std::max(1, std::max(F.n_levels, B.n_levels));
  if (rows_per_level <= 24.0 && !std::getenv("S35_WAVE_GROUPS") && !std::getenv("S35_WAVE_GROUP_ROWS"))
    {
      WavefrontOptions narrow = opt;
      narrow.groups = 5;
      narrow.group_rows = 32;
      try
        {
          WavefrontPlan B2 = wavefront_pack(n, A.h_rowptr.data(), A.h_col.data(), diag.data(), false, nullptr, narrow);
          WavefrontPlan F2 = wavefront_pack(n, A.h_rowptr.data(), A.h_col.data(), diag.data(), true, B2.rhs_slot.data(), narrow);
          if (force || (double)(F2.far_entries + B2.far_entries) <= 0.02 * (double)(F2.entries + B2.entries + 1))
            {
              B = std::move(B2);
              F = std::move(F2);
              opt = narrow;
            }
        }
      catch (const std::invalid_argument &)
        {
        }
    }
  w.groups = opt.groups;
  w.group_rows = opt.group_rows;
  w.window = opt.window;
  w.ring_bytes = opt.ring_bytes;
  w.rhs_slot.upload(F.rhs_slot, s);
  WavefrontPlan *plans[2] = {&F, &B};
  for (int d = 0; d < 2; ++d)
    {
      WavefrontPlan &P = *plans[d];
      w.n_tiles[d] = P.n_tiles;
      w.n_levels[d] = P.n_levels;
      w.n_values[d] = (int64_t)P.value_dst.size();
      w.stream_bytes[d] = (int64_t)P.blob.size();
      w.far_fraction[d] = P.entries ? (double)P.far_entries / (double)P.entries : 0.0;
      w.fresh_fraction[d] = P.entries ? (double)P.fresh_entries / (double)P.entries : 0.0;
      w.blob[d].upload(P.blob, s);
      w.directory[d].upload(P.directory, s);
      w.rhs_dir[d].upload(P.rhs_dir, s);
      w.rhs[d].alloc((size_t)P.rhs_len);
      w.rhs[d].zero(s);
      w.value_dst[d].upload(P.value_dst, s);
      w.value_src[d].upload(P.value_src, s);
      w.by_position[d].alloc(n);
      w.by_position[d].zero(s);
    }
  S35_CUDA(cudaFuncSetAttribute(k_wave_sweep<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  S35_CUDA(cudaFuncSetAttribute(k_wave_sweep<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  if (const char *f = std::getenv("S35_WAVE_TRACE"))
    if (*f && n > 100000) // the large factor of a run only
      {
        w.trace.alloc((size_t)24 * w.n_tiles[0]);
        w.trace.zero(s);
      }
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
  // Which sweep.  The wavefront mode (one block, levels handed over inside an SM) wins where the levels are narrow
  // and the sweep's values stay in its shared-memory window: factors with short rows (the Q1 pressure matrices:
  // 2.13 -> 1.27 ms per application at r1), and factors with long rows whose bandwidth is below half the window (the
  // Q2 velocity matrices of the coarse meshes: cylinder3d r0 21.5 -> 16.0 ms per diffusion step).  On the large
  // velocity factors a sixth of the values has left the window and one SM cannot stream 600 MB per sweep: they keep
  // the wide sweep.  S35_ILU_MODE = wide | wavefront overrides the choice, S35_WAVE_MAX_ROW moves the row-length bound.
  wave = Wave();
  const char *mode = std::getenv("S35_ILU_MODE");
  const bool force_wide = mode && std::strcmp(mode, "wide") == 0, force_wave = mode && std::strcmp(mode, "wavefront") == 0;
  const double mean_row = n ? (double)A.nnz / n : 0.0;
  int32_t bandwidth = 0;
  for (int32_t r = 0; r < n; ++r)
    if (rp[r + 1] > rp[r])
      bandwidth = std::max(bandwidth, std::max(r - cl[rp[r]], cl[rp[r + 1] - 1] - r));
  const bool suits = mean_row <= (double)env_int("S35_WAVE_MAX_ROW", 40) || bandwidth <= env_int("S35_WAVE_WINDOW", 8192) / 2;
  if (!force_wide && (force_wave || suits))
    wave_create(*this, A, s, force_wave);
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
          S35_LAUNCH(k_wave_fill_values, grid, 256, 0, s, wave.n_values[d], wave.value_dst[d].ptr, wave.value_src[d].ptr, lu.ptr,
                     reinterpret_cast<double *>(wave.blob[d].ptr));
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
