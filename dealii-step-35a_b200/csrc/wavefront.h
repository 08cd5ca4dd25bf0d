// Level-sequential ("wavefront") schedule of a sparse triangular sweep for ONE thread block, host side: the
// packed streams the kernel k_wave_sweep (ilu.cu) consumes, and an executor of the same streams on the CPU with
// which the packing is tested without a GPU (tests/cpp/test_wavefront.cpp).  No CUDA in this file.
//
// Why: the substitutions of SparseILU::vmult have thousands of dependency levels of ~100 rows each.  The wide
// sweep (k_ilu_sweep) pays a round trip through L2 per level.  Here one block walks the levels in order and keeps
// the most recent `window` solution values in shared memory, so that a level costs shared-memory latency plus a
// block barrier; the factor arrives through bulk copies that run several tiles ahead.  The arithmetic per row is
// that of the serial algorithm: the row's entries in ascending column order, one accumulator.
//
// Order of the rows ("positions"): by dependency level; inside a level by decreasing number of entries (so that
// a tile can be stored as jagged diagonals without padding), then by row index.  A TILE is a run of at most
// tile_rows rows of ONE level with at most tile_entries entries; thread i of the block owns row i of the tile.
//
// Streams (one pair per sweep direction), per tile:
//   meta (int32, length a multiple of 4):
//     [0] n_rows  [1] n_entries  [2] max_len  [3] position of the tile's first row
//     off[max_len rounded up to 4]      entry j of row i sits at  off[j] + i  (jagged diagonal j)
//     row[n_rows r.u. 4]                natural row index (where the result goes)
//     rhs[n_rows r.u. 4]                index of the row's right-hand side in the caller's vector
//     len[n_rows r.u. 4]                number of entries of the row
//     src[n_entries r.u. 4]             where the solution value of the entry's column is read:
//                                       >= 0: slot of the shared-memory window;  < 0: ~position, from global memory
//   values (double, length a multiple of 2):
//     val[n_entries r.u. 2]             factor entries, same indexing as src
//     piv[n_rows r.u. 2]                inverse pivots (backward sweep only)
// A value is "near" (read from the window) if its position is less than `window` behind the END of the reading
// tile: then no row of that tile overwrites its slot.  Everything else is "far" and read from the position-ordered
// copy of the solution in global memory, where it has been for a long time.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <stdexcept>
#include <vector>

namespace s35 {

struct WavefrontOptions
{
  int tile_rows = 256;     // = consumer threads of the block
  int tile_entries = 1536; // bound on a tile's entries (stage size in shared memory)
  int window = 16384;      // solution values kept in shared memory; a power of two
};

struct WavefrontPlan
{
  int32_t n = 0;
  bool forward = true;
  int32_t window = 0;
  int32_t n_tiles = 0, n_levels = 0;
  std::vector<int32_t> meta;       // all tiles' meta blocks
  std::vector<int32_t> directory;  // per tile: meta offset (ints), meta length (ints), value offset (doubles), value length
  std::vector<int64_t> value_source; // per slot of the value stream: index into the factor's value array, -1 = 0.0
  std::vector<int32_t> position;   // position of every row in this sweep's order
  int64_t n_values = 0;            // length of the value stream
  int32_t max_meta = 0, max_values = 0; // largest tile (ints / doubles): the stage size
  int64_t entries = 0, far_entries = 0;
};

inline int32_t round_up(int32_t x, int32_t m) { return (x + m - 1) / m * m; }

#if defined(__CUDACC__)
#define S35_WAVE_INLINE __host__ __device__ __forceinline__
#else
#define S35_WAVE_INLINE inline
#endif

// A tile as it lies in the two streams (or in a stage of shared memory).  The kernel and the CPU walk below use
// these two functions, so the index arithmetic and the order of the operations of the kernel are what the CPU
// test exercises.
struct WaveTile
{
  int32_t rows, entries, max_len, first_pos;
  const int32_t *off, *row, *rhs, *len, *src;
  const double *val, *piv;
};

S35_WAVE_INLINE WaveTile wave_tile(const int32_t *m, const double *v)
{
  WaveTile t;
  t.rows = m[0];
  t.entries = m[1];
  t.max_len = m[2];
  t.first_pos = m[3];
  const int32_t rows_pad = (t.rows + 3) & ~3;
  t.off = m + 4;
  t.row = t.off + ((t.max_len + 3) & ~3);
  t.rhs = t.row + rows_pad;
  t.len = t.rhs + rows_pad;
  t.src = t.len + rows_pad;
  t.val = v;
  t.piv = v + ((t.entries + 1) & ~1);
  return t;
}

// reads a solution value that has left the window from the position-ordered array; on the device through L2
// (the value was stored by another thread of the block, long ago)
struct WaveFarLoad
{
  const double *by_position;
  S35_WAVE_INLINE double operator()(int32_t pos) const
  {
#if defined(__CUDA_ARCH__)
    double v;
    asm volatile("ld.global.cg.f64 %0, [%1];" : "=d"(v) : "l"(by_position + pos) : "memory");
    return v;
#else
    return by_position[pos];
#endif
  }
};

// Row i of the tile: rhs - sum_j val_j * solution(column_j), times the inverse pivot in the backward sweep.  Eight
// entries at a time, all their loads issued before the first product is needed; the products are added in
// ascending column order into one accumulator.  far(position) reads a value that has left the window.
S35_WAVE_INLINE double wave_row(const WaveTile &t, int32_t i, double rhs, const double *window, WaveFarLoad far, bool forward)
{
  const int32_t n_mine = t.len[i];
  double acc = 0.0;
  for (int32_t j0 = 0; j0 < n_mine; j0 += 8)
    {
      int32_t where[8];
      double factor[8], solved[8];
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (j0 + u < n_mine)
          {
            const int32_t e = t.off[j0 + u] + i;
            where[u] = t.src[e];
            factor[u] = t.val[e];
          }
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (j0 + u < n_mine)
          solved[u] = where[u] >= 0 ? window[where[u]] : far(~where[u]);
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (j0 + u < n_mine)
          acc = fma(factor[u], solved[u], acc);
    }
  double result = rhs - acc;
  if (!forward)
    result *= t.piv[i];
  return result;
}

// rowptr/col: CSR with ascending columns; diag[r]: position of the diagonal entry of row r.
// forward: the strictly lower part (unit diagonal); backward: the strictly upper part and the pivot at diag[r].
// rhs_index: for every row the index of its right-hand side in the vector the sweep reads (nullptr: the row index).
inline WavefrontPlan wavefront_pack(int32_t n, const int32_t *rowptr, const int32_t *col, const int32_t *diag, bool forward,
                                    const int32_t *rhs_index, const WavefrontOptions &opt)
{
  if (opt.window <= 0 || (opt.window & (opt.window - 1)) != 0)
    throw std::invalid_argument("wavefront: the window must be a power of two");
  if (opt.tile_rows <= 0 || opt.tile_entries <= 0)
    throw std::invalid_argument("wavefront: bad tile bounds");
  if (opt.tile_rows > opt.window)
    throw std::invalid_argument("wavefront: a tile must not have more rows than the window has slots"); // two rows, one slot
  WavefrontPlan P;
  P.n = n;
  P.forward = forward;
  P.window = opt.window;
  auto first = [&](int32_t r) { return forward ? rowptr[r] : diag[r] + 1; };
  auto last = [&](int32_t r) { return forward ? diag[r] : rowptr[r + 1]; };

  // dependency levels
  std::vector<int32_t> level(n, 0);
  int32_t n_levels = n ? 1 : 0;
  for (int32_t s = 0; s < n; ++s)
    {
      const int32_t r = forward ? s : n - 1 - s;
      int32_t lv = 0;
      for (int32_t k = first(r); k < last(r); ++k)
        lv = std::max(lv, level[col[k]] + 1);
      level[r] = lv;
      n_levels = std::max(n_levels, lv + 1);
    }
  P.n_levels = n_levels;

  // rows by (level, more entries first, row index)
  std::vector<int32_t> order(n);
  for (int32_t r = 0; r < n; ++r)
    order[r] = r;
  std::sort(order.begin(), order.end(), [&](int32_t a, int32_t b) {
    if (level[a] != level[b])
      return level[a] < level[b];
    const int32_t la = last(a) - first(a), lb = last(b) - first(b);
    if (la != lb)
      return la > lb;
    return a < b;
  });
  P.position.assign(n, -1);
  for (int32_t p = 0; p < n; ++p)
    P.position[order[p]] = p;

  // tiles
  int32_t p = 0;
  while (p < n)
    {
      const int32_t lv = level[order[p]];
      int32_t rows = 0, entries = 0;
      while (p + rows < n && level[order[p + rows]] == lv && rows < opt.tile_rows)
        {
          const int32_t r = order[p + rows];
          const int32_t len = last(r) - first(r);
          if (len > opt.tile_entries)
            throw std::invalid_argument("wavefront: a row has more entries than a tile may hold");
          if (rows > 0 && entries + len > opt.tile_entries)
            break;
          entries += len;
          ++rows;
        }
      const int32_t tile_end = p + rows;
      const int32_t max_len = last(order[p]) - first(order[p]); // rows are sorted by decreasing length
      const int32_t off_pad = round_up(max_len, 4), rows_pad4 = round_up(rows, 4), ent_pad4 = round_up(entries, 4);
      const int32_t ent_pad2 = round_up(entries, 2), rows_pad2 = forward ? 0 : round_up(rows, 2);
      const int64_t meta_begin = (int64_t)P.meta.size(), val_begin = P.n_values;
      const int32_t meta_len = 4 + off_pad + 3 * rows_pad4 + ent_pad4, val_len = ent_pad2 + rows_pad2;
      if (meta_begin + meta_len > INT32_MAX || val_begin + val_len > INT32_MAX)
        throw std::invalid_argument("wavefront: streams exceed 32-bit offsets");
      P.meta.resize(meta_begin + meta_len, 0);
      P.value_source.resize(val_begin + val_len, -1);
      int32_t *m = P.meta.data() + meta_begin;
      m[0] = rows;
      m[1] = entries;
      m[2] = max_len;
      m[3] = p;
      int32_t *off = m + 4, *row = off + off_pad, *rhs = row + rows_pad4, *len = rhs + rows_pad4, *src = len + rows_pad4;
      // jagged diagonals: cnt[j] rows have more than j entries; they are the first cnt[j] rows of the tile
      {
        int32_t running = 0;
        for (int32_t j = 0; j < max_len; ++j)
          {
            off[j] = running;
            int32_t cnt = 0;
            while (cnt < rows && last(order[p + cnt]) - first(order[p + cnt]) > j)
              ++cnt;
            running += cnt;
          }
        if (running != entries)
          throw std::logic_error("wavefront: jagged diagonal sizes do not add up");
      }
      for (int32_t i = 0; i < rows; ++i)
        {
          const int32_t r = order[p + i];
          row[i] = r;
          rhs[i] = rhs_index ? rhs_index[r] : r;
          len[i] = last(r) - first(r);
          for (int32_t j = 0; j < len[i]; ++j)
            {
              const int32_t k = first(r) + j, q = P.position[col[k]];
              if (q < 0 || q >= p)
                throw std::logic_error("wavefront: a dependency is not in an earlier tile");
              const int32_t e = off[j] + i;
              const bool near = q >= tile_end - opt.window;
              src[e] = near ? (q & (opt.window - 1)) : ~q;
              P.far_entries += near ? 0 : 1;
              P.value_source[val_begin + e] = k;
            }
          if (!forward)
            P.value_source[val_begin + ent_pad2 + i] = diag[r];
        }
      P.entries += entries;
      P.directory.push_back((int32_t)meta_begin);
      P.directory.push_back(meta_len);
      P.directory.push_back((int32_t)val_begin);
      P.directory.push_back(val_len);
      P.n_values += val_len;
      P.max_meta = std::max(P.max_meta, meta_len);
      P.max_values = std::max(P.max_values, val_len);
      ++P.n_tiles;
      p = tile_end;
    }
  return P;
}

// value stream of a plan from the factor's value array
inline std::vector<double> wavefront_values(const WavefrontPlan &P, const double *lu)
{
  std::vector<double> v(P.n_values, 0.0);
  for (int64_t s = 0; s < P.n_values; ++s)
    if (P.value_source[s] >= 0)
      v[s] = lu[P.value_source[s]];
  return v;
}

// The kernel's walk over the streams, on the CPU.  rhs: the vector the rows' right-hand sides are read from
// (through the rhs index of the plan); by_position: the solution in position order (also the source of far
// values); natural: the solution by row index, or nullptr.  The rows of a tile are visited last to first and
// write the window at once, so a window slot that a tile overwrites while it still needs it shows up as a
// wrong result.
inline void wavefront_apply_host(const WavefrontPlan &P, const double *values, const double *rhs, double *by_position, double *natural)
{
  std::vector<double> window(P.window, 0.0);
  for (int32_t t = 0; t < P.n_tiles; ++t)
    {
      const WaveTile tile = wave_tile(P.meta.data() + P.directory[4 * t], values + P.directory[4 * t + 2]);
      for (int32_t i = tile.rows - 1; i >= 0; --i)
        {
          const double res = wave_row(tile, i, rhs[tile.rhs[i]], window.data(), WaveFarLoad{by_position}, P.forward);
          const int32_t pos = tile.first_pos + i;
          window[pos & (P.window - 1)] = res;
          by_position[pos] = res;
          if (natural)
            natural[tile.row[i]] = res;
        }
    }
}

} // namespace s35
