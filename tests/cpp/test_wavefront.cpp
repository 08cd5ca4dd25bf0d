// CPU test of the wavefront schedule (dealii-step-35a_b200/csrc/wavefront.h): packs the forward and the backward
// sweep of random factors on several patterns, group counts and tile / window / ring sizes, walks the packed streams
// with the kernel's timing made adversarial (values not yet guaranteed to exist are NaN), and compares with plain
// substitutions over the CSR arrays.
//
//   g++ -std=c++17 -O1 -I dealii-step-35a_b200/csrc tests/cpp/test_wavefront.cpp -o test_wavefront
#include "wavefront.h"

#include <cstdio>
#include <cstring>
#include <random>
#include <set>

namespace {

struct Csr
{
  int32_t n = 0;
  std::vector<int32_t> rowptr, col, diag;
  std::vector<double> val;
};

// pattern from a neighbour predicate on a box grid, rows renumbered by `perm`
Csr grid_pattern(int nx, int ny, int nz, int reach, const std::vector<int32_t> &perm, std::mt19937 &rng)
{
  const int32_t n = nx * ny * nz;
  std::vector<std::set<int32_t>> rows(n);
  for (int z = 0; z < nz; ++z)
    for (int y = 0; y < ny; ++y)
      for (int x = 0; x < nx; ++x)
        for (int dz = -reach; dz <= reach; ++dz)
          for (int dy = -reach; dy <= reach; ++dy)
            for (int dx = -reach; dx <= reach; ++dx)
              {
                const int X = x + dx, Y = y + dy, Z = z + dz;
                if (X < 0 || Y < 0 || Z < 0 || X >= nx || Y >= ny || Z >= nz)
                  continue;
                rows[perm[(z * ny + y) * nx + x]].insert(perm[(Z * ny + Y) * nx + X]);
              }
  Csr A;
  A.n = n;
  A.rowptr.assign(1, 0);
  A.diag.resize(n);
  std::uniform_real_distribution<double> u(-1.0, 1.0);
  for (int32_t r = 0; r < n; ++r)
    {
      for (int32_t c : rows[r])
        {
          if (c == r)
            A.diag[r] = (int32_t)A.col.size();
          A.col.push_back(c);
          A.val.push_back(c == r ? 1.0 / (4.0 + u(rng)) : 0.3 * u(rng)); // inverse pivot on the diagonal
        }
      A.rowptr.push_back((int32_t)A.col.size());
    }
  return A;
}

// random band pattern with rows of very different lengths (some hold only their diagonal)
Csr random_band(int32_t n, int reach, int max_per_row, std::mt19937 &rng)
{
  Csr A;
  A.n = n;
  A.rowptr.assign(1, 0);
  A.diag.resize(n);
  std::uniform_real_distribution<double> u(-1.0, 1.0);
  for (int32_t r = 0; r < n; ++r)
    {
      std::set<int32_t> cols{r};
      const int want = (int)(rng() % (unsigned)(max_per_row + 1));
      for (int k = 0; k < want; ++k)
        {
          const int32_t c = r - reach + (int32_t)(rng() % (unsigned)(2 * reach + 1));
          if (c >= 0 && c < n)
            cols.insert(c);
        }
      for (int32_t c : cols)
        {
          if (c == r)
            A.diag[r] = (int32_t)A.col.size();
          A.col.push_back(c);
          A.val.push_back(c == r ? 1.0 / (3.0 + u(rng)) : 0.2 * u(rng));
        }
      A.rowptr.push_back((int32_t)A.col.size());
    }
  return A;
}

bool same_bits(const std::vector<double> &a, const std::vector<double> &b)
{
  return a.size() == b.size() && std::memcmp(a.data(), b.data(), a.size() * sizeof(double)) == 0;
}

// largest difference relative to the largest entry
double rel_diff(const std::vector<double> &a, const std::vector<double> &b)
{
  double d = 0.0, m = 0.0;
  for (size_t i = 0; i < a.size(); ++i)
    {
      d = std::max(d, std::fabs(a[i] - b[i]));
      m = std::max(m, std::fabs(b[i]));
    }
  return a.size() == b.size() ? d / std::max(m, 1e-300) : 1.0;
}

int check(const char *name, const Csr &A, const s35::WavefrontOptions &opt, std::mt19937 &rng)
{
  const int32_t n = A.n;
  std::vector<double> b(n), y(n), x(n);
  std::uniform_real_distribution<double> u(-1.0, 1.0);
  for (double &v : b)
    v = u(rng);
  for (int32_t r = 0; r < n; ++r)
    {
      double acc = 0.0;
      for (int32_t k = A.rowptr[r]; k < A.diag[r]; ++k)
        acc = std::fma(A.val[k], y[A.col[k]], acc);
      y[r] = b[r] - acc;
    }
  for (int32_t r = n - 1; r >= 0; --r)
    {
      double acc = 0.0;
      for (int32_t k = A.diag[r] + 1; k < A.rowptr[r + 1]; ++k)
        acc = std::fma(A.val[k], x[A.col[k]], acc);
      x[r] = (y[r] - acc) * A.val[A.diag[r]];
    }

  // the backward sweep first: the forward sweep writes its results straight into the backward sweep's rhs layout
  s35::WavefrontPlan B = s35::wavefront_pack(n, A.rowptr.data(), A.col.data(), A.diag.data(), false, nullptr, opt);
  s35::WavefrontPlan F = s35::wavefront_pack(n, A.rowptr.data(), A.col.data(), A.diag.data(), true, B.rhs_slot.data(), opt);
  s35::wavefront_fill_values(F, A.val.data());
  s35::wavefront_fill_values(B, A.val.data());
  const double junk = std::nan("");
  std::vector<double> b_layout(F.rhs_len, junk), y_layout(B.rhs_len, junk);
  for (int32_t r = 0; r < n; ++r)
    b_layout[F.rhs_slot[r]] = b[r]; // what the permutation kernel does before the forward sweep
  std::vector<double> y_pos(n, junk), x_pos(n, junk), x_nat(n, junk);
  s35::wavefront_apply_host(F, b_layout.data(), y_pos.data(), nullptr, y_layout.data());
  s35::wavefront_apply_host(B, y_layout.data(), x_pos.data(), x_nat.data(), nullptr);

  // the sums are re-associated (old entries first), so the comparison is to rounding, not bit for bit; a schedule
  // error reads a NaN or a stale value and is far outside this bound
  int bad = 0;
  std::vector<double> y_back(n), x_back(n);
  for (int32_t r = 0; r < n; ++r)
    {
      y_back[r] = y_pos[F.position[r]];
      x_back[r] = x_pos[B.position[r]];
    }
  const double tol = 1e-11;
  if (!(rel_diff(y_back, y) < tol))
    bad |= 1;
  if (!(rel_diff(x_nat, x) < tol))
    bad |= 2;
  if (!same_bits(x_back, x_nat))
    bad |= 4;
  // every tile respects its bounds, lies 128-byte aligned in stream and ring, and waits for an earlier tile only
  int64_t post_max = 0;
  for (const s35::WavefrontPlan *P : {&F, &B})
    {
      for (int32_t t = 0; t < P->n_tiles; ++t)
        {
          const int32_t *d = P->directory.data() + 4 * (size_t)t;
            const int32_t rows = reinterpret_cast<const int32_t *>(P->blob.data() + (size_t)d[0] * 16)[1];
          const int32_t *rd = P->rhs_dir.data() + 2 * (size_t)t;
          if (rows < 1 || rows > opt.group_rows || d[2] + rd[1] > opt.ring_bytes / 2 || d[2] % 128 || d[0] % 8 || d[1] % 8 || rd[1] % 128 ||
              rd[0] % 16 || rd[1] < 8 * rows || d[1] * 16 + d[2] + rd[1] > opt.ring_bytes || d[3] >= t || d[3] < t - opt.slots)
            bad |= 8;
        }
      post_max = std::max<int64_t>(post_max, P->max_post_chunks);
    }
  std::printf("%-28s n=%6d  levels %5d/%5d  tiles %6d/%6d  far %5.2f%% / %5.2f%%  fresh %4.1f%%  stored/real %.2f  chunks after the wait: max %d  %s\n",
              name, n, F.n_levels, B.n_levels, F.n_tiles, B.n_tiles, 100.0 * F.far_entries / std::max<int64_t>(1, F.entries),
              100.0 * B.far_entries / std::max<int64_t>(1, B.entries), 100.0 * F.fresh_entries / std::max<int64_t>(1, F.entries),
              (double)F.stored_entries / std::max<int64_t>(1, F.entries), (int)post_max, bad ? "FAILED" : "ok");
  return bad;
}

std::vector<int32_t> identity(int32_t n)
{
  std::vector<int32_t> p(n);
  for (int32_t i = 0; i < n; ++i)
    p[i] = i;
  return p;
}

} // namespace

// a pattern written by the Python test (int32: n, nnz, rowptr[n + 1], col[nnz]): the matrices of a real mesh
Csr read_pattern(const char *file, std::mt19937 &rng)
{
  Csr A;
  FILE *f = std::fopen(file, "rb");
  int32_t head[2] = {0, 0};
  if (!f || std::fread(head, 4, 2, f) != 2)
    throw std::runtime_error("cannot read the pattern file");
  A.n = head[0];
  A.rowptr.resize(A.n + 1);
  A.col.resize(head[1]);
  if (std::fread(A.rowptr.data(), 4, A.rowptr.size(), f) != A.rowptr.size() || std::fread(A.col.data(), 4, A.col.size(), f) != A.col.size())
    throw std::runtime_error("short pattern file");
  std::fclose(f);
  A.diag.assign(A.n, -1);
  A.val.resize(A.col.size());
  std::uniform_real_distribution<double> u(-1.0, 1.0);
  for (int32_t r = 0; r < A.n; ++r)
    {
      const int32_t len = A.rowptr[r + 1] - A.rowptr[r];
      for (int32_t k = A.rowptr[r]; k < A.rowptr[r + 1]; ++k)
        {
          if (A.col[k] == r)
            A.diag[r] = k;
          A.val[k] = A.col[k] == r ? 1.0 / (1.0 + 0.1 * u(rng)) : u(rng) / (2.0 * len); // inverse pivots near 1, small couplings
        }
    }
  return A;
}

int main(int argc, char **argv)
{
  std::mt19937 rng(35);
  int bad = 0;
  if (argc > 1)
    {
      // the shapes the library chooses: the default, and five one-warp groups for narrow levels
      s35::WavefrontOptions standard, narrow;
      narrow.groups = 5;
      narrow.group_rows = 32;
      for (int k = 1; k < argc; ++k)
        {
          const Csr A = read_pattern(argv[k], rng);
          for (const s35::WavefrontOptions *o : {&standard, &narrow})
            try
              {
                bad |= check(argv[k], A, *o, rng);
              }
            catch (const std::invalid_argument &e) // "not for this sweep" is an answer; a logic_error is a fault
              {
                std::printf("  %s: declined (%s)\n", argv[k], e.what());
              }
        }
      std::printf(bad ? "wavefront: FAILED (%d)\n" : "wavefront: all schedules reproduce the substitutions\n", bad);
      return bad ? 1 : 0;
    }
  s35::WavefrontOptions standard, tiny_window, tiny_tiles, two_groups;
  tiny_window.window = 256; // = (groups + 1) * group_rows: every older value is far
  tiny_window.ring_bytes = 64 * 1024;
  tiny_window.max_tile_bytes = 24 * 1024;
  tiny_window.group_rows = 32;
  tiny_window.groups = 7;
  tiny_tiles.group_rows = 64;
  tiny_tiles.groups = 3;
  tiny_tiles.max_tile_bytes = 15 * 1024; // one warp, or two with short rows
  tiny_tiles.ring_bytes = 32 * 1024;    // and two or three tiles in the ring: every load waits for an earlier tile
  tiny_tiles.window = 512;
  tiny_tiles.slots = 4;
  two_groups.groups = 2;
  two_groups.group_rows = 64;
  two_groups.ring_bytes = 64 * 1024;
  two_groups.max_tile_bytes = 24 * 1024;
  {
    const Csr A = grid_pattern(60, 50, 1, 1, identity(3000), rng); // 9-point grid, natural order: levels are diagonals
    bad |= check("grid 60x50, 9-point", A, standard, rng);
    bad |= check("  window 256, 7 groups", A, tiny_window, rng);
    bad |= check("  tiny tiles and ring", A, tiny_tiles, rng);
    bad |= check("  two groups", A, two_groups, rng);
  }
  {
    const Csr A = grid_pattern(14, 12, 10, 1, identity(1680), rng); // 27-point box
    bad |= check("box 14x12x10, 27-point", A, standard, rng);
    bad |= check("  window 256, 7 groups", A, tiny_window, rng);
    bad |= check("  tiny tiles and ring", A, tiny_tiles, rng);
    bad |= check("  two groups", A, two_groups, rng);
  }
  {
    std::vector<int32_t> perm = identity(2000);
    std::shuffle(perm.begin(), perm.end(), rng); // random numbering: few, wide levels and long-range dependencies
    const Csr A = grid_pattern(20, 10, 10, 1, perm, rng);
    bad |= check("box 20x10x10, shuffled", A, standard, rng);
    bad |= check("  window 256, 7 groups", A, tiny_window, rng);
    bad |= check("  tiny tiles and ring", A, tiny_tiles, rng);
    bad |= check("  two groups", A, two_groups, rng);
  }
  {
    const Csr A = grid_pattern(40, 1, 1, 2, identity(40), rng); // a chain: one row per level
    bad |= check("chain of 40, reach 2", A, standard, rng);
    bad |= check("  tiny tiles and ring", A, tiny_tiles, rng);
    const Csr D = grid_pattern(50, 1, 1, 0, identity(50), rng); // diagonal matrix: one level, no entries
    bad |= check("diagonal of 50", D, tiny_tiles, rng);
  }
  for (int round = 0; round < 3; ++round)
    {
      const Csr A = random_band(1500 + 371 * round, 150 + 400 * round, 6 + 10 * round, rng);
      bad |= check("random band", A, standard, rng);
      bad |= check("  window 256, 7 groups", A, tiny_window, rng);
      bad |= check("  tiny tiles and ring", A, tiny_tiles, rng);
      bad |= check("  two groups", A, two_groups, rng);
    }
  try
    {
      s35::WavefrontOptions wrong;
      wrong.window = 512; // fewer slots than the rows of groups + 1 tiles (4 x 256)
      const Csr A = grid_pattern(10, 10, 1, 1, identity(100), rng);
      s35::wavefront_pack(A.n, A.rowptr.data(), A.col.data(), A.diag.data(), false, nullptr, wrong);
      bad |= 16;
    }
  catch (const std::invalid_argument &)
    {
    }
  if (bad)
    {
      std::printf("wavefront: FAILED (%d)\n", bad);
      return 1;
    }
  std::printf("wavefront: all schedules reproduce the substitutions\n");
  return 0;
}
