// CPU test of the wavefront schedule (dealii-step-35a_b200/csrc/wavefront.h): packs the forward and the backward
// sweep of random factors on several patterns and tile / window sizes, walks the packed streams the way the kernel
// does, and compares with plain substitutions over the CSR arrays -- bit for bit, since the order of the operations
// per row is the same.
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

  const s35::WavefrontPlan F = s35::wavefront_pack(n, A.rowptr.data(), A.col.data(), A.diag.data(), true, nullptr, opt);
  // the backward sweep reads its right-hand sides from the forward sweep's position-ordered result
  const s35::WavefrontPlan B = s35::wavefront_pack(n, A.rowptr.data(), A.col.data(), A.diag.data(), false, F.position.data(), opt);
  const std::vector<double> fv = s35::wavefront_values(F, A.val.data()), bv = s35::wavefront_values(B, A.val.data());
  std::vector<double> y_pos(n, -7.0), y_nat(n, -7.0), x_pos(n, -7.0), x_nat(n, -7.0);
  s35::wavefront_apply_host(F, fv.data(), b.data(), y_pos.data(), y_nat.data());
  s35::wavefront_apply_host(B, bv.data(), y_pos.data(), x_pos.data(), x_nat.data());

  int bad = 0;
  if (!same_bits(y_nat, y))
    bad |= 1;
  if (!same_bits(x_nat, x))
    bad |= 2;
  for (int32_t r = 0; r < n; ++r)
    if (y_pos[F.position[r]] != y[r] || x_pos[B.position[r]] != x[r])
      bad |= 4;
  // every tile respects its bounds, and the streams are aligned for 16-byte bulk copies
  for (const s35::WavefrontPlan *P : {&F, &B})
    for (int32_t t = 0; t < P->n_tiles; ++t)
      {
        const int32_t *m = P->meta.data() + P->directory[4 * t];
        if (m[0] < 1 || m[0] > opt.tile_rows || (m[1] > opt.tile_entries && m[0] > 1) || P->directory[4 * t] % 4 || P->directory[4 * t + 1] % 4 ||
            P->directory[4 * t + 2] % 2 || P->directory[4 * t + 3] % 2)
          bad |= 8;
      }
  std::printf("%-28s n=%6d  levels %5d/%5d  tiles %6d/%6d  far %5.2f%% / %5.2f%%  %s\n", name, n, F.n_levels, B.n_levels, F.n_tiles,
              B.n_tiles, 100.0 * F.far_entries / std::max<int64_t>(1, F.entries), 100.0 * B.far_entries / std::max<int64_t>(1, B.entries),
              bad ? "FAILED" : "ok");
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

int main()
{
  std::mt19937 rng(35);
  int bad = 0;
  s35::WavefrontOptions standard, tiny_window, tiny_tiles;
  tiny_window.window = 64;
  tiny_window.tile_rows = 48;
  tiny_tiles.tile_rows = 8;
  tiny_tiles.tile_entries = 40;
  tiny_tiles.window = 256;
  {
    const Csr A = grid_pattern(60, 50, 1, 1, identity(3000), rng); // 9-point grid, natural order: levels are diagonals
    bad |= check("grid 60x50, 9-point", A, standard, rng);
    bad |= check("  window 64", A, tiny_window, rng);
    bad |= check("  tiles 8 rows / 40 entries", A, tiny_tiles, rng);
  }
  {
    const Csr A = grid_pattern(14, 12, 10, 1, identity(1680), rng); // 27-point box
    bad |= check("box 14x12x10, 27-point", A, standard, rng);
    bad |= check("  window 64", A, tiny_window, rng);
    bad |= check("  tiles 8 rows / 40 entries", A, tiny_tiles, rng);
  }
  {
    std::vector<int32_t> perm = identity(2000);
    std::shuffle(perm.begin(), perm.end(), rng); // random numbering: few, wide levels and long-range dependencies
    const Csr A = grid_pattern(20, 10, 10, 1, perm, rng);
    bad |= check("box 20x10x10, shuffled", A, standard, rng);
    bad |= check("  window 64", A, tiny_window, rng);
    bad |= check("  tiles 8 rows / 40 entries", A, tiny_tiles, rng);
  }
  {
    const Csr A = grid_pattern(40, 1, 1, 2, identity(40), rng); // a chain: one row per level
    bad |= check("chain of 40, reach 2", A, standard, rng);
    const Csr D = grid_pattern(50, 1, 1, 0, identity(50), rng); // diagonal matrix: one level, no entries
    bad |= check("diagonal of 50", D, tiny_tiles, rng);
  }
  for (int round = 0; round < 3; ++round)
    {
      const Csr A = random_band(1500 + 371 * round, 150 + 400 * round, 6 + 10 * round, rng);
      bad |= check("random band", A, standard, rng);
      bad |= check("  window 64", A, tiny_window, rng);
      bad |= check("  tiles 8 rows / 40 entries", A, tiny_tiles, rng);
    }
  try
    {
      s35::WavefrontOptions wrong;
      wrong.window = 128; // fewer slots than rows of a tile
      const Csr A = grid_pattern(10, 10, 1, 1, identity(100), rng);
      s35::wavefront_pack(A.n, A.rowptr.data(), A.col.data(), A.diag.data(), true, nullptr, wrong);
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
  std::printf("wavefront: all schedules reproduce the substitutions bit for bit\n");
  return 0;
}
