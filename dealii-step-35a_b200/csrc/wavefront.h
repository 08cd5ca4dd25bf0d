// Level-sequential ("wavefront") schedule of a sparse triangular sweep for ONE thread block, host side: the packed
// stream the kernel k_wave_sweep (ilu.cu) consumes, and an executor of the same stream on the CPU with which the
// packing is tested without a GPU (tests/cpp/test_wavefront.cpp).  No CUDA in this file.
//
// Why: the substitutions of SparseILU::vmult have thousands of dependency levels of ~100 rows each.  A sweep that
// spreads the rows of a level over many SMs pays a trip through L2 per level (measured on B200: ~490 cycles per
// hand-off, tools/microbench/latency.cu).  Here one block walks the levels in order, keeps the most recent `window`
// solution values in shared memory, and hands a level over with a named barrier (~110 cycles).
//
// Order of the rows ("positions"): by dependency level; inside a level by decreasing numbers of entry chunks, then by
// row index.  A TILE is a run of at most group_rows rows of ONE level whose packed size is at most max_tile_bytes.  The
// block has `groups` warp groups of group_rows threads; tile t belongs to group t % groups, lane l of warp w of the
// group owns row 32 w + l of the tile.
//
// Pipelining: a row's entries are split by the age of the value they read.  OLD entries read values of tiles
// <= t - groups - 1; those are complete when the group starts on tile t (it has just finished tile t - groups and
// waited for tile t - groups - 1 then), so they can be summed ahead of time, off the critical path.  FRESH entries
// (tiles t - groups .. t - 1) are added after the wait for tile t - 1.  Per row the sum runs over the old entries
// in ascending column order, then over the fresh ones in ascending column order, in one accumulator: a fixed
// order, so results do not depend on timing; it is a re-association of the serial algorithm's sum.
//
// What the format is shaped by (all measured on B200, profiles/ of round 2):
//  * one SM is bound by instruction issue long before bandwidth: a thread's entries therefore come in CHUNKS of four
//    (two 16-byte value loads and one 8-byte slot load serve four products), and the chunks of a WARP's 32 rows lie
//    side by side -- chunk c of the warp is one 1280-byte block (32 x 8 bytes of slots, 32 x 16 of values 0-1, 32 x
//    16 of values 2-3) -- so that every address is "warp base + lane * size + constant" and the chunk counts are the
//    same for the whole warp: no per-row index arithmetic, no divergence;
//  * a bulk copy whose ends are not 128-byte aligned is slow, and only > 100 KB in flight keep the stream going, so
//    tiles are 128-byte aligned and travel back to back through a byte ring (not through equal stages).
// Per warp, chunks [0, n_pre) hold the rows' old entries and are summed ahead; chunks [n_pre, n_chunks) hold their
// fresh entries; rows with fewer entries than the warp's counts are padded (value 0.0, slot 0; the window starts as
// zeros, so a padding product is 0 * finite).
// The few old entries whose value has left the window ("far": its position is more than `window` behind the END
// of the reading tile, so a concurrently running row may overwrite its slot) are kept out of the chunks, in a list
// per tile that the owning thread walks ahead of time.
//
//  * a global load in a consumer warp stalls it for the load's whole latency at the next shared-memory result it
//    needs (the loads share scoreboard slots), however far ahead the load was issued: the consumers therefore read
//    NOTHING from global memory.  The right-hand sides travel through the ring like the factor: they are kept in
//    "rhs layout" (tile after tile, each tile's rows in order, padded to 128 bytes) and copied in behind the tile.
//    The forward sweep's input is brought into that layout by a small kernel before the sweep; the forward sweep
//    writes its results straight into the backward sweep's layout.
//
// Packed tile (byte offsets from the tile's start; tile start and size are multiples of 128):
//   header   16 x int32   [0] first position  [1] rows  [2] warps  [3] n_far  [12] size of the packed tile = offset of
//                         the right-hand sides in the ring      [4 + w] warp w: (block offset / 128) << 16 | n_pre << 8 | n_chunks
//   far      n_far x 16 bytes   int32 row in tile, int32 position, double value
//   warp block (128-byte aligned), for each warp:
//       forward sweep:  32 int32: where the row's result goes in the backward sweep's rhs layout
//       backward sweep: 32 doubles inverse pivots, 32 int32 natural row indices
//       n_chunks x { 32 x 4 uint16 slots (byte offsets into the window) | 32 x 2 doubles | 32 x 2 doubles }
// In the ring a tile is followed by its right-hand sides (rows doubles, padded to a multiple of 16).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <vector>

namespace s35 {

struct WavefrontOptions
{
  int group_rows = 192;           // threads of a warp group = most rows of a tile; a multiple of 32, at most 256
  int groups = 4;                 // warp groups (>= 2)
  int window = 8192;              // solution values kept in shared memory; a power of two, <= 8192 (16-bit byte offsets)
  int ring_bytes = 160 * 1024;    // shared-memory ring the tiles travel through; a multiple of 128, <= 256 KB: all that
                                  // is left beside the window (measured, pressure factor r1: 112 KB 1.34 ms per application,
                                  // 128 KB 1.23, 144 KB 1.19, 160 KB 1.15 -- the consumers wait for data a fifth of the time)
  int max_tile_bytes = 48 * 1024; // bound on a packed tile (several tiles must fit the ring)
  int slots = 32;                 // tiles in flight at most (one pair of mbarriers each)
};

struct WavefrontPlan
{
  int32_t n = 0;
  bool forward = true;
  WavefrontOptions opt;
  int32_t n_tiles = 0, n_levels = 0;
  std::vector<unsigned char> blob;   // all tiles back to back
  std::vector<int32_t> directory;    // per tile: blob offset / 16, ring offset / 16, bytes, tile that must be finished before
                                     // this one may be loaded over it (-1: none)
  std::vector<int32_t> rhs_dir;      // per tile: offset (in doubles) and size (in bytes) of its right-hand sides in rhs layout
  std::vector<int32_t> rhs_slot;     // per ROW: where its right-hand side sits in this sweep's rhs layout
  int32_t rhs_len = 0;               // length of a vector in rhs layout
  std::vector<int32_t> value_dst;    // blob offset / 8 of every factor value in the stream ...
  std::vector<int32_t> value_src;    // ... and its index in the factor's value array
  std::vector<int32_t> position;     // position of every row in this sweep's order
  int64_t entries = 0, far_entries = 0, fresh_entries = 0, stored_entries = 0, post_chunks = 0, warp_tiles = 0;
  int32_t max_tile = 0, max_post_chunks = 0; // largest tile in bytes; most chunks a warp sums after the wait
};

inline int32_t round_up(int32_t x, int32_t m) { return (x + m - 1) / m * m; }

constexpr int32_t kWaveChunkBytes = 1280;      // one chunk of one warp
constexpr int32_t kWaveBackwardHead = 384;     // pivots and natural row indices of one warp (backward sweep)
constexpr int32_t kWaveForwardHead = 128;      // where the rows' results go in the backward sweep's rhs layout

struct WaveFar
{
  int32_t row, position;
  double value;
};

// rowptr/col: CSR with ascending columns; diag[r]: position of the diagonal entry of row r.
// forward: the strictly lower part (unit diagonal); backward: the strictly upper part and the pivot at diag[r].
// next_slot (forward sweep only): for every row, where its result goes in the rhs layout of the sweep that follows
// (the backward plan's rhs_slot; pack the backward sweep first).
inline WavefrontPlan wavefront_pack(int32_t n, const int32_t *rowptr, const int32_t *col, const int32_t *diag, bool forward,
                                    const int32_t *next_slot, const WavefrontOptions &opt)
{
  if (opt.window <= 0 || (opt.window & (opt.window - 1)) != 0 || opt.window > 8192)
    throw std::invalid_argument("wavefront: the window must be a power of two, at most 8192");
  if (opt.group_rows <= 0 || opt.group_rows % 32 != 0 || opt.group_rows > 256 || opt.groups < 2)
    throw std::invalid_argument("wavefront: a warp group is a multiple of 32 threads (at most 256), and there are at least two groups");
  if ((int64_t)opt.group_rows * (opt.groups + 1) > opt.window)
    throw std::invalid_argument("wavefront: the window must hold the rows of groups + 1 tiles"); // fresh values are near
  if (opt.ring_bytes % 128 != 0 || opt.ring_bytes > 256 * 1024 || opt.max_tile_bytes + 8 * opt.group_rows > opt.ring_bytes / 2 || opt.slots < opt.groups + 1)
    throw std::invalid_argument("wavefront: bad ring / tile / slot sizes");
  WavefrontPlan P;
  P.n = n;
  P.forward = forward;
  P.opt = opt;
  auto first = [&](int32_t r) { return forward ? rowptr[r] : diag[r] + 1; };
  auto last = [&](int32_t r) { return forward ? diag[r] : rowptr[r + 1]; };
  auto len_of = [&](int32_t r) { return last(r) - first(r); };
  const int32_t warp_head = forward ? kWaveForwardHead : kWaveBackwardHead;
  if (forward && !next_slot && n > 0)
    throw std::invalid_argument("wavefront: the forward sweep needs the layout of the sweep that reads its result");

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

  // Rows by (level, chunks of old entries, chunks of fresh entries, row index), larger first: a warp's 32 rows are
  // padded to the warp's largest counts, so rows with equal counts should sit together.  Old and fresh are not known
  // exactly before the tiles are cut; here an entry counts as fresh if its value is produced less than
  // (groups + 1) * group_rows / 2 rows' worth of levels back -- an estimate that only affects the padding.
  std::vector<int32_t> order(n);
  for (int32_t r = 0; r < n; ++r)
    order[r] = r;
  std::vector<int32_t> level_rows(n_levels + 1, 0);
  for (int32_t r = 0; r < n; ++r)
    ++level_rows[level[r] + 1];
  for (int32_t l = 0; l < n_levels; ++l)
    level_rows[l + 1] += level_rows[l]; // rows in levels < l + 1
  std::vector<int32_t> key(n);
  {
    // the oldest level that can still hold fresh values for a row of level lv: walk back over about `groups` tiles
    std::vector<int32_t> fresh_from(n_levels, 0);
    for (int32_t lv = 0; lv < n_levels; ++lv)
      {
        int32_t l = lv, tiles = 0;
        while (l > 0 && tiles < opt.groups)
          {
            --l;
            tiles += (level_rows[l + 1] - level_rows[l] + opt.group_rows - 1) / opt.group_rows;
          }
        fresh_from[lv] = l;
      }
    for (int32_t r = 0; r < n; ++r)
      {
        int32_t n_fresh = 0;
        for (int32_t k = first(r); k < last(r); ++k)
          n_fresh += level[col[k]] >= fresh_from[level[r]];
        key[r] = (((len_of(r) - n_fresh + 3) / 4) << 8) | ((n_fresh + 3) / 4);
      }
  }
  std::sort(order.begin(), order.end(), [&](int32_t a, int32_t b) {
    if (level[a] != level[b])
      return level[a] < level[b];
    if (key[a] != key[b])
      return key[a] > key[b];
    return a < b;
  });
  P.position.assign(n, -1);
  for (int32_t p = 0; p < n; ++p)
    P.position[order[p]] = p;

  // tile boundaries first (an entry is old or fresh by the TILE of the value it reads); the size bound assumes the
  // worst: every entry in a chunk of its own warp's longest row, plus a far record
  std::vector<int32_t> tile_begin; // positions; one more entry at the end
  {
    int32_t p = 0;
    while (p < n)
      {
        const int32_t lv = level[order[p]];
        int32_t rows = 0, bytes = 64;
        while (p + rows < n && level[order[p + rows]] == lv && rows < opt.group_rows)
          {
            const int32_t r = order[p + rows];
            int32_t add = 0;
            for (int32_t k = first(r); k < last(r); ++k) // far records, at most (the tile ends at p + group_rows at the latest)
              add += P.position[col[k]] < p + opt.group_rows - opt.window ? 16 : 0;
            if (rows % 32 == 0) // a new warp: room for its longest row, old and fresh entries rounded up separately
              {
                int32_t longest = 0;
                for (int32_t k2 = p + rows; k2 < std::min(n, p + rows + 32) && level[order[k2]] == lv; ++k2)
                  longest = std::max(longest, len_of(order[k2]));
                add += 128 + warp_head + kWaveChunkBytes * ((longest + 3) / 4 + 1);
              }
            if (round_up(bytes + add, 128) > opt.max_tile_bytes)
              {
                if (rows == 0)
                  throw std::invalid_argument("wavefront: a row does not fit a tile");
                break;
              }
            bytes += add;
            ++rows;
          }
        tile_begin.push_back(p);
        p += rows;
      }
    tile_begin.push_back(n);
  }
  P.n_tiles = (int32_t)tile_begin.size() - 1;
  P.rhs_slot.assign(n, 0);
  std::vector<int32_t> tile_of(n);
  for (int32_t t = 0; t < P.n_tiles; ++t)
    for (int32_t p = tile_begin[t]; p < tile_begin[t + 1]; ++p)
      tile_of[p] = t;

  int32_t ring_cursor = 0;
  std::vector<int32_t> ring_begin(P.n_tiles), ring_end(P.n_tiles);
  int32_t oldest_live = 0; // tiles before this one have been overwritten in the ring by earlier decisions
  struct Entry
  {
    int32_t k, slot; // factor position; byte offset in the window
  };
  std::vector<std::vector<Entry>> near(opt.group_rows); // per row of the tile: old entries first, far ones left out
  std::vector<int32_t> n_old(opt.group_rows);
  std::vector<WaveFar> far;
  std::vector<int32_t> far_k;
  constexpr int32_t kMaxPlaces = 4 * 255;
  std::vector<std::vector<int32_t>> place_of(opt.group_rows);              // per row of the tile: place of entry j
  std::vector<char> placed((size_t)(opt.group_rows + 32) * kMaxPlaces, 0); // per row and place: taken by a real entry
  std::vector<int32_t> bank_load((size_t)kMaxPlaces * 16);                 // per place and bank pair: different values so far
  std::vector<std::vector<int32_t>> place_slots(kMaxPlaces);
  for (int32_t t = 0; t < P.n_tiles; ++t)
    {
      const int32_t p = tile_begin[t], tile_end = tile_begin[t + 1], rows = tile_end - p, warps = (rows + 31) / 32;
      far.clear();
      far_k.clear();
      for (int32_t i = 0; i < rows; ++i)
        {
          const int32_t r = order[p + i];
          P.rhs_slot[r] = P.rhs_len + i;
          near[i].clear();
          for (int pass = 0; pass < 2; ++pass) // old entries, then fresh ones; ascending columns in both
            {
              for (int32_t k = first(r); k < last(r); ++k)
                {
                  const int32_t q = P.position[col[k]];
                  if (q < 0 || q >= p)
                    throw std::logic_error("wavefront: a dependency is not in an earlier tile");
                  const bool old = tile_of[q] <= t - opt.groups - 1;
                  if (old != (pass == 0))
                    continue;
                  if (q < tile_end - opt.window)
                    {
                      if (!old)
                        throw std::logic_error("wavefront: a fresh value outside the window");
                      far.push_back(WaveFar{i, q, 0.0});
                      far_k.push_back(k);
                    }
                  else
                    near[i].push_back(Entry{k, 8 * (q & (opt.window - 1))});
                }
              if (pass == 0)
                n_old[i] = (int32_t)near[i].size();
            }
          P.entries += len_of(r);
          P.fresh_entries += (int64_t)near[i].size() - n_old[i];
        }
      P.far_entries += (int64_t)far.size();
      // warp blocks
      int32_t cursor = round_up(64 + 16 * (int32_t)far.size(), 128);
      int32_t warp_word[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      int32_t warp_at[8];
      for (int32_t w = 0; w < warps; ++w)
        {
          // old entries fill the warp's first n_pre chunks, fresh ones the n_post chunks after them (each row padded
          // with zeros to the warp's counts): after the wait a warp sums fresh entries only
          int32_t n_pre = 0, n_post = 0;
          const int32_t lanes = std::min(32, rows - 32 * w);
          for (int32_t l = 0; l < lanes; ++l)
            {
              n_pre = std::max(n_pre, (n_old[32 * w + l] + 3) / 4);
              n_post = std::max(n_post, ((int32_t)near[32 * w + l].size() - n_old[32 * w + l] + 3) / 4);
            }
          if (n_pre + n_post > 255)
            throw std::invalid_argument("wavefront: a row has more than 255 chunks");
          warp_at[w] = cursor;
          warp_word[w] = ((cursor / 128) << 16) | (n_pre << 8) | (n_pre + n_post);
          cursor = round_up(cursor + warp_head + kWaveChunkBytes * (n_pre + n_post), 128);
          // Where a row's entries sit in its chunks.  The 16 lanes of a half-warp read the window together, one place at
          // a time; two different values in the same pair of banks cost an extra trip.  Old entries may take any place
          // of the first n_pre chunks, fresh ones any place of the chunks after them: lane by lane, entry by entry, the
          // free place with the fewest clashes so far.  (The order of a row's sum is the order of its places: fixed here.)
          if (4 * (n_pre + n_post) > kMaxPlaces)
            throw std::invalid_argument("wavefront: a row has too many chunks");
          for (int32_t half = 0; half < 2; ++half)
            {
              std::fill(bank_load.begin(), bank_load.end(), 0);
              for (auto &v : place_slots)
                v.clear();
              for (int32_t l = 16 * half; l < std::min(lanes, 16 * half + 16); ++l)
                {
                  const int32_t i = 32 * w + l;
                  place_of[i].assign(near[i].size(), -1);
                  std::fill(placed.begin() + (size_t)i * kMaxPlaces, placed.begin() + (size_t)(i + 1) * kMaxPlaces, (char)0);
                  for (size_t j = 0; j < near[i].size(); ++j)
                    {
                      const bool old = (int32_t)j < n_old[i];
                      const int32_t lo = old ? 0 : 4 * n_pre, hi = old ? 4 * n_pre : 4 * (n_pre + n_post);
                      const int32_t slot = near[i][j].slot, bank = (slot / 8) % 16;
                      int32_t best = -1, best_cost = 1 << 30;
                      for (int32_t place = lo; place < hi; ++place)
                        {
                          if (placed[(size_t)i * kMaxPlaces + place])
                            continue;
                          const auto &seen = place_slots[place];
                          const int32_t cost = std::find(seen.begin(), seen.end(), slot) != seen.end() ? 0 : bank_load[place * 16 + bank];
                          if (cost < best_cost)
                            {
                              best_cost = cost;
                              best = place;
                            }
                        }
                      if (best < 0)
                        throw std::logic_error("wavefront: no place left for an entry");
                      place_of[i][j] = best;
                      placed[(size_t)i * kMaxPlaces + best] = 1;
                      auto &seen = place_slots[best];
                      if (std::find(seen.begin(), seen.end(), slot) == seen.end())
                        {
                          seen.push_back(slot);
                          ++bank_load[best * 16 + bank];
                        }
                    }
                }
            }
          for (int32_t l = lanes; l < 32; ++l) // lanes without a row: nothing placed
            std::fill(placed.begin() + (size_t)(32 * w + l) * kMaxPlaces, placed.begin() + (size_t)(32 * w + l + 1) * kMaxPlaces, (char)0);
          P.post_chunks += n_post;
          P.max_post_chunks = std::max(P.max_post_chunks, n_post);
          P.stored_entries += (int64_t)128 * (n_pre + n_post);
          ++P.warp_tiles;
        }
      // The cut above assumed that a warp's longest run of old entries and its longest run of fresh entries belong to
      // the same row; when they do not, the tile comes out larger than max_tile_bytes.  That is fine as long as two
      // tiles still fit the ring; beyond that this factor is not for the wavefront sweep.
      const int32_t bytes = round_up(cursor, 128);
      if (bytes + 8 * round_up(rows, 16) > opt.ring_bytes / 2)
        throw std::invalid_argument("wavefront: a tile does not fit half the ring");
      const int64_t blob_begin = (int64_t)P.blob.size();
      if (blob_begin + bytes > (int64_t)INT32_MAX * 8)
        throw std::invalid_argument("wavefront: stream exceeds 32-bit offsets");
      P.blob.resize(blob_begin + bytes, 0);
      unsigned char *base = P.blob.data() + blob_begin;
      int32_t *h = reinterpret_cast<int32_t *>(base);
      h[0] = p;
      h[1] = rows;
      h[2] = warps;
      h[3] = (int32_t)far.size();
      h[12] = bytes;
      for (int32_t w = 0; w < 8; ++w)
        h[4 + w] = warp_word[w];
      for (size_t f = 0; f < far.size(); ++f)
        {
          std::memcpy(base + 64 + 16 * f, &far[f], 16);
          P.value_dst.push_back((int32_t)((blob_begin + 64 + 16 * (int64_t)f + 8) / 8));
          P.value_src.push_back(far_k[f]);
        }
      for (int32_t w = 0; w < warps; ++w)
        {
          unsigned char *wb = base + warp_at[w];
          const int32_t lanes = std::min(32, rows - 32 * w);
          for (int32_t l = 0; l < lanes; ++l)
            {
              const int32_t i = 32 * w + l, r = order[p + i];
              if (forward)
                reinterpret_cast<int32_t *>(wb)[l] = next_slot[r];
              else
                {
                  reinterpret_cast<int32_t *>(wb + 256)[l] = r;
                  P.value_dst.push_back((int32_t)((blob_begin + warp_at[w] + 8 * l) / 8));
                  P.value_src.push_back(diag[r]);
                }
              const int32_t n_pre = (warp_word[w] >> 8) & 0xff;
              for (size_t j = 0; j < near[i].size(); ++j)
                {
                  const int32_t place = place_of[i][j];
                  const int32_t c = place / 4, u = place % 4;
                  unsigned char *cb = wb + warp_head + kWaveChunkBytes * c;
                  reinterpret_cast<uint16_t *>(cb)[4 * l + u] = (uint16_t)near[i][j].slot;
                  const int32_t o_val = warp_at[w] + warp_head + kWaveChunkBytes * c + (u < 2 ? 256 : 768) + 16 * l + 8 * (u & 1);
                  P.value_dst.push_back((int32_t)((blob_begin + o_val) / 8));
                  P.value_src.push_back(near[i][j].k);
                }
              (void)n_pre;
            }
          // a padding entry (value 0.0) reads what a real entry at the same place of its half-warp reads: a broadcast,
          // no extra trip through the banks; where the half-warp has no real entry there, all its lanes read slot 0
          const int32_t n_places = 4 * (warp_word[w] & 0xff);
          for (int32_t half = 0; half < 2; ++half)
            for (int32_t place = 0; place < n_places; ++place)
              {
                uint16_t *slots = reinterpret_cast<uint16_t *>(wb + warp_head + kWaveChunkBytes * (place / 4));
                int32_t shared_slot = -1;
                for (int32_t l = 16 * half; l < 16 * half + 16 && shared_slot < 0; ++l)
                  if (placed[(32 * w + l) * kMaxPlaces + place])
                    shared_slot = slots[4 * l + place % 4];
                if (shared_slot >= 0)
                  for (int32_t l = 16 * half; l < 16 * half + 16; ++l)
                    if (!placed[(32 * w + l) * kMaxPlaces + place])
                      slots[4 * l + place % 4] = (uint16_t)shared_slot;
              }
        }
      P.max_tile = std::max(P.max_tile, bytes);

      // place tile and right-hand sides in the ring; the tiles they overwrite must be finished before the load
      const int32_t rhs_bytes = 8 * round_up(rows, 16);
      if (ring_cursor + bytes + rhs_bytes > opt.ring_bytes)
        ring_cursor = 0;
      ring_begin[t] = ring_cursor;
      ring_end[t] = ring_cursor + bytes + rhs_bytes;
      ring_cursor += bytes + rhs_bytes;
      P.rhs_dir.push_back(P.rhs_len);
      P.rhs_dir.push_back(rhs_bytes);
      P.rhs_len += rhs_bytes / 8;
      int32_t need = -1;
      for (int32_t u = oldest_live; u < t; ++u)
        if (ring_begin[u] < ring_end[t] && ring_begin[t] < ring_end[u])
          need = u;
      if (need >= 0)
        oldest_live = need + 1;
      need = std::max(need, t - opt.slots); // and its pair of barriers must be free
      P.directory.push_back((int32_t)(blob_begin / 16));
      P.directory.push_back(ring_begin[t] / 16);
      P.directory.push_back(bytes);
      P.directory.push_back(need);
    }
  return P;
}

// writes the factor's values into the stream
inline void wavefront_fill_values(WavefrontPlan &P, const double *lu)
{
  double *slots = reinterpret_cast<double *>(P.blob.data());
  for (size_t s = 0; s < P.value_dst.size(); ++s)
    slots[P.value_dst[s]] = lu[P.value_src[s]];
}

// The kernel's walk over the stream, on the CPU, with the kernel's timing made adversarial: a tile is copied into
// the ring (a real buffer of ring_bytes) as EARLY as the directory allows, the part of tile u that is summed ahead
// is summed as early as the kernel may do it -- when only tiles <= u - groups - 1 are finished; every window slot
// written by a younger tile is poisoned (NaN) for the duration -- and a tile must not read its own slots.  A wrong
// `need`, a window slot overwritten too early or an entry classified old too early shows up as a wrong result.
// rhs: the right-hand sides in this plan's rhs layout; by_position should come in as NaN; next_rhs (forward sweep):
// the vector in the next sweep's rhs layout that receives the results; natural (backward sweep): results by row.
inline void wavefront_apply_host(const WavefrontPlan &P, const double *rhs, double *by_position, double *natural, double *next_rhs)
{
  const WavefrontOptions &o = P.opt;
  const int32_t warp_head = P.forward ? kWaveForwardHead : kWaveBackwardHead;
  std::vector<double> window(o.window, 0.0);
  std::vector<unsigned char> ring(o.ring_bytes, 0xA5);
  const int G = o.groups;
  std::vector<double> acc_old((size_t)(G + 1) * o.group_rows);
  std::vector<int32_t> first_pos(P.n_tiles + 1, P.n);
  {
    int32_t p = 0;
    for (int32_t t = 0; t < P.n_tiles; ++t)
      {
        first_pos[t] = p;
        p += reinterpret_cast<const int32_t *>(P.blob.data() + (size_t)P.directory[4 * (size_t)t] * 16)[1];
      }
  }
  int32_t next_load = 0, next_prepare = 0;
  auto tile_at = [&](int32_t t) { return ring.data() + (size_t)P.directory[4 * (size_t)t + 1] * 16; };
  // chunks [c0, c1) of lane l of a warp block added to acc
  auto chunks = [&](const unsigned char *wb, int32_t l, int32_t c0, int32_t c1, double acc, const std::vector<double> &win) {
    for (int32_t c = c0; c < c1; ++c)
      {
        const unsigned char *cb = wb + warp_head + kWaveChunkBytes * c;
        for (int u = 0; u < 4; ++u)
          {
            const uint32_t off = reinterpret_cast<const uint16_t *>(cb)[4 * l + u];
            const double v = reinterpret_cast<const double *>(cb + (u < 2 ? 256 : 768))[2 * l + (u & 1)];
            if (v != 0.0) // a padding entry (0.0 * slot 0) contributes nothing; the kernel's window holds no NaN to spoil it
              acc = std::fma(v, win[off / 8], acc);
          }
      }
    return acc;
  };
  auto progress = [&](int32_t finished) { // tiles < finished are complete
    while (next_load < P.n_tiles && P.directory[4 * (size_t)next_load + 3] < finished)
      {
        const int32_t *d = P.directory.data() + 4 * (size_t)next_load, *rd = P.rhs_dir.data() + 2 * (size_t)next_load;
        std::memcpy(ring.data() + (size_t)d[1] * 16, P.blob.data() + (size_t)d[0] * 16, d[2]);
        std::memcpy(ring.data() + (size_t)d[1] * 16 + d[2], rhs + rd[0], rd[1]);
        ++next_load;
      }
    while (next_prepare < next_load && next_prepare <= finished + G)
      {
        const int32_t u = next_prepare;
        // what the kernel may rely on: tiles <= u - G - 1.  Poison everything younger that is already in the window.
        std::vector<double> poisoned = window;
        for (int32_t q = first_pos[std::max(0, u - G)]; q < first_pos[finished]; ++q)
          poisoned[q & (o.window - 1)] = std::nan("");
        const unsigned char *T = tile_at(u);
        const int32_t *h = reinterpret_cast<const int32_t *>(T);
        const int32_t rows = h[1], n_far = h[3];
        if (h[0] != first_pos[u] || h[12] != P.directory[4 * (size_t)u + 2])
          throw std::logic_error("wavefront: directory and tile header disagree");
        for (int32_t i = 0; i < rows; ++i)
          {
            const int32_t word = h[4 + i / 32];
            const unsigned char *wb = T + (size_t)((uint32_t)word >> 16) * 128;
            double acc = chunks(wb, i % 32, 0, (word >> 8) & 0xff, 0.0, poisoned);
            for (int32_t f = 0; f < n_far; ++f)
              {
                WaveFar e;
                std::memcpy(&e, T + 64 + 16 * (size_t)f, 16);
                if (e.row == i)
                  acc = std::fma(e.value, by_position[e.position], acc);
              }
            acc_old[(size_t)(u % (G + 1)) * o.group_rows + i] = acc;
          }
        ++next_prepare;
      }
  };
  progress(0);
  for (int32_t t = 0; t < P.n_tiles; ++t)
    {
      if (next_prepare <= t)
        throw std::logic_error("wavefront: the schedule is stuck (a tile waits for ring space held by a later one)");
      const unsigned char *T = tile_at(t);
      const int32_t *h = reinterpret_cast<const int32_t *>(T);
      const int32_t rows = h[1];
      const double *tile_rhs = reinterpret_cast<const double *>(T + h[12]);
      // a value of this tile's own slots must not be read by the tile: poison them first
      std::vector<double> win = window;
      for (int32_t i = 0; i < rows; ++i)
        win[(h[0] + i) & (o.window - 1)] = std::nan("");
      std::vector<double> res(rows);
      for (int32_t i = 0; i < rows; ++i)
        {
          const int32_t word = h[4 + i / 32];
          const unsigned char *wb = T + (size_t)((uint32_t)word >> 16) * 128;
          const double acc = chunks(wb, i % 32, (word >> 8) & 0xff, word & 0xff, acc_old[(size_t)(t % (G + 1)) * o.group_rows + i], win);
          res[i] = tile_rhs[i] - acc;
          if (!P.forward)
            res[i] *= reinterpret_cast<const double *>(wb)[i % 32];
        }
      for (int32_t i = 0; i < rows; ++i)
        {
          const int32_t pos = h[0] + i;
          const unsigned char *wb = T + (size_t)((uint32_t)h[4 + i / 32] >> 16) * 128;
          window[pos & (o.window - 1)] = res[i];
          by_position[pos] = res[i];
          if (P.forward)
            next_rhs[reinterpret_cast<const int32_t *>(wb)[i % 32]] = res[i];
          else
            natural[reinterpret_cast<const int32_t *>(wb + 256)[i % 32]] = res[i];
        }
      progress(t + 1);
    }
}

} // namespace s35
