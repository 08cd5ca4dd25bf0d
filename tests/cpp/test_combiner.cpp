// CPU test of the collective combiner's protocol (dealii-step-35a_b200/csrc/combiner.h): threads stand in for
// ranks and for the component solvers of each rank; the mock transport checks that every rank issues the same
// operations in the same order and computes the sums.  Exit code 0 = all runs consistent, no dead-lock.
//
//   g++ -std=c++17 -O1 -pthread -I dealii-step-35a_b200/csrc tests/cpp/test_combiner.cpp -o test_combiner
#include "combiner.h"

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <random>
#include <thread>

namespace {

struct Signature
{
  int party, kind;
  long tag;
  bool operator==(const Signature &o) const { return party == o.party && kind == o.kind && tag == o.tag; }
};

// the "network": the i-th operation issued by every rank meets the i-th of all others
struct World
{
  explicit World(int n) : n_ranks(n) {}
  int n_ranks;
  std::mutex m;
  std::condition_variable cv;
  struct Meeting
  {
    int arrived = 0, departed = 0;
    double sum = 0.0;
    Signature first{};
  };
  std::map<long, Meeting> meetings;
  std::atomic<bool> failed{false};

  double meet(long index, const Signature &sig, double value)
  {
    std::unique_lock<std::mutex> lock(m);
    Meeting &mt = meetings[index];
    if (mt.arrived == 0)
      mt.first = sig;
    else if (!(mt.first == sig))
      {
        failed = true;
        cv.notify_all();
        throw std::runtime_error("ranks issued different operations at the same position");
      }
    mt.sum += value;
    ++mt.arrived;
    cv.notify_all();
    if (!cv.wait_for(lock, std::chrono::seconds(20), [&] { return mt.arrived == n_ranks || failed.load(); }) || failed)
      {
        failed = true;
        cv.notify_all();
        throw std::runtime_error("dead-lock: an operation was not matched by all ranks");
      }
    const double result = mt.sum;
    if (++mt.departed == n_ranks)
      meetings.erase(index);
    return result;
  }
};

struct Op
{
  int party = 0, kind = 0;
  long tag = 0;
  double *value = nullptr;
};

struct MockBackend
{
  using Stream = int;
  using Event = int;
  using Op = ::Op;
  World &world;
  long issued = 0;
  int open_groups = 0;
  long rounds = 0, largest_round = 0, in_this_round = 0;
  explicit MockBackend(World &w) : world(w) {}
  Stream comm_stream() { return 99; }
  Event party_event(int party) { return party; }
  Event round_event(int parity) { return 10 + parity; }
  void record(Event, Stream) {}
  void wait(Stream, Event) {}
  void group_begin()
  {
    if (open_groups++ != 0)
      throw std::runtime_error("nested group");
    in_this_round = 0;
  }
  void issue(const Op &op, Stream s)
  {
    if (open_groups != 1 || s != 99)
      throw std::runtime_error("operation issued outside a group or on a wrong stream");
    *op.value = world.meet(issued++, Signature{op.party, op.kind, op.tag}, *op.value);
    ++in_this_round;
  }
  void group_end()
  {
    --open_groups;
    ++rounds;
    largest_round = std::max(largest_round, in_this_round);
  }
};

void nap(std::mt19937 &rng, int max_us)
{
  const int us = (int)(rng() % (unsigned)(max_us + 1));
  if (us > 0)
    std::this_thread::sleep_for(std::chrono::microseconds(us));
}

// one rank: `steps` times { all parties enter; each runs its own number of iterations; leaves }
void run_rank(World &world, int rank, int n_parties, int steps, const std::vector<std::vector<int>> &iterations, unsigned seed,
              std::atomic<long> &mismatches, long &rounds, long &largest)
{
  MockBackend be(world);
  s35::Combiner<MockBackend> comb(be, n_parties);
  for (int step = 0; step < steps; ++step)
    {
      comb.enter_all();
      std::vector<std::thread> parties;
      for (int p = 0; p < n_parties; ++p)
        parties.emplace_back([&, p, step] {
          std::mt19937 rng(seed * 7919u + rank * 104729u + p * 1299709u + step);
          try
            {
              for (int it = 0; it < iterations[step][p]; ++it)
                for (int j = 0; j < it % 4 + 2; ++j) // a ghost exchange, then a growing number of sums
                  {
                    nap(rng, 200);
                    const long tag = ((long)step << 40) | ((long)it << 20) | j;
                    double v = (rank + 1) * 1000.0 + p * 10.0 + j;
                    comb.submit(p, Op{p, j == 0 ? 1 : 0, tag, &v}, /*stream*/ p);
                    double expect = 0.0;
                    for (int r = 0; r < world.n_ranks; ++r)
                      expect += (r + 1) * 1000.0 + p * 10.0 + j;
                    if (v != expect)
                      ++mismatches;
                  }
              nap(rng, 300);
              comb.leave(p);
            }
          catch (const std::exception &e)
            {
              std::fprintf(stderr, "rank %d party %d: %s\n", rank, p, e.what());
              ++mismatches;
              world.failed = true;
              world.cv.notify_all();
            }
        });
      for (auto &t : parties)
        t.join();
      if (world.failed)
        break;
    }
  rounds = be.rounds;
  largest = be.largest_round;
}

} // namespace

int main(int argc, char **argv)
{
  const int runs = argc > 1 ? std::atoi(argv[1]) : 20;
  long total_rounds = 0;
  for (int run = 0; run < runs; ++run)
    {
      const int n_ranks = 2 + run % 3, n_parties = 3, steps = 3;
      std::mt19937 rng(1234u + run);
      // iteration counts differ between the parties (they leave at different times) but not between the ranks
      std::vector<std::vector<int>> iterations(steps, std::vector<int>(n_parties));
      for (auto &s : iterations)
        for (int &n : s)
          n = (int)(rng() % 7); // 0 iterations: the party leaves at once
      World world(n_ranks);
      std::atomic<long> mismatches{0};
      std::vector<long> rounds(n_ranks, 0), largest(n_ranks, 0);
      std::vector<std::thread> ranks;
      for (int r = 0; r < n_ranks; ++r)
        ranks.emplace_back(run_rank, std::ref(world), r, n_parties, steps, std::cref(iterations), (unsigned)run, std::ref(mismatches),
                           std::ref(rounds[r]), std::ref(largest[r]));
      for (auto &t : ranks)
        t.join();
      if (world.failed || mismatches != 0)
        {
          std::fprintf(stderr, "run %d FAILED (%ld mismatches)\n", run, mismatches.load());
          return 1;
        }
      for (int r = 1; r < n_ranks; ++r)
        if (rounds[r] != rounds[0])
          {
            std::fprintf(stderr, "run %d: ranks formed different numbers of rounds\n", run);
            return 1;
          }
      if (largest[0] > n_parties)
        return 1;
      total_rounds += rounds[0];
    }
  std::printf("combiner: %d runs consistent, %ld rounds per rank in total\n", runs, total_rounds);
  return 0;
}
