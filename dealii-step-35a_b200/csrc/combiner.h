// Collective combiner: lets several solver threads of ONE rank share ONE communicator.
//
// The velocity components are solved concurrently, each by a host thread with a stream of its own.  On a mesh
// partition every one of them needs collectives (ghost exchanges, sums of Krylov scalars).  A communicator serves
// one issue order only, and separate communicators per component can block each other on the device in different
// orders on different ranks.  The combiner keeps one communicator and one issue order:
//
//   * a party (component thread) hands its next collective to submit() and waits;
//   * when every party that is still active has handed one in, the thread that completed the set issues all of
//     them together -- between group_begin() / group_end() on the communication stream, in party order -- and
//     wakes the others; every party's stream then waits for the round's completion event;
//   * a party that has finished its solve calls leave(); it is not waited for any more.
//
// Why every rank forms the same rounds: a party's sequence of collectives and the point where it leaves depend
// only on quantities that are identical on all ranks (iteration counts follow from all-reduced residuals).  A
// round consists of the next collective of every party that has not left; whether a leave() is seen before or
// after the others have submitted changes who issues the round, not what is in it.
//
// The class knows nothing about CUDA or NCCL: a Backend supplies streams, events and the issue of an operation,
// so that the protocol is tested on the CPU with threads standing in for ranks (tests/cpp/test_combiner.cpp).
//
//   struct Backend {
//     using Stream = ...;  using Event = ...;  using Op = ...;      // copyable handles
//     Stream comm_stream();
//     Event party_event(int party);     // one reusable event per party
//     Event round_event(int parity);    // two reusable events for round completion, used alternately
//     void record(Event, Stream);  void wait(Stream, Event);
//     void group_begin();  void issue(const Op &, Stream);  void group_end();
//   };
#pragma once
#include <condition_variable>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

namespace s35 {

template <class Backend>
class Combiner
{
public:
  using Stream = typename Backend::Stream;
  using Event = typename Backend::Event;
  using Op = typename Backend::Op;

  Combiner(Backend &backend, int n_parties) : be(backend), slots(n_parties) {}

  int n_parties() const { return (int)slots.size(); }

  // All parties become active.  Called by the thread that starts the parties, before it starts them.
  void enter_all()
  {
    std::lock_guard<std::mutex> lock(m);
    for (Slot &s : slots)
      {
        if (s.submitted)
          throw std::logic_error("combiner: enter_all() while a round is being collected");
        s.active = true;
      }
  }

  // Hands in the party's next collective; `s` is the stream the party's work before and after it runs on.
  // Returns when the round holding the operation has been issued and `s` has been told to wait for it.
  void submit(int party, const Op &op, Stream s)
  {
    be.record(be.party_event(party), s); // what the operation reads has been enqueued on s up to here
    std::unique_lock<std::mutex> lock(m);
    Slot &me = slots.at(party);
    if (!me.active || me.submitted)
      throw std::logic_error("combiner: submit() from a party that is not active or already waits");
    me.op = op;
    me.submitted = true;
    const long round = generation;
    if (round_complete())
      issue_round();
    else
      cv.wait(lock, [&] { return generation != round; });
    if (!failure.empty())
      throw std::runtime_error(failure);
    const Event done = be.round_event((int)(round & 1));
    lock.unlock();
    be.wait(s, done);
  }

  // The party takes no further part (its solve is over).  May complete the round the others wait for.
  void leave(int party)
  {
    std::lock_guard<std::mutex> lock(m);
    Slot &me = slots.at(party);
    if (me.submitted)
      throw std::logic_error("combiner: leave() while the party waits for a round");
    me.active = false;
    if (round_complete())
      issue_round();
  }

private:
  struct Slot
  {
    bool active = false, submitted = false;
    Op op{};
  };

  // called with the lock held
  bool round_complete() const
  {
    bool any = false;
    for (const Slot &s : slots)
      {
        if (s.active && !s.submitted)
          return false;
        any = any || s.submitted;
      }
    return any;
  }

  // called with the lock held, by the thread that completed the round
  void issue_round()
  {
    try
      {
        const Stream cs = be.comm_stream();
        for (size_t p = 0; p < slots.size(); ++p)
          if (slots[p].submitted)
            be.wait(cs, be.party_event((int)p));
        be.group_begin();
        for (Slot &s : slots)
          if (s.submitted)
            be.issue(s.op, cs);
        be.group_end();
        be.record(be.round_event((int)(generation & 1)), cs);
      }
    catch (const std::exception &e)
      {
        failure = e.what(); // every waiting party rethrows it
      }
    for (Slot &s : slots)
      s.submitted = false;
    ++generation;
    cv.notify_all();
  }

  Backend &be;
  std::mutex m;
  std::condition_variable cv;
  std::vector<Slot> slots;
  long generation = 0;
  std::string failure;
};

} // namespace s35
