// host_pack.h — host half of the PACKED ingest (peng_cuda_ingest_codes with host buffers).
//
// SequenceSet::readFASTA / Sequence hand the path one byte per base (src/shared/SequenceSet.cpp:285-447,
// src/shared/Alphabet.cpp:33-41).  Sent as they are, those bytes make the ingest PCIe-bound (1 B/base at ~46 GB/s).  The
// host threads therefore write the device layout of core_seq.cuh directly — 2-bit codes + 1-bit validity, two separator
// positions behind every sequence — into pinned staging buffers, and only 0.375 B/base cross the bus.
//
// Plain C++ interface (the AVX-512 body lives in host_pack.cpp, compiled by g++: nvcc's front end does not see it).
#pragma once
#include <stddef.h>
#include <stdint.h>

#include <vector>

namespace peng_hostpack {

// bits of a 64-position block that a worker covered only in part (range boundaries): OR-ed together by the caller
struct Partial {
  uint64_t blk, lo, hi, v;
};

// Implied validity: without letters other than ACGT the validity bits of a block follow from the offsets alone (the device
// rebuilds them, see k_valid_implied), so a phase only has to send the blocks where that is not true — those holding another
// letter, and the blocks at share boundaries (merged by the caller, listed unconditionally).
struct ValidExc {
  uint64_t blk, v;
};

// One worker's share: sequences [s_begin, s_end) of the set, written as whole 64-position blocks
//   codes64[2 * (blk - blk0)], codes64[2 * (blk - blk0) + 1]  (position q at bits 2 (q & 63) of the pair)
//   valid64[blk - blk0]                                        (position q at bit q & 63)
// for every block the share covers completely; incomplete first / last blocks go to `partials`.
// Stream position of base i of sequence s: front_pad + offsets[s] + 2 s + i.
// lead_zero : the share starts at position 0 with the front pad (first share of the set)
// q_end_pad : != 0: after the last sequence, zeros are appended up to this position (last share of the set)
struct Share {
  const uint8_t* src;
  const uint64_t* offsets;
  uint64_t s_begin, s_end;
  bool ascii;
  uint32_t front_pad;
  bool lead_zero;
  uint64_t q_end_pad;
  uint64_t blk0;
  uint64_t* codes64;
  uint64_t* valid64;
  std::vector<Partial>* partials;
  const uint8_t* const* seq_ptrs = nullptr;  // != NULL: sequence s is read from seq_ptrs[s] (per-sequence arrays) instead of src + offsets[s]
  std::vector<ValidExc>* exc = nullptr;      // != NULL: whole blocks whose validity word is not the implied one are listed here (AVX-512 path only)
};
void pack_share(const Share& sh);
bool have_avx512();
bool lists_exceptions();  // the path pack_share takes fills Share::exc (AVX-512 present and not switched off)

class Pool;
// Half-way packing for the part of a set the host does not pack completely: letter codes (0 = other, 1..4 = ACGT, anything
// else counts as other) as NIBBLES, dst[k] = nib(src[2 k]) | nib(src[2 k + 1]) << 4 — four vector instructions per 64 bases
// (near memcpy speed) and half the bytes over PCIe; the device unpacks them in k_pack<2> (core_seq.cuh: pack_word).
// n even or the last byte's high nibble 0.  Runs on the pool's threads.
void nibble_pack(Pool& pool, const uint8_t* src, uint64_t n, uint8_t* dst);

// minimal persistent worker pool: run(n, fn) calls fn(i) for i = 0 .. n-1 on the pool's threads (the caller takes part)
class Pool {
 public:
  explicit Pool(int n_threads);
  ~Pool();
  int size() const { return n_; }
  void run(int n_tasks, void (*fn)(int, void*), void* arg);
  // the same in two steps: start() wakes the workers and returns, finish() lets the caller take part and waits for the last
  // task — what the caller does in between runs beside the workers.  One job at a time: a second thread's start() waits
  // until the first has called finish(), so every start() must be followed by finish() on the same thread, and a thread
  // must not start a second job (or call run()) before it has finished its first
  void start(int n_tasks, void (*fn)(int, void*), void* arg);
  void finish();

 private:
  struct Impl;
  Impl* p_;
  int n_;
};
Pool& global_pool();  // sized by PENG_HOST_THREADS, else hardware threads / LOCAL_WORLD_SIZE, at most 32


// One phase of an ingest: sequences [s_begin, s_end) of the set are packed by the pool's threads into staging buffers that
// hold the blocks [blk0, blk1) (blk0 = first block the phase touches, blk1 = block holding the first position behind
// the phase, or the block count of the whole stream for the last phase).  A block shared with the previous phase arrives
// as `carry` (has_carry) and is merged; the incomplete block blk1 leaves the same way.
struct Phase {
  const uint8_t* src;
  const uint64_t* offsets;
  uint64_t s_begin, s_end;
  bool ascii;
  uint32_t front_pad;
  bool first, last;
  uint64_t q_total;  // last phase: zeros up to this position (multiple of 64)
  uint64_t* codes64;
  uint64_t* valid64;
  const uint8_t* const* seq_ptrs = nullptr;  // see Share
  // != NULL and the AVX-512 path runs: receives every block of [blk0, blk1) whose validity word differs from the implied one
  // (or may: the merged boundary blocks), sorted by block; *exc_complete says whether the list was made at all
  std::vector<ValidExc>* exc = nullptr;
  bool* exc_complete = nullptr;
};
void phase_blocks(const Phase& ph, uint64_t* blk0, uint64_t* blk1);
// end of the phase that starts with sequence s_begin: the largest s_end whose blocks (the carried one included) fit
// max_blocks; returns s_begin if not even one sequence fits (the caller falls back to the unpacked ingest)
uint64_t plan_phase(const uint64_t* offsets, uint64_t n_seq, uint64_t s_begin, uint32_t front_pad, bool first, uint64_t max_blocks);
void pack_phase(Pool& pool, const Phase& ph, bool* has_carry, Partial* carry);
// the same in two steps (the Phase must stay alive in between): begin() cuts the shares and starts the pool, end() joins the
// packing, merges the boundary blocks and hands the carry on.  Between the two the caller can queue the copies of the
// previous phase.
struct PhaseJob;
PhaseJob* pack_phase_begin(Pool& pool, const Phase& ph);
void pack_phase_end(Pool& pool, PhaseJob* job, bool* has_carry, Partial* carry);

}  // namespace peng_hostpack
