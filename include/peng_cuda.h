/* peng_cuda.h — C ABI of libpeng_cuda.so, the B200 (sm_100a) replacement for PEnG-motif's
 * data-parallel hot path.
 *
 * The reference (soedinglab/PEnG-motif) has no plugin / FFI interface: the path sits behind in-process
 * C++ class seams called from Peng::process (src/peng.cpp:322-435).  Every entry point below names the
 * reference seam it replaces (file:line relative to the reference tree); INTEGRATION.md shows the
 * binding a reference maintainer would add at each seam.
 *
 * Conventions
 *   - every function returns PENG_OK (0) or a negative PENG_E_* code and never exits the process;
 *     peng_cuda_last_error() gives the message (the reference prints to stderr and calls exit()).
 *   - the opaque handle owns all device memory; host buffers passed in/out are caller-owned.
 *   - one host thread per handle; one handle per GPU; all work of a handle is ordered on one CUDA stream.
 *   - encodings are the reference's own (src/base_pattern.h:24-29, src/iupac_pattern.h:26-30):
 *       sequence codes 0 = other, 1..4 = A,C,G,T (src/shared/Alphabet.cpp:33-41)
 *       base-pattern id = sum c_p * 4^p, position 0 least significant, A,C,G,T = 0..3
 *       IUPAC id        = sum c_p * 11^p, A,C,G,T,S,W,R,Y,M,K,N = 0..10
 *   - `precision`: PENG_F32REF reproduces the reference's float32 arithmetic operation by operation
 *     (bit-identical tables, sequential float32 accumulation order included); PENG_F64 computes the same
 *     quantities in double (compared at 1e-9 relative against the FP64 build of the reference).
 *   - there is no CPU fallback: without a CUDA device every call fails with PENG_E_CUDA.
 */
#ifndef PENG_CUDA_H_
#define PENG_CUDA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct peng_cuda_ctx* peng_handle_t;

enum {
  PENG_OK = 0,
  PENG_E_INVALID = -1, /* bad argument / call order */
  PENG_E_CUDA = -2,    /* CUDA runtime error (message holds cudaGetErrorString) */
  PENG_E_IO = -3,      /* cannot open / read FASTA (reference: exit(1), src/shared/SequenceSet.cpp:430-433) */
  PENG_E_FORMAT = -4,  /* malformed FASTA (reference: exit(1), src/shared/SequenceSet.cpp:361-373) */
  PENG_E_NOMEM = -5
};

enum { PENG_STRAND_PLUS = 0, PENG_STRAND_BOTH = 1 };          /* Strand, src/Global.h:13-15 */
enum { PENG_F32REF = 0, PENG_F64 = 1 };
enum {                                                         /* tables of BasePattern, src/base_pattern.h:116-120 */
  PENG_TABLE_BGPROB = 0,   /* pattern_bg_probabilities[order] */
  PENG_TABLE_EXPECTED = 1, /* expected_counts */
  PENG_TABLE_ZSCORE = 2,   /* pattern_zscore */
  PENG_TABLE_LOGP = 3      /* pattern_logp; PENG_F32REF: the one table that is NOT bit-identical to the reference — its
                            * log() is the device's, not glibc's (<= 4 differing float entries per 10^5, checked at float
                            * resolution); unused on the default path (MUTUAL_INFO), the host driver recomputes the logp of
                            * the selected seeds with glibc */
};

/* ---- lifetime ------------------------------------------------------------------------------------- */
int peng_cuda_device_count(void);
int peng_cuda_create(int device, peng_handle_t* out);
void peng_cuda_destroy(peng_handle_t h);
const char* peng_cuda_last_error(peng_handle_t h); /* h may be NULL: last error of a failed create */
/* number of kernels this handle has launched so far (bench.py's gpu_launches) */
uint64_t peng_cuda_launch_count(peng_handle_t h);
/* run the handle's work on a caller-owned stream (cudaStream_t passed as void*), NULL = handle's own */
int peng_cuda_set_stream(peng_handle_t h, void* cuda_stream);
int peng_cuda_synchronize(peng_handle_t h);

/* Device-side timing of the phases of the last peng_cuda_count[_local] call (CUDA events on the handle's stream,
 * recorded only while enabled): phase 0 = k-mer start marking (fused into stage P on the partition path),
 * 1 = the counting kernel(s), 2 = fold; partition path only: 3 = stage P, 4 = slow units, 5 = stage H, 6 = stage P2
 * (deep path of W >= 12, else 0) of the first chunk of <= 2^30 stream positions (the only one for sets up to 1 Gbp);
 * 7 = the last peng_cuda_count_exchange32 kernel (its wait for the slowest rank included).
 * This is how bench.py measures the dominant kernel's duration for the roofline without a profiler. */
int peng_cuda_enable_timing(peng_handle_t h, int on);
int peng_cuda_get_timing(peng_handle_t h, int phase, float* milliseconds);

/* ---- (a) ingest: SequenceSet::readFASTA + Sequence (src/shared/SequenceSet.cpp:285-447,
 *      src/shared/Sequence.cpp:4-35) -> packed 2-bit device buffer + validity mask ------------------- */
/* Read a FASTA file with the reference's line rules and pack it on the device. */
int peng_cuda_ingest_fasta(peng_handle_t h, const char* path, uint64_t* n_seq, uint64_t* total_bases,
                           uint32_t* min_len, uint32_t* max_len);
/* Shard `part` of `n_parts` of the same file for a multi-GPU run (one handle per GPU): contiguous runs of records cut by
 * their position in the file, i.e. balanced by bases; every record lands in exactly one shard, a shard may be empty.
 * Every shard applies the reference's line rules to the whole file, so errors are reported by all of them. */
int peng_cuda_ingest_fasta_shard(peng_handle_t h, const char* path, int part, int n_parts, uint64_t* n_seq,
                                 uint64_t* total_bases, uint32_t* min_len, uint32_t* max_len);
/* Same for sequences the caller already holds in the reference's encoding: what a binding inside the
 * reference passes from Sequence::getSequence()/getL() (src/shared/Sequence.h:18-19). codes = all sequences
 * concatenated, offsets[n_seq+1]. */
int peng_cuda_ingest_codes(peng_handle_t h, const uint8_t* codes, const uint64_t* offsets, uint64_t n_seq);
/* Same for one array PER SEQUENCE — exactly what the reference holds: SequenceSet::getSequences()[s]->getSequence() and
 * ->getL() (src/shared/SequenceSet.h, src/shared/Sequence.h:18-19), in ordinary pageable memory.  The host threads pack
 * the sequences from where they lie into the 2-bit device layout (0.375 B/base cross PCIe); nothing is concatenated or
 * pinned by the caller. */
int peng_cuda_ingest_sequences(peng_handle_t h, const uint8_t* const* seqs, const uint32_t* lengths, uint64_t n_seq);
/* Same, codes/offsets already resident on this handle's device (no host copy). */
/* How host buffers travel (peng_cuda_ingest_codes): 0 = automatic — with >= 4 host threads and AVX-512 the first part
 * of the set is packed into the 2-bit device layout BY THE HOST THREADS (0.375 B/base over PCIe) while the DMA engine
 * moves the raw bytes of the rest for the device to pack; 1 = raw bytes only; 2 = host packing only.  host_fraction
 * in [0, 1] fixes the packed share of the bases (< 0: derived from the thread count).  Threads: PENG_HOST_THREADS,
 * else hardware threads / LOCAL_WORLD_SIZE.  The device buffers are bit-identical in every mode. */
int peng_cuda_set_ingest_mode(peng_handle_t h, int mode, double host_fraction);
uint64_t peng_cuda_last_ingest_host_packed(peng_handle_t h); /* sequences packed by the host in the last ingest */
uint64_t peng_cuda_last_ingest_h2d_bytes(peng_handle_t h);   /* bytes the last ingest from host buffers sent to the device */
int peng_cuda_ingest_codes_device(peng_handle_t h, const void* d_codes, const uint64_t* offsets, uint64_t n_seq);
/* Count prefetch: BasePattern's constructor counts right after the sequences exist (src/base_pattern.cpp:36-80), so a
 * binding knows W and the strand before it hands the sequences over.  Armed with them, peng_cuda_ingest_codes runs the
 * first stage of that count (stage P of the W = 10 partition path) unit range by unit range as the data lands on the
 * device, under the DMA of the rest; the next peng_cuda_count[_local / _local32] with the same (W, strand) only runs what is
 * left (slow units, stage H, fold).  Same table bit for bit.  Other pattern lengths / small sets: the hint is ignored.
 * W = 0 disarms. */
int peng_cuda_set_count_prefetch(peng_handle_t h, int W, int strand);
/* Synthetic ChIP-seq-like set generated on the device (BASELINE.json configs[2..4]): n_seq sequences of
 * seq_len i.i.d. uniform bases from a counter-based generator (seed, base index), motif (base codes 0..3,
 * motif_len <= 32, may be NULL) planted in ~plant_fraction of the sequences at a uniform offset;
 * first_seq offsets the counter so that ranks generate disjoint shards of one global set. */
int peng_cuda_ingest_synthetic(peng_handle_t h, uint64_t n_seq, uint32_t seq_len, uint64_t seed, uint64_t first_seq,
                               const uint8_t* motif, uint32_t motif_len, double plant_fraction, double n_fraction);
/* Copy the ingested sequences back in the reference's encoding (tests, CPU baseline). */
int peng_cuda_sequences_info(peng_handle_t h, uint64_t* n_seq, uint64_t* total_bases, uint32_t* min_len,
                             uint32_t* max_len);
int peng_cuda_export_codes(peng_handle_t h, uint8_t* codes /* total_bases */, uint64_t* offsets /* n_seq+1 */);

/* ---- background model: BackgroundModel ctor (src/shared/BackgroundModel.cpp:3-92, 490-530) ----------- */
/* Count the (k+1)-mers of the ingested set on the device (N quirks of src/shared/Sequence.cpp:28-33
 * included) and derive the interpolated conditional probabilities v[0..K].
 * v_out: v[0] (4) | v[1] (16) | ... | v[K] (4^(K+1)), float for PENG_F32REF, double for PENG_F64. */
int peng_cuda_background_model(peng_handle_t h, int K, const float* alpha /* K+1 */, int interpolate, int precision,
                               void* v_out, int64_t* kmer_counts_out /* may be NULL, same layout */);
/* calculateV (src/shared/BackgroundModel.cpp:490-530) on (k+1)-mer counts summed over the shards of a multi-GPU run
 * (kmer_counts_out of peng_cuda_background_model, added element-wise by the caller). */
int peng_cuda_background_from_counts(int K, const float* alpha, int interpolate, int precision, const int64_t* kmer_counts,
                                     void* v_out);
/* Raw (k+1)-mer counts only, left on the device for a cross-GPU sum. */
int peng_cuda_background_counts(peng_handle_t h, int K);
int peng_cuda_background_finish(peng_handle_t h, int K, const float* alpha, int interpolate, int precision,
                                void* v_out);

/* ---- (b) counting: BasePattern::count_patterns / count_patterns_single_strand
 *      (src/base_pattern.cpp:331-441) ------------------------------------------------------------------ */
/* Builds the 4^W table on the device: greedy non-overlapping occurrences per (canonical) id, the
 * "resume two behind a non-ACGT letter" rule, BOTH: mirror copy at the reverse-complement id.
 * ltot = number of W-mers looked at (BasePattern::getLtot). */
int peng_cuda_count(peng_handle_t h, int W, int strand, uint64_t* ltot);
int peng_cuda_get_counts(peng_handle_t h, uint64_t* counts /* 4^W */); /* BasePattern::getPatternCounter */
/* The same table as uint32 (half the bytes over PCIe) — for sets of fewer than 2^32 positions, where no bin can exceed
 * 32 bits (the reference's own position counter is `unsigned int`, src/base_pattern.cpp:335-336); fails otherwise. */
int peng_cuda_get_counts32(peng_handle_t h, uint32_t* counts /* 4^W */);
/* Counting kernel family for W >= 10: 0 = automatic (partition path from 2^22 / 2^25 / 2^24 / 2^28 stream positions on for W = 10 / 12 / 14 / 16), 1 = one L2 atomic
 * per counted W-mer (k_count_global), 2 = partition path (W = 10: k_count_part / k_count_slow / k_hist_part, shared-
 * memory histograms per id range; W = 12, 14, 16: the deep path k_count_part / k_count_slow / k_part2 / k_hist_leaf),
 * 3 = as 2 with the second stage of the deep path replaced by one L2 RED per id (k_hist_l2; a cross-check for tests).
 * All produce the same table bit for bit; shorter pattern lengths ignore the setting. */
int peng_cuda_set_count_path(peng_handle_t h, int path);
int peng_cuda_last_count_path(peng_handle_t h); /* 1 or 2: what the last peng_cuda_count[_local] ran */
/* Multi-GPU: peng_cuda_count_local leaves the UNMIRRORED local histogram + ltot on the device;
 * the caller sums `n_u64` uint64 at `ptr` over ranks (ncclAllReduce), then
 * peng_cuda_count_finish applies the mirror step and publishes ltot. */
int peng_cuda_count_local(peng_handle_t h, int W, int strand);
int peng_cuda_count_buffer(peng_handle_t h, void** device_ptr, uint64_t* n_u64);
int peng_cuda_count_finish(peng_handle_t h, uint64_t* ltot);
/* The same with a 32-bit exchange, for sets of fewer than 2^32 stream positions IN TOTAL over all ranks and at most 2^30
 * per rank (no uint32 bin, nor ltot, can overflow in the sum): _count_local32 leaves the UNFOLDED uint32 launch table with
 * ltot in word 4^W; the caller sums `n_u32` uint32 at `ptr` over ranks — half the bytes of the 64-bit exchange —;
 * _count_finish32 folds (reverse-complement pairs, mirror) once, after the sum, and publishes ltot. */
int peng_cuda_count_local32(peng_handle_t h, int W, int strand);
int peng_cuda_count_buffer32(peng_handle_t h, void** device_ptr, uint64_t* n_u32);
int peng_cuda_count_finish32(peng_handle_t h, uint64_t* ltot);
/* Peer group: the multi-GPU counting step without a library collective in the data path.  After peng_cuda_count_local32
 * on every rank, peng_cuda_count_exchange32 runs ONE kernel per rank that sums the ranks' uint32 launch tables over NVLink
 * peer memory (each rank reduces the units it owns with 16-byte peer loads), folds reverse-complement pairs and writes the
 * finished 64-bit table — mirror copy included (src/base_pattern.cpp:387-392) — into every rank's table with peer stores;
 * arrival / departure of the ranks are flag words in peer memory, nothing passes through the host.  Replaces
 * count_buffer32 -> ncclAllReduce -> count_finish32 (kept: NCCL remains the portable path and carries the 64-bit exchange).
 *   peng_cuda_peer_export : allocates the tables of pattern length W (8 .. 16) and writes PENG_PEER_BLOB_BYTES describing
 *                           them (CUDA IPC handles; raw pointers for handles of the same process);
 *   peng_cuda_peer_attach : blobs = the exports of all `world` (<= 8) ranks in rank order — gathered by the caller with
 *                           whatever it has (torch.distributed / MPI / threads of one process);
 *   peng_cuda_peer_detach : before the pattern length changes or the handle goes away (destroy detaches).
 * ltot == NULL in peng_cuda_count_exchange32 (and in peng_cuda_count): no read-back and no synchronisation — the table is
 * complete in stream order and peng_cuda_count_finish(h, &ltot), or the first call that needs the table, fetches ltot. */
#define PENG_PEER_BLOB_BYTES 256
int peng_cuda_peer_export(peng_handle_t h, int W, void* blob);
int peng_cuda_peer_attach(peng_handle_t h, int rank, int world, const void* blobs);
int peng_cuda_peer_detach(peng_handle_t h);
int peng_cuda_count_exchange32(peng_handle_t h, uint64_t* ltot);
/* Where the last exchange kernel of this rank spent its time (device clock, ns since its start): [0] all ranks arrived,
 * [1] the units of CTA 0 done, [2] CTA 0's peer stores visible system-wide, [3] all CTAs done, [4] all ranks departed. */
int peng_cuda_peer_trace(peng_handle_t h, double* out5);
/* Id-sliced tables for the long pattern lengths on several GPUs (BASELINE configs[4]; the reference's tables are one
 * float array per order plus expected / z / logp over all 4^W ids, src/base_pattern.h:116-120 — at W = 16 more than one
 * GPU holds).  Rank `part` of `n_parts` keeps the ids [lo, hi) (whole tiles of 1024 ids, peng_cuda_table_slice) of every
 * table peng_cuda_stats builds; peng_cuda_get_table returns the slice, _gather_table / _gather_counts accept ids of the
 * slice, peng_cuda_iupac_aggregate (PENG_F64) and _iupac_count return this rank's PARTIAL sums (the caller adds them
 * over the ranks), the EM session must be opened with the same (part, n_parts).  The float32 arm's sequential sums and
 * seed selection need whole tables and refuse.  n_parts = 1 switches slicing off.
 * peng_cuda_count_finish32_sliced (PLUS strand) is the matching finish of the 32-bit exchange: the caller reduce-scatters
 * the 4^W uint32 bins in place (ncclReduceScatter with recvbuff = sendbuff + lo) and all-reduces word 4^W (ltot); only
 * the slice of the 64-bit table is then written and valid. */
int peng_cuda_set_table_slice(peng_handle_t h, int part, int n_parts);
int peng_cuda_table_slice(peng_handle_t h, uint64_t* lo, uint64_t* hi); /* after a count: this rank's id range */
int peng_cuda_count_finish32_sliced(peng_handle_t h, uint64_t* ltot);
/* Sparse read-back for tables that are too large to copy (W >= 14): the bins at `ids`, and the sum of all bins. */
int peng_cuda_gather_counts(peng_handle_t h, const uint64_t* ids, uint64_t n, uint64_t* out);
int peng_cuda_counts_total(peng_handle_t h, uint64_t* total);
/* Load an externally produced table instead of counting (tests; reference binding that keeps its own
 * count_patterns). counts must already be mirrored for BOTH. */
int peng_cuda_set_counts(peng_handle_t h, int W, int strand, const uint64_t* counts, uint64_t ltot);

/* ---- (c) background probabilities of all W-mers, expected counts, z-scores, log p-values
 *      calculate_bg_probabilities / aggregate_double_strand_background / calculate_expected_counts /
 *      calculate_zscores / calculate_log_pvalues (src/base_pattern.cpp:231-325) ------------------------- */
/* v: conditional probabilities as returned by BackgroundModel::getV() (src/shared/BackgroundModel.h:85-87)
 * flattened v[0]|v[1]|...|v[K]; both precisions are taken so that both arms can be produced.
 * k = order used for expected counts (BasePattern::k), max_k = highest order kept (BasePattern::max_k). */
int peng_cuda_stats(peng_handle_t h, int K, const float* v32, const double* v64, int k, int max_k);
int peng_cuda_get_table(peng_handle_t h, int table, int order, int precision, void* out /* 4^W elems */);
/* Sparse read-back: out[i] = table[ids[i]] (float for PENG_F32REF, double for PENG_F64) — what
 * BasePattern::print_patterns and getOptimizationScore read for the selected seeds (src/base_pattern.cpp:180-224,517-532). */
int peng_cuda_gather_table(peng_handle_t h, int table, int order, int precision, const uint64_t* ids, uint64_t n, void* out);

/* select_base_patterns (src/base_pattern.cpp:443-515).  Runs on the HOST on the float32 z table copied
 * back from the device: the result depends on the order libstdc++'s unstable std::sort leaves exact ties
 * in (every reverse-complement pair ties), which only the same algorithm reproduces. */
int peng_cuda_select_base_patterns(peng_handle_t h, float z_threshold, uint64_t count_threshold,
                                   int single_stranded, int filter_neighbors, uint64_t* out, uint64_t cap,
                                   uint64_t* n_selected);

/* ---- (d) IUPAC degenerate sums: IUPACPattern::aggregate_attributes_from_basepatterns /
 *      count_combined_occurences (src/iupac_pattern.cpp:410-473, 806-833) ------------------------------ */
/* For each of n IUPAC ids: n_sites = sum counts, bg_p = sum bgprob[k], expected = sum expected over the
 * distinct (BOTH: canonical) base patterns of its expansion.  PENG_F32REF accumulates the float sums
 * sequentially in the reference's list order (BOTH: ascending canonical id; PLUS: expansion order). */
int peng_cuda_iupac_aggregate(peng_handle_t h, const uint64_t* iupac_ids, int n, int precision, uint64_t* n_sites,
                              void* bg_p, void* expected);
int peng_cuda_iupac_count(peng_handle_t h, const uint64_t* iupac_ids, int n, uint64_t* counts_out);

/* ---- (e) EM: Peng::em_optimize_pwms + calculate_prob_odds (src/peng.cpp:48-197) ------------------------ */
/* n_pwms PWMs of W x 4 (row = position, col = A,C,G,T), all refined in one batch against the count table
 * and bgprob[bg_order]; each PWM stops when its L1 change <= threshold or after max_iterations.
 * pwm_out holds the final PWMs after the extra normalisation of IUPACPattern(ori, pwm)
 * (src/iupac_pattern.cpp:44-72).  float for PENG_F32REF (bit-identical to the reference), double for F64. */
int peng_cuda_em(peng_handle_t h, int n_pwms, const void* pwm_in, int precision, double saturation, double threshold,
                 int max_iterations, int bg_order, void* pwm_out, int* iterations_out);
/* Multi-GPU EM: rank `part` of `n_parts` accumulates its slice of the id range; the raw (unnormalised)
 * accumulators (n_pwms*W*4 doubles on the device at *acc_ptr) are summed over ranks by the caller between
 * _em_accumulate and _em_update.  F64 only. */
int peng_cuda_em_begin(peng_handle_t h, int n_pwms, const double* pwm_in, double saturation, double threshold,
                       int max_iterations, int bg_order, int part, int n_parts, void** acc_ptr);
int peng_cuda_em_accumulate(peng_handle_t h);
int peng_cuda_em_update(peng_handle_t h, int* n_active);
int peng_cuda_em_end(peng_handle_t h, double* pwm_out, int* iterations_out);

#ifdef __cplusplus
}
#endif
#endif /* PENG_CUDA_H_ */
