// abi_shim.cpp — TEST INFRASTRUCTURE ONLY.  The subset of include/peng_cuda.h that the host driver
// (peng-motif_b200/host) calls, answered by the CPU oracle (oracle/peng_oracle.h), so that the HOST logic of the
// driver — hill-climb bookkeeping, filters, merging, writers, number formatting — can be checked against the
// reference's recorded console report and MEME file on a machine without a GPU (tests/test_host_driver.py builds
// tests/sim/peng_motif_sim from the driver's sources + this file).  Never linked into the product: the shipped
// peng_motif links libpeng_cuda.so and stops without a CUDA device.
#include <atomic>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/peng_cuda.h"
#include "../../oracle/peng_oracle.h"
#include "../../peng-motif_b200/csrc/host_ingest.h"  // calculate_v (host arithmetic of the product, no device code)

struct peng_cuda_ctx {
  std::vector<uint8_t> codes;
  std::vector<uint64_t> offsets;
  uint64_t n_seq = 0;
  int W = 0, strand = 0, k = 0, max_k = 0, K = 0;
  uint64_t ltot = 0;
  std::vector<uint64_t> counts;
  std::vector<std::vector<float>> bgprob;  // per order
  std::vector<float> expected, z, logp, v;
  std::string err;
  // --gpus N on the shim: the ranks are threads of this process; the "exchange" adds the ranks' tables behind a barrier
  std::vector<peng_cuda_ctx*> group;
  std::vector<uint64_t> local;
  uint64_t local_ltot = 0;
  std::atomic<int> arrived{0}, departed{0};
};
static std::string g_err;

extern "C" {

int peng_cuda_create(int, peng_handle_t* out) {
  *out = new peng_cuda_ctx();
  return PENG_OK;
}
void peng_cuda_destroy(peng_handle_t h) { delete h; }
const char* peng_cuda_last_error(peng_handle_t h) { return h ? h->err.c_str() : g_err.c_str(); }

int peng_cuda_ingest_fasta(peng_handle_t h, const char* path, uint64_t* n_seq, uint64_t* total, uint32_t* min_len,
                           uint32_t* max_len) {
  uint64_t n = 0, t = 0;
  uint32_t a = 0, b = 0;
  if (po_fasta_scan(path, &n, &t, &a, &b)) {
    h->err = std::string("Error: Cannot read FASTA file: ") + path;
    return PENG_E_IO;
  }
  h->codes.assign(t ? t : 1, 0);
  h->offsets.assign(n + 1, 0);
  po_fasta_read(path, h->codes.data(), h->offsets.data());
  h->n_seq = n;
  if (n_seq) *n_seq = n;
  if (total) *total = t;
  if (min_len) *min_len = a;
  if (max_len) *max_len = b;
  return PENG_OK;
}

int peng_cuda_background_model(peng_handle_t h, int K, const float* alpha, int interpolate, int precision, void* v_out,
                               int64_t* counts_out) {
  const size_t nv = (((size_t)1 << (2 * (K + 2))) - 4) / 3;
  std::vector<int64_t> n(nv);
  if (precision == PENG_F32REF) {
    po_background_model_f32(h->codes.data(), h->offsets.data(), h->n_seq, K, alpha, interpolate, (float*)v_out, n.data());
  } else {
    std::vector<double> a(alpha, alpha + K + 1);
    po_background_model_f64(h->codes.data(), h->offsets.data(), h->n_seq, K, a.data(), interpolate, (double*)v_out, n.data());
  }
  if (counts_out) memcpy(counts_out, n.data(), nv * 8);
  return PENG_OK;
}

int peng_cuda_count(peng_handle_t h, int W, int strand, uint64_t* ltot) {
  h->W = W;
  h->strand = strand;
  h->counts.assign((size_t)1 << (2 * W), 0);
  po_count(h->codes.data(), h->offsets.data(), h->n_seq, W, strand == PENG_STRAND_BOTH, h->counts.data(), &h->ltot);
  if (ltot) *ltot = h->ltot;
  return PENG_OK;
}

int peng_cuda_stats(peng_handle_t h, int K, const float* v32, const double*, int k, int max_k) {
  const size_t n = (size_t)1 << (2 * h->W);
  const size_t nv = (((size_t)1 << (2 * (K + 2))) - 4) / 3;
  h->v.assign(v32, v32 + nv);
  h->K = K;
  h->k = k;
  h->max_k = max_k < k ? k : max_k;
  h->bgprob.assign((size_t)h->max_k + 1, std::vector<float>(n));
  for (int o = 0; o <= h->max_k; o++) {
    po_bg_probabilities_f32(v32, h->W, o, h->bgprob[(size_t)o].data());
    if (h->strand == PENG_STRAND_BOTH) po_aggregate_double_strand_f32(h->bgprob[(size_t)o].data(), h->W);
  }
  h->expected.assign(n, 0);
  h->z.assign(n, 0);
  h->logp.assign(n, 0);
  po_stats_f32(h->counts.data(), h->bgprob[(size_t)k].data(), h->ltot, h->W, h->expected.data(), h->z.data(), h->logp.data());
  return PENG_OK;
}

int peng_cuda_select_base_patterns(peng_handle_t h, float z_threshold, uint64_t count_threshold, int single_stranded,
                                   int filter_neighbors, uint64_t* out, uint64_t cap, uint64_t* n_selected) {
  const uint64_t n = po_select_f32(h->z.data(), h->counts.data(), h->W, z_threshold, count_threshold, single_stranded,
                                   filter_neighbors, out, cap);
  if (n_selected) *n_selected = n;
  return PENG_OK;
}

int peng_cuda_gather_counts(peng_handle_t h, const uint64_t* ids, uint64_t n, uint64_t* out) {
  for (uint64_t i = 0; i < n; i++) out[i] = h->counts[ids[i]];
  return PENG_OK;
}

int peng_cuda_gather_table(peng_handle_t h, int table, int order, int precision, const uint64_t* ids, uint64_t n, void* out) {
  if (precision != PENG_F32REF) return PENG_E_INVALID;
  const std::vector<float>* t = table == PENG_TABLE_BGPROB ? &h->bgprob[(size_t)order]
                                : table == PENG_TABLE_EXPECTED ? &h->expected
                                : table == PENG_TABLE_ZSCORE   ? &h->z
                                                               : &h->logp;
  for (uint64_t i = 0; i < n; i++) ((float*)out)[i] = (*t)[ids[i]];
  return PENG_OK;
}

int peng_cuda_iupac_aggregate(peng_handle_t h, const uint64_t* ids, int n, int precision, uint64_t* n_sites, void* bg_p,
                              void* expected) {
  if (precision != PENG_F32REF) return PENG_E_INVALID;
  for (int i = 0; i < n; i++) {
    float out5[5];
    po_iupac_aggregate_f32(ids[i], h->W, h->strand == PENG_STRAND_BOTH, h->counts.data(), h->bgprob[(size_t)h->k].data(),
                           h->expected.data(), h->n_seq, &n_sites[i], out5);
    ((float*)bg_p)[i] = out5[0];
    ((float*)expected)[i] = out5[1];
  }
  return PENG_OK;
}

int peng_cuda_iupac_count(peng_handle_t h, const uint64_t* ids, int n, uint64_t* out) {
  for (int i = 0; i < n; i++) out[i] = po_count_combined(ids[i], h->W, h->strand == PENG_STRAND_BOTH, h->counts.data());
  return PENG_OK;
}

int peng_cuda_em(peng_handle_t h, int n_pwms, const void* pwm_in, int precision, double saturation, double threshold,
                 int max_iterations, int bg_order, void* pwm_out, int* iterations_out) {
  if (precision != PENG_F32REF) return PENG_E_INVALID;
  const int W = h->W;
  for (int i = 0; i < n_pwms; i++) {
    const int it = po_em_f32((const float*)pwm_in + (size_t)i * W * 4, W, h->counts.data(), h->bgprob[(size_t)bg_order].data(),
                             (float)saturation, (float)threshold, max_iterations, (float*)pwm_out + (size_t)i * W * 4);
    if (iterations_out) iterations_out[i] = it;
  }
  return PENG_OK;
}

int peng_cuda_device_count(void) { return 8; }
int peng_cuda_synchronize(peng_handle_t) { return PENG_OK; }
int peng_cuda_count_finish(peng_handle_t h, uint64_t* ltot) {
  if (ltot) *ltot = h->ltot;
  return PENG_OK;
}

int peng_cuda_ingest_fasta_shard(peng_handle_t h, const char* path, int part, int n_parts, uint64_t* n_seq, uint64_t* total,
                                 uint32_t* min_len, uint32_t* max_len) {
  uint64_t n = 0, t = 0;
  uint32_t a = 0, b = 0;
  const int rc = peng_cuda_ingest_fasta(h, path, &n, &t, &a, &b);
  if (rc != PENG_OK) return rc;
  // contiguous runs of records (the product cuts by file position; any cut gives the same sums)
  const uint64_t s0 = n * (uint64_t)part / (uint64_t)n_parts, s1 = n * (uint64_t)(part + 1) / (uint64_t)n_parts;
  std::vector<uint8_t> codes(h->codes.begin() + (long)h->offsets[s0], h->codes.begin() + (long)h->offsets[s1]);
  std::vector<uint64_t> offs;
  uint32_t mn = 0xFFFFFFFFu, mx = 0;
  for (uint64_t s = s0; s <= s1; s++) offs.push_back(h->offsets[s] - h->offsets[s0]);
  for (uint64_t s = s0; s < s1; s++) {
    const uint32_t L = (uint32_t)(h->offsets[s + 1] - h->offsets[s]);
    mn = L < mn ? L : mn;
    mx = L > mx ? L : mx;
  }
  if (codes.empty()) codes.push_back(0);
  h->codes.swap(codes);
  h->offsets.swap(offs);
  h->n_seq = s1 - s0;
  if (n_seq) *n_seq = s1 - s0;
  if (total) *total = h->offsets.back();
  if (min_len) *min_len = s1 > s0 ? mn : 0;
  if (max_len) *max_len = mx;
  return PENG_OK;
}

int peng_cuda_background_from_counts(int K, const float* alpha, int interpolate, int precision, const int64_t* kmer_counts, void* v_out) {
  const size_t nv = (((size_t)1 << (2 * (K + 2))) - 4) / 3;
  std::vector<uint64_t> n(kmer_counts, kmer_counts + nv);
  if (precision == PENG_F32REF) peng_host::calculate_v<float>(n.data(), K, alpha, interpolate != 0, (float*)v_out);
  else peng_host::calculate_v<double>(n.data(), K, alpha, interpolate != 0, (double*)v_out);
  return PENG_OK;
}

int peng_cuda_get_counts(peng_handle_t h, uint64_t* counts) {
  memcpy(counts, h->counts.data(), h->counts.size() * 8);
  return PENG_OK;
}
int peng_cuda_set_counts(peng_handle_t h, int W, int strand, const uint64_t* counts, uint64_t ltot) {
  h->W = W;
  h->strand = strand;
  h->counts.assign(counts, counts + ((size_t)1 << (2 * W)));
  h->ltot = ltot;
  return PENG_OK;
}

int peng_cuda_peer_export(peng_handle_t h, int W, void* blob) {
  h->W = W;
  memset(blob, 0, PENG_PEER_BLOB_BYTES);
  memcpy(blob, &h, sizeof h);
  return PENG_OK;
}
int peng_cuda_peer_attach(peng_handle_t h, int, int world, const void* blobs) {
  h->group.clear();
  for (int p = 0; p < world; p++) {
    peng_cuda_ctx* q;
    memcpy(&q, (const unsigned char*)blobs + (size_t)p * PENG_PEER_BLOB_BYTES, sizeof q);
    h->group.push_back(q);
  }
  return PENG_OK;
}
int peng_cuda_peer_detach(peng_handle_t h) {
  h->group.clear();
  return PENG_OK;
}
int peng_cuda_count_local32(peng_handle_t h, int W, int strand) {
  h->W = W;
  h->strand = strand;
  h->local.assign((size_t)1 << (2 * W), 0);
  po_count(h->codes.data(), h->offsets.data(), h->n_seq, W, strand == PENG_STRAND_BOTH, h->local.data(), &h->local_ltot);
  return PENG_OK;
}
int peng_cuda_count_exchange32(peng_handle_t h, uint64_t* ltot) {
  const int world = (int)h->group.size();
  for (auto* q : h->group) q->arrived++;             // "my table is complete", told to every rank
  while (h->arrived.load() < world) std::this_thread::yield();
  h->counts.assign(h->local.size(), 0);
  h->ltot = 0;
  for (auto* q : h->group) {
    for (size_t i = 0; i < h->counts.size(); i++) h->counts[i] += q->local[i];
    h->ltot += q->local_ltot;
  }
  for (auto* q : h->group) q->departed++;            // "I have read your table"
  while (h->departed.load() < world) std::this_thread::yield();
  h->arrived = 0;
  h->departed = 0;
  if (ltot) *ltot = h->ltot;
  return PENG_OK;
}

}  // extern "C"
