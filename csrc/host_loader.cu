// Host-side minibatch assembly (plain C++, no device code): gathers the CSR rows of a shuffled
// minibatch from the corpus CSR into caller-provided (pinned) staging buffers.
// Ref: the reference's torch DataLoader over a dense [n_docs, V] array (federated_model.py:82-86).
#include "../include/gfedntm_kernels.h"
#include <string.h>
#include <thread>
#include <vector>

extern "C" int gf_host_gather_csr(const int* indptr, const int* indices, const float* data,
                                  const long long* rows, int n_rows, int* out_indptr, int* out_indices,
                                  float* out_data, long long out_capacity, int n_threads) {
  if (n_rows < 0) return 2;
  out_indptr[0] = 0;
  for (int i = 0; i < n_rows; ++i) {
    const long long r = rows[i];
    out_indptr[i + 1] = out_indptr[i] + (indptr[r + 1] - indptr[r]);
  }
  if ((long long)out_indptr[n_rows] > out_capacity) return 2;
  auto work = [&](int lo, int hi) {
    for (int i = lo; i < hi; ++i) {
      const long long r = rows[i];
      const int n = indptr[r + 1] - indptr[r];
      memcpy(out_indices + out_indptr[i], indices + indptr[r], sizeof(int) * (size_t)n);
      memcpy(out_data + out_indptr[i], data + indptr[r], sizeof(float) * (size_t)n);
    }
  };
  if (n_threads <= 1 || n_rows < 256) {
    work(0, n_rows);
  } else {
    std::vector<std::thread> th;
    const int per = (n_rows + n_threads - 1) / n_threads;
    for (int t = 0; t < n_threads; ++t) {
      const int lo = t * per, hi = lo + per < n_rows ? lo + per : n_rows;
      if (lo < hi) th.emplace_back(work, lo, hi);
    }
    for (auto& t : th) t.join();
  }
  return 0;
}

// Gather rows of a dense row-major float matrix (contextual embeddings).
extern "C" int gf_host_gather_rows(const float* src, long long row_len, const long long* rows, int n_rows,
                                   float* out) {
  for (int i = 0; i < n_rows; ++i)
    memcpy(out + (size_t)i * row_len, src + (size_t)rows[i] * row_len, sizeof(float) * (size_t)row_len);
  return 0;
}
