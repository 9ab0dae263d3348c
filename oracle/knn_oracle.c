/*
 * ORACLE — TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference k-NN for parity checks and
 * for the timed CPU baseline.  Nothing on the product path may call into this file.
 *
 * Follows (reference repo paths):
 *   distances  lib/knn/src/cuda/knn.cu:88-91   ssd += (a-b)*(a-b), dims in order; nvcc contracts
 *                                              this to fma(tmp,tmp,ssd)           -> contract=1
 *              lib/knn/src/cpu/knn_cpu.cpp:8-16 same sum, g++ x86-64 (no FMA)     -> contract=0
 *   selection  lib/knn/src/cuda/knn.cu:127-173  running top-k, strict '<' (lowest index wins ties)
 *              == lib/knn/src/cpu/knn_cpu.cpp:27-43 (stable bubble sort, then first k)
 *   indices    1-based (knn.cu:128,147,171; knn_cpu.cpp:25)
 *   layout     ref [dim][n_ref], query [dim][n_query], idx [k][n_query]  (knn.h:33-34)
 *
 * Pinned against the reference's own knn_cpu (oracle/_ref, built from the reference sources where
 * they lie) in tests/test_oracle_knn.py.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

static inline float pair_ssd(const float *ref, int n_ref, int j, const float *query, int n_query,
                             int q, int dim, int contract) {
  float ssd = 0.0f;
  for (int d = 0; d < dim; ++d) {
    float t = ref[(size_t)d * n_ref + j] - query[(size_t)d * n_query + q];
    if (contract) {
      ssd = fmaf(t, t, ssd);
    } else {
      volatile float p = t * t; /* force a rounded product: no contraction */
      ssd = ssd + p;
    }
  }
  return ssd;
}

/* k nearest references of every query.  idx [k][n_query] 1-based; dist [k][n_query] or NULL.
 * second [n_query] or NULL: for k==1 the smallest distance among the OTHER references
 * (+inf when n_ref==1) — used by the tests for the "best/second-best gap" rule. */
int knn_oracle(const float *ref, int n_ref, const float *query, int n_query, int dim, int k,
               int contract, int64_t *idx, float *dist, float *second) {
  if (k <= 0 || n_ref <= 0 || dim <= 0) return -1;
  int kk = k < n_ref ? k : n_ref;
#pragma omp parallel
  {
    float *td = (float *)malloc(sizeof(float) * (size_t)k);
    int64_t *ti = (int64_t *)malloc(sizeof(int64_t) * (size_t)k);
#pragma omp for schedule(static)
    for (int q = 0; q < n_query; ++q) {
      int cnt = 0;
      float sec = INFINITY;
      for (int j = 0; j < n_ref; ++j) {
        float d = pair_ssd(ref, n_ref, j, query, n_query, q, dim, contract);
        if (k == 1) {
          /* knn.cu:127: max_dist = dist[0]; then `curr_dist < max_dist` for the rest */
          if (cnt == 0) { td[0] = d; ti[0] = 0; cnt = 1; }
          else if (d < td[0]) { sec = td[0]; td[0] = d; ti[0] = j; }
          else if (d < sec) sec = d;
          continue;
        }
        if (cnt < kk) {
          int p = cnt++;
          while (p > 0 && td[p - 1] > d) { td[p] = td[p - 1]; ti[p] = ti[p - 1]; --p; }
          td[p] = d; ti[p] = j;
        } else if (d < td[kk - 1]) {
          int p = kk - 1;
          while (p > 0 && td[p - 1] > d) { td[p] = td[p - 1]; ti[p] = ti[p - 1]; --p; }
          td[p] = d; ti[p] = j;
        }
      }
      for (int i = 0; i < k; ++i) {
        idx[(size_t)i * n_query + q] = i < cnt ? ti[i] + 1 : 0;
        if (dist) dist[(size_t)i * n_query + q] = i < cnt ? td[i] : INFINITY;
      }
      if (second) second[q] = sec;
    }
    free(td);
    free(ti);
  }
  return 0;
}

/* single-thread pair rate probe helper: number of (query, ref) pairs evaluated */
double knn_oracle_pairs(int n_ref, int n_query) { return (double)n_ref * (double)n_query; }
