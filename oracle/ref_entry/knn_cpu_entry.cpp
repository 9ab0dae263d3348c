// C entry point around the reference's knn_cpu (declared in the reference's own header) so that
// ctypes can call the UNMODIFIED reference CPU implementation.  Test infrastructure only.
#include "cpu/vision.h"
#include <stdlib.h>

extern "C" int ref_knn_cpu(float *ref, int n_ref, float *query, int n_query, int dim, int k,
                           long *idx) {
  // scratch exactly as lib/knn/src/knn.h:49-50 allocates it
  float *dist = (float *)malloc(sizeof(float) * (size_t)n_ref * (size_t)n_query);
  long *ind_buf = (long *)malloc(sizeof(long) * (size_t)n_ref);
  if (!dist || !ind_buf) return -1;
  knn_cpu(ref, n_ref, query, n_query, dim, k, dist, idx, ind_buf);
  free(dist);
  free(ind_buf);
  return 0;
}
