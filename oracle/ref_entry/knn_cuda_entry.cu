// C entry point around the reference's OWN CUDA k-NN (lib/knn/src/cuda/knn.cu, compiled unmodified
// where it lies under /root/reference) so that ctypes can call it on the B200.  It replaces the
// THC-dependent glue of lib/knn/src/knn.h:12-42 (THCudaMalloc / THError are gone from torch >= 1.11)
// with the same sequence on plain CUDA runtime calls.  Test / baseline infrastructure only: the
// product never links this.
#include <cuda_runtime.h>

// lib/knn/src/cuda/vision.h:5-7 (that header also pulls in torch + THC, so the prototype is restated)
void knn_device(float *ref_dev, int ref_width, float *query_dev, int query_width, int height, int k,
                float *dist_dev, long *ind_dev, cudaStream_t stream);

extern "C" {

// bytes of the ref_nb x query_nb distance matrix knn.h:29 allocates per call
size_t ref_knn_cuda_scratch_bytes(int n_ref, int n_query) {
  return sizeof(float) * (size_t)n_ref * (size_t)n_query;
}

// knn.h:29-41 for device pointers: loop over the batch re-using one scratch matrix.  `dist_scratch` may be
// NULL (allocated and freed inside, like the reference) or a caller-held buffer of ref_knn_cuda_scratch_bytes
// (what torch's caching allocator amounts to after the first call).  After the call the first k rows of the
// scratch hold the (squared) distances of the LAST batch element (knn.cu:113-176 sorts them in place).
// Returns 0, or the cudaError_t.
int ref_knn_cuda(float *ref, float *query, long *idx, int batch, int dim, int n_ref, int n_query, int k,
                 float *dist_scratch, void *stream) {
  float *dist = dist_scratch;
  if (!dist) {
    cudaError_t e = cudaMalloc((void **)&dist, ref_knn_cuda_scratch_bytes(n_ref, n_query));
    if (e != cudaSuccess) return (int)e;
  }
  for (int b = 0; b < batch; b++)
    knn_device(ref + (size_t)b * dim * n_ref, n_ref, query + (size_t)b * dim * n_query, n_query, dim, k, dist,
               idx + (size_t)b * k * n_query, (cudaStream_t)stream);
  cudaError_t err = cudaGetLastError();
  if (!dist_scratch) {
    cudaStreamSynchronize((cudaStream_t)stream);
    cudaFree(dist);
  }
  return (int)err;
}

}  // extern "C"
