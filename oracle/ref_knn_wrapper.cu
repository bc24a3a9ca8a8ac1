// C entry point around the REFERENCE's own SimpleKNN::knn (compiled from
// /root/reference/submodules/simple-knn/simple_knn.cu where it lies; see build_ref.sh).
// Test infrastructure only: used to pin street_crafter_b200's distCUDA2 against the reference
// kernels on the GPU box.
#include <cuda_runtime.h>
#include "simple_knn.h"

extern "C" int ref_knn(int P, float* points, float* mean_dists) {
    SimpleKNN::knn(P, (float3*)points, mean_dists);
    return (int)cudaDeviceSynchronize();
}
