// C wrapper around the UNMODIFIED reference simple-knn (submodules/simple-knn/simple_knn.cu, compiled from where it
// lies by oracle/Makefile). TEST INFRASTRUCTURE ONLY: the golden implementation for distCUDA2 on the GPU box.
#include <cuda_runtime.h>

#include "simple_knn.h"

extern "C" int ref_knn(int P, const float* points_dev, float* mean_dist2_dev) {
    SimpleKNN::knn(P, (float3*)points_dev, mean_dist2_dev);   // legacy default stream, like the reference
    return (int)cudaDeviceSynchronize();
}
