// How many clusters of a 1-CTA-per-SM kernel (225 KB dynamic shared memory, 192 threads: the footprint of
// ssi::gemm_kernel) can be co-resident on this GPU, by cluster size.  Backs the DESIGN.md note on why K1 does not
// multicast W tiles across a 4-CTA cluster: GPCs with an odd number of TPCs strand one TPC per GPC.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/cluster_probe.bin tools/cluster_probe.cu
#include <cuda_runtime.h>
#include <cstdio>

__global__ void __launch_bounds__(192, 1) probe_kernel(int* out) {
  extern __shared__ unsigned char smem[];
  if (out && threadIdx.x == 0 && blockIdx.x == 0) out[0] = smem[0];
}

int main() {
  const int smem = 230656;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) return 1;
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"max_active_clusters\": {", prop.name, prop.multiProcessorCount);
  const int sizes[] = {1, 2, 4, 8, 16};
  for (int i = 0; i < 5; ++i) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(prop.multiProcessorCount / sizes[i] * sizes[i]);
    cfg.blockDim = dim3(192);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = sizes[i], at[0].val.clusterDim.y = 1, at[0].val.clusterDim.z = 1;
    cfg.attrs = at, cfg.numAttrs = 1;
    int n = -1;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&n, probe_kernel, &cfg);
    printf("%s\"%d\": %d", i ? ", " : "", sizes[i], e == cudaSuccess ? n : -static_cast<int>(e));
  }
  printf("}}\n");
  return 0;
}
