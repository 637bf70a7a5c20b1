#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(int* p) { extern __shared__ int s[]; if (p) p[0] = s[0]; }
int main() {
  for (int smem_kb : {100, 145, 190, 220}) {
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_kb * 1024);
    cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    printf("smem %d KB:", smem_kb);
    for (int C = 1; C <= 16; C++) {
      cudaLaunchConfig_t cfg{}; cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = C; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.attrs = at; cfg.numAttrs = 1; cfg.blockDim = dim3(256); cfg.gridDim = dim3(C * 64); cfg.dynamicSmemBytes = smem_kb * 1024;
      int n = 0; cudaError_t e = cudaOccupancyMaxActiveClusters(&n, k, &cfg);
      if (e != cudaSuccess) { n = -1; cudaGetLastError(); }
      printf(" C%d:%d(%d)", C, n, n * C);
    }
    printf("\n");
  }
}
