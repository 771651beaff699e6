// How many clusters of a given size / shared-memory footprint are co-resident on this GPU?
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(int* p) { extern __shared__ float s[]; if (p) p[0] = (int)s[0]; }
int main() {
  for (int cs : {2, 4, 8, 16}) for (int kb : {64, 100, 128, 190, 210, 224}) for (int threads : {256, 512}) {
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, kb * 1024);
    if (cs > 8) cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3(cs * 64); cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = kb * 1024;
    cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int n = -1; cudaError_t e = cudaOccupancyMaxActiveClusters(&n, k, &cfg);
    printf("cluster %2d smem %3d KB threads %d: max active clusters %d (%s)\n", cs, kb, threads, n, cudaGetErrorString(e));
    cudaGetLastError();
  }
  return 0;
}
