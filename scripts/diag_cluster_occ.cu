// How many clusters of a given shape fit the device at once (diagnostic for hipe_cluster.cu).
//   nvcc -gencode arch=compute_100a,code=sm_100a -o /tmp/occ scripts/diag_cluster_occ.cu && /tmp/occ
#include <cuda_runtime.h>
#include <stdio.h>
__global__ void __maxnreg__(112) k_mr(int* p) { extern __shared__ int s[]; if (p) p[0] = s[0]; }
__global__ void __launch_bounds__(576, 1) k_lb(int* p) { extern __shared__ int s[]; if (p) p[0] = s[0]; }
template <typename K>
static void probe(const char* name, K kern, int threads, int smem) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    printf("%s threads=%d smem=%d: setattr %s\n", name, threads, smem, cudaGetErrorString(e));
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, kern);
    printf("   regs %d maxThreadsPerBlock %d\n", fa.numRegs, fa.maxThreadsPerBlock);
    for (int C = 1; C <= 16; C == 8 ? C = 16 : ++C) {
        if (C > 8) cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        cudaLaunchConfig_t cfg = {};
        cfg.blockDim = dim3(threads);
        cfg.gridDim = dim3(C);
        cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute a[1];
        a[0].id = cudaLaunchAttributeClusterDimension;
        a[0].val.clusterDim.x = C; a[0].val.clusterDim.y = 1; a[0].val.clusterDim.z = 1;
        cfg.attrs = a; cfg.numAttrs = 1;
        int n = -1;
        e = cudaOccupancyMaxActiveClusters(&n, kern, &cfg);
        printf("   C=%2d: %s -> %d clusters (%d SMs)\n", C, cudaGetErrorString(e), n, n * C);
        (void)cudaGetLastError();
    }
}
int main() {
    cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
    printf("%s SMs %d smem/SM %zu optin %zu reserved %zu regs/SM %d\n", pr.name, pr.multiProcessorCount,
           pr.sharedMemPerMultiprocessor, pr.sharedMemPerBlockOptin, pr.reservedSharedMemPerBlock, pr.regsPerMultiprocessor);
    probe("maxnreg112", k_mr, 576, 232208);
    probe("maxnreg112", k_mr, 576, 200000);
    probe("launch_bounds576", k_lb, 576, 232208);
    probe("launch_bounds576", k_lb, 640, 232208);
    return 0;
}
