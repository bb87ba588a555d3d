// Instantiates the VM kernels of one geometry (see qnn_vm_kernel.cuh); one translation unit per geometry so
// that they compile in parallel.
#include "qnn_vm_kernel.cuh"

namespace vm {
int launch_g2(bool bwd, bool state, const LaunchCfg& lc, const VmArgs& a) { return launch_geom<2, false>(bwd, state, lc, a); }

// How many clusters of `cluster` CTAs can be resident at once (GPC boundaries strand some SMs, so this is
// less than 148 / cluster); a persistent cluster beyond that number would run as a second wave.
int max_active_clusters_g2(int cluster, size_t smem) {
  static int cache[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  if (cluster < 1 || cluster > 8) return 1;
  if (cache[cluster] > 0) return cache[cluster];
  int n = 148 / cluster;
  auto kern = k_vm<true, false, 2, false>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) == cudaSuccess) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(148 / cluster * cluster); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr{};
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = cluster; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
    cfg.attrs = &attr; cfg.numAttrs = 1;
    int q = 0;
    if (cudaOccupancyMaxActiveClusters(&q, kern, &cfg) == cudaSuccess && q > 0) n = std::min(n, q);
    else (void)cudaGetLastError();
  } else {
    (void)cudaGetLastError();
  }
  cache[cluster] = n;
  return n;
}
}  // namespace vm
