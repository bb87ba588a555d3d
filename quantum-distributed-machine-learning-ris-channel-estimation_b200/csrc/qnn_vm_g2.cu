// Instantiates the VM kernels of one geometry (see qnn_vm_kernel.cuh); one translation unit per geometry so
// that they compile in parallel.
#include <map>
#include <mutex>
#include <tuple>

#include "qnn_vm_kernel.cuh"

namespace vm {
int launch_g2(bool bwd, bool state, const LaunchCfg& lc, const VmArgs& a) { return launch_geom<2, false>(bwd, state, lc, a); }

// How many clusters of `cluster` CTAs can be resident at once (GPC boundaries strand some SMs, so this is
// less than num_sm / cluster); a persistent cluster beyond that number would run as a second wave.  The
// answer depends on the device and on the shared-memory footprint, so the cache is keyed by all three.
int max_active_clusters_g2(int cluster, size_t smem, int device, int num_sm, bool bwd) {
  if (cluster < 1 || cluster > 8) return 1;
  static std::mutex mu;
  static std::map<std::tuple<int, int, size_t, bool>, int> cache;
  std::lock_guard<std::mutex> lock(mu);
  const auto key = std::make_tuple(device, cluster, smem, bwd);
  const auto hit = cache.find(key);
  if (hit != cache.end()) return hit->second;
  int n = num_sm / cluster * (bwd ? 1 : 2);
  auto kern = bwd ? k_vm<true, false, 2, false> : k_vm<false, true, 2, false>;
  if (qnn_allow_smem(reinterpret_cast<const void*>(kern), 227 * 1024) == QNN_OK) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(num_sm / cluster * cluster * (bwd ? 1 : 2)); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = smem;
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
  cache[key] = n;
  return n;
}
}  // namespace vm
