// Instantiates the VM kernels of one geometry (see qnn_vm_kernel.cuh); one translation unit per geometry so
// that they compile in parallel.
#include "qnn_vm_kernel.cuh"

namespace vm {
int launch_g0(bool bwd, bool state, const LaunchCfg& lc, const VmArgs& a) { return launch_geom<0, false>(bwd, state, lc, a); }
}  // namespace vm
