// Instantiates the VM kernels of one geometry (see qnn_vm_kernel.cuh) with 2^4 amplitudes per thread; one translation
// unit per variant so that they compile in parallel.
#include "qnn_vm_kernel.cuh"

namespace vm {
int launch_g0_k4(bool bwd, bool state, const LaunchCfg& lc, const VmArgs& a) { return launch_geom<0, false, 4>(bwd, state, lc, a); }
}  // namespace vm
