#!/bin/bash
# Register / spill summary of every VM kernel from the ptxas logs of the last build.
for f in qnn_vm_g*.o.ptxas.log; do
  awk -v F=$f '/Compiling entry function/{n=$7} /spill stores/{sp=$0} /Used [0-9]+ registers/{print F, n, "|", $5, "regs |", sp}' $f
done | sed "s/_ZN2vm4k_vm//; s/EEvNS_6VmArgsE//; s/ptxas info    ://; s/  */ /g"
