#!/bin/bash
# SASS opcode histogram of one kernel of lut_ldpc_b200/csrc/build/kernels.o: tools/sass_hist.sh k_vn_fusedILi8 [object]
K=${1:-k_vn_fusedILi8}; O=${2:-lut_ldpc_b200/csrc/build/kernels.o}
cuobjdump -sass $O | awk -v k="$K" '/Function : /{f=index($0,k)>0} f' | grep -E "^\s+/\*[0-9a-f]{4}\*/" \
  | sed -E 's/^\s+\/\*[0-9a-f]+\*\/\s+//; s/\s*\/\*.*//; s/^@!?U?P[0-9T] //' | awk '{print $1}' | sed 's/\..*//' | sort | uniq -c | sort -rn | head -${3:-20}
