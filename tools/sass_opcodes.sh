#!/bin/bash
# Per-kernel SASS opcode histogram of the SHIPPED library (evidence that K1 is TMA-fed VABSDIFF4, K2 is tcgen05 / TMEM,
# K1f is packed FP32, K8 is unfused FP64, and how many spill instructions each kernel carries).
#   tools/sass_opcodes.sh [spectavi_b200/libspectavi.so] > profiles/r02_sass_opcodes.txt
SO=${1:-spectavi_b200/libspectavi.so}
echo "# $(basename $SO): $(stat -c %s $SO) bytes; $(cuobjdump --version | tail -1)"
echo "# kernel | SASS instructions | counts of the opcodes that identify the design (STL / LDL = spills)"
cuobjdump -sass $SO | python3 "$(dirname $0)/sass_opcodes.py"
