#!/bin/bash
# k_ps variants (PS_T x PS_RPT, built as lib_<name>.so with make OUT= EXTRA=) timed with the probe
for lib in ${LIBS:-libb200ba.so}; do
  B200BA_LIB=$PWD/jaxopt-3d-reconstruction_b200/$lib python tools/cs_warp_probe.py ${SHAPES:-venice1778} 0 0 2>&1 | grep task_rows
done
