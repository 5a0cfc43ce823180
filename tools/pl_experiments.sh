#!/bin/bash
# build experiment variants of the library (diagnostics of the pseudo-label kernel's critical path); run on the GPU box with
#   for v in 1 2 4 3 7; do TSS_LIB=$PWD/variants/libtss_exp$v.so TSS_NVCC_EXTRA="-DTSS_PL_EXP=$v" python tools/bench_stages.py --only pl_low_cfg2; done
mkdir -p variants
for v in ${@:-1 2 4 3 7}; do
  TSS_LIB=$PWD/variants/libtss_exp$v.so TSS_NVCC_EXTRA="-DTSS_PL_EXP=$v" python -c "import sys; sys.path.insert(0,'.'); from b200seg import _build; print(_build.ensure_current())"
done
