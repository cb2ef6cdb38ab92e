#!/bin/bash
# Compile the reference's own alt_cuda_corr extension, UNMODIFIED, from where its sources lie
# (/root/reference/alt_cuda_corr/{correlation.cpp, correlation_kernel.cu}), into oracle/_ref/
# (git-ignored; travels to the GPU box).  Test infrastructure: the GPU tests compare this library's
# rc_alt_fwd / rc_alt_bwd with the reference extension's own output, and tools/ time it as the
# "reference kernel on B200" arm.  Plain nvcc/g++ on the two files, not the reference's setup.py.
# Takes ~6 minutes (torch/extension.h).  Usage: bash oracle/build_ref_ext.sh [--if-missing]
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${RAFT_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
SO="$OUT/alt_cuda_corr_ref.so"
if [ "$1" = "--if-missing" ] && [ -f "$SO" ]; then exit 0; fi
if [ ! -f "$REF/alt_cuda_corr/correlation_kernel.cu" ]; then echo "reference checkout not found at $REF"; exit 0; fi
mkdir -p "$OUT/build"
PY="${PYTHON:-python}"
TORCH_INC=$($PY -c "import torch.utils.cpp_extension as c; print(' '.join('-I'+p for p in c.include_paths()))" 2>/dev/null)
TORCH_LIB=$($PY -c "import torch.utils.cpp_extension as c; print(c.library_paths()[0])" 2>/dev/null)
PY_INC=$($PY -c "import sysconfig; print('-I'+sysconfig.get_paths()['include'])")
EXT=$($PY -c "import sysconfig; print(sysconfig.get_config_var('EXT_SUFFIX'))")
FLAGS="-O3 -std=c++17 -DTORCH_EXTENSION_NAME=alt_cuda_corr_ref -DTORCH_API_INCLUDE_EXTENSION_H -D_GLIBCXX_USE_CXX11_ABI=1 $TORCH_INC $PY_INC"
nvcc -gencode arch=compute_100a,code=sm_100a $FLAGS -Xcompiler -fPIC -w -c "$REF/alt_cuda_corr/correlation_kernel.cu" -o "$OUT/build/correlation_kernel.o" &
g++ $FLAGS -fPIC -w -I/usr/local/cuda/include -c "$REF/alt_cuda_corr/correlation.cpp" -o "$OUT/build/correlation.o" &
wait
g++ -shared "$OUT/build/correlation.o" "$OUT/build/correlation_kernel.o" -L"$TORCH_LIB" -ltorch -ltorch_cpu -ltorch_cuda -lc10 -lc10_cuda -ltorch_python \
    -L/usr/local/cuda/lib64 -lcudart -Wl,-rpath,"$TORCH_LIB" -o "$OUT/alt_cuda_corr_ref$EXT"
ln -sf "alt_cuda_corr_ref$EXT" "$SO"
echo "built $OUT/alt_cuda_corr_ref$EXT"
