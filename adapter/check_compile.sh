#!/bin/sh
# Compile-check of the adapter against the reference headers (needs /root/reference).
# It proves the signatures and PD_REGISTER_KERNEL registrations type-check; linking needs libphi.
set -e
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=${OUT:-/tmp/phi_b200_kernels.o}
OUT2=${OUT2:-${OUT%.o}_functors.o}
for pair in "phi_b200_kernels.cu:$OUT" "phi_b200_functor_kernels.cu:$OUT2"; do
  src=${pair%%:*}
  obj=${pair#*:}
  /usr/local/cuda/bin/nvcc -std=c++17 -gencode arch=compute_100a,code=sm_100a -DPADDLE_WITH_CUDA -DEIGEN_USE_GPU \
    --expt-relaxed-constexpr --expt-extended-lambda -w \
    -I"$HERE/shim" -I"$HERE/../include" -I"$REF" -I"$REF/third_party/eigen3" \
    -c "$HERE/$src" -o "$obj"
  echo "adapter compiles: $obj"
done
# third TU: the auto-parallel hook needs the NCCL comm context (system nccl.h)
OUT3=${OUT3:-${OUT%.o}_dist.o}
/usr/local/cuda/bin/nvcc -std=c++17 -gencode arch=compute_100a,code=sm_100a -DPADDLE_WITH_CUDA -DPADDLE_WITH_NCCL -DEIGEN_USE_GPU \
  --expt-relaxed-constexpr --expt-extended-lambda -w \
  -I"$HERE/shim" -I"$HERE/../include" -I"$REF" -I"$REF/third_party/eigen3" \
  -c "$HERE/phi_b200_dist_hook.cu" -o "$OUT3"
echo "adapter compiles: $OUT3"
