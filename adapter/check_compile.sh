#!/bin/sh
# Compile-check of the adapter against the reference headers (needs /root/reference).
# It proves the signatures and PD_REGISTER_KERNEL registrations type-check; linking needs libphi.
#   1-3  the adapter TUs (kernel bodies + registrations, functor plug-ins + grads, dist hook, the
#        Partial->Replicate reshard function)
#   4    the template entry points (adapter/b200_funcs_shim.h through the shadow headers in
#        adapter/funcs_shim) under three UNMODIFIED reference clients
#   5    the host-side unit test of the shim's functor classification (built AND run)
# The compiles run side by side; any failure fails the script.
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=${OUT:-/tmp/phi_b200_kernels.o}
OUT2=${OUT2:-${OUT%.o}_functors.o}
OUT3=${OUT3:-${OUT%.o}_dist.o}
LOGDIR=$(mktemp -d)
NVCC="/usr/local/cuda/bin/nvcc -std=c++17 -gencode arch=compute_100a,code=sm_100a -DPADDLE_WITH_CUDA -DEIGEN_USE_GPU --expt-relaxed-constexpr --expt-extended-lambda -w"
INC="-I$HERE/shim -I$HERE/../include -I$REF -I$REF/third_party/eigen3"
SHIM_INC="-I$HERE/funcs_shim -I$HERE $INC"

run() {  # run <tag> <message> <command...>
  tag=$1; msg=$2; shift 2
  ( "$@" > "$LOGDIR/$tag.log" 2>&1 && echo "$msg" || { echo "FAILED: $tag"; cat "$LOGDIR/$tag.log"; touch "$LOGDIR/failed"; } ) &
}

run kernels "adapter compiles: $OUT" $NVCC $INC -c "$HERE/phi_b200_kernels.cu" -o "$OUT"
run functors "adapter compiles: $OUT2" $NVCC $INC -c "$HERE/phi_b200_functor_kernels.cu" -o "$OUT2"
# the auto-parallel hook needs the NCCL comm context (system nccl.h)
run dist "adapter compiles: $OUT3" $NVCC -DPADDLE_WITH_NCCL $INC -c "$HERE/phi_b200_dist_hook.cu" -o "$OUT3"
OUT4=${OUT4:-${OUT%.o}_reshard.o}
run reshard "adapter compiles: $OUT4" $NVCC -DPADDLE_WITH_NCCL -DPADDLE_WITH_DISTRIBUTE $INC -c "$HERE/phi_b200_reshard.cu" -o "$OUT4"
run shim_p_norm "shim client compiles: ${OUT%.o}_shim_p_norm.o" $NVCC $SHIM_INC -c "$REF/paddle/phi/kernels/gpu/p_norm_kernel.cu" -o "${OUT%.o}_shim_p_norm.o"
run shim_compare "shim client compiles: ${OUT%.o}_shim_compare.o" $NVCC $SHIM_INC -c "$REF/paddle/phi/kernels/kps/compare_kernel.cu" -o "${OUT%.o}_shim_compare.o"
run shim_grad "shim client compiles: ${OUT%.o}_shim_elementwise_grad.o" $NVCC $SHIM_INC -c "$HERE/shim_client_elementwise_grad.cu" -o "${OUT%.o}_shim_elementwise_grad.o"
# (unrelated phi symbols that the reference headers' unused inline functions mention stay
# unresolved on purpose: the probe only runs the classification templates)
PROBE=${OUT%.o}_shim_probe
run probe "shim probe builds: $PROBE" $NVCC -O1 -Xcompiler -ffunction-sections -Xcompiler -fdata-sections \
  -Xlinker --gc-sections -Xlinker --unresolved-symbols=ignore-all $SHIM_INC "$HERE/shim_probe_test.cu" -o "$PROBE"
wait
if [ -e "$LOGDIR/failed" ]; then rm -rf "$LOGDIR"; exit 1; fi
rm -rf "$LOGDIR"
"$PROBE" || exit 1
