#!/bin/bash
# Builds a second copy of the engine library with extra nvcc flags (A/B runs): tools/build_variant.sh NAME -DFLAG ...
# -> aws-k8s-ansible-provisioner_b200/libb2engine_NAME.so ; use with B2_LIB_PATH=...
set -e
cd "$(dirname "$0")/.."
name=$1; shift
out=build/variant_$name
mkdir -p $out
objs=""
for f in elementwise elementwise_opt gemm_simt gemm_tc gemm_pair gemm_gateup gemm_resid gemm_mc attn_decode attn_prefill attn_prefill_tc sampling chain chain_cluster engine capi; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -Xcudafe --diag_suppress=177 "$@" \
    -c aws-k8s-ansible-provisioner_b200/csrc/$f.cu -o $out/$f.o &
  objs="$objs $out/$f.o"
done
wait
nvcc -shared -o aws-k8s-ansible-provisioner_b200/libb2engine_$name.so $objs -Xcompiler -fPIC -lcudart_static -ldl -lpthread -lrt
echo aws-k8s-ansible-provisioner_b200/libb2engine_$name.so
