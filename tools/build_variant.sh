#!/bin/bash
# usage: tools/build_variant.sh NAME "-DFOO=1 ..."  -> su2_b200/variants/libsu2b200_NAME.so (for A/B kernel tuning)
set -e
name=$1; shift
cd "$(dirname "$0")/../su2_b200/csrc"
out=../variants; mkdir -p $out/obj_$name
for f in capi.cu capi_flow.cu flow_kernels.cu edge_kernels.cu bc_kernels.cu linalg.cu precond.cu precond_poll.cu precond_factor.cu krylov.cu comm.cu turb_sa.cu turb_sst.cu mesh_host.cu restart_host.cu; do
  ( /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr "$@" -c $f -o $out/obj_$name/${f%.cu}.o ) &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/libsu2b200_$name.so $out/obj_$name/*.o -lcudart -ldl
rm -rf $out/obj_$name
echo built $out/libsu2b200_$name.so
