#!/bin/bash
# Build a variant of libhkbcn.so with extra -D flags for kernel experiments: tools/build_variant.sh NAME -DFOO=1 ...
# -> build/variants/libhkbcn_NAME.so (use with HKBCN_LIB=...).  Only bcn_decode.cu / bc7_encode.cu / bc6h_encode.cu are rebuilt.
set -e
cd "$(dirname "$0")/../hork-source_b200/csrc"
name=$1; shift
out=../../build/variants; mkdir -p $out/$name
FL="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden --fmad=false"
objs=""
for s in hkbcn_api bc45_encode bc123_encode bc7_encode bc6h_encode bcn_decode mip_resize normal_pack envprobe; do
  case " ${VARIANT_SRCS:-bcn_decode} " in
    *" $s "*) nvcc $FL "$@" -c $s.cu -o $out/$name/$s.o & objs="$objs $out/$name/$s.o";;
    *) objs="$objs build/$s.o";;
  esac
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/libhkbcn_$name.so $objs -cudart static
echo $out/libhkbcn_$name.so
