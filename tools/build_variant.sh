#!/bin/sh
# tools/build_variant.sh NAME [-DFLAG ...]: experiment builds of libvr5.so into volume-rendering_b200/build/var/NAME.so
# (git-ignored; travels to the GPU box, where a sweep copies one over libvr5.so per run)
set -e
cd "$(dirname "$0")/../volume-rendering_b200/csrc"
name=$1; shift
mkdir -p ../build/var/$name.o
F="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -ftz=false -prec-div=true -prec-sqrt=true -Xcompiler -fPIC $*"
for f in vr5_render.cu vr5_synth.cu vr5_prep.cu vr5_peer.cu vr5_hostring.cu vr5_scene.cpp vr5_volio.cpp; do nvcc $F -c $f -o ../build/var/$name.o/$f.o & done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../build/var/$name.so ../build/var/$name.o/*.o -lcudart
echo built $name
