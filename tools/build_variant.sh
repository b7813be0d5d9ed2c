#!/bin/bash
# tools/build_variant.sh <name> [-DFLAG ...] : kg_omega.cu compiled with extra flags, linked with the tree's other objects into
# _variants/libkgcuda_<name>.so (git-ignored; travels to the GPU box). Run a variant with KGCUDA_LIB=_variants/libkgcuda_<name>.so.
set -e
name=$1; shift
cd "$(dirname "$0")/.."
mkdir -p _variants
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC "$@" -c -o _variants/kg_omega_$name.o kineticgas_b200/csrc/kg_omega.cu
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o _variants/libkgcuda_$name.so _variants/kg_omega_$name.o \
    kineticgas_b200/csrc/kg_transport.o kineticgas_b200/csrc/kg_genpot.o kineticgas_b200/csrc/kg_table.o kineticgas_b200/csrc/kg_eos.o
echo built _variants/libkgcuda_$name.so
