#!/bin/bash
# build the library of a git revision (default HEAD) into tools/prev.bin for same-box A/B runs (tools/gpu_ab.sh)
rev=${1:-HEAD}
d=$(mktemp -d)
mkdir -p $d/grl_b200/csrc $d/include
for f in $(git ls-tree --name-only $rev grl_b200/csrc/ | grep -E "\.(cu|cuh|h)$"); do git show $rev:$f > $d/$f; done
git show $rev:include/grlcu.h > $d/include/grlcu.h
(cd $d/grl_b200/csrc && nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -lineinfo -std=c++17 -Xcompiler -fPIC -shared -o "$OLDPWD/tools/prev.bin" grlcu.cu)
rc=$?
rm -rf $d
exit $rc
