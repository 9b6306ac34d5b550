#!/bin/bash
# SASS evidence for profiles/: which instructions the built library contains (bulk copies, mbarrier waits, warp reductions, ...).
# usage: bash scripts/sass_counts.sh > profiles/sass_rNN.txt
SO=headlamp_intel_gpu_plugin_b200/csrc/libigpu_rollup.so
echo "# cuobjdump -sass $SO | grep -c <mnemonic>   ($(date -u +%F), nvcc $(nvcc --version | grep -o 'release [0-9.]*'))"
cuobjdump -lelf $SO | sed 's/^/# /'
cuobjdump -sass $SO > /tmp/sass_all.txt
for m in UBLKCP UBLKPF SYNCS REDUX "LDS.128" "LDG.E.128" ATOMS "REDG.E" "ATOMG" "STG.E.64" DADD "F2F.F64.F32" UTMALDG tcgen05; do
  printf "%-14s %s\n" "$m" "$(grep -c -- "$m" /tmp/sass_all.txt)"
done
echo "# per kernel (Function : name, then counts of UBLKCP / SYNCS / REDUX / ATOMS / RED)"
awk '/Function :/ {if (name) printf "%-110s UBLKCP %3d SYNCS %3d REDUX %3d ATOMS %3d RED %3d\n", name, u, s, r, a, d; name=$3; u=s=r=a=d=0}
     /UBLKCP/ {u++} /SYNCS/ {s++} /REDUX/ {r++} /ATOMS/ {a++} /REDG/ {d++}
     END {printf "%-110s UBLKCP %3d SYNCS %3d REDUX %3d ATOMS %3d RED %3d\n", name, u, s, r, a, d}' /tmp/sass_all.txt | sed 's/_ZN[0-9]*_GLOBAL__N__[0-9a-f_]*[a-z_]*cu_[0-9a-f]*//' | cut -c1-190
