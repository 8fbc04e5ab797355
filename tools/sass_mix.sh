#!/bin/bash
# tools/sass_mix.sh <T> <G> <K>: compile one fill instantiation and print the instruction mix of its hot loop
set -e
T=$1; G=$2; K=$3
cd "$(dirname "$0")/../mgl_b200/csrc"
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xptxas -v -DINST_T=$T -DINST_G=$G -DINST_K=$K -c -o /tmp/fill_$T_$G_$K.o sw_fill_inst.cu 2>&1 | grep -E "registers|spill|error" || true
cuobjdump -sass /tmp/fill_$T_$G_$K.o | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed -E 's/^\s+\/\*[0-9a-f]+\*\/\s+//; s/\s+\/\*.*//' > /tmp/fill.sass
first=$(grep -n -m1 "SHFL.UP PT" /tmp/fill.sass | cut -d: -f1)
last=$(grep -n "WARPSYNC.ALL" /tmp/fill.sass | head -1 | cut -d: -f1)
echo "hot loop: lines $first..$last of /tmp/fill.sass ($((last-first)) instructions static)"
sed -n "${first},${last}p" /tmp/fill.sass | awk '{ if ($1 ~ /^@/) print $2; else print $1 }' | sed 's/;//' | sort | uniq -c | sort -rn | head -${4:-16}
