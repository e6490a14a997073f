#!/bin/bash
# VQ role profile (per-role wait / work cycle counters of a -DVK_ROLE_PROF build), then the sanitizer suite
set -u
mkdir -p gpurun_out
T=${1:-r02k}
echo "== vq role profile"; VAEKIT_LIB=$PWD/pytorch-vae_b200/vaekit/_lib/libvaekit_prof.so timeout 120 python scripts/prof_vq.py 2>&1 | grep -vE "^trace" | sort | uniq -c | sort -rn | head -12 | tee gpurun_out/role_prof_vq_$T.log
SAN_TIMEOUT=420 bash scripts/sanitize.sh $T
for tool in memcheck racecheck synccheck; do tail -4 gpurun_out/sanitize_${tool}_$T.log | cut -c1-200; done
