#!/bin/bash
# SASS evidence for the TMA-fed DMMA GEMMs: per kernel instantiation the count of the mnemonics that prove the data
# path (DMMA.8x8x4 = FP64 tensor op, UTMALDG = TMA tensor load, SYNCS.* = mbarrier), then the K-loop body of the Yvo kernel.
#   bash tools/sass_excerpt.sh > profiles/sass_gemm_tma.txt
SO=pyscf_mpcc_b200/csrc/libmpcc_b200.so
echo "# cuobjdump -sass $SO  ($(nvcc --version | tail -2 | head -1))"
cuobjdump -sass $SO | awk '
/Function :/ {f=$3}
f ~ /gemm_tma_kernel/ && /\/\*[0-9a-f]+\*\// {
  for (i=2;i<=NF;i++) if ($i !~ /^@/ ) { m=$i; break }
  sub(/;$/,"",m)
  if (m ~ /^(DMMA|UTMALDG|UTMASTG|SYNCS|UTCHMMA|UTCQMMA|LDTM|LDS|STS|STG|LDG|BAR|DFMA|DADD|DMUL|MUFU)/) c[f" "m]++
}
END { for (k in c) print c[k], k }' | sort -k2,2 -k3,3 | c++filt | awk '{n=$1; $1=""; printf "%6d %s\n", n, $0}'
echo
echo "# K-loop body of gemm_tma_kernel<0,4> (Yvo): first DMMA to the loop-closing branch"
cuobjdump -sass $SO | awk '/Function :/ {on = ($3 ~ /gemm_tma_kernelILi0ELi4E/)} on && /\/\*[0-9a-f][0-9a-f][0-9a-f][0-9a-f]\*\// {print}' \
  | sed -e 's#/\* 0x[0-9a-f]* \*/##' | awk '/SYNCS.PHASECHK/ && !s {s=1} s {print} /DMMA/ {d++} d>=64 && /BRA/ {exit}' | head -150
