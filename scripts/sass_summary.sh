#!/bin/bash
# Per-kernel counts of the SASS mnemonics that show which hardware paths the shipped
# library uses (sm_100a only): FP64 tensor cores (DMMA.8x8x4 -- tcgen05 has no FP64 kind),
# TMA bulk copies shared -> global (UBLKCP, committed / waited with UTMACMDFLUSH / DEPBAR),
# 16-byte streaming stores (STG.E.EF.128).   scripts/sass_summary.sh > profiles/r02_sass.txt
LIB=mpc_interface_b200/csrc/libmpci.so
echo "# cuobjdump -sass $LIB (built by csrc/Makefile: -gencode arch=compute_100a,code=sm_100a -lineinfo)"
cuobjdump -lelf $LIB | sed 's/^/# /'
echo "# kernel | instructions | DMMA.8x8x4 | UBLKCP (bulk S2G) | UTMACMDFLUSH | STG.E.EF.128 | STG.E.EF.64 | LDS.128 | SHFL"
cuobjdump -sass $LIB | awk '
  /Function :/ { if (name != "") print name " | " n " | " dmma " | " ublk " | " flush " | " st128 " | " st64 " | " lds128 " | " shfl;
                 name=$3; n=0; dmma=0; ublk=0; flush=0; st128=0; st64=0; lds128=0; shfl=0; next }
  /^ +\/\*[0-9a-f]+\*\/ / { n++ }
  /DMMA\.8x8x4/ { dmma++ } /UBLKCP/ { ublk++ } /UTMACMDFLUSH/ { flush++ }
  /STG\.E\.EF\.128/ { st128++ } /STG\.E\.EF\.64/ { st64++ } /LDS\.128/ { lds128++ } /SHFL/ { shfl++ }
  END { print name " | " n " | " dmma " | " ublk " | " flush " | " st128 " | " st64 " | " lds128 " | " shfl }' | c++filt | sed 's/(anonymous namespace):://g'
echo "# excerpt: the DMMA loop of hessian_rows / hessian_scan in lti_assemble_kernel<1, 0> (first 12 DMMA lines with context)"
cuobjdump -sass $LIB | awk '/Function :/ {f = ($0 ~ /lti_assemble_kernelILi1ELi0/)} f' | grep -E "DMMA|UBLKCP|STG.E.EF|LDS.128|SHFL.DOWN" | head -24
