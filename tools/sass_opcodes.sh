#!/bin/bash
# SASS opcode summary of the in-tree library (evidence that the GEMM is tcgen05 / TMEM / TMA): bash tools/sass_opcodes.sh > profiles/rNN_sass_opcodes.txt
LIB=oc_cleanrl_b200/csrc/libocppo.so
T=$(mktemp); cuobjdump -sass $LIB > $T 2>/dev/null
echo "# SASS opcode evidence of $LIB (sm_100a)"
echo "# command: cuobjdump -sass $LIB | grep -cE '(^|[^A-Z])<mnemonic>'   (B200_PROFILING.md: the mnemonics that prove tcgen05 / TMEM / TMA)"
for m in UTCHMMA UTCHMMA.2CTA UTCQMMA LDTM STTM UTMALDG UTMALDG.2D UTMALDG.3D UTMASTG UTMASTG.2D UTMAPF UTCBAR UTCBAR.2CTA.MULTICAST UTCATOMSWS SYNCS LDGSTS HMMA IMMA FFMA DFMA; do
  printf "%-24s %s\n" "$m" "$(grep -cE "(^|[^A-Z])${m//./\\.}" $T)"
done
echo
echo "# tcgen05.mma (UTCHMMA) instructions per kernel:"
awk '/Function :/{f=$3} /UTCHMMA/{c[f]++} END{for(k in c) print c[k], k}' $T | sort -rn | sed 's/_ZN5ocppo[0-9a-zA-Z_]*tc_gemm_kernelILi\([0-9]\)ELb\([01]\)ELi\([0-9]\)E.*/tc_gemm_kernel<MODE=\1 (0 FWD, 1 DGRAD, 2 WGRAD), SPLIT=\2, CORR=\3>/'
rm -f $T
