#!/bin/bash
# profiles/<round>_sass_excerpts.txt from the built library: bash profiles/sass_excerpts.sh r02
R=${1:-r02}; SO=acclimate_b200/libacclimate_b200.so; OUT=profiles/${R}_sass_excerpts.txt; TMP=$(mktemp)
cuobjdump -sass $SO > $TMP
{
echo "# cuobjdump -sass $SO (sm_100a), excerpts -- round ${R#r}, last build of the round"
echo "# seller_kernel<SellerProduction>: the demand book is staged by the copy engine: mbarrier init (SYNCS.EXCH), expected bytes"
echo "# (SYNCS.ARRIVE.TRANS64), four bulk copies global -> shared (UBLKCP.S.G = cp.async.bulk.shared::cluster.global.mbarrier), wait"
echo "# (SYNCS.PHASECHK.TRANS64.TRYWAIT); UBLKPF.L2 = cp.async.bulk.prefetch.L2 (the book of the seller half a wave further on)"
awk '/Function : .*seller_kernelINS_16SellerProduction/{f=1; print; next} /Function :/{f=0} f && /SYNCS|UBLKCP|UBLKPF/' $TMP
echo
echo "# block_smem_kernel<Purchase<0,1,1>,96>: the L2 prefetch of the next wave's connection lists (UBLKPF), peer stores of the demand"
echo "# request as 16-byte stores (ST.E.128), barriers and the software divisions (MUFU.RCP64H) by count"
awk '/Function : .*PurchaseILb0ELb1ELb1EEELi96/{f=1; next} /Function :/{f=0} f' $TMP > $TMP.p
grep -E "UBLKPF|ST.E.128" $TMP.p | head -12
for pat in "BAR.SYNC" "MUFU.RCP64H" "DADD" "DMUL" "LDS" "STS" "LDG" "LDC" "LDL" "STL"; do echo "$(grep -c "$pat" $TMP.p) $pat"; done
echo
echo "# cross_barrier_kernel: the cross-part barrier over peer memory (system-scope release store / acquire load)"
awk '/Function : .*cross_barrier_kernel/{f=1; print; next} /Function :/{f=0} f && /\.SYS|MEMBAR|ST\.E|LD\.E|ERRBAR/' $TMP | head -12
echo
echo "# instruction counts per kernel (static SASS lines)"
awk '/Function :/{name=$3} /^ +\/\*[0-9a-f]+\*\/ /{n[name]++} END{for (k in n) print n[k], k}' $TMP | sort -rn | head -24 | cut -c1-150
} > $OUT
rm -f $TMP $TMP.p
wc -l $OUT
