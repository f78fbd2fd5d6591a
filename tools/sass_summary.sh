#!/bin/bash
# Per-kernel SASS evidence for the shipped library: which kernels issue FP64 tensor-core MMAs (DMMA), 1-D bulk TMA
# (UBLKCP) with mbarriers (SYNCS), cp.async (LDGSTS), tensor-map TMA (UTMALDG) or tcgen05 (UTC*MMA / LDTM — none: tcgen05
# has no f64 kind).  Usage: tools/sass_summary.sh [lib] > profiles/<round>_sass_summary.txt
LIB=${1:-pls_b200/libplsb.so}
echo "# $(basename $LIB): instruction counts per kernel (cuobjdump -sass, sm_100a); registers from cuobjdump -res-usage"
echo "# kernel | REG | DMMA | DFMA | UBLKCP | SYNCS | LDGSTS | UTMALDG | UTCMMA+LDTM | total"
cuobjdump -res-usage "$LIB" 2>/dev/null | awk '/Function/ {f=$2; sub(/:$/,"",f)} /REG:/ {for(i=1;i<=NF;i++) if ($i ~ /^REG:/) {r=$i; sub("REG:","",r); print f, r}}' > /tmp/plsb_regs.$$
cuobjdump -sass "$LIB" | awk '
  /Function :/ { if (f != "") out(); f=$3; d=df=u=s=l=t=tc=n=0; next }
  /^[ \t]+\/\*[0-9a-f][0-9a-f][0-9a-f][0-9a-f][0-9a-f]*\*\// { n++; if ($0 ~ /DMMA/) d++; if ($0 ~ / DFMA/) df++; if ($0 ~ /UBLKCP/) u++; if ($0 ~ /SYNCS/) s++;
                           if ($0 ~ /LDGSTS/) l++; if ($0 ~ /UTMALDG/) t++; if ($0 ~ /UTC[A-Z]*MMA|LDTM/) tc++ }
  function out() { print f, d, df, u, s, l, t, tc, n }
  END { if (f != "") out() }' | while read f d df u s l t tc n; do
    r=$(grep -m1 "^$f " /tmp/plsb_regs.$$ | awk '{print $2}')
    echo "$(echo $f | c++filt | sed 's/(.*//') | ${r:-?} | $d | $df | $u | $s | $l | $t | $tc | $n"
done | sort
rm -f /tmp/plsb_regs.$$
