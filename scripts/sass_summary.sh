#!/bin/bash
# Per-kernel SASS evidence of the Blackwell-native instructions (tcgen05.mma = UTC*MMA, tcgen05.ld = LDTM,
# TMA = UTMALDG / UTMASTG (tensor) and UBLKCP (cp.async.bulk, 1-D), tcgen05.commit = UTCBAR, mma.sync = HMMA) and the ptxas register / spill report.
# Usage: scripts/sass_summary.sh > profiles/sass_summary.txt   (CPU only: cuobjdump / cuobjdump --dump-resource-usage)
set -e
SO=${1:-self-attention-tacotron_b200/libsat.so}
echo "# cuobjdump -sass $SO : instruction counts per kernel"
printf "%-70s %8s %6s %8s %7s %7s %6s %7s\n" kernel UTCxMMA LDTM UTMALDG UBLKCP UTCBAR HMMA total
cuobjdump -sass "$SO" | awk '
  /Function :/ { if (name != "") flush(); name=$3; mma=ld=tma=blk=bar=h=tot=0; next }
  /^ +\/\*[0-9a-f]+\*\/ / { tot++;
     if ($0 ~ /UTC[A-Z]*MMA/) mma++; if ($0 ~ /LDTM/) ld++; if ($0 ~ /UTMALDG/) tma++; if ($0 ~ /UBLKCP/) blk++;
     if ($0 ~ /UTCBAR/) bar++; if ($0 ~ / HMMA/) h++; }
  function flush() { cmd="echo " name " | c++filt | cut -c1-68"; cmd | getline dn; close(cmd);
     printf "%-70s %8d %6d %8d %7d %7d %6d %7d\n", dn, mma, ld, tma, blk, bar, h, tot }
  END { flush() }' | sort -k8 -n -r
echo
echo "# cuobjdump --dump-resource-usage $SO : registers / shared / stack (spills) per kernel"
cuobjdump --dump-resource-usage "$SO" | awk '/Function/ {f=$2} /REG:/ {cmd="echo " f " | tr -d : | c++filt | cut -c1-68"; cmd | getline dn; close(cmd); print dn, $0}'
