#!/bin/bash
# counts of the SASS mnemonics that prove the tensor-core / TMA / TMEM / async-copy / reduction paths, per kernel
cuobjdump -sass harp_b200/libharp_b200.so 2>/dev/null | awk '
/Function :/ {fn=$3}
/UTCHMMA|UTCBAR|UTMALDG|UBLKCP|LDTM|UTCATOMSWS|REDG|LDGSTS|SYNCS\.|ATOMG|MEMBAR|ERRBAR|LDGMC|STG.*MC|UCGABAR/ {
  for (i=1;i<=NF;i++) if ($i ~ /^(UTCHMMA|UTCBAR|UTMALDG|UBLKCP|LDTM|REDG|LDGSTS|ATOMG|MEMBAR|ERRBAR|LDGMC|UCGABAR|STG\.[A-Z0-9.]*MC)/) {split($i,a,";"); print fn, a[1]; break}
}' | sort | uniq -c | sort -k2,2 -k3,3
