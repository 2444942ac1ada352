#!/bin/bash
# Tensor-core / TMA / TMEM instructions in the SASS of the in-tree library, per kernel:
#   bash scripts/sass_evidence.sh > profiles/r2_sass_tensor_ops.txt
so=dba-fusion_b200/csrc/libdbaf_b200.so
echo "# cuobjdump -sass $so : counts of UTCHMMA/UTCQMMA (tcgen05.mma), LDTM (tcgen05.ld), UTMALDG/UTMASTG (TMA load/store),"
echo "# UTCBAR (tcgen05.commit), LDGSTS (cp.async), per kernel (kernels without any of them are omitted)"
cuobjdump -sass $so | awk '
  /Function :/ { fn=$3 }
  { for (i = 1; i <= NF; i++) if ($i ~ /^(UTCHMMA|UTCQMMA|UTCIMMA|UTCMMA|LDTM|STTM|UTMALDG|UTMASTG|UTCBAR|LDGSTS|UTCATOMSWS)/) { split($i, a, "."); c[fn " " a[1]]++ } }
  END { for (k in c) print c[k], k }' | sort -k2,2 -k3,3 | while read n fn op; do printf "%-28s %-10s %5d\n" "$(echo $fn | grep -o 'k_[a-z0-9_]*' | head -1)" $op $n; done
