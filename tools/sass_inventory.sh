#!/bin/bash
# SASS inventory of libflagg.so: per kernel family (template instantiations summed), the Blackwell-native
# instructions that prove the tensor-core / TMA / TMEM path (B200_PROFILING.md): UTCHMMA (tcgen05.mma), LDTM / STTM
# (tcgen05.ld / st), UTMALDG (TMA tensor loads; .GATHER4 = tile::gather4), LDGSTS (cp.async), FMNMX (K1 networks).
# No GPU needed:  tools/sass_inventory.sh > profiles/r02_sass_inventory.txt
SO=${1:-multimodal-fl-security_b200/flagg/libflagg.so}
cuobjdump -sass "$SO" 2>/dev/null | awk '
/Function : / { fn=$3; if (match(fn, /[a-z_0-9]*_kernel/)) fn=substr(fn, RSTART, RLENGTH); inst[fn]++; next }
/UTCHMMA/ {a[fn]++} /LDTM/ {b[fn]++} /STTM/ {c[fn]++} /UTMALDG/ {d[fn]++} /GATHER4/ {g[fn]++} /LDGSTS/ {e[fn]++} /FMNMX/ {f[fn]++}
END { printf "%-34s %6s %8s %6s %6s %8s %8s %7s %8s\n","kernel family","inst.","UTCHMMA","LDTM","STTM","UTMALDG","GATHER4","LDGSTS","FMNMX";
      for (k in inst) printf "%-34s %6d %8d %6d %6d %8d %8d %7d %8d\n", k, inst[k], a[k],b[k],c[k],d[k],g[k],e[k],f[k] }' | (read -r hdr; echo "$hdr"; sort)
