#!/bin/bash
# profiles/<tag>_sass.txt: per-kernel counts of the SASS mnemonics that prove tcgen05 / TMEM / TMA (B200_PROFILING.md)
cd "$(dirname "$0")/.."
tag=${1:-r02}
so=nldsc_b200/libnldsc_ld.so
out=profiles/${tag}_sass.txt
{
  echo "# cuobjdump -sass $so ($(date -u +%F)): per-kernel counts; UTCIMMA = tcgen05.mma.kind::i8, UTCOMMA = tcgen05.mma.kind::mxf4 (fp4, block-scaled), LDTM / STTM = tcgen05.ld / st,"
  echo "# UTMALDG = cp.async.bulk.tensor (TMA load), UTCBAR = tcgen05.commit, SYNCS = mbarrier, DFMA/DADD/DMUL = FP64 epilogue"
  cuobjdump -sass $so | awk '
    /Function : /{name=$3}
    /UTCIMMA/{c[name,"UTCIMMA"]++} /UTCOMMA/{c[name,"UTCOMMA"]++} /STTM/{c[name,"STTM"]++} /LDTM/{c[name,"LDTM"]++} /UTMALDG/{c[name,"UTMALDG"]++} /UTCBAR/{c[name,"UTCBAR"]++}
    /SYNCS/{c[name,"SYNCS"]++} /DFMA|DADD|DMUL/{c[name,"FP64"]++} /IDP.4A|IDP4A/{c[name,"IDP4A"]++} /STG|LDG/{c[name,"LDG_STG"]++}
    /Function : /{names[name]=1}
    END{for(n in names){printf "%-70s", n; split("UTCIMMA UTCOMMA LDTM STTM UTMALDG UTCBAR SYNCS FP64 IDP4A LDG_STG",k," "); for(i=1;i<=10;i++) printf " %s=%d", k[i], c[n,k[i]]+0; printf "\n"}}' | sort
  echo
  echo "# variants of the MMA / TMA instructions in the band kernels"
  cuobjdump -sass $so | grep -oE "UTCIMMA[.A-Z0-9_]*|UTCOMMA[.A-Z0-9_]*|STTM[.a-z0-9A-Z_]*|UTMALDG[.A-Z0-9_]*|UTCBAR[.A-Z0-9_]*|LDTM[.a-z0-9A-Z_]*" | sort | uniq -c
  echo
  echo "# resource usage (cuobjdump -res-usage)"
  cuobjdump -res-usage $so 2>/dev/null | grep -A1 "Function" | grep -v "^--" | paste - - | sed 's/ Fatbin.*//' | awk '{$1=$1};1' | cut -c1-220
} > $out
cat $out
