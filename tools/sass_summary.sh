#!/bin/bash
# SASS evidence per kernel (the built .so is git-ignored): counts of the Blackwell-native instructions
#   UTCHMMA  tcgen05.mma        LDTM  tcgen05.ld (TMEM -> registers)     UTCBAR  tcgen05.commit
#   UTMALDG  cp.async.bulk.tensor (TMA tiled load)     UBLKCP  cp.async.bulk (1-D bulk copy)     SYNCS  mbarrier ops
# Usage: tools/sass_summary.sh > profiles/rNN_sass_summary.txt
LIB=${1:-localai_tfp_b200/libtfpstore.so}
echo "# cuobjdump -sass $LIB ($(date -u +%Y-%m-%d)); nvcc $(nvcc --version | grep release | sed 's/.*release //')"
echo "# kernel | registers | UTCHMMA | UTCHMMA.2CTA | LDTM | UTCBAR(.MULTICAST) | UTMALDG | UBLKCP | SYNCS | DADD | FADD | HMMA/IMMA/legacy mma"
cuobjdump -sass "$LIB" 2>/dev/null | awk '
  /Function : / { if (name != "") flush(); name = $3; delete c; next }
  { for (k in pat) if ($0 ~ pat[k]) c[k]++ }
  function flush() {
      printf "%s | - | %d | %d | %d | %d(%d) | %d | %d | %d | %d | %d | %d\n", name, c["mma"], c["mma2"], c["ldtm"], c["bar"], c["barm"], c["tma"], c["blk"], c["syncs"], c["dadd"], c["fadd"], c["old"]
  }
  BEGIN { pat["mma"]="UTCHMMA"; pat["mma2"]="UTCHMMA\\.2CTA"; pat["ldtm"]="LDTM"; pat["bar"]="UTCBAR"; pat["barm"]="UTCBAR.*MULTICAST";
          pat["tma"]="UTMALDG"; pat["blk"]="UBLKCP"; pat["syncs"]="SYNCS"; pat["dadd"]=" DADD"; pat["fadd"]=" FADD"; pat["old"]=" (HMMA|IMMA|DMMA|HGMMA)" }
  END { if (name != "") flush() }' | c++filt | sort | grep -v "^_ZN3cub\|cub::" 
echo "# registers / shared memory per kernel (cuobjdump --dump-resource-usage)"
cuobjdump --dump-resource-usage "$LIB" 2>/dev/null | grep -A1 "Function" | grep -E "Function|REG" | paste - - | \
  sed -E 's/ *Function ([^:]*):.*REG:([0-9]+).*SHARED:([0-9]+).*/\1 REG=\2 STATIC_SHARED=\3/' | c++filt | grep -v "cub::" | sort
