#!/bin/bash
# Count the SASS mnemonics that prove tcgen05 / TMA / TMEM / mbarrier / multimem use, per kernel file.
# Usage: scripts/sass_mnemonics.sh > profiles/sass_mnemonics.txt   (after python build_native.py; no GPU needed)
for f in conv_sm100 wgrad_sm100 stem_sm100 allreduce; do
  echo "== csrc/$f.cu"
  cuobjdump -sass build/obj/$f.o | grep -oE "UTCHMMA[A-Z0-9_.]*|UTMALDG[A-Z0-9_.]*|UTCBAR[A-Z0-9_.]*|UTCATOMSWS[A-Z0-9_.]*|LDTM[A-Za-z0-9_.]*|SYNCS[A-Z0-9_.]*|REDG[A-Z0-9_.]*|LDGMC[A-Z0-9_.a-z]*|LDG\.E\.[A-Z0-9_.]*STRONG\.SYS|STG\.E\.[A-Z0-9_.]*STRONG\.SYS|HMMA[A-Z0-9_.]*|HGMMA[A-Z0-9_.]*" \
    | sort | uniq -c | sort -rn
done
