#!/bin/bash
# SASS opcode histogram per object file of libb2ba.so (sm_100a), written under profiles/ (VERDICT r1 "What's weak" 13)
cd "$(dirname "$0")/.."
R=${1:-r2}
out=profiles/${R}_sass_opcodes.txt
: > $out
for o in visualodometry_b200/build/*.o; do
  echo "== $(basename $o) (arch $(cuobjdump -lelf $o | head -1 | sed 's/.*\.sm_/sm_/;s/\.elf.*//'))" >> $out
  cuobjdump -sass $o | grep -E "^\s+/\*[0-9a-f]{4,5}\*/" | sed 's#^\s*/\*[0-9a-f]*\*/\s*##' | sed 's/^@!\?U\?P[0-9T]* *//' | awk '{print $1}' | sed 's/;$//' | cut -d. -f1-2 | sort | uniq -c | sort -rn | awk '{printf "%s:%s  ", $2, $1} END {print ""}' | fold -w 150 >> $out
done
echo "-- tensor / bulk-copy mnemonics over all objects" >> $out
for m in DMMA UTCHMMA UTCQMMA UTMALDG LDGSTS HMMA IMMA; do
  n=$(for o in visualodometry_b200/build/*.o; do cuobjdump -sass $o; done | grep -c "\b$m")
  echo "$m $n" >> $out
done
echo $out
