#!/usr/bin/env bash
# Tally of the SASS mnemonics that characterise libb200mt.so (which memory instructions the kernels use, that the
# Adam-family math is MUFU-based, that no tensor-core instruction appears -- by design: <= 3 flop/byte).
#   bash tools/sass_tally.sh > profiles/sass_r02.txt
set -euo pipefail
LIB=${1:-rocm-apex_b200/libb200mt.so}
TMP=$(mktemp)
cuobjdump -sass "$LIB" > "$TMP"
echo "# cuobjdump -sass $LIB  ($(date -u +%Y-%m-%dT%H:%MZ), $(nvcc --version | tail -1))"
echo "# arch: $(grep -m1 -o 'arch = sm_[0-9a]*' "$TMP")   kernels (Function :): $(grep -c 'Function :' "$TMP")"
echo "# count  mnemonic   (STG.E.*.STRONG.SYS = multimem.st: the multicast address makes the switch replicate it)"
for pat in 'LDG\.E\.NA\.128' 'STG\.E\.NA\.128' 'LDG\.E\.NA\.ENL2\.256' 'STG\.E\.NA\.ENL2\.256' 'LDG\.E\.NA\.64' 'STG\.E\.NA\.64' \
           'LDGMC\.E\.HPADD\.[A-Z0-9]*' 'STG\.E\.(64|128)\.STRONG\.SYS' 'UBLKCP' 'SYNCS' 'PREEXIT' 'ACQBULK' \
           'MUFU\.RCP' 'MUFU\.SQRT' 'MUFU\.RSQ' 'VOTE' 'SHFL' 'HMMA' 'UTC[A-Z]*MMA' 'IMMA' 'WGMMA|HGMMA' 'STL' 'LDL'; do
  printf "%7d  %s\n" "$(grep -cE "$pat" "$TMP" || true)" "$pat"
done
echo "# distinct multicast / bulk-copy opcodes:"
grep -oE '(LDGMC|UBLKCP|STG\.E\.[A-Z0-9.]*MC[A-Z0-9.]*|ST\.E\.[A-Z0-9.]*MC[A-Z0-9.]*)[A-Z0-9.]*' "$TMP" | sort | uniq -c | sort -rn | head -12
rm -f "$TMP"
