#!/bin/bash
# Full SASS listing of every kernel object (cuobjdump -sass, hex encodings stripped, gzip) plus a per-kernel
# table of the mnemonics that prove the Blackwell-native paths (B200_PROFILING.md "What proves a
# Blackwell-native kernel").  Runs on the CPU box: the objects are cross-compiled for sm_100a.
#   bash scripts/dump_sass.sh        -> profiles/sass/<object>.sass.gz, profiles/sass_evidence.txt
set -eu
cd "$(dirname "$0")/.."
python -m pipe_b200.ops._build > /dev/null
mkdir -p profiles/sass
out=profiles/sass_evidence.txt
{
  echo "# SASS evidence per kernel (cuobjdump -sass of pipe_b200/_native/*.o, sm_100a; full listings: profiles/sass/*.sass.gz)"
  echo "# columns: kernel | UTC*MMA (tcgen05.mma) | LDTM/STTM (tcgen05.ld/st) | UTMALDG/UTMASTG (TMA) | HMMA (mma.sync) | LDGMC/multimem | total instructions"
} > $out
for obj in pipe_b200/_native/*.o; do
  name=$(basename $obj .o)
  cuobjdump -sass $obj | sed -E 's@[[:space:]]*/\* 0x[0-9a-f]+ \*/@@' > /tmp/$name.sass
  gzip -9 -c /tmp/$name.sass > profiles/sass/$name.sass.gz
  python - "$name" /tmp/$name.sass >> $out <<'PY'
import re, sys
name, path = sys.argv[1], sys.argv[2]
cur, rows = None, {}
for line in open(path):
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); rows[cur] = dict(mma=0, tm=0, tma=0, hmma=0, mc=0, n=0); continue
    if cur is None: continue
    m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if not m: continue
    op = m.group(1); r = rows[cur]; r["n"] += 1
    if re.match(r"UTC[A-Z]*MMA", op): r["mma"] += 1
    if op.startswith(("LDTM", "STTM")): r["tm"] += 1
    if op.startswith(("UTMALDG", "UTMASTG", "UBLKCP")): r["tma"] += 1
    if op.startswith("HMMA"): r["hmma"] += 1
    if "LDGMC" in op or "MULTIMEM" in op or op.startswith(("STGMC", "REDGMC")): r["mc"] += 1
import subprocess
for k, r in rows.items():
    try:
        dem = subprocess.run(["cu++filt", k], capture_output=True, text=True).stdout.strip().split("(")[0]
    except Exception:
        dem = k
    print(f"{name}.o | {dem} | {r['mma']} | {r['tm']} | {r['tma']} | {r['hmma']} | {r['mc']} | {r['n']}")
PY
done
cat $out
