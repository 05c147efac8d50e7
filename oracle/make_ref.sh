#!/usr/bin/env bash
# TEST INFRASTRUCTURE.  Installs the UNMODIFIED reference (a pure-Python package) from
# /root/reference into oracle/_ref/ (git-ignored; it travels to the GPU box with the
# gpurun snapshot) so that parity tests, golden-vector generation and the reference arm of
# bench.py can execute the reference's own torch path.  /root/reference is read-only and
# pip needs to write egg-info, so the install goes through a scratch copy in /tmp.
#
# One byte-level repair is applied to the INSTALLED copy (never to /root/reference):
# `vmas/simulator/utils.py:61` re-defines `Color.YELLOW`, which is a TypeError on
# `import vmas` under any Python ("'YELLOW' already defined"); the duplicate line is
# dropped.  `gym` (not in this image) is supplied by oracle/gym_stub at run time.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
SRC="${1:-/root/reference}"
if [ ! -d "$SRC/vmas" ]; then
  echo "make_ref: $SRC not present (GPU box?) - keeping whatever oracle/_ref holds" >&2
  exit 0
fi
TMP="$(mktemp -d /tmp/vmas_src.XXXXXX)"
cp -r "$SRC/." "$TMP/"
chmod -R u+w "$TMP"
rm -rf "$HERE/_ref"
python -m pip install --quiet --no-index --no-build-isolation --no-deps \
  --find-links /opt/wheelhouse --target "$HERE/_ref" "$TMP"
rm -rf "$TMP"
python - "$HERE/_ref/vmas/simulator/utils.py" <<'PY'
import sys, re
p = sys.argv[1]
lines = open(p).read().split("\n")
seen = False
out = []
for ln in lines:
    if re.match(r"\s+YELLOW\s*=", ln):
        if seen:
            continue            # drop the duplicate enum member (utils.py:61)
        seen = True
    out.append(ln)
open(p, "w").write("\n".join(out))
PY
echo "make_ref: installed reference into $HERE/_ref"
