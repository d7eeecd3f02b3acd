#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O /tmp/d
python - <<'PY'
import numpy as np
z=np.load('tests/golden/synth_600_x200.npz')
z['ref_file'].tofile('/tmp/d/ref.fa'); z['coll_file'].tofile('/tmp/d/coll.fa')
open('/tmp/d/list.txt','w').write('/tmp/d/ref.fa\n/tmp/d/coll.fa\n')
z['bwt'].tofile('/tmp/d/want.bwt')
PY
L=$O/diag_cli.log; : > $L
for opt in "" "--no-stream" "--no-host-sa" "--no-host-sa --no-stream"; do
  echo "=== $opt" >> $L
  cms_bwt_b200/cms_bwt $opt -o /tmp/d/out /tmp/d/list.txt 2>&1 | tail -2 >> $L
  cmp /tmp/d/out.bwt /tmp/d/want.bwt >> $L 2>&1 && echo same >> $L
  rm -f /tmp/d/out.bwt
done
echo "=== sanitizer (no-stream)" >> $L
timeout 120 compute-sanitizer --report-api-errors all --show-backtrace yes cms_bwt_b200/cms_bwt --no-stream -o /tmp/d/out /tmp/d/list.txt 2>&1 | grep -v "^====\s*$" | grep -A25 "Program hit" | head -80 >> $L
echo "=== sanitizer (stream)" >> $L
timeout 120 compute-sanitizer --report-api-errors all --show-backtrace yes cms_bwt_b200/cms_bwt -o /tmp/d/out /tmp/d/list.txt 2>&1 | grep -A25 "Program hit" | head -80 >> $L
tail -5 $L
