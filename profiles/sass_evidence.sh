#!/bin/bash
# profiles/sass_evidence.sh — which Blackwell / NVLink mechanisms the built objects really contain (run after the build):
# counts of the telling SASS mnemonics per object file. python -c 'import __graft_entry__ as g; g.build()'; bash profiles/sass_evidence.sh
cd "$(dirname "$0")/../flex-gpu-scheduler_b200/csrc" || exit 1
for o in eval_narrow.o kernels.o kernels_narrow.o sequential.o peer_merge.o preempt.o; do
  echo "== $o"
  cuobjdump -sass $o | grep -oE "UBLKCP[.A-Z0-9]*|SYNCS[.A-Z0-9_]*|LDGMC[.A-Z0-9]*|CREDUX[.A-Z0-9]*|UCGABAR[_A-Z]*|ST\.E[.0-9]*\.STRONG\.(GPU|SYS)|LD\.E[.0-9]*\.STRONG\.(GPU|SYS)|MEMBAR[.A-Z]*|IMAD\.HI\.U32" \
    | sort | uniq -c | sort -rn
done
