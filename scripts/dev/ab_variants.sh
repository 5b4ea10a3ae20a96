#!/bin/bash
# A/B of QP-kernel build variants on the GPU box: build/variants/libihm2b_<name>.so copied over the in-tree library in turn
# (the box copy of the repo is scratch), scripts/profile_qp.py timed on each; CMD (optional) runs after the timing.
for v in "$@"; do
  cp build/variants/libihm2b_$v.so ihm2_b200/lib/libihm2b.so
  echo "== $v"; python scripts/profile_qp.py 4096 5 2>&1 | grep -E "solve|ipm|phase" | tail -3
  if [ -n "$CMD" ]; then bash -c "$CMD"; fi
done
