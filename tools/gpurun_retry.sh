#!/bin/sh
# usage: tools/gpurun_retry.sh [gpurun options] -- 'command'
# Re-submits while the pod answers "busy / draining" (nothing is charged for those).
n=0
while [ $n -lt 30 ]; do
  /usr/local/graft/bin/gpurun "$@" > /tmp/gpurun_retry.$$ 2>&1
  rc=$?
  if grep -q "status=transient\|retry in a few minutes\|no box or slot" /tmp/gpurun_retry.$$; then
    n=$((n + 1)); sleep 75; continue
  fi
  cat /tmp/gpurun_retry.$$; rm -f /tmp/gpurun_retry.$$
  exit $rc
done
cat /tmp/gpurun_retry.$$; rm -f /tmp/gpurun_retry.$$
exit 3
