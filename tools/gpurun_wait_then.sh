#!/bin/bash
# usage: tools/gpurun_wait_then.sh <logfile> <gpurun args...>: waits for a call still in flight, then runs with retries
while /usr/local/graft/bin/gpurun --status 2>/dev/null | grep -q '"in_flight": 1'; do sleep 20; done
exec "$(dirname "$0")/gpurun_retry.sh" "$@"
