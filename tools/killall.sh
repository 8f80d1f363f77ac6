#!/usr/bin/env bash
# Stop the ranks started by scripts/launch.sh -- by the PIDs it recorded, never by process name
# (the reference's tools/killall.sh runs `killall python` on every host).
PIDS="$(dirname "$0")/pids"
while read -r host pid; do [ -n "$pid" ] && ssh "$host" "kill -- -$pid 2>/dev/null || kill $pid"; done < "$PIDS"
