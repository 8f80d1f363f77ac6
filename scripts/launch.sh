#!/usr/bin/env bash
# ssh fan-out over the hosts listed in tools/hosts (one per line), like the reference's src/launch.sh.
# Records the remote PIDs in tools/pids so tools/killall.sh can stop exactly those processes.
MASTER_IP="$1"; HOSTS="${HOSTS_FILE:-$(dirname "$0")/../tools/hosts}"; DIR="$(cd "$(dirname "$0")/.." && pwd)"
N=$(grep -c . "$HOSTS"); i=0; : > "$DIR/tools/pids"
while read -r host; do
  [ -z "$host" ] && continue
  pid=$(ssh "$host" "cd $DIR && nohup bash scripts/run_pytorch_dist.sh $i $N $MASTER_IP >/dev/null 2>&1 & echo \$!")
  echo "$host $pid" >> "$DIR/tools/pids"; i=$((i+1))
done < "$HOSTS"
