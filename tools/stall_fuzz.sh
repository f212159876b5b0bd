#!/usr/bin/env bash
# Random small federations through tools/stall_hunt.py: nodes 4-11, any train-set size, 2-4 rounds, ring / full / star / line,
# FedAvg / FedMedian / Krum / Scaffold, test or fast preset.  Usage: tools/stall_fuzz.sh [scenarios=14] [runs per scenario=6]
cd "$(dirname "$0")/.."
for i in $(seq 1 "${1:-14}"); do
  n=$((RANDOM % 8 + 4)); ts=$((RANDOM % n + 1)); r=$((RANDOM % 3 + 2))
  topo=$(shuf -n1 -e ring full star line); agg=$(shuf -n1 -e fedavg fedmedian krum scaffold)
  if [ $((RANDOM % 2)) -eq 0 ]; then fast="--fast"; else fast=""; fi
  printf "n=%s train_set=%s rounds=%s %s %s %s: " "$n" "$ts" "$r" "$topo" "$agg" "$fast"
  timeout 900 python tools/stall_hunt.py --nodes "$n" --rounds "$r" --topology "$topo" --train-set "$ts" --aggregator "$agg" --iters "${2:-6}" $fast --stall 8 2>&1 | tail -n 1 | cut -c1-200
done
