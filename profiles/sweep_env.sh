#!/bin/bash
# A/B runs of the bench workload under different environment switches: `bash profiles/sweep_env.sh "A=1" "B=2 C=3" ...`
# prints the engine's GPU timeline line (EMRF_DEBUG_TIMERS) of the last of three fits for each setting.
for e in "$@"; do
  echo "== $e"
  env $e EMRF_DEBUG_TIMERS=1 timeout 120 python profiles/profile_fit.py 3 2>&1 | grep -E "GPU timeline" | tail -1
done
