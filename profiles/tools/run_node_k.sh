#!/bin/bash
# node-loop exit policy (GPT_WF_NODE_K) on the mid-size and large configs; K=0 = the library's own choice
cd "$(dirname "$0")/../.."
for k in 0 3 6 10 16; do
  echo "== K=$k"
  if [ $k = 0 ]; then unset GPT_WF_NODE_K; else export GPT_WF_NODE_K=$k; fi
  python profiles/tools/configs_bench.py "$@" 2>/dev/null | python profiles/tools/summ_configs.py
done
