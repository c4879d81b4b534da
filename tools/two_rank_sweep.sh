#!/bin/bash
# Diagnostic: the headline stream leg on two GPUs at once (disjoint halves of the cores), for several
# socket buffer sizes.  Usage: tools/two_rank_sweep.sh  (on a 2-GPU box)
n=$(nproc); h=$((n/2))
for buf in 4194304 1048576 262144 65536; do
  for tr in tcp ipc; do
    export BT_PUSH_SNDBUF=$buf BT_PULL_RCVBUF=$buf
    python tools/e2e_sweep.py --gpu 0 --cpus 0-$((h-1)) --split "$1" --transport $tr --tag "buf$buf" 2>/dev/null &
    python tools/e2e_sweep.py --gpu 1 --cpus $h-$((n-1)) --split "$1" --transport $tr --tag "buf$buf" 2>/dev/null
    wait
  done
done
