#!/bin/bash
# start every node's script with your launcher (ssh / srun / qsub); NODES = host list, node 0 first
i=0
for h in ${NODES:?set NODES="host0 host1 ..."}; do
  s=$([ $i -eq 0 ] && echo main.sh || echo sub$i.sh)
  ${LAUNCH:-ssh} $h "cd $(pwd) && MASTER=$MASTER DATA=$DATA bash $s" &
  i=$((i+1))
done
wait
