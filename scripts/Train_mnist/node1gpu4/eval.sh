#!/bin/bash
# evaluate the last checkpoint of node1gpu4 on the MNIST test split
DIST=${DIST:-$(cd "$(dirname "$0")/../../.." && pwd)}
python3 ${DIST}/hetseq_9cme_b200/eval_mnist.py --model_ckpt node1gpu4/checkpoint_last.pt \
  --mnist_dir ${DATA:-$DIST/mnist_data} 2>&1 | tee node1gpu4_eval.log
