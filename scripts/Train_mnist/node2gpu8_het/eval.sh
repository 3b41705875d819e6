#!/bin/bash
# evaluate the last checkpoint of node2gpu8_het on the MNIST test split
DIST=${DIST:-$(cd "$(dirname "$0")/../../.." && pwd)}
python3 ${DIST}/hetseq_9cme_b200/eval_mnist.py --model_ckpt node2gpu8_het/checkpoint_last.pt \
  --mnist_dir ${DATA:-$DIST/mnist_data} 2>&1 | tee node2gpu8_het_eval.log
