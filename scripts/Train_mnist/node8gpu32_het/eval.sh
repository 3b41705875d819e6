#!/bin/bash
# evaluate the last checkpoint of node8gpu32_het on the MNIST test split
DIST=${DIST:-$(cd "$(dirname "$0")/../../.." && pwd)}
python3 ${DIST}/hetseq_9cme_b200/eval_mnist.py --model_ckpt node8gpu32_het/checkpoint_last.pt \
  --mnist_dir ${DATA:-$DIST/mnist_data} 2>&1 | tee node8gpu32_het_eval.log
