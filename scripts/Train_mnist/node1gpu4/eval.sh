#!/bin/bash
DIST=${DIST:-$(cd "$(dirname "$0")/../../.." && pwd)}
python3 ${DIST}/hetseq_9cme_b200/eval_mnist.py --model_ckpt mnist_node1gpu4/checkpoint_last.pt --mnist_dir ${DATA:-$DIST/mnist_data}
