#!/bin/bash
# The reference's loss x model grid (sweeps/experiment_synthetic.sh: model=small,medium,large x loss=mse,nll,combined x
# trainer=fast,slow) on the B200 path: trainer=b200 with the clip / epoch settings of `fast` and `slow`, one job per GPU.
python3 sweeps/run_sweep.py --gpus "${B200_SWEEP_GPUS:-8}" -- \
  datamodule=synthetic \
  model=small,medium,large \
  loss=mse,nll,combined \
  trainer=b200 \
  trainer.max_epochs=4,32 "$@"
