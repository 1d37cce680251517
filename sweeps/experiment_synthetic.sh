#!/bin/bash
# The reference's loss x model grid (sweeps/experiment_synthetic.sh: model=small,medium,large x loss=mse,nll,combined x
# trainer=fast,slow) on the B200 path: trainer=b200_fast,b200_slow carry the epochs, clip value and validation cadence of
# configs/trainer/fast.yaml and slow.yaml (4 epochs / clip 5.0; 32 epochs / clip 2.0 / validation every 4 epochs), one job per GPU.
python3 sweeps/run_sweep.py --gpus "${B200_SWEEP_GPUS:-8}" -- \
  datamodule=synthetic \
  model=small,medium,large \
  loss=mse,nll,combined \
  trainer=b200_fast,b200_slow "$@"
