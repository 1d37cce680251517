#!/bin/bash
# Shape of the reference's sweeps/example.sh (model=large, 3 learning rates x 2 epoch budgets).  The reference runs it
# on datamodule=real; the Fama-French CSVs are not shipped, so this runs the same grid on the synthetic datamodule.
python3 sweeps/run_sweep.py --gpus "${B200_SWEEP_GPUS:-8}" -- \
  datamodule=synthetic \
  model=large \
  model.learning_rate=0.001,0.0001,0.00001 \
  trainer=b200 \
  trainer.max_epochs=100,200 "$@"
