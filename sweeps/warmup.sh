#!/bin/bash
# The thesis' warm-up recipe (tex/diplomski_rad.tex:1134-1147): train with the combined objective on the synthetic
# datamodule, then continue from the best checkpoint on the real datamodule.  The reference keeps a `checkpoint: null` key
# in configs/config.yaml for this but its train.py never reads it; this entry point does (parameters only).
# Stage 2 needs data/real/stocks.pt + market.pt (or the Fama-French CSVs in data/real_raw, which are not shipped).
set -e
MODEL="${1:-small}"
python3 train.py datamodule=synthetic model="$MODEL" loss=combined trainer=b200_slow
CKPT=$(ls -t logs/FinancialLstm/synthetic/combined_"$MODEL"_lr*_b200_slow/checkpoints/best_*.ckpt | head -1)
for LOSS in mse nll combined; do
  python3 train.py datamodule=real model="$MODEL" loss="$LOSS" trainer=b200_slow checkpoint="$CKPT"
done
