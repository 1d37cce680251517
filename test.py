"""Drop-in for the reference's evaluation script (test.py:147-218): `python test.py checkpoint=<path> [group=name ...]`.

Loads a checkpoint written by `train.py trainer=b200` (or by the reference: same keys), runs the model over the test
split on the device next to the look-back OLS baseline and logs the reference's result families into the run's
TensorBoard directory.  Figures need matplotlib, which is not installed here: histograms, scalars and
`test_results.npz` are written instead (masters-thesis_b200/evaluate.py).
"""
import sys
from pathlib import Path

import torch

import masters_thesis_b200 as b200
from masters_thesis_b200.config import compose_config

CONFIG_DIR = Path(__file__).resolve().parent / "configs"


def run(cfg):
    if not cfg.checkpoint:
        print("No checkpoint provided")
        return None
    print("\n" + "=" * 70)
    print("EXPERIMENT CONFIGURATION")
    print("=" * 70)
    print(f"Datamodule: {cfg.datamodule.name}")
    print(f"Model: {cfg.model.name}")
    print(f"Checkpoint: {cfg.checkpoint}")
    print("=" * 70 + "\n")

    tcfg = cfg.trainer
    seed = int(tcfg.get("seed", 1234))
    dm = b200.FinancialLstmDataModule(
        data_dir=Path(cfg.datamodule.data_dir), lookback_window=cfg.datamodule.lookback_window,
        target_window=cfg.datamodule.target_window, stride=cfg.datamodule.stride,
        prediction_task=cfg.datamodule.prediction_task, interaction_only=cfg.datamodule.interaction_only,
        batch_size=cfg.datamodule.batch_size, seed=seed, n_samples=int(tcfg.get("n_samples", 1_000_000)))

    ModelClass = b200.get_model_class(cfg.loss.module_class)
    model = ModelClass.load_from_checkpoint(cfg.checkpoint, map_location='cpu')
    print(model)

    checkpoint = Path(cfg.checkpoint)
    run_dir = checkpoint.parent.parent if checkpoint.parent.name == "checkpoints" else checkpoint.parent   # test.py:182-185

    print("\n" + "=" * 50)
    print("Starting Testing...")
    print("=" * 50 + "\n")
    dm.prepare_data()
    dm.setup('test')
    mode = b200.trainer._precision_to_mode(tcfg.get("precision", "32-true"), tcfg.get("engine", "auto"))
    hp = dm.hparams
    model.eval()
    model.attach(dm.device, mode, hp["lookback_window"], hp["target_window"], dm.n_stocks)
    results = b200.evaluate.collect_test_results(model, dm, int(tcfg.get("eval_batch_windows", 256)),
                                                 tcfg.get("limit_test_windows"))
    writer = None
    try:
        from torch.utils.tensorboard import SummaryWriter
        writer = SummaryWriter(str(run_dir))
    except Exception:
        pass
    summary = b200.evaluate.write_results(results, run_dir, writer)
    for k, v in summary.items():
        print(f"{k}: {v:.6f}")
    print("\n" + "=" * 50)
    print("Testing Complete!  (figures: matplotlib not installed; see test_results.npz and the TensorBoard histograms)")
    print("=" * 50)
    return summary


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    return run(compose_config(CONFIG_DIR, "config", [a for a in argv if a not in ("-m", "--multirun")]))


if __name__ == '__main__':
    main()
