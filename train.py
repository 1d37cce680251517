"""Drop-in entry point: `python train.py [group=option ...] [a.b=value ...] [-m]`.

Same command line, config groups (configs/**) and printed report as the reference's train.py
(/root/reference/train.py:70-220).  A trainer option from the b200 family (`trainer=b200`, `b200_fast`, `b200_slow`,
`b200_slowest`: the files of configs/trainer/ that carry an `engine:` key) selects the B200-native path: the device
batch generator, fused LSTM/objective/optimiser kernels and B200Trainer of `masters-thesis_b200`.

Any other trainer option (`fast`, `slow`, `slowest` -- `fast` is the default of configs/config.yaml) is the
reference's own Lightning path and is NOT re-implemented here: the untouched command line is handed to the
reference's own train.py (`reference_fallthrough`) when a reference checkout is reachable ($B200_REFERENCE_ROOT, or a
`src/` package next to this file) and lightning + hydra are importable.  Neither is part of this image, so here the
default `python train.py` stops with a message that says so and names `trainer=b200`.

Differences from the reference entry point, all forced by the environment:
  * no import-time dataset bootstrap: the synthetic series is a pure function of (seed, stock, day)
    evaluated on the device, and the real-data CSV bootstrap (train.py:19-28) is skipped because the raw
    files are not shipped;
  * configuration: Hydra composes the job when it is importable (`hydra_main`); otherwise
    masters_thesis_b200.config composes the same groups, overrides, interpolations and `-m` grids on PyYAML;
  * `-m` grids run one job after the other in this process (the reference's joblib launcher is
    configured with n_jobs: 1); sweeps/run_sweep.py spreads a grid over the 8 GPUs of a box.
"""
from __future__ import annotations

import sys
import os
from pathlib import Path

import torch

import masters_thesis_b200 as b200
from masters_thesis_b200.config import compose_config, expand_multirun, split_cli, to_container

CONFIG_DIR = Path(__file__).resolve().parent / "configs"


def banner(title, width=50):
    print("\n" + "=" * width)
    print(title)
    print("=" * width)


def run(cfg):
    if int(os.environ.get("RANK", "0")) != 0:           # like Lightning's rank_zero_only printing
        global print
        print = lambda *a, **k: None
    banner("EXPERIMENT CONFIGURATION", 70)
    print(f"Datamodule: {cfg.datamodule.name}")
    print(f"Model: {cfg.model.name} (hidden={cfg.model.hidden_size}, layers={cfg.model.num_layers})")
    print(f"Loss: {cfg.loss.name} (class={cfg.loss.module_class})")
    if cfg.loss.name == 'combined':
        print(f"  └─ MSE Weight: {cfg.loss.mse_weight}")
    print(f"Trainer: {cfg.trainer.name} (epochs={cfg.trainer.max_epochs})")
    print(f"Learning Rate: {cfg.model.learning_rate}")
    print(f"Logger: {cfg.logger.name}/{cfg.logger.version}")
    print("=" * 70 + "\n")

    tcfg = cfg.trainer
    seed = int(tcfg.get("seed", 1234))
    torch.manual_seed(seed)

    dm = b200.FinancialLstmDataModule(
        data_dir=Path(cfg.datamodule.data_dir), lookback_window=cfg.datamodule.lookback_window,
        target_window=cfg.datamodule.target_window, stride=cfg.datamodule.stride,
        prediction_task=cfg.datamodule.prediction_task, interaction_only=cfg.datamodule.interaction_only,
        batch_size=cfg.datamodule.batch_size, seed=seed, streaming=bool(tcfg.get("streaming", False)),
        steps_per_epoch=tcfg.get("steps_per_epoch"), n_samples=int(tcfg.get("n_samples", 1_000_000)))

    ModelClass = b200.get_model_class(cfg.loss.module_class)
    model_params = dict(input_size=cfg.model.input_size, hidden_size=cfg.model.hidden_size,
                        num_layers=cfg.model.num_layers, dropout=cfg.model.dropout,
                        learning_rate=cfg.model.learning_rate)
    if cfg.loss.name == 'combined':
        model_params['mse_weight'] = cfg.loss.mse_weight
    model = ModelClass(**model_params)
    if cfg.get("checkpoint"):
        # Warm start (configs/config.yaml:15 `checkpoint: null`; thesis tex/diplomski_rad.tex:1134-1147 "zagrijavanje"): continue
        # training -- typically on datamodule=real -- from the parameters of a model trained on the synthetic datamodule.
        # Parameters only: optimiser state, learning-rate schedule and epoch counter start fresh, as in a new run.
        state = torch.load(cfg.checkpoint, map_location="cpu", weights_only=False)
        model.load_state_dict(state["state_dict"] if "state_dict" in state else state)
        print(f"Warm start from checkpoint: {cfg.checkpoint}")
    model.compile()
    print(model)
    print(f"\nUsing Lightning Module: {ModelClass.__name__}\n")

    run_dir = Path(cfg.logger.save_dir) / cfg.logger.name / cfg.logger.version
    extra = {}
    if "check_val_every_n_epoch" in tcfg:
        extra["check_val_every_n_epoch"] = tcfg.check_val_every_n_epoch
    trainer = b200.B200Trainer(
        max_epochs=tcfg.max_epochs, gradient_clip_val=tcfg.gradient_clip_val, precision=tcfg.precision,
        enable_progress_bar=tcfg.enable_progress_bar, enable_model_summary=tcfg.enable_model_summary,
        max_steps=tcfg.get("max_steps", -1), seed=seed, bucket_mb=tcfg.get("bucket_mb", 0.05),
        engine=tcfg.get("engine", "auto"), eval_batch_windows=tcfg.get("eval_batch_windows", 512),
        micro_batch_windows=tcfg.get("micro_batch_windows", 4096),
        checkpoint_dir=str(run_dir / "checkpoints"), log_dir=str(run_dir), **extra)

    banner("Starting Training...")
    trainer.fit(model, datamodule=dm)

    banner("Training Complete!")
    print(f"Best model path: {trainer.checkpoint_callback.best_model_path}")
    print(f"Best validation loss: {trainer.checkpoint_callback.best_model_score:.6f}")
    print("=" * 50 + "\n")

    trainer.test(model, datamodule=dm)

    config_dict = to_container(cfg)        # plain dict (resolved) from either composer
    config_dict.pop('logger', None)
    final = {'test/mae': model.test_results['test_loss']['mae'], 'test/nll': model.test_results['test_loss']['nll'],
             'test/best_val_loss': trainer.checkpoint_callback.best_model_score.item()}
    writer = trainer._writer()
    if writer:
        flat = {}

        def walk(prefix, node):
            for k, v in node.items():
                if isinstance(v, dict):
                    walk(f"{prefix}{k}/", v)
                else:
                    flat[f"{prefix}{k}"] = v if isinstance(v, (int, float, str, bool)) else str(v)
        walk("", config_dict)
        writer.add_hparams(flat, final, run_name=".")
        writer.flush()

    banner("FINAL TEST RESULTS")
    print(f"Test MAE: {model.test_results['test_loss']['mae']:.6f}")
    print(f"Test NLL: {model.test_results['test_loss']['nll']:.6f}")
    print("=" * 50 + "\n")
    return trainer.checkpoint_callback.best_model_score.item()


def is_b200_trainer(cfg):
    return "engine" in cfg.trainer


def reference_root():
    """A checkout of the reference this entry point can hand non-b200 trainers to, or None."""
    cands = [os.environ.get("B200_REFERENCE_ROOT"), str(Path(__file__).resolve().parent)]
    for c in cands:
        if c and (Path(c) / "src" / "model.py").exists() and (Path(c) / "train.py").exists() \
                and Path(c, "train.py").resolve() != Path(__file__).resolve():
            return Path(c)
    return None


def reference_fallthrough(argv, trainer_name):
    """trainer=fast|slow|slowest: the reference's own Lightning path (/root/reference/train.py:110-187), run unmodified with
    the same command line.  Needs the reference checkout plus lightning, hydra and torchmetrics."""
    import importlib.util
    import runpy
    missing = [m for m in ("lightning", "hydra", "omegaconf", "torchmetrics") if importlib.util.find_spec(m) is None]
    root = reference_root()
    if missing or root is None:
        why = []
        if missing:
            why.append("not installed: " + ", ".join(missing))
        if root is None:
            why.append("no reference checkout found (set B200_REFERENCE_ROOT)")
        raise SystemExit(f"trainer={trainer_name} is the reference's own Lightning path, which this entry point hands to the "
                         f"reference's train.py unchanged -- unavailable here ({'; '.join(why)}).\n"
                         "Use the B200-native path instead:  python train.py trainer=b200   "
                         "(b200_fast / b200_slow / b200_slowest carry the epochs, clip value and validation cadence of "
                         "fast / slow / slowest).")
    os.chdir(root)
    sys.path.insert(0, str(root))
    sys.argv = [str(root / "train.py"), *argv]
    runpy.run_path(str(root / "train.py"), run_name="__main__")
    return None


def hydra_main(argv):
    """Composition by Hydra itself (used when hydra is importable): same config directory, same resolver
    (/root/reference/train.py:39-42, 70)."""
    import hydra
    from omegaconf import OmegaConf
    if not OmegaConf.has_resolver("input_size_from_interaction"):
        OmegaConf.register_new_resolver("input_size_from_interaction", lambda interaction_only: 3 if interaction_only else 5)
    out = []

    @hydra.main(version_base=None, config_path=str(CONFIG_DIR), config_name="config")
    def _main(cfg):
        if not is_b200_trainer(cfg):
            return reference_fallthrough(argv, cfg.trainer.name)
        out.append(run(cfg))
        return out[-1]

    sys.argv = [sys.argv[0], *argv]
    _main()
    return out[0] if len(out) == 1 else out


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    try:
        import hydra  # noqa: F401
        have_hydra = True
    except ImportError:
        have_hydra = False
    if have_hydra:
        return hydra_main(argv)
    multirun, overrides = split_cli(argv)
    jobs = expand_multirun(overrides) if multirun else [overrides]
    results = []
    for i, ov in enumerate(jobs):
        if multirun:
            print(f"[multirun] job {i + 1}/{len(jobs)}: {' '.join(ov)}")
        cfg = compose_config(CONFIG_DIR, "config", ov)
        if not is_b200_trainer(cfg):
            return reference_fallthrough(argv, cfg.trainer.name)
        results.append(run(cfg))
    return results[0] if len(results) == 1 else results


if __name__ == '__main__':
    main()
