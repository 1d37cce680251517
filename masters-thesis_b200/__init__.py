"""B200-native training hot path of the masters-thesis trainer.

Host-side mirror of the reference's operator surface (model / datamodule / trainer) over the C-ABI
CUDA library libb200lstm.so.  The directory name carries a hyphen; import it as
``masters_thesis_b200`` (shim module at the repo root) or through importlib.
"""
from . import _lib, data, evaluate
from ._lib import B200Error, LOSS_KINDS, MODES
from .config import compose_config
from .data import ChunkedBatch, DeviceBatch, FinancialLstmDataModule, HostBatchFeeder, SeriesBatchSource
from .engine import LstmEngine
from .model import FinancialLstm, FinancialLstmCombined, FinancialLstmMse, FinancialLstmNll, get_model_class
from .trainer import B200Trainer, LossMailbox

__all__ = ["B200Error", "B200Trainer", "DeviceBatch", "FinancialLstm", "FinancialLstmCombined",
           "FinancialLstmDataModule", "FinancialLstmMse", "HostBatchFeeder", "FinancialLstmNll", "LOSS_KINDS", "LossMailbox", "LstmEngine",
           "MODES", "SeriesBatchSource", "compose_config", "data", "evaluate", "get_model_class", "_lib"]
