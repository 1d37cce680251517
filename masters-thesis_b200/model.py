"""Host-side mirror of the reference Lightning modules (src/model.py:72-331).

Same class names, constructor keyword arguments, hparams, state_dict keys and `test_results`
structure, so train.py drives them exactly as it drives the reference classes; the arithmetic is in
libb200lstm.so.  Initial parameters are drawn through torch.nn's own default initialisers in the
reference's construction order (nn.LSTM, alpha_head, beta_head -- src/model.py:88-97), so the same
torch.manual_seed gives the same initial state_dict as the reference.
"""
from __future__ import annotations

from collections import OrderedDict

import torch

from .engine import LstmEngine


def _param_names(num_layers):
    names = []
    for l in range(num_layers):
        names += [f"lstm.weight_ih_l{l}", f"lstm.weight_hh_l{l}", f"lstm.bias_ih_l{l}", f"lstm.bias_hh_l{l}"]
    return names + ["alpha_head.weight", "alpha_head.bias", "beta_head.weight", "beta_head.bias"]


def _param_shapes(input_size, hidden, num_layers):
    shapes = []
    for l in range(num_layers):
        i = input_size if l == 0 else hidden
        shapes += [(4 * hidden, i), (4 * hidden, hidden), (4 * hidden,), (4 * hidden,)]
    return shapes + [(1, hidden), (1,), (1, hidden), (1,)]


class FinancialLstm:
    """Base mirror of src/model.py:72-173."""
    loss_kind = "mse"

    def __init__(self, input_size: int = 3, hidden_size: int = 64, num_layers: int = 2, dropout: float = 0.2,
                 learning_rate: float = 1e-4, weight_decay: float = 1e-5):
        self.hparams = OrderedDict(input_size=input_size, hidden_size=hidden_size, num_layers=num_layers,
                                   dropout=dropout, learning_rate=learning_rate, weight_decay=weight_decay)
        # default initialisation with torch's own RNG stream, in the reference's module order
        lstm = torch.nn.LSTM(input_size=input_size, hidden_size=hidden_size, num_layers=num_layers,
                             dropout=dropout if num_layers > 1 else 0, batch_first=True)
        alpha_head = torch.nn.Linear(hidden_size, 1)
        beta_head = torch.nn.Linear(hidden_size, 1)
        with torch.no_grad():
            self._host_params = torch.cat([p.reshape(-1) for p in
                                           list(lstm.parameters()) + list(alpha_head.parameters()) + list(beta_head.parameters())])
        self.names = _param_names(num_layers)
        self.shapes = _param_shapes(input_size, hidden_size, num_layers)
        self.effective_dropout = float(dropout) if num_layers > 1 else 0.0     # src/model.py:92
        self.engine: LstmEngine | None = None
        self.test_results = None
        self.training = True

    # ---- module-like surface used by train.py -------------------------------------------------
    def compile(self, *a, **k):      # train.py:137 -- kernels are already native code
        return self

    def train(self, mode=True):
        self.training = mode
        return self

    def eval(self):
        return self.train(False)

    def __repr__(self):
        h = self.hparams
        return (f"{type(self).__name__}(\n  (lstm): LSTM({h['input_size']}, {h['hidden_size']}, num_layers={h['num_layers']}, "
                f"batch_first=True, dropout={self.effective_dropout})\n  (alpha_head): Linear(in_features={h['hidden_size']}, "
                f"out_features=1, bias=True)\n  (beta_head): Linear(in_features={h['hidden_size']}, out_features=1, bias=True)\n"
                f"  [engine: libb200lstm.so sm_100a, {self.num_parameters()} parameters]\n)")

    def num_parameters(self):
        return int(self._host_params.numel())

    def attach(self, device="cuda:0", mode="ffma", lookback=60, target=30, n_stocks=100):
        """Create the device engine and upload the parameters."""
        h = self.hparams
        self.engine = LstmEngine(h["input_size"], h["hidden_size"], h["num_layers"], self.effective_dropout, device, mode,
                                 lookback, target, n_stocks)
        self.engine.load_params(self._host_params)
        return self.engine

    def flat_parameters(self) -> torch.Tensor:
        if self.engine is not None:
            return self.engine.params[: self.engine.P].detach().cpu()
        return self._host_params.clone()

    def state_dict(self):
        flat, out, off = self.flat_parameters(), OrderedDict(), 0
        for name, shape in zip(self.names, self.shapes):
            n = 1
            for s in shape:
                n *= s
            out[name] = flat[off:off + n].reshape(shape).clone()
            off += n
        return out

    def load_state_dict(self, sd):
        flat = torch.cat([sd[name].reshape(-1).to(torch.float32) for name in self.names])
        self._host_params = flat.clone()
        if self.engine is not None:
            self.engine.load_params(flat)

    @classmethod
    def load_from_checkpoint(cls, checkpoint_path, map_location=None, **overrides):
        """Lightning's classmethod as test.py:178 uses it: hyper-parameters and weights from a checkpoint written by
        B200Trainer (or by the reference's ModelCheckpoint: same `state_dict` / `hyper_parameters` keys)."""
        ck = torch.load(checkpoint_path, map_location="cpu", weights_only=False)
        hp = dict(ck.get("hyper_parameters", {}))
        hp.update(overrides)
        import inspect
        accepted = inspect.signature(cls.__init__).parameters
        model = cls(**{k: v for k, v in hp.items() if k in accepted})
        model.load_state_dict(ck["state_dict"])
        return model

    def forward(self, lstm_input: torch.Tensor):
        """(N, T, F) -> alpha (N,1), beta (N,1); src/model.py:103-109."""
        from .data import DeviceBatch
        assert self.engine is not None, "call attach() first"
        e = self.engine
        N, T, F = lstm_input.shape
        x = lstm_input.to(e.device, torch.float32).permute(1, 0, 2).contiguous()
        # a forward-only call needs no targets: treat the N rows as N single-stock windows
        S_saved = e.S
        try:
            e.S = 1
            e._cap = 0
            b = DeviceBatch(x, torch.zeros(N, e.n, device=e.device), torch.zeros(N, e.n, device=e.device),
                            torch.ones(N, 2, device=e.device), torch.ones(N, device=e.device), N, 1)
            e.forward(b, training=False)
            e.heads(b, "mse", want_grad=False, windows_per_group=1)
            return e.alpha[:N].clone().unsqueeze(1), e.beta[:N].clone().unsqueeze(1)
        finally:
            e.S = S_saved
            e._cap = 0

    __call__ = forward

    def on_test_start(self):
        self.test_results = {'test_loss': {'mae': None, 'nll': None}}

    def configure_optimizers(self):
        """Description of the optimiser/schedule the trainer applies (src/model.py:149-173)."""
        return {"optimizer": {"name": "Adam", "lr": self.hparams["learning_rate"], "weight_decay": self.hparams["weight_decay"],
                              "betas": (0.9, 0.999), "eps": 1e-8},
                "lr_scheduler": {"name": "ReduceLROnPlateau", "mode": "min", "factor": 0.5, "patience": 2,
                                 "monitor": "loss/total/val", "interval": "epoch"}}


class FinancialLstmMse(FinancialLstm):
    loss_kind = "mse"


class FinancialLstmNll(FinancialLstm):
    loss_kind = "nll"


class FinancialLstmCombined(FinancialLstm):
    loss_kind = "combined"

    def __init__(self, input_size: int = 3, hidden_size: int = 64, num_layers: int = 2, dropout: float = 0.2,
                 learning_rate: float = 1e-4, weight_decay: float = 1e-5, mse_weight: float = 1e2):
        super().__init__(input_size, hidden_size, num_layers, dropout, learning_rate, weight_decay)
        self.hparams["mse_weight"] = mse_weight


def get_model_class(module_class_name: str):
    """train.py:45-67."""
    classes = {'FinancialLstmMse': FinancialLstmMse, 'FinancialLstmNll': FinancialLstmNll,
               'FinancialLstmCombined': FinancialLstmCombined}
    if module_class_name not in classes:
        raise ValueError(f"Unknown module class: {module_class_name}. Available: {list(classes.keys())}")
    return classes[module_class_name]
