"""Early stopping on a validation loss.
Ref: src/models/base/utils/early_stopping/pytorchtools.py:4-55 (same constructor and call protocol)."""
from __future__ import annotations

import numpy as np


class EarlyStopping:
    """Stops training when the validation loss has not improved by ``delta`` for ``patience`` calls.
    On every improvement ``model.save(path)`` is invoked (when the model has a ``save`` method)."""

    def __init__(self, patience: int = 7, verbose: bool = False, delta: float = 0, path="checkpoint.pt",
                 trace_func=print) -> None:
        self.patience = patience
        self.verbose = verbose
        self.delta = delta
        self.path = path
        self.trace_func = trace_func
        self.counter = 0
        self.best_score = None
        self.early_stop = False
        self.val_loss_min = np.inf

    def __call__(self, val_loss: float, model) -> None:
        score = -val_loss
        if self.best_score is None or score >= self.best_score + self.delta:
            if self.best_score is not None:
                self.counter = 0
            self.best_score = score
            self.save_checkpoint(val_loss, model)
        else:
            self.counter += 1
            if self.verbose:
                self.trace_func(f"EarlyStopping counter: {self.counter} out of {self.patience}")
            if self.counter >= self.patience:
                self.early_stop = True

    def save_checkpoint(self, val_loss: float, model) -> None:
        if self.verbose:
            self.trace_func(f"Validation loss decreased ({self.val_loss_min:.6f} --> {val_loss:.6f}).  Saving model ...")
        if hasattr(model, "save") and self.path is not None:
            model.save(self.path)
        self.val_loss_min = val_loss
