"""Metric tracking -- mirrors the parts of ``ReSIDE/util.py`` on the eval path (MetricsTracker :12-92,
AverageMeter :110-134, Timer :198-247).  MetricsTracker.update is ONE fused kernel + ONE 128-byte D2H
instead of ~9 `.item()` syncs; formulas (including the log10 batch-size quirk, util.py:68) unchanged."""
import datetime
import math

import torch

from .. import loss as _loss


class AverageMeter:
    """Running average that ignores NaN/inf updates (util.py:110-134)."""

    def __init__(self):
        self._sum, self._count = 0, 0

    @property
    def value(self):
        return self._sum / self._count if self._count else float("nan")

    def update(self, value, num_elements=1):
        if not math.isnan(value) and not math.isinf(value):
            self._sum += value
            self._count += num_elements

    def __str__(self):
        return str(self.value)

    def __format__(self, spec):
        return f"{self.value:{spec}}"


class MetricsTracker:
    """Tracks running averages of MAE/MSE/RMSE/REL/log10/delta1-3 (util.py:12-92)."""
    _KEYS = ("mae", "mse", "abs_rel", "log10", "delta1", "delta2", "delta3")

    def __init__(self):
        for k in self._KEYS:
            setattr(self, k, AverageMeter())
        self.rmse = 0

    def to_dict(self):
        return {k: (v.value if isinstance(v, AverageMeter) else v) for k, v in self.__dict__.items()}

    def update(self, outputs: torch.Tensor, labels: torch.Tensor):
        self.update_from_sums(_loss.loss_metric_sums(outputs, labels).cpu().tolist())

    def update_from_sums(self, s):
        """`s` = the 16 sums of ede_sobel_loss_metrics for ONE batch (possibly all-reduced over ranks)."""
        nv, b = s[11], int(s[13])
        if nv == 0:
            return
        # util.py:57-77: value = batch_size * sum / num_valid, weighted by batch_size in the meter
        self.mae.update(b * s[4] / nv, b)
        self.mse.update(b * s[5] / nv, b)
        self.rmse = math.sqrt(self.mse.value) if self.mse.value == self.mse.value else float("nan")
        self.abs_rel.update(b * s[6] / nv, b)
        self.log10.update(s[7] / nv, b)                       # quirk: no batch_size factor (util.py:68)
        self.delta1.update(s[8] / nv * b, b)
        self.delta2.update(s[9] / nv * b, b)
        self.delta3.update(s[10] / nv * b, b)

    def __str__(self):
        return (f"ABS_REL: {self.abs_rel:.3f} - MAE: {self.mae:.3f} - MSE: {self.mse:.3f} - RMSE: {self.rmse:.3f} - "
                f"LOG10: {self.log10:.3f} - DELTA1: {self.delta1:.3f} - DELTA2: {self.delta2:.3f} - "
                f"DELTA3: {self.delta3:.3f}        ")


class Timer:
    """Wall-clock timer / context manager (util.py:198-247)."""

    def __init__(self):
        self._start_time = datetime.datetime.fromtimestamp(0)
        self._stop_time = None

    start_time = property(lambda self: self._start_time)
    stop_time = property(lambda self: self._stop_time)

    @property
    def elapsed(self):
        return (self._stop_time or datetime.datetime.now()) - self._start_time

    def start(self):
        self._start_time, self._stop_time = datetime.datetime.now(), None

    def stop(self):
        self._stop_time = datetime.datetime.now()

    def __enter__(self):
        self.start()

    def __exit__(self, *exc):
        self.stop()
