"""Import stub (see README.md): nerf/utils.py imports ExponentialMovingAverage at module scope; the reference arm passes
ema_decay=None (the legacy Trainer only updates the EMA once per epoch and uses it for evaluation), so it is never constructed."""


class ExponentialMovingAverage:
    def __init__(self, *a, **k):
        raise RuntimeError("stub torch_ema.ExponentialMovingAverage: pass ema_decay=None")
