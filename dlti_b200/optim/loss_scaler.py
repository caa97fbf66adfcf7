"""fp16 loss scaling: DeepSpeed `DynamicLossScaler` state machine, restated.

Config block (reference configs/ds_config_zero{1,2,3}.json `fp16`): loss_scale 0 => dynamic,
initial_scale_power 16 => 65536, loss_scale_window 1000, hysteresis 2, min_loss_scale 1.
Overflow = any inf/nan in any rank's gradients; here that is `not isfinite(global sum of squares)`
(detected by the same kernel that produces the gradient norm).
"""
from __future__ import annotations


class StaticLossScaler:
    dynamic = False

    def __init__(self, scale: float = 1.0):
        self.cur_scale = float(scale)

    def update(self, overflow: bool) -> None:
        pass

    def state_dict(self) -> dict:
        return {"cur_scale": self.cur_scale}

    def load_state_dict(self, sd: dict) -> None:
        self.cur_scale = float(sd["cur_scale"])


class DynamicLossScaler:
    dynamic = True

    def __init__(self, initial_scale_power: int = 16, scale_window: int = 1000, hysteresis: int = 2,
                 min_scale: float = 1.0, scale_factor: float = 2.0, consecutive_hysteresis: bool = False):
        self.cur_scale = float(2 ** initial_scale_power)
        self.cur_iter = 0
        self.last_overflow_iter = -1
        self.scale_factor = scale_factor
        self.scale_window = scale_window
        self.min_scale = min_scale
        self.delayed_shift = hysteresis
        self.cur_hysteresis = hysteresis
        self.consecutive_hysteresis = consecutive_hysteresis

    def update(self, overflow: bool) -> None:
        if overflow:
            if self.delayed_shift == 1 or self.cur_hysteresis == 1:
                self.cur_scale = max(self.cur_scale / self.scale_factor, self.min_scale)
            else:
                self.cur_hysteresis -= 1
            self.last_overflow_iter = self.cur_iter
        else:
            if self.consecutive_hysteresis:
                self.cur_hysteresis = self.delayed_shift
            if (self.cur_iter - self.last_overflow_iter) % self.scale_window == 0:
                if not self.consecutive_hysteresis:
                    self.cur_hysteresis = self.delayed_shift
                self.cur_scale *= self.scale_factor
        self.cur_iter += 1

    def state_dict(self) -> dict:
        return {k: getattr(self, k) for k in ("cur_scale", "cur_iter", "last_overflow_iter",
                                              "cur_hysteresis")}

    def load_state_dict(self, sd: dict) -> None:
        for k, v in sd.items():
            setattr(self, k, v)


def grad_scaler_like() -> DynamicLossScaler:
    """torch.amp.GradScaler defaults, which is what the reference's baseline runs under (HF Trainer fp16=True
    without DeepSpeed, reference training/train_baseline.py): initial scale 2^16, halve immediately on the first
    overflow (no hysteresis), double after 2000 clean steps, no lower bound to speak of."""
    return DynamicLossScaler(initial_scale_power=16, scale_window=2000, hysteresis=1, min_scale=2.0 ** -24)


def build_loss_scaler(fp16_cfg: dict | None):
    """fp16 block of a ds_config -> scaler. None / disabled => static 1.0 (bf16 / fp32)."""
    if not fp16_cfg or not fp16_cfg.get("enabled", False):
        return StaticLossScaler(1.0)
    ls = fp16_cfg.get("loss_scale", 0)
    if ls and ls > 0:
        return StaticLossScaler(float(ls))
    return DynamicLossScaler(initial_scale_power=int(fp16_cfg.get("initial_scale_power", 16)),
                             scale_window=int(fp16_cfg.get("loss_scale_window", 1000)),
                             hysteresis=int(fp16_cfg.get("hysteresis", 2)),
                             min_scale=float(fp16_cfg.get("min_loss_scale", 1)),
                             consecutive_hysteresis=bool(fp16_cfg.get("consecutive_hysteresis", False)))
