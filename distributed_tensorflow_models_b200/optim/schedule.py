"""Learning-rate schedule: ``tf.train.exponential_decay`` (e.g. cifarnet/cifar10_cifarnet_bsp.py:70-74,
inception/imagenet_inception_bsp.py:127-133).  Pure host arithmetic: the value is passed to
the fused update kernel as a scalar argument."""
from __future__ import annotations

import math


def exponential_decay(learning_rate: float, global_step: int, decay_steps: int, decay_rate: float,
                      staircase: bool = False) -> float:
    """lr * rate^(step/decay_steps), exponent floored when ``staircase``."""
    if decay_steps <= 0:
        return float(learning_rate)
    p = global_step / float(decay_steps)
    if staircase:
        p = math.floor(p)
    return float(learning_rate) * (decay_rate ** p)


def ema_decay(decay: float, num_updates=None) -> float:
    """tf.train.ExponentialMovingAverage effective decay: min(decay, (1+n)/(10+n)) when ``num_updates``
    is given (cnn/cifar10_cnn_bsp.py:57-64 passes global_step)."""
    if num_updates is None:
        return decay
    n = float(num_updates)
    return min(decay, (1.0 + n) / (10.0 + n))
