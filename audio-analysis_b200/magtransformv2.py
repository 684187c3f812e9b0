"""Drop-in for src/magtransformv2.py: same y = x ** sigmoid(a) with `a` of shape [1] initialised to
-1.0; the MinMaxNorm(-2, 1) constraint (:13-15) acts on the weight after training updates only, so
it is offered as `constrain` and not applied to a loaded value."""
import numpy as np

from .magtransform import MagTransform as _Base


class MagTransform(_Base):
    INIT_A = -1.0

    @staticmethod
    def constrain(a, min_value=-2.0, max_value=1.0):
        """tf.keras.constraints.MinMaxNorm(min, max, rate=1.0, axis=-1) on a one-element weight:
        w * clip(|w|, min, max) / (eps + |w|) with eps = 1e-7 (Keras backend epsilon)."""
        a = np.float32(a)
        norm = np.float32(abs(a))
        return np.float32(a * (np.clip(norm, min_value, max_value) / (np.float32(1e-7) + norm)))
