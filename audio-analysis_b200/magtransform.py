"""Drop-in for the reference's MagTransform Keras layer (src/magtransform.py, duplicated at
src/identify_tracks.py:26-40): y = x ** sigmoid(a) with a learned scalar `a` (init 0.0).  The
layer here is a plain object (no Keras): `a` is set from the checkpoint by the caller, `call`
runs the elementwise kernel (cac_magtransform_run) on numpy arrays or CUDA tensors."""
import numpy as np


class MagTransform:
    INIT_A = 0.0

    def __init__(self, a=None, **kwargs):
        self.kwargs = kwargs                      # Keras layer kwargs (name, dtype, ...) are accepted and kept
        self.a = np.float32(self.INIT_A if a is None else np.asarray(a, dtype=np.float32).reshape(-1)[0])

    def call(self, inputs):
        from . import engine
        return engine.magtransform(inputs, float(self.a))

    __call__ = call

    def get_config(self):
        return dict(self.kwargs)
