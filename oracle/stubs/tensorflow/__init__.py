"""numpy-backed stand-in for the handful of tensorflow ops the reference's feature
path touches (TEST INFRASTRUCTURE ONLY): expand_dims / reduce_mean / repeat / concat
(src/identify_tracks.py:267,283-287), tf.math.pow / sigmoid and the Keras layer
scaffolding used by MagTransform (src/magtransform.py, src/magtransformv2.py,
src/identify_tracks.py:26-40)."""
import types
import numpy as np


def expand_dims(x, axis):
    return np.expand_dims(np.asarray(x), axis)


def reduce_mean(x, axis=None):
    return np.mean(np.asarray(x), axis=axis, dtype=np.asarray(x).dtype)


def repeat(x, repeats, axis=None):
    return np.repeat(np.asarray(x), repeats, axis=axis)


def concat(values, axis):
    return np.concatenate([np.asarray(v) for v in values], axis=axis)


class _Math(types.SimpleNamespace):
    @staticmethod
    def pow(x, y):
        return np.power(np.asarray(x, dtype=np.float32), np.float32(y)).astype(np.float32)

    @staticmethod
    def sigmoid(x):
        x = np.asarray(x, dtype=np.float32)
        return (np.float32(1.0) / (np.float32(1.0) + np.exp(-x))).astype(np.float32)


math = _Math()


class _Constant:
    def __init__(self, value=0.0):
        self.value = value


class _MinMaxNorm:
    def __init__(self, min_value=0.0, max_value=1.0, rate=1.0, axis=0):
        self.min_value, self.max_value, self.rate, self.axis = min_value, max_value, rate, axis


class _Layer:
    def __init__(self, **kwargs):
        self._kwargs = kwargs

    def add_weight(self, initializer=None, name=None, dtype="float32", shape=(), trainable=True,
                   constraint=None):
        v = getattr(initializer, "value", 0.0)
        return np.full(shape, v, dtype=np.float32) if shape not in ((), None) else np.float32(v)

    def __call__(self, inputs):
        return self.call(inputs)


def _register_keras_serializable(package="Custom", name=None):
    def deco(cls):
        return cls
    return deco


keras = types.SimpleNamespace(
    utils=types.SimpleNamespace(register_keras_serializable=_register_keras_serializable),
    layers=types.SimpleNamespace(Layer=_Layer),
    initializers=types.SimpleNamespace(Constant=_Constant),
    constraints=types.SimpleNamespace(MinMaxNorm=_MinMaxNorm),
    models=types.SimpleNamespace(load_model=None),
)
image = types.SimpleNamespace(resize_with_pad=None)
