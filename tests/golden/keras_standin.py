"""A minimal stand-in for the `tensorflow` / `tensorflow.keras` names that the reference's pointnet/PointNet.py touches,
backed by PyTorch-CPU fp32, so that the UNMODIFIED reference model code (class PointNet, TNet, ConvLayer, DenseLayer,
PointCloudNormalization: layer table, dataflow, T-Net algebra, regulariser, freeze/thaw logic) executes in this
container, where TensorFlow 2.20 / Keras 3.10 cannot be installed.  GOLDEN-VECTOR TOOLING ONLY: used by
tests/golden/make_network_golden.py; nothing in the package, the tests proper, smoke() or bench.py imports it.

What is the reference's own code when the generator runs: everything in PointNet.py.  What is restated here from the
published Keras 3 / TF 2 semantics (and therefore NOT pinned by the golden vectors) are the primitives only:
  Conv2D 1x1 'same' stride 1      y[b,n,1,f] = sum_c x[b,n,1,c] k[0,0,c,f] (+ bias)           keras.layers.Conv2D
  Dense                           y = x @ kernel (+ bias)                                      keras.layers.Dense
  BatchNormalization              eps 1e-3, momentum m; `training and trainable` -> biased batch moments over all axes
                                  but the last and moving <- moving*m + batch*(1-m); otherwise the moving statistics
  Dropout                         identity unless training (rate 0 in the generator)
  tf.nn.l2_loss = sum(x^2)/2, tf.nn.softmax over the last axis, tf.reduce_* / expand_dims / squeeze / tile / concat /
  reshape / transpose / eye / matmul with TF argument names.
Initial values (GlorotUniform etc.) are placeholders: the generator overwrites every variable by name."""
import sys
import types

import numpy as np
import torch

float32 = torch.float32


# ----------------------------------------------------------------------------------------------- tf.* functions
def _t(x):
    return x if isinstance(x, torch.Tensor) else torch.as_tensor(np.asarray(x), dtype=torch.float32)


def matmul(a, b):
    return torch.matmul(_t(a), _t(b))


def expand_dims(x, axis):
    return torch.unsqueeze(x, axis)


def squeeze(x, axis=None):
    return torch.squeeze(x) if axis is None else torch.squeeze(x, axis)


def reduce_max(x, axis=None, keepdims=False):
    return torch.amax(x, dim=axis, keepdim=keepdims)


def reduce_mean(x, axis=None, keepdims=False):
    return torch.mean(x, dim=axis, keepdim=keepdims)


def reduce_sum(x, axis=None, keepdims=False):
    return torch.sum(x, dim=axis, keepdim=keepdims)


def sqrt(x):
    return torch.sqrt(x)


def square(x):
    return x * x


def maximum(a, b):
    return torch.maximum(_t(a), _t(b))


def tile(x, multiples):
    return x.repeat(*[int(m) for m in multiples])


def concat(values, axis):
    return torch.cat(list(values), dim=axis)


def reshape(x, shape):
    return torch.reshape(x, tuple(int(s) for s in shape))


def transpose(x, perm=None):
    return x.permute(*perm) if perm is not None else x.T


def eye(num_rows, batch_shape=None):
    e = torch.eye(int(num_rows))
    if batch_shape is not None:
        e = e.expand(*[int(b) for b in batch_shape], -1, -1).clone()
    return e


def shape(x):
    return tuple(x.shape)


def constant(value, dtype=None):
    return torch.as_tensor(np.asarray(value), dtype=dtype or torch.float32)


def zeros_initializer(shape, dtype=None):
    return torch.zeros(tuple(shape))


class _NN:
    @staticmethod
    def relu(x):
        return torch.relu(x)

    @staticmethod
    def softmax(x, axis=-1):
        return torch.softmax(x, dim=axis)

    @staticmethod
    def l2_loss(x):
        return (x * x).sum() / 2


class _Debugging:
    @staticmethod
    def check_numerics(x, message):
        if not torch.isfinite(x).all():
            raise ValueError(message)
        return x


# ----------------------------------------------------------------------------------------------- keras.initializers
class GlorotUniform:
    def __init__(self, seed=None):
        self.seed = seed

    def __call__(self, shape, fan_in, fan_out):
        lim = float(np.sqrt(6.0 / (fan_in + fan_out)))
        g = torch.Generator().manual_seed(int(self.seed) if isinstance(self.seed, int) else 0)
        return (torch.rand(tuple(shape), generator=g) * 2 - 1) * lim


def _initial(initializer, shape):
    if isinstance(initializer, str):
        if initializer == "identity":
            return torch.eye(shape[0], shape[1])
        if initializer == "zeros":
            return torch.zeros(tuple(shape))
        if initializer == "ones":
            return torch.ones(tuple(shape))
        raise NotImplementedError(initializer)
    if isinstance(initializer, GlorotUniform):
        rf = int(np.prod(shape[:-2])) if len(shape) > 2 else 1
        return initializer(shape, shape[-2] * rf, shape[-1] * rf)
    if callable(initializer):
        return initializer(shape)
    raise NotImplementedError(repr(initializer))


# ----------------------------------------------------------------------------------------------- keras.layers
class Layer:
    def __init__(self, name=None, trainable=True, **kwargs):
        if kwargs:
            raise TypeError(f"stand-in Layer got unexpected arguments {sorted(kwargs)}")
        object.__setattr__(self, "_children", [])
        object.__setattr__(self, "_variables", [])
        object.__setattr__(self, "_losses", [])
        self.name = name or type(self).__name__.lower()
        self.trainable = trainable
        self.built = False

    def __setattr__(self, key, value):
        if isinstance(value, Layer) and value not in self._children:
            self._children.append(value)
        object.__setattr__(self, key, value)

    # keras API touched by the reference
    def build(self, input_shape):
        self.built = True

    def add_weight(self, shape, initializer="zeros", trainable=True, name=None):
        v = _initial(initializer, tuple(int(s) for s in shape)).clone().requires_grad_(bool(trainable))
        self._variables.append((name, v, bool(trainable)))
        return v

    def add_loss(self, value):
        self._losses.append(value)

    def get_config(self):
        return {"name": self.name, "trainable": self.trainable}

    def compute_output_shape(self, input_shape):
        return tuple(input_shape)

    def __call__(self, *args, **kwargs):
        if not self.built:
            x = args[0]
            self.build(tuple(x.shape))
            self.built = True
        if getattr(self, "_is_outermost", False):
            self._clear_losses()
        return self.call(*args, **kwargs)

    # helpers for the generator
    def _clear_losses(self):
        self._losses.clear()
        for c in self._children:
            c._clear_losses()

    @property
    def losses(self):
        out = list(self._losses)
        for c in self._children:
            out += c.losses
        return out


class Model(Layer):
    pass


class Conv2D(Layer):
    def __init__(self, filters, kernel_size, strides=(1, 1), padding="valid", activation=None, use_bias=True, name=None,
                 kernel_initializer=None):
        super().__init__(name=name)
        assert tuple(kernel_size) == (1, 1) and tuple(strides) == (1, 1) and activation is None, "stand-in covers the 1x1 case the reference uses"
        self.filters, self.use_bias, self.kernel_initializer = filters, use_bias, kernel_initializer or GlorotUniform()
        self.kernel = self.bias = None

    def build(self, input_shape):
        if self.kernel is None:
            self.kernel = self.add_weight((1, 1, int(input_shape[-1]), self.filters), self.kernel_initializer, True, "kernel")
            if self.use_bias:
                self.bias = self.add_weight((self.filters,), "zeros", True, "bias")
        self.built = True

    def compute_output_shape(self, input_shape):
        return tuple(input_shape[:-1]) + (self.filters,)

    def call(self, x):
        y = torch.matmul(x, self.kernel[0, 0])
        return y + self.bias if self.use_bias else y


class Dense(Layer):
    def __init__(self, units, activation=None, use_bias=True, name=None, kernel_initializer=None):
        super().__init__(name=name)
        assert activation is None
        self.units, self.use_bias, self.kernel_initializer = units, use_bias, kernel_initializer or GlorotUniform()
        self.kernel = self.bias = None

    def build(self, input_shape):
        if self.kernel is None:
            self.kernel = self.add_weight((int(input_shape[-1]), self.units), self.kernel_initializer, True, "kernel")
            if self.use_bias:
                self.bias = self.add_weight((self.units,), "zeros", True, "bias")
        self.built = True

    def call(self, x):
        y = torch.matmul(x, self.kernel)
        return y + self.bias if self.use_bias else y


class BatchNormalization(Layer):
    def __init__(self, momentum=0.99, epsilon=1e-3, name=None):
        super().__init__(name=name)
        self.momentum, self.epsilon = momentum, epsilon
        self.gamma = None

    def build(self, input_shape):
        if self.gamma is None:
            c = int(input_shape[-1])
            self.gamma = self.add_weight((c,), "ones", True, "gamma")
            self.beta = self.add_weight((c,), "zeros", True, "beta")
            self.moving_mean = self.add_weight((c,), "zeros", False, "moving_mean")
            self.moving_variance = self.add_weight((c,), "ones", False, "moving_variance")
        self.built = True

    def call(self, x, training=False):
        if training and self.trainable:
            axes = tuple(range(x.dim() - 1))
            mean = x.mean(dim=axes)
            var = ((x - mean) ** 2).mean(dim=axes)
            with torch.no_grad():
                m = self.momentum
                self.moving_mean.copy_(self.moving_mean * m + mean * (1 - m))
                self.moving_variance.copy_(self.moving_variance * m + var * (1 - m))
        else:
            mean, var = self.moving_mean, self.moving_variance
        return (x - mean) * (self.gamma * torch.rsqrt(var + self.epsilon)) + self.beta


class Dropout(Layer):
    def __init__(self, rate, name=None):
        super().__init__(name=name)
        self.rate = rate

    def call(self, x, training=False):
        if training and self.rate > 0:
            raise NotImplementedError("the generator runs with dropout_rate = 0 (TF's dropout stream cannot be reproduced)")
        return x


def Input(*a, **k):
    raise NotImplementedError


# ----------------------------------------------------------------------------------------------- module assembly
def install():
    """Put `tensorflow`, `tensorflow.keras`, `tensorflow.keras.layers` (and an empty plotly) into sys.modules."""
    tf = types.ModuleType("tensorflow")
    for name in ("matmul", "expand_dims", "squeeze", "reduce_max", "reduce_mean", "reduce_sum", "sqrt", "square", "maximum", "tile",
                 "concat", "reshape", "transpose", "eye", "shape", "constant", "zeros_initializer", "float32"):
        setattr(tf, name, globals()[name])
    tf.nn, tf.debugging = _NN, _Debugging
    keras = types.ModuleType("tensorflow.keras")
    layers = types.ModuleType("tensorflow.keras.layers")
    for name in ("Layer", "Conv2D", "BatchNormalization", "Dense", "Dropout", "Input"):
        setattr(layers, name, globals()[name])
    keras.layers, keras.Model = layers, Model
    keras.saving = types.SimpleNamespace(register_keras_serializable=lambda package=None, name=None: (lambda cls: cls))
    keras.initializers = types.SimpleNamespace(GlorotUniform=GlorotUniform)
    tf.keras = keras
    sys.modules.update({"tensorflow": tf, "tensorflow.keras": keras, "tensorflow.keras.layers": layers})
    for name in ("plotly", "plotly.graph_objects", "plotly.subplots"):
        sys.modules.setdefault(name, types.ModuleType(name))
    return tf
