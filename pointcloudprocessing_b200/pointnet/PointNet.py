"""Drop-in mirror of the reference's `pointnet/PointNet.py` (PointNet, TNet, ConvLayer, DenseLayer,
PointCloudNormalization) over libpcp_b200.so.

Same class names, constructor arguments, layer names and call semantics as
/root/reference/point_cloud_analysis/pointnet/PointNet.py (cited as PointNet.py:line below), but no
TensorFlow: `PointNet.__call__` hands device pointers to `pcp_model_forward` and every FLOP runs in
hand-written sm_100a CUDA.  There is no CPU fallback — constructing a model without a CUDA device
raises.  Inputs may be CUDA torch tensors (zero-copy) or NumPy arrays (host path through
`pcp_model_predict_host`, H2D/D2H inside the library).
"""
from __future__ import annotations

import ctypes as C
import json
from pathlib import Path

import numpy as np
import torch

from .. import _ffi

TNET_WIDTHS = [64, 128, 1024, 512, 256]     # PointNet.py:384


def _glorot(rng: np.random.Generator, cin: int, cout: int) -> np.ndarray:
    """keras GlorotUniform (PointNet.py:532,627)."""
    lim = np.sqrt(6.0 / (cin + cout))
    return rng.uniform(-lim, lim, size=(cin, cout)).astype(np.float32)


class _LayerBase:
    """Shared freeze/thaw/is_trainable contract (PointNet.py:585-594, :670-679)."""

    def __init__(self, name: str):
        self.name = name
        self.trainable = True

    def freeze(self):
        self.trainable = False

    def thaw(self):
        self.trainable = True

    def is_trainable(self):
        return self.trainable


class ConvLayer(_LayerBase):
    """1x1 Conv2D (no bias when BN) -> BatchNormalization -> activation (PointNet.py:493-594).
    A descriptor: the arithmetic runs fused inside the owning PointNet's CUDA chain."""

    def __init__(self, filters: int, name: str, kernel_size=(1, 1), strides=(1, 1), padding="same", activation=None,
                 apply_bn: bool = True, bn_momentum: float = 0.99, random_seed=None, **kwargs):
        super().__init__(f"{name}_convolution_layer")          # PointNet.py:529
        if tuple(kernel_size) != (1, 1) or tuple(strides) != (1, 1):
            raise ValueError("only the reference's 1x1 / stride-1 shared-MLP convolution is supported")
        self.filters, self.kernel_size, self.strides, self.padding = filters, tuple(kernel_size), tuple(strides), padding
        self.activation, self.apply_bn, self.bn_momentum, self.seed = activation, apply_bn, bn_momentum, random_seed
        self.cin = None

    def build(self, input_shape):
        self.cin = int(input_shape[-1])

    def get_config(self):
        return {"filters": self.filters, "name": self.name, "kernel_size": self.kernel_size, "strides": self.strides,
                "padding": self.padding, "activation": self.activation, "apply_bn": self.apply_bn,
                "bn_momentum": self.bn_momentum, "random_seed": self.seed}


class DenseLayer(_LayerBase):
    """Dense (bias only without BN) -> BatchNormalization -> activation (PointNet.py:597-679)."""

    def __init__(self, units: int, name: str, activation=None, apply_bn: bool = False, bn_momentum: float = 0.99,
                 random_seed=None, **kwargs):
        super().__init__(f"{name}_dense_layer")                 # PointNet.py:625
        self.units, self.activation, self.apply_bn, self.bn_momentum, self.seed = units, activation, apply_bn, bn_momentum, random_seed
        self.cin = None

    def build(self, input_shape):
        self.cin = int(input_shape[-1])

    def get_config(self):
        return {"units": self.units, "name": self.name, "activation": self.activation, "apply_bn": self.apply_bn,
                "bn_momentum": self.bn_momentum, "random_seed": self.seed}


class PointCloudNormalization(_LayerBase):
    """Centre on the centroid, scale to the unit sphere (PointNet.py:681-712)."""

    def __init__(self, name: str = "input_normalization", **kwargs):
        super().__init__(name)
        self.trainable = False

    def __call__(self, inp):
        from .. import ops
        return ops.normalize(inp, return_stats=True)

    call = __call__

    def get_config(self):
        return {"name": self.name}

    def is_trainable(self):
        return False                                             # PointNet.py:711-712


class TNet(_LayerBase):
    """Transformation network descriptor (PointNet.py:379-490): conv 64-128-1024, max-pool, dense 512-256,
    256 x K^2 matrix `w` + identity-initialised `b`."""

    def __init__(self, name: str, add_regularization: bool = False, bn_momentum: float = 0.99,
                 layer_widths: list = TNET_WIDTHS, random_seed=None, **kwargs):
        super().__init__(name)
        if list(layer_widths) != TNET_WIDTHS:
            raise ValueError("only the reference's T-Net widths [64,128,1024,512,256] are supported")
        self.add_regularization, self.bn_momentum, self.layer_widths, self.seed = add_regularization, bn_momentum, list(layer_widths), random_seed
        w = self.layer_widths
        self.conv_layer_1 = ConvLayer(w[0], f"{name}_convolution_layer_1", activation="relu", bn_momentum=bn_momentum, random_seed=random_seed)
        self.conv_layer_2 = ConvLayer(w[1], f"{name}_convolution_layer_2", activation="relu", bn_momentum=bn_momentum, random_seed=random_seed)
        self.conv_layer_3 = ConvLayer(w[2], f"{name}_convolution_layer_3", activation="relu", bn_momentum=bn_momentum, random_seed=random_seed)
        self.dense_layer_1 = DenseLayer(w[3], f"{name}_dense_layer_1", activation="relu", apply_bn=True, bn_momentum=bn_momentum, random_seed=random_seed)
        self.dense_layer_2 = DenseLayer(w[4], f"{name}_dense_layer_2", activation="relu", apply_bn=True, bn_momentum=bn_momentum, random_seed=random_seed)
        self.K = None
        self._last_predicted = None

    def sublayers(self):
        return [self.conv_layer_1, self.conv_layer_2, self.conv_layer_3, self.dense_layer_1, self.dense_layer_2]

    def build(self, input_shape):
        self.K = int(input_shape[-1])                            # PointNet.py:414
        cin = self.K
        for l in self.sublayers():
            l.build((None, cin))
            cin = l.filters if isinstance(l, ConvLayer) else l.units

    def freeze(self):                                            # PointNet.py:468-475
        self.trainable = False
        for l in self.sublayers():
            l.freeze()

    def thaw(self):                                              # PointNet.py:477-484
        self.trainable = True
        for l in self.sublayers():
            l.thaw()

    def is_trainable(self):                                      # PointNet.py:486-490
        return all(l.is_trainable() for l in self.sublayers())

    def get_last_predicted_transformation(self):
        return self._last_predicted

    def get_config(self):
        return {"name": self.name, "add_regularization": self.add_regularization, "bn_momentum": self.bn_momentum,
                "layer_widths": self.layer_widths, "random_seed": self.seed}


class PointNet:
    """PointNet classification + part segmentation + input-transform output (PointNet.py:84-376).

    model(pc, training=False) -> [classification (b,C) probabilities, segmentation (b,n,P) probabilities, R (b,3,3)].
    `precision`: 'bf16' (tcgen05 tensor cores, 2e-2 contract), 'fp32' (CUDA-core FMA, 1e-3 contract) or 'fp32_tc' (the 1e-3
    contract on the tensor cores: every layer product as three bf16 products of hi / lo halves, ~16 mantissa bits; inference —
    its training steps run the fp32 path).
    """

    def __init__(self, classification_output_width: int, segmentation_output_width: int, dropout_rate: float,
                 random_seed: int, debugging: bool = False, vanilla: bool = False,
                 regularize_input_transform: bool = False, regularize_feature_transform: bool = False,
                 precision: str = "bf16", **kwargs):
        self._classification_output_width = int(classification_output_width)
        self._segmentation_output_width = int(segmentation_output_width)
        self._dropout_rate = float(dropout_rate)
        self._random_seed = random_seed
        self._debugging = bool(debugging)
        self._vanilla = bool(vanilla)
        self._regularize_input_transform = bool(regularize_input_transform)
        self._regularize_feature_transform = bool(regularize_feature_transform)
        self._precision = precision
        self.name = kwargs.get("name", "point_net")
        self.input_names = ["pointnet_input"]                                      # PointNet.py:113
        self.output_names = ["classification_output", "segmentation_output", "se3"]  # PointNet.py:114
        self.stop_training = False
        self.built = False
        self._handle = None
        self._lib = _ffi.load()
        seed = self._random_seed
        self.normalize_input = PointCloudNormalization(name="input_normalization")
        self.input_transform = None if self._vanilla else TNet(name="input_transform", add_regularization=regularize_input_transform, random_seed=seed)
        self.mlp_1_1 = ConvLayer(64, "s1_l1_64", activation="relu", random_seed=seed)
        self.mlp_1_2 = ConvLayer(64, "s1_l2_64", activation="relu", random_seed=seed)
        self.feature_transform = None if self._vanilla else TNet(name="feature_transform", add_regularization=regularize_feature_transform, random_seed=seed)
        self.mlp_2_1 = ConvLayer(64, "s2_l1_64", activation="relu", random_seed=seed)
        self.mlp_2_2 = ConvLayer(128, "s2_l2_128", activation="relu", random_seed=seed)
        self.mlp_2_3 = ConvLayer(1024, "s2_l3_1024", activation="relu", random_seed=seed)
        self.mlp_cls_1 = DenseLayer(512, "s3_l1_512", activation="relu", apply_bn=True)
        self.mlp_cls_2 = DenseLayer(256, "s3_l2_256", activation="relu", apply_bn=True)
        self.mlp_cls_3 = DenseLayer(self._classification_output_width, "output", activation="softmax")
        self.mlp_seg_1 = ConvLayer(512, "seg_l1_512", activation="relu", random_seed=seed)
        self.mlp_seg_2 = ConvLayer(256, "seg_l2_256", activation="relu", random_seed=seed)
        self.mlp_seg_3 = ConvLayer(128, "seg_l3_128", activation="relu", random_seed=seed)
        self.mlp_seg_4 = ConvLayer(128, "seg_l4_128", activation="relu", random_seed=seed)
        self.mlp_seg_5 = ConvLayer(self._segmentation_output_width, "seg_l5_output", activation="softmax", apply_bn=False, random_seed=seed)
        self._custom_layers = [self.normalize_input]                               # PointNet.py:143-158
        if not self._vanilla:
            self._custom_layers.append(self.input_transform)
        self._custom_layers += [self.mlp_1_1, self.mlp_1_2]
        if not self._vanilla:
            self._custom_layers.append(self.feature_transform)
        self._custom_layers += [self.mlp_2_1, self.mlp_2_2, self.mlp_2_3, self.mlp_cls_1, self.mlp_cls_2, self.mlp_cls_3,
                                self.mlp_seg_1, self.mlp_seg_2, self.mlp_seg_3, self.mlp_seg_4, self.mlp_seg_5]
        self._compiled = None
        self._step = 0

    # ------------------------------------------------------------------ lifecycle
    def _create_handle(self):
        prec = {"fp32": _ffi.PRECISION_FP32, "bf16": _ffi.PRECISION_BF16, "fp32_tc": _ffi.PRECISION_FP32_TC}[self._precision]
        desc = _ffi.ModelDesc(self._classification_output_width, self._segmentation_output_width, int(self._vanilla),
                              int(self._regularize_input_transform), int(self._regularize_feature_transform),
                              self._dropout_rate, prec)
        h = C.c_void_p()
        _ffi.check(self._lib.pcp_model_create(C.byref(desc), C.byref(h)))
        self._handle = h

    def __del__(self):
        try:
            if getattr(self, "_handle", None):
                self._lib.pcp_model_destroy(self._handle)
                self._handle = None
        except Exception:
            pass

    def build(self, input_shape):
        """PointNet.build (PointNet.py:161-195): declare every weight; Glorot-uniform kernels."""
        self._input_width = input_shape[1]
        if self._handle is None:
            self._create_handle()
        # q4: every seeded conv layer draws from an identically seeded generator, so
        # equal-shaped kernels start identical; unseeded cls-head layers draw fresh entropy.
        for name, count in self.tensor_table():
            if name.endswith("/kernel") or name.endswith("/w"):
                layer = name.rsplit("/", 1)[0]
                unseeded_layers = ("s3_l1_512_dense_layer", "s3_l2_256_dense_layer", "output_dense_layer")  # PointNet.py:130-134
                seeded = isinstance(self._random_seed, int) and layer not in unseeded_layers
                rng = np.random.default_rng(self._random_seed if seeded else None)
                cin, cout = self._shape_of(name, count)
                self.set_tensor(name, _glorot(rng, cin, cout))
        self._sync_trainable()
        self._prepare()
        self.built = True
        return self

    def _shape_of(self, name, count):
        C_, P_ = self._classification_output_width, self._segmentation_output_width
        if name.endswith("/w"):
            return 256, count // 256
        widths = {
            "s1_l1_64_convolution_layer": (3, 64), "s1_l2_64_convolution_layer": (64, 64),
            "s2_l1_64_convolution_layer": (64, 64), "s2_l2_128_convolution_layer": (64, 128),
            "s2_l3_1024_convolution_layer": (128, 1024), "s3_l1_512_dense_layer": (1024, 512),
            "s3_l2_256_dense_layer": (512, 256), "output_dense_layer": (256, C_),
            "seg_l1_512_convolution_layer": (1088, 512), "seg_l2_256_convolution_layer": (512, 256),
            "seg_l3_128_convolution_layer": (256, 128), "seg_l4_128_convolution_layer": (128, 128),
            "seg_l5_output_convolution_layer": (128, P_),
        }
        layer = name.rsplit("/", 1)[0]
        if layer in widths:
            return widths[layer]
        for tn, k in (("input_transform", 3), ("feature_transform", 64)):
            w = TNET_WIDTHS
            sub = {f"{tn}_convolution_layer_1_convolution_layer": (k, w[0]), f"{tn}_convolution_layer_2_convolution_layer": (w[0], w[1]),
                   f"{tn}_convolution_layer_3_convolution_layer": (w[1], w[2]), f"{tn}_dense_layer_1_dense_layer": (w[2], w[3]),
                   f"{tn}_dense_layer_2_dense_layer": (w[3], w[4])}
            if layer in sub:
                return sub[layer]
        raise KeyError(name)

    def _prepare(self):
        _ffi.check(self._lib.pcp_model_prepare(self._handle))

    # ------------------------------------------------------------------ tensors
    def tensor_table(self):
        n = self._lib.pcp_model_num_tensors(self._handle)
        out = []
        buf = C.create_string_buffer(256)
        cnt = C.c_size_t()
        grp = C.c_int()
        for i in range(n):
            _ffi.check(self._lib.pcp_model_tensor_info(self._handle, i, buf, 256, C.byref(cnt), C.byref(grp)))
            out.append((buf.value.decode(), cnt.value))
        return out

    def set_tensor(self, name: str, value: np.ndarray):
        a = np.ascontiguousarray(np.asarray(value, np.float32).reshape(-1))
        _ffi.check(self._lib.pcp_model_set_tensor(self._handle, name.encode(), _ffi.fptr(a), a.size))

    def get_tensor(self, name: str, count: int) -> np.ndarray:
        a = np.empty(count, np.float32)
        _ffi.check(self._lib.pcp_model_get_tensor(self._handle, name.encode(), _ffi.fptr(a), a.size))
        return a

    def set_weights_dict(self, params: dict):
        """Load a {reference tensor name: array} dict (conv kernels (Cin,Cout) or Keras (1,1,Cin,Cout))."""
        if self._handle is None:
            self._create_handle()
        names = dict(self.tensor_table())
        for k, v in params.items():
            if k not in names:
                raise KeyError(f"unknown tensor {k}")
            self.set_tensor(k, v)
        self._sync_trainable()
        self._prepare()
        self.built = True

    def get_weights_dict(self) -> dict:
        return {name: self.get_tensor(name, cnt) for name, cnt in self.tensor_table()}

    def save(self, path: str, input_width=None):
        """`*.keras`: the reference's archive format (pointnet_train.py:207-214, :280-295) — zip of config.json / metadata.json /
        model.weights.h5 written without Keras or h5py (pointnet/keras_io.py, pointnet/hdf5_min.py).  Anything else: our own
        container, `.npz` keyed by the reference's variable names + the config JSON."""
        p = Path(path)
        if p.suffix == ".keras":
            from . import keras_io
            keras_io.save_keras(str(p), self.get_weights_dict(), self.get_config(), input_width)
            return
        np.savez(p.with_suffix(".npz") if p.suffix != ".npz" else p, __config__=json.dumps(self.get_config()), **self.get_weights_dict())

    def export_onnx(self, path: str, input_width: int) -> str:
        """The opset-13 inference graph `tf2onnx.convert.from_keras(model, [(None, input_width, 3) f32], opset=13)` would
        give (pointnet_train.py:237-247), written without onnx / tf2onnx — see pointnet/onnx_export.py."""
        from . import onnx_export
        return onnx_export.export_onnx(path, self.get_weights_dict(), self._classification_output_width, self._segmentation_output_width,
                                       int(input_width), vanilla=self._vanilla)

    @classmethod
    def load(cls, path: str, **overrides):
        if str(path).endswith(".keras"):                          # tf.keras.models.load_model(path, custom_objects=...) (pointnet_train.py:280-295)
            from . import keras_io
            weights, cfg = keras_io.load_keras(str(path))
            cfg.update(overrides)
            m = cls.from_config(cfg)
            m.set_weights_dict(weights)
            return m
        z = np.load(path, allow_pickle=False)
        cfg = json.loads(str(z["__config__"]))
        cfg.update(overrides)
        m = cls.from_config(cfg)
        m.set_weights_dict({k: z[k] for k in z.files if k != "__config__"})
        return m

    # ------------------------------------------------------------------ forward
    def __call__(self, pc, training: bool = False, heads=("cls", "seg")):
        """PointNet.call (PointNet.py:197-292)."""
        if training:
            raise NotImplementedError("training-mode forward runs inside fit()/train_on_batch()")
        if not self.built:
            raise RuntimeError("call build() or set_weights_dict() first")
        mask = (_ffi.HEAD_CLS if "cls" in heads else 0) | (_ffi.HEAD_SEG if "seg" in heads else 0)
        Cw, Pw = self._classification_output_width, self._segmentation_output_width
        if isinstance(pc, torch.Tensor):
            if not pc.is_cuda:
                raise TypeError("torch inputs must live on the GPU (pass NumPy for the host path)")
            x = pc.to(torch.float32).contiguous()
            b, n, _ = x.shape
            if self._debugging and not bool(torch.isfinite(x).all()):
                raise FloatingPointError("Input point cloud contains nan values")        # PointNet.py:199
            cls = torch.empty((b, Cw), device=x.device, dtype=torch.float32) if mask & 1 else None
            seg = torch.empty((b, n, Pw), device=x.device, dtype=torch.float32) if mask & 2 else None
            R = torch.empty((b, 3, 3), device=x.device, dtype=torch.float32)
            _ffi.check(self._lib.pcp_model_forward(self._handle, _ffi.ptr(x), b, n, mask, _ffi.ptr(cls), _ffi.ptr(seg),
                                                   _ffi.ptr(R), None, None, _ffi.stream_ptr()))
        else:
            x = np.ascontiguousarray(np.asarray(pc, np.float32))
            b, n, _ = x.shape
            if self._debugging and not np.isfinite(x).all():
                raise FloatingPointError("Input point cloud contains nan values")
            # outputs come from torch's caching PINNED host allocator (the NumPy arrays are views that keep their block alive and
            # hand it back to the cache when dropped): the b x n x P segmentation output is 50 MB at 1024 x 1024 x 12 — fresh
            # pageable memory costs first-touch page faults plus a driver-staged copy (12.6 ms end to end, 4 ms pinned)
            host = lambda *shape: torch.empty(shape, dtype=torch.float32, pin_memory=True).numpy()
            cls = host(b, Cw) if mask & 1 else None
            seg = host(b, n, Pw) if mask & 2 else None
            R = host(b, 3, 3)
            _ffi.check(self._lib.pcp_model_predict_host(self._handle, _ffi.fptr(x), b, n, mask, _ffi.fptr(cls), _ffi.fptr(seg),
                                                        _ffi.fptr(R), _ffi.stream_ptr() if torch.cuda.is_available() else None))
        if self.input_transform is not None:
            self.input_transform._last_predicted = R
        return [cls, seg, R]

    call = __call__

    def forward_with_logits(self, x: torch.Tensor, heads=("cls", "seg")):
        """Device forward that also returns the pre-softmax logits (tolerance checks)."""
        x = x.to(torch.float32).contiguous()
        b, n, _ = x.shape
        Cw, Pw = self._classification_output_width, self._segmentation_output_width
        mask = (_ffi.HEAD_CLS if "cls" in heads else 0) | (_ffi.HEAD_SEG if "seg" in heads else 0)
        mk = lambda *s: torch.empty(s, device=x.device, dtype=torch.float32)
        out = {"R": mk(b, 3, 3)}
        if mask & 1:
            out["cls"], out["cls_logits"] = mk(b, Cw), mk(b, Cw)
        if mask & 2:
            out["seg"], out["seg_logits"] = mk(b, n, Pw), mk(b, n, Pw)
        _ffi.check(self._lib.pcp_model_forward(self._handle, _ffi.ptr(x), b, n, mask, _ffi.ptr(out.get("cls")), _ffi.ptr(out.get("seg")),
                                               _ffi.ptr(out["R"]), _ffi.ptr(out.get("cls_logits")), _ffi.ptr(out.get("seg_logits")),
                                               _ffi.stream_ptr()))
        return out

    def pooled_features(self, which: int, b: int) -> np.ndarray:
        a = np.empty((b, 1024), np.float32)
        _ffi.check(self._lib.pcp_model_get_pooled(self._handle, which, b, _ffi.fptr(a)))
        return a

    def predict(self, x, batch_size: int = 32, verbose: int = 0):
        """keras Model.predict: batches of `batch_size`, outputs concatenated as NumPy."""
        x = np.asarray(x, np.float32)
        outs = [[], [], []]
        for i in range(0, x.shape[0], batch_size):
            o = self(x[i:i + batch_size])
            for j in range(3):
                outs[j].append(o[j])
        return [np.concatenate(o, 0) for o in outs]

    # ------------------------------------------------------------------ training (pointnet_train.py:221-228, :310-351)
    def compile(self, optimizer=None, loss=None, loss_weights=None, metrics=None, sync_bn=False, **kwargs):
        """keras Model.compile as the reference driver calls it (pointnet_train.py:334-351): Adam (+ ExponentialDecay),
        SCCE on the two softmax outputs, MSE on `se3`, per-output loss weights.  Resets the optimizer state (a fresh
        optimizer is compiled for every training profile, pointnet_train.py:310-319).
        sync_bn=True (data parallel only): BatchNorm batch statistics are all-reduced over the ranks inside the step, so
        N ranks x b/N clouds compute exactly the reference's single-device step on b clouds."""
        from . import keras_compat as K
        opt = optimizer if optimizer is not None else K.Adam()
        if not isinstance(opt, K.Adam):
            raise TypeError("only Adam (the reference's optimizer) is implemented on the device")
        lw = {"classification_output": 1.0, "segmentation_output": 1.0, "se3": 1.0}
        lw.update(loss_weights or {})
        self._compiled = {"optimizer": opt, "loss_weights": lw}
        self._set_sync_bn(bool(sync_bn))
        self._step = 0
        self._dropout_seed = (self._random_seed if isinstance(self._random_seed, int) else 0) * 7919 + 12345
        self._adam_reset = True
        return self

    def _set_sync_bn(self, on: bool):
        """Register / clear the BatchNorm all-reduce callback of the training step (pcp_model_set_bn_allreduce)."""
        import torch.distributed as dist
        if not on or not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
            if getattr(self, "_sync_cb", None) is not None:
                _ffi.check(self._lib.pcp_model_set_bn_allreduce(self._handle, None, None))
            self._sync_cb = None
            return
        dev = torch.cuda.current_device()

        def _allreduce(user, ptr, count, dtype, stream):          # called by the library on this thread, mid-step
            try:
                class _Dev:
                    __cuda_array_interface__ = {"shape": (int(count),), "typestr": "<f8" if dtype == 1 else "<f4", "data": (int(ptr), False), "version": 2}
                dist.all_reduce(torch.as_tensor(_Dev(), device=torch.device("cuda", dev)), op=dist.ReduceOp.SUM)
                return 0
            except Exception:                                      # never let an exception cross the C boundary
                import traceback
                traceback.print_exc()
                return 1
        self._sync_cb = _ffi.ALLREDUCE_FN(_allreduce)              # keep the thunk alive as long as it is registered
        _ffi.check(self._lib.pcp_model_set_bn_allreduce(self._handle, C.cast(self._sync_cb, C.c_void_p), None))

    def _train_config(self, explicit_masks=False, sample_offset=0, skip_unweighted_heads=False):
        if self._compiled is None:
            raise RuntimeError("call compile() before training")
        opt, lw = self._compiled["optimizer"], self._compiled["loss_weights"]
        lr0, steps, rate = opt.schedule()
        return _ffi.TrainConfig(lw["classification_output"], lw["segmentation_output"], lw["se3"], lr0, rate, steps,
                                opt.beta_1, opt.beta_2, opt.epsilon, (self._dropout_seed + self._step) & (2**64 - 1), sample_offset,
                                int(explicit_masks), int(skip_unweighted_heads))

    def _grad_tensor(self) -> torch.Tensor:
        """The library's flat fp32 gradient buffer as a zero-copy torch tensor (one NCCL bucket)."""
        if getattr(self, "_grads", None) is None:
            p, cnt = C.c_void_p(), C.c_size_t()
            _ffi.check(self._lib.pcp_model_grad_buffer(self._handle, C.byref(p), C.byref(cnt)))

            class _Dev:
                __cuda_array_interface__ = {"shape": (cnt.value,), "typestr": "<f4", "data": (p.value, False), "version": 2}
            self._grads = torch.as_tensor(_Dev(), device=torch.device("cuda", torch.cuda.current_device()))
        return self._grads

    @staticmethod
    def _labels_to_device(y, dev):
        yc = y.get("classification_output") if y else None
        ys = y.get("segmentation_output") if y else None
        yr = y.get("se3") if y else None
        conv = lambda a, dt: None if a is None else torch.as_tensor(np.asarray(a) if not isinstance(a, torch.Tensor) else a).to(device=dev, dtype=dt).contiguous()
        return conv(yc, torch.int32), conv(ys, torch.int32), conv(yr, torch.float32)

    METRIC_KEYS = ["loss", "classification_output_loss", "segmentation_output_loss", "se3_loss",
                   "classification_output_sparse_categorical_accuracy", "segmentation_output_sparse_categorical_accuracy",
                   "se3_root_mean_squared_error", "regularization_loss"]

    def train_on_batch(self, x, y, dropout_masks=None, global_batch=None, sample_offset=0, skip_unweighted_heads=False,
                       apply=True, return_dict=True, allreduce=True):
        """One optimisation step: training-mode forward, losses, backward, [gradient all-reduce], Adam.
        `x` (b,n,3) and labels may be NumPy (copied H2D here) or CUDA tensors.  Under torch.distributed the flat gradient
        bucket is all-reduced (SUM; the loss already divides by the global batch) before the identical Adam step."""
        import torch.distributed as dist
        dev = torch.device("cuda", torch.cuda.current_device())
        xd = (x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(np.asarray(x, np.float32)))).to(dev, torch.float32).contiguous()
        b, n, _ = xd.shape
        yc, ys, yr = self._labels_to_device(y, dev)
        world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        gb = int(global_batch) if global_batch is not None else b * world
        m1 = m2 = None
        if dropout_masks is not None:
            m1 = torch.as_tensor(np.asarray(dropout_masks[0], np.float32)).to(dev).contiguous()
            m2 = torch.as_tensor(np.asarray(dropout_masks[1], np.float32)).to(dev).contiguous()
        cfg = self._train_config(dropout_masks is not None, sample_offset, skip_unweighted_heads)
        met = np.zeros(8, np.float32)
        st = _ffi.stream_ptr()
        _ffi.check(self._lib.pcp_model_loss_backward(self._handle, _ffi.ptr(xd), b, n, _ffi.ptr(yc), _ffi.ptr(ys), _ffi.ptr(yr), C.byref(cfg), gb,
                                                     _ffi.ptr(m1), _ffi.ptr(m2), _ffi.fptr(met), st))
        if world > 1 and allreduce:
            dist.all_reduce(self._grad_tensor(), op=dist.ReduceOp.SUM)       # NCCL over NVLink: one 16.9 MB bucket
        if apply:
            if getattr(self, "_adam_reset", False):
                self._reset_adam_state()
            _ffi.check(self._lib.pcp_model_adam_step(self._handle, C.byref(cfg), self._step, st))
            self._step += 1
        return dict(zip(self.METRIC_KEYS, map(float, met))) if return_dict else met

    def _reset_adam_state(self):
        _ffi.check(self._lib.pcp_model_reset_optimizer(self._handle))
        self._adam_reset = False

    def test_on_batch(self, x, y, return_dict=True):
        dev = torch.device("cuda", torch.cuda.current_device())
        xd = (x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(np.asarray(x, np.float32)))).to(dev, torch.float32).contiguous()
        b, n, _ = xd.shape
        yc, ys, yr = self._labels_to_device(y, dev)
        cfg = self._train_config()
        met = np.zeros(8, np.float32)
        _ffi.check(self._lib.pcp_model_evaluate(self._handle, _ffi.ptr(xd), b, n, _ffi.ptr(yc), _ffi.ptr(ys), _ffi.ptr(yr), C.byref(cfg),
                                                _ffi.fptr(met), _ffi.stream_ptr()))
        return dict(zip(self.METRIC_KEYS, map(float, met))) if return_dict else met

    def evaluate(self, x, steps=None, verbose=0, return_dict=True):
        """keras Model.evaluate over an iterator of (x, labels) batches: metric means over `steps` batches."""
        tot, cnt = np.zeros(8, np.float64), 0
        for i, (xb, yb) in enumerate(x):
            if steps is not None and i >= steps:
                break
            tot += self.test_on_batch(xb, yb, return_dict=False)
            cnt += 1
        mean = tot / max(cnt, 1)
        return dict(zip(self.METRIC_KEYS, map(float, mean))) if return_dict else list(map(float, mean))

    def fit(self, x=None, validation_data=None, epochs=1, verbose=1, callbacks=None, steps_per_epoch=None, validation_steps=None, **kwargs):
        """keras Model.fit as the reference driver calls it (pointnet_train.py:221-228): `x` / `validation_data` are
        (infinite) iterators of (points (b,n,3), {'classification_output','segmentation_output','se3'}) batches.
        Returns a History whose `.history` has the reference's 14 keys (7 metrics + their val_ twins)."""
        from . import keras_compat as K
        if steps_per_epoch is None:
            raise ValueError("steps_per_epoch is required (the reference passes count // batch_size)")
        hist = K.History()
        cbs = [hist] + list(callbacks or [])
        for cb in cbs:
            cb.set_model(self)
            cb.on_train_begin()
        self.stop_training = False
        it = iter(x)
        keys = self.METRIC_KEYS[:7]
        import torch.distributed as dist
        from .. import parallel
        rank, world = parallel.world()
        for epoch in range(epochs):
            tot = np.zeros(8, np.float64)
            for _ in range(steps_per_epoch):
                xb, yb = next(it)
                if world > 1:        # every rank draws the same global batch and trains on its own contiguous shard
                    gb = len(xb)
                    xb, yb, off = parallel.shard_batch(xb, yb)
                    tot += self.train_on_batch(xb, yb, global_batch=gb, sample_offset=off, return_dict=False)
                else:
                    tot += self.train_on_batch(xb, yb, return_dict=False)
            logs_vec = tot[:7] / steps_per_epoch
            if world > 1:
                self._average_bn_statistics()
            val_vec = None
            if validation_data is not None and validation_steps:
                val = self.evaluate(iter(validation_data), steps=validation_steps)
                val_vec = np.array([val[k] for k in keys], np.float64)
            if world > 1:            # identical logs on every rank => identical callback decisions (no divergent early stop)
                dev = torch.device("cuda", torch.cuda.current_device())
                t = torch.from_numpy(np.concatenate([logs_vec, val_vec if val_vec is not None else np.zeros(0)])).to(dev)
                dist.all_reduce(t, op=dist.ReduceOp.SUM)
                t = (t / world).cpu().numpy()
                logs_vec, val_vec = t[:7], (t[7:] if val_vec is not None else None)
            logs = {k: float(v) for k, v in zip(keys, logs_vec)}
            if val_vec is not None:
                logs.update({f"val_{k}": float(v) for k, v in zip(keys, val_vec)})
            if verbose and rank == 0:
                print(f"Epoch {epoch + 1}/{epochs} - " + " - ".join(f"{k}: {v:.4f}" for k, v in logs.items()))
            for cb in cbs:
                cb.on_epoch_end(epoch, logs)
            if self.stop_training:
                break
        for cb in cbs:
            cb.on_train_end()
        return hist

    def _average_bn_statistics(self):
        """Data parallel: BN moving statistics are accumulated per shard; average them across ranks once per epoch so
        every rank validates (and checkpoints) the same model."""
        import torch.distributed as dist
        dev = torch.device("cuda", torch.cuda.current_device())
        w = self.get_weights_dict()
        names = [k for k in w if "/bn/moving_" in k]
        flat = torch.from_numpy(np.concatenate([w[k] for k in names])).to(dev)
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        flat = (flat / dist.get_world_size()).cpu().numpy()
        o = 0
        for k in names:
            self.set_tensor(k, flat[o:o + w[k].size])
            o += w[k].size

    def gradients_dict(self) -> dict:
        """Every gradient tensor of the last loss_backward (tests)."""
        out = {}
        for name, cnt in self.tensor_table():
            a = np.empty(cnt, np.float32)
            _ffi.check(self._lib.pcp_model_get_grad(self._handle, name.encode(), _ffi.fptr(a), cnt))
            out[name] = a
        return out

    def train_output(self, which: int, count: int) -> np.ndarray:
        a = np.empty(count, np.float32)
        _ffi.check(self._lib.pcp_model_get_train_output(self._handle, which, _ffi.fptr(a), count))
        return a

    # ------------------------------------------------------------------ freeze / thaw (PointNet.py:294-342)
    def _sync_trainable(self):
        if self._handle is None:
            return
        groups = {
            _ffi.GROUP_SHARED: self.mlp_1_1.is_trainable(),
            _ffi.GROUP_INPUT_TRANSFORM: self.input_transform.is_trainable() if self.input_transform is not None else False,
            _ffi.GROUP_CLS_HEAD: self.mlp_cls_1.is_trainable(),
            _ffi.GROUP_SEG_HEAD: self.mlp_seg_1.is_trainable(),
        }
        for g, f in groups.items():
            _ffi.check(self._lib.pcp_model_set_trainable(self._handle, g, int(f)))

    def freeze_input_transform(self) -> None:
        if self._vanilla:
            print("PointNet:  No input transorm available to freeze.")
        else:
            self.input_transform.freeze()
        self._sync_trainable()

    def thaw_input_transform(self) -> None:
        if self._vanilla:
            print("PointNet:  No input transorm available to thaw.")
        else:
            self.input_transform.thaw()
        self._sync_trainable()

    def _shared_layers(self):
        # q1: the reference dereferences the T-Nets unconditionally and crashes in vanilla
        # mode (PointNet.py:302-318); absent T-Nets are skipped here.
        ls = [self.input_transform, self.mlp_1_1, self.mlp_1_2, self.feature_transform, self.mlp_2_1, self.mlp_2_2, self.mlp_2_3]
        return [l for l in ls if l is not None]

    def freeze_shared_network(self) -> None:
        for l in self._shared_layers():
            l.freeze()
        self._sync_trainable()

    def thaw_shared_network(self) -> None:
        for l in self._shared_layers():
            l.thaw()
        self._sync_trainable()

    def freeze_segmentation_head(self) -> None:
        for l in (self.mlp_seg_1, self.mlp_seg_2, self.mlp_seg_3, self.mlp_seg_4, self.mlp_seg_5):
            l.freeze()
        self._sync_trainable()

    def thaw_segmentation_head(self) -> None:
        for l in (self.mlp_seg_1, self.mlp_seg_2, self.mlp_seg_3, self.mlp_seg_4, self.mlp_seg_5):
            l.thaw()
        self._sync_trainable()

    def freeze_classification_head(self) -> None:
        for l in (self.mlp_cls_1, self.mlp_cls_2, self.mlp_cls_3):
            l.freeze()
        self._sync_trainable()

    def thaw_classification_head(self) -> None:
        for l in (self.mlp_cls_1, self.mlp_cls_2, self.mlp_cls_3):
            l.thaw()
        self._sync_trainable()

    def get_layer_trainability(self) -> dict:                     # PointNet.py:344-349
        return {l.name: l.is_trainable() for l in self._custom_layers}

    def get_config(self):                                         # PointNet.py:351-364 (+ vanilla, q2)
        return {"classification_output_width": self._classification_output_width,
                "segmentation_output_width": self._segmentation_output_width,
                "dropout_rate": self._dropout_rate, "random_seed": self._random_seed, "debugging": self._debugging,
                "vanilla": self._vanilla, "regularize_input_transform": self._regularize_input_transform,
                "regularize_feature_transform": self._regularize_feature_transform, "precision": self._precision}

    @classmethod
    def from_config(cls, config):                                 # PointNet.py:366-373
        return cls(**config)

    def get_last_predicted_dcm(self):                             # PointNet.py:375-376
        return None if self.input_transform is None else self.input_transform.get_last_predicted_transformation()
