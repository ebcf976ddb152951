"""ONNX (opset 13) export of the inference graph — the artefact `pointnet_train.py:237-247` writes next to every
checkpoint with `tf2onnx.convert.from_keras(model, input_signature=[(None, n, 3) f32], opset=13)`.

Neither `onnx` nor `tf2onnx` exists in this environment, so the file is serialised straight from the ONNX schema: the
subset of `onnx.proto` (IR version 7 — the one that introduced opset 13) that a feed-forward graph needs is declared below
with its published field numbers and built into message classes by `google.protobuf` (a plugin-free, dependency-free
writer; any ONNX runtime reads the result).  Graph = `PointNet.call` (PointNet.py:197-292) in inference mode:

* input `pointnet_input` (N, n, 3) f32; outputs `classification_output` (N, C), `segmentation_output` (N, n, P), `se3`
  (N, 3, 3) — the names of PointNet.py:113-114;
* `PointCloudNormalization` (:691-706) op for op: ReduceMean, Sub, Mul, ReduceSum, Sqrt, ReduceMax, Max(1e-7), Div;
* every ConvLayer / DenseLayer (:554-566, :642-654) as MatMul -> Mul, Add (BatchNormalization in inference mode, folded to
  the per-channel affine the way tf2onnx's optimiser leaves it) -> Relu; 1x1 convolutions on (N, n, 1, Cin) are MatMuls on
  (N, n, Cin);
* T-Net (:418-454): ReduceMax over points, two dense layers, MatMul with `w`, Reshape to (-1, K, K), Add `b`;
* the segmentation head's `concat([X_64, tile(global)])` (:268-270) followed by the 1088 -> 512 convolution is written as
  X_64 . W[:64] + (global . W[64:]) broadcast over the points — the same sum without a shape-dependent Tile;
* Dropout is the identity in inference and is left out, as tf2onnx does.
"""
from __future__ import annotations

import numpy as np
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

BN_EPS = 1e-3                      # keras BatchNormalization default
TNET_WIDTHS = (64, 128, 1024, 512, 256)
_F = descriptor_pb2.FieldDescriptorProto

# ---- the onnx.proto subset (message, [(field name, number, type, label, type_name)]) ----
_SCHEMA = {
    "OperatorSetIdProto": [("domain", 1, _F.TYPE_STRING, 1, None), ("version", 2, _F.TYPE_INT64, 1, None)],
    "StringStringEntryProto": [("key", 1, _F.TYPE_STRING, 1, None), ("value", 2, _F.TYPE_STRING, 1, None)],
    "TensorProto": [("dims", 1, _F.TYPE_INT64, 3, None), ("data_type", 2, _F.TYPE_INT32, 1, None), ("float_data", 4, _F.TYPE_FLOAT, 3, None),
                    ("int64_data", 7, _F.TYPE_INT64, 3, None), ("name", 8, _F.TYPE_STRING, 1, None), ("raw_data", 9, _F.TYPE_BYTES, 1, None)],
    "Dimension": [("dim_value", 1, _F.TYPE_INT64, 1, None), ("dim_param", 2, _F.TYPE_STRING, 1, None)],
    "TensorShapeProto": [("dim", 1, _F.TYPE_MESSAGE, 3, "Dimension")],
    "TensorTypeProto": [("elem_type", 1, _F.TYPE_INT32, 1, None), ("shape", 2, _F.TYPE_MESSAGE, 1, "TensorShapeProto")],
    "TypeProto": [("tensor_type", 1, _F.TYPE_MESSAGE, 1, "TensorTypeProto")],
    "ValueInfoProto": [("name", 1, _F.TYPE_STRING, 1, None), ("type", 2, _F.TYPE_MESSAGE, 1, "TypeProto")],
    "AttributeProto": [("name", 1, _F.TYPE_STRING, 1, None), ("f", 2, _F.TYPE_FLOAT, 1, None), ("i", 3, _F.TYPE_INT64, 1, None),
                       ("s", 4, _F.TYPE_BYTES, 1, None), ("t", 5, _F.TYPE_MESSAGE, 1, "TensorProto"), ("floats", 7, _F.TYPE_FLOAT, 3, None),
                       ("ints", 8, _F.TYPE_INT64, 3, None), ("type", 20, _F.TYPE_INT32, 1, None)],
    "NodeProto": [("input", 1, _F.TYPE_STRING, 3, None), ("output", 2, _F.TYPE_STRING, 3, None), ("name", 3, _F.TYPE_STRING, 1, None),
                  ("op_type", 4, _F.TYPE_STRING, 1, None), ("attribute", 5, _F.TYPE_MESSAGE, 3, "AttributeProto"), ("domain", 7, _F.TYPE_STRING, 1, None)],
    "GraphProto": [("node", 1, _F.TYPE_MESSAGE, 3, "NodeProto"), ("name", 2, _F.TYPE_STRING, 1, None), ("initializer", 5, _F.TYPE_MESSAGE, 3, "TensorProto"),
                   ("doc_string", 10, _F.TYPE_STRING, 1, None), ("input", 11, _F.TYPE_MESSAGE, 3, "ValueInfoProto"),
                   ("output", 12, _F.TYPE_MESSAGE, 3, "ValueInfoProto")],
    "ModelProto": [("ir_version", 1, _F.TYPE_INT64, 1, None), ("producer_name", 2, _F.TYPE_STRING, 1, None), ("producer_version", 3, _F.TYPE_STRING, 1, None),
                   ("domain", 4, _F.TYPE_STRING, 1, None), ("model_version", 5, _F.TYPE_INT64, 1, None), ("doc_string", 6, _F.TYPE_STRING, 1, None),
                   ("graph", 7, _F.TYPE_MESSAGE, 1, "GraphProto"), ("opset_import", 8, _F.TYPE_MESSAGE, 3, "OperatorSetIdProto"),
                   ("metadata_props", 14, _F.TYPE_MESSAGE, 3, "StringStringEntryProto")],
}
ATTR_FLOAT, ATTR_INT, ATTR_INTS = 1, 2, 7          # AttributeProto.AttributeType
DT_FLOAT, DT_INT64 = 1, 7                          # TensorProto.DataType
_PKG = "pcp_onnx_subset"
_classes = None


def messages():
    """{message name: class} for the schema subset (built once; proto2 like onnx.proto, so scalar presence is explicit)."""
    global _classes
    if _classes is None:
        fd = descriptor_pb2.FileDescriptorProto(name=f"{_PKG}.proto", package=_PKG, syntax="proto2")
        for mname, fields in _SCHEMA.items():
            m = fd.message_type.add(name=mname)
            for fname, num, ftype, label, tname in fields:
                f = m.field.add(name=fname, number=num, type=ftype, label=label)
                if tname:
                    f.type_name = f".{_PKG}.{tname}"
                if label == 3 and fname.endswith("_data"):
                    f.options.packed = True                  # onnx.proto: float_data / int64_data are [packed = true]; dims, ints, floats are not
        pool = descriptor_pool.DescriptorPool()
        pool.Add(fd)
        _classes = {n: message_factory.GetMessageClass(pool.FindMessageTypeByName(f"{_PKG}.{n}")) for n in _SCHEMA}
    return _classes


class _Graph:
    def __init__(self):
        self.M = messages()
        self.g = self.M["GraphProto"](name="pointnet")
        self._n = 0

    def const(self, name, arr, dtype=np.float32):
        a = np.ascontiguousarray(arr, dtype)
        t = self.g.initializer.add(name=name, data_type=DT_FLOAT if dtype == np.float32 else DT_INT64, raw_data=a.astype(a.dtype.newbyteorder("<")).tobytes())
        t.dims.extend(a.shape)
        return name

    def node(self, op, inputs, out=None, **attrs):
        self._n += 1
        out = out or f"{op.lower()}_{self._n}"
        nd = self.g.node.add(op_type=op, name=f"{op}_{self._n}")
        nd.input.extend(inputs)
        nd.output.append(out)
        for k, v in attrs.items():
            a = nd.attribute.add(name=k)
            if isinstance(v, float):
                a.type, a.f = ATTR_FLOAT, v
            elif isinstance(v, int):
                a.type, a.i = ATTR_INT, v
            else:
                a.type = ATTR_INTS
                a.ints.extend(int(i) for i in v)
        return out

    def value_info(self, field, name, shape):
        vi = field.add(name=name)
        vi.type.tensor_type.elem_type = DT_FLOAT
        for d in shape:
            dim = vi.type.tensor_type.shape.dim.add()
            if isinstance(d, str):
                dim.dim_param = d
            else:
                dim.dim_value = int(d)


def _tnet_names(name):
    return [f"{name}_convolution_layer_{i}_convolution_layer" for i in (1, 2, 3)] + [f"{name}_dense_layer_{i}_dense_layer" for i in (1, 2)]


def build_onnx(params: dict, num_classes: int, num_parts: int, input_width: int, vanilla: bool = False, producer_version: str = "") -> bytes:
    """Serialised ModelProto of the inference graph.  `params`: {reference variable name: array} as `get_weights_dict()` /
    `oracle.pointnet_numpy.make_params` give them (conv kernels (Cin, Cout) or Keras' (1, 1, Cin, Cout))."""
    G = _Graph()
    p = {k: np.asarray(v, np.float32) for k, v in params.items()}

    def kernel(name, cin, cout):
        return p[f"{name}/kernel"].reshape(cin, cout)

    def layer(x, name, cin, cout, relu=True, w=None):
        """ConvLayer / DenseLayer: MatMul (+ bias) -> folded inference BN -> Relu."""
        y = G.node("MatMul", [x, G.const(f"{name}/kernel", kernel(name, cin, cout) if w is None else w)])
        return finish(y, name, cout, relu)

    def finish(y, name, cout, relu=True):
        if f"{name}/bias" in p:
            y = G.node("Add", [y, G.const(f"{name}/bias", p[f"{name}/bias"].reshape(cout))])
        if f"{name}/bn/gamma" in p:
            s = (p[f"{name}/bn/gamma"] / np.sqrt(p[f"{name}/bn/moving_variance"] + np.float32(BN_EPS))).astype(np.float32).reshape(cout)
            t = (p[f"{name}/bn/beta"].reshape(cout) - p[f"{name}/bn/moving_mean"].reshape(cout) * s).astype(np.float32)
            y = G.node("Add", [G.node("Mul", [y, G.const(f"{name}/bn/scale", s)]), G.const(f"{name}/bn/shift", t)])
        return G.node("Relu", [y]) if relu else y

    def tnet(x, name, k, out=None):
        names = _tnet_names(name)
        widths = (k,) + TNET_WIDTHS
        h = x
        for i in range(3):
            h = layer(h, names[i], widths[i], widths[i + 1])
        h = G.node("ReduceMax", [h], axes=[1], keepdims=0)                        # PointNet.py:429
        for i in (3, 4):
            h = layer(h, names[i], widths[i], widths[i + 1])
        m = G.node("MatMul", [h, G.const(f"{name}/w", p[f"{name}/w"].reshape(256, k * k))])
        m = G.node("Reshape", [m, G.const(f"{name}/shape", np.array([-1, k, k]), np.int64)])
        return G.node("Add", [m, G.const(f"{name}/b", p[f"{name}/b"].reshape(k, k))], out=out)

    x = "pointnet_input"
    # ---- PointCloudNormalization (PointNet.py:691-706) ----
    c = G.node("Sub", [x, G.node("ReduceMean", [x], axes=[1], keepdims=1)])
    d = G.node("Sqrt", [G.node("ReduceSum", [G.node("Mul", [c, c]), G.const("norm/axis", np.array([2]), np.int64)], keepdims=1)])
    scale = G.node("Max", [G.node("ReduceMax", [d], axes=[1], keepdims=1), G.const("norm/eps", np.array(1e-7, np.float32))])
    pc = G.node("Div", [c, scale])
    # ---- trunk (PointNet.py:205-248) ----
    if not vanilla:
        R = tnet(pc, "input_transform", 3, out="se3")
        X = G.node("MatMul", [pc, R])
    else:
        R = G.node("Add", [G.node("Mul", [G.node("ReduceSum", [pc, G.const("eye/axes", np.array([1, 2]), np.int64)], keepdims=1),
                                          G.const("eye/zero", np.zeros((1, 1, 1), np.float32))]),
                           G.const("eye/I", np.eye(3, dtype=np.float32).reshape(1, 3, 3))], out="se3")       # tile(eye(3), (N, 1, 1)), :211
        X = pc
    X = layer(X, "s1_l1_64_convolution_layer", 3, 64)
    X = layer(X, "s1_l2_64_convolution_layer", 64, 64)
    X64 = G.node("MatMul", [X, tnet(X, "feature_transform", 64)]) if not vanilla else X
    X = layer(X64, "s2_l1_64_convolution_layer", 64, 64)
    X = layer(X, "s2_l2_128_convolution_layer", 64, 128)
    X = layer(X, "s2_l3_1024_convolution_layer", 128, 1024)
    gf = G.node("ReduceMax", [X], axes=[1], keepdims=0)
    # ---- classification head (PointNet.py:252-263) ----
    h = layer(gf, "s3_l1_512_dense_layer", 1024, 512)
    h = layer(h, "s3_l2_256_dense_layer", 512, 256)
    G.node("Softmax", [layer(h, "output_dense_layer", 256, num_classes, relu=False)], out="classification_output", axis=-1)
    # ---- segmentation head (PointNet.py:268-290) ----
    name = "seg_l1_512_convolution_layer"
    w1 = kernel(name, 1088, 512)
    local = G.node("MatMul", [X64, G.const(f"{name}/kernel_local", w1[:64])])
    glob = G.node("Unsqueeze", [G.node("MatMul", [gf, G.const(f"{name}/kernel_global", w1[64:])]), G.const("seg/axis", np.array([1]), np.int64)])
    s = finish(G.node("Add", [local, glob]), name, 512)
    s = layer(s, "seg_l2_256_convolution_layer", 512, 256)
    s = layer(s, "seg_l3_128_convolution_layer", 256, 128)
    s = layer(s, "seg_l4_128_convolution_layer", 128, 128)
    G.node("Softmax", [layer(s, "seg_l5_output_convolution_layer", 128, num_parts, relu=False)], out="segmentation_output", axis=-1)

    G.value_info(G.g.input, "pointnet_input", ("N", input_width, 3))
    G.value_info(G.g.output, "classification_output", ("N", num_classes))
    G.value_info(G.g.output, "segmentation_output", ("N", input_width, num_parts))
    G.value_info(G.g.output, "se3", ("N", 3, 3))
    model = G.M["ModelProto"](ir_version=7, producer_name="pointcloudprocessing_b200", producer_version=producer_version, graph=G.g)
    model.opset_import.add(domain="", version=13)
    return model.SerializeToString()


def export_onnx(path, params: dict, num_classes: int, num_parts: int, input_width: int, vanilla: bool = False) -> str:
    data = build_onnx(params, num_classes, num_parts, input_width, vanilla)
    with open(path, "wb") as f:
        f.write(data)
    return str(path)
