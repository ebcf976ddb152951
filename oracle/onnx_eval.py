"""TEST INFRASTRUCTURE (checker only — nothing under pointcloudprocessing_b200/ imports this).

A minimal ONNX reader + NumPy interpreter for the operator subset the exported PointNet graph uses, so that the file
`pointnet/onnx_export.py` writes (the artefact of pointnet_train.py:237-247) can be executed here, where neither `onnx`
nor `onnxruntime` is installed.  The reader decodes the protobuf wire format itself (varint / 64-bit / length-delimited /
32-bit records) with the field numbers of onnx.proto written out again below — it shares no code with the writer.
Operator semantics follow the ONNX operator specification at opset 13 (ReduceSum takes its axes as an input, ReduceMean /
ReduceMax as an attribute; Softmax over a single axis; Unsqueeze axes as an input; MatMul = numpy.matmul).
"""
from __future__ import annotations

import struct

import numpy as np


def _varint(buf, i):
    v = s = 0
    while True:
        b = buf[i]
        i += 1
        v |= (b & 0x7F) << s
        if b < 0x80:
            return v, i
        s += 7


def _fields(buf):
    """[(field number, wire type, value)] of one message; value = int (varint), bytes (len-delimited) or raw 4/8 bytes."""
    out, i = [], 0
    while i < len(buf):
        key, i = _varint(buf, i)
        num, wt = key >> 3, key & 7
        if wt == 0:
            v, i = _varint(buf, i)
        elif wt == 1:
            v, i = bytes(buf[i:i + 8]), i + 8
        elif wt == 2:
            ln, i = _varint(buf, i)
            v, i = bytes(buf[i:i + ln]), i + ln
        elif wt == 5:
            v, i = bytes(buf[i:i + 4]), i + 4
        else:
            raise ValueError(f"wire type {wt}")
        out.append((num, wt, v))
    return out


def _s64(v):
    return v - (1 << 64) if v >= (1 << 63) else v


def _packed_ints(wt, v):
    if wt == 0:
        return [_s64(v)]
    out, i = [], 0
    while i < len(v):
        x, i = _varint(v, i)
        out.append(_s64(x))
    return out


def _tensor(buf):
    dims, dtype, raw, name, floats, ints = [], 0, None, "", [], []
    for num, wt, v in _fields(buf):
        if num == 1: dims += _packed_ints(wt, v)
        elif num == 2: dtype = v
        elif num == 4: floats += list(struct.unpack(f"<{len(v) // 4}f", v)) if wt == 2 else [struct.unpack("<f", v)[0]]
        elif num == 7: ints += _packed_ints(wt, v)
        elif num == 8: name = v.decode()
        elif num == 9: raw = v
    np_t = {1: np.float32, 7: np.int64}[dtype]
    a = np.frombuffer(raw, np.dtype(np_t).newbyteorder("<")).astype(np_t) if raw is not None else np.array(floats if dtype == 1 else ints, np_t)
    return name, a.reshape(dims)


def _attr(buf):
    name, val, ints = "", None, []
    for num, wt, v in _fields(buf):
        if num == 1: name = v.decode()
        elif num == 2: val = struct.unpack("<f", v)[0]
        elif num == 3: val = _s64(v)
        elif num == 8: ints += _packed_ints(wt, v)
    return name, (ints if ints else val)


def load(data: bytes):
    """-> dict(ir_version, opset, inputs [(name, shape)], outputs [(name, shape)], initializers {name: array},
    nodes [(op, [inputs], [outputs], {attr: value})])."""
    model = {"opset": {}, "inputs": [], "outputs": [], "initializers": {}, "nodes": []}

    def value_info(buf):
        name, shape = "", []
        for num, _, v in _fields(buf):
            if num == 1: name = v.decode()
            elif num == 2:
                for n2, _, tt in _fields(v):
                    if n2 != 1: continue
                    for n3, _, sh in _fields(tt):
                        if n3 != 2: continue
                        for n4, _, dim in _fields(sh):
                            if n4 != 1: continue
                            d = None
                            for n5, _, dv in _fields(dim):
                                d = _s64(dv) if n5 == 1 else dv.decode()
                            shape.append(d)
        return name, shape

    for num, _, v in _fields(data):
        if num == 1: model["ir_version"] = v
        elif num == 8:
            dom, ver = "", 0
            for n2, _, x in _fields(v):
                if n2 == 1: dom = x.decode()
                elif n2 == 2: ver = x
            model["opset"][dom] = ver
        elif num == 7:
            for n2, _, g in _fields(v):
                if n2 == 1:
                    ins, outs, op, attrs = [], [], "", {}
                    for n3, _, x in _fields(g):
                        if n3 == 1: ins.append(x.decode())
                        elif n3 == 2: outs.append(x.decode())
                        elif n3 == 4: op = x.decode()
                        elif n3 == 5:
                            k, val = _attr(x)
                            attrs[k] = val
                    model["nodes"].append((op, ins, outs, attrs))
                elif n2 == 5:
                    name, a = _tensor(g)
                    model["initializers"][name] = a
                elif n2 == 11: model["inputs"].append(value_info(g))
                elif n2 == 12: model["outputs"].append(value_info(g))
    return model


def _softmax(z, axis):
    z = z - z.max(axis=axis, keepdims=True)
    e = np.exp(z)
    return e / e.sum(axis=axis, keepdims=True)


_OPS = {
    "Sub": lambda a, i: i[0] - i[1], "Add": lambda a, i: i[0] + i[1], "Mul": lambda a, i: i[0] * i[1], "Div": lambda a, i: i[0] / i[1],
    "Sqrt": lambda a, i: np.sqrt(i[0]), "Max": lambda a, i: np.maximum(i[0], i[1]), "Relu": lambda a, i: np.maximum(i[0], 0),
    "MatMul": lambda a, i: np.matmul(i[0], i[1]),
    "ReduceMean": lambda a, i: i[0].mean(axis=tuple(a["axes"]), keepdims=bool(a.get("keepdims", 1))),
    "ReduceMax": lambda a, i: i[0].max(axis=tuple(a["axes"]), keepdims=bool(a.get("keepdims", 1))),
    "ReduceSum": lambda a, i: i[0].sum(axis=tuple(int(x) for x in i[1]), keepdims=bool(a.get("keepdims", 1))),
    "Reshape": lambda a, i: i[0].reshape([int(x) for x in i[1]]),
    "Unsqueeze": lambda a, i: np.expand_dims(i[0], tuple(int(x) for x in i[1])),
    "Softmax": lambda a, i: _softmax(i[0], a.get("axis", -1)),
}


def run(model, feeds: dict) -> dict:
    """Execute the graph in file order (the exporter emits nodes topologically sorted, as ONNX requires)."""
    env = dict(model["initializers"])
    env.update({k: np.asarray(v, np.float32) for k, v in feeds.items()})
    for op, ins, outs, attrs in model["nodes"]:
        env[outs[0]] = np.asarray(_OPS[op](attrs, [env[i] for i in ins]), np.float32)
    return {name: env[name] for name, _ in model["outputs"]}
