"""ONNX export (pointnet_train.py:237-247): the file written without `onnx` / `tf2onnx` is decoded by an independent reader
and executed in NumPy (oracle/onnx_eval.py); its outputs must be the oracle's inference outputs."""
import numpy as np
import pytest

from oracle import onnx_eval
from oracle import pointnet_numpy as onp
from pointcloudprocessing_b200.pointnet import onnx_export
from tests import synth


@pytest.mark.parametrize("vanilla", [False, True])
def test_exported_graph_reproduces_the_oracle(vanilla, tmp_path):
    C, P, n = 23, 12, 256
    p = onp.make_params(C, P, seed=7, vanilla=vanilla)
    path = onnx_export.export_onnx(tmp_path / "m.onnx", p, C, P, n, vanilla=vanilla)
    model = onnx_eval.load(open(path, "rb").read())
    assert model["ir_version"] == 7 and model["opset"] == {"": 13}                      # opset = 13 (pointnet_train.py:244)
    assert model["inputs"] == [("pointnet_input", ["N", n, 3])]                          # input_signature (None, n, 3) f32 (:238-240)
    assert [o[0] for o in model["outputs"]] == ["classification_output", "segmentation_output", "se3"]   # PointNet.py:114
    assert model["outputs"][1][1] == ["N", n, P] and model["outputs"][0][1] == ["N", C]
    x = synth.gaussian_clouds(11, 3, n)
    got = onnx_eval.run(model, {"pointnet_input": x})
    ref = onp.forward(x, p, vanilla=vanilla)
    for k_onnx, k_ref in (("classification_output", "cls"), ("segmentation_output", "seg"), ("se3", "R")):
        assert got[k_onnx].shape == ref[k_ref].shape
        err = np.abs(got[k_onnx] - ref[k_ref]).max() / np.abs(ref[k_ref]).max()
        assert err <= 1e-5, (k_onnx, err)
    assert np.array_equal(got["classification_output"].argmax(-1), ref["cls"].argmax(-1))


def test_every_weight_of_the_layer_table_is_in_the_file():
    C, P = 5, 4
    p = onp.make_params(C, P, seed=3)
    model = onnx_eval.load(onnx_export.build_onnx(p, C, P, 64))
    init = model["initializers"]
    for kind, name, cin, cout, has_bn in onp.layer_table(C, P):
        if name == "seg_l1_512_convolution_layer":                                       # split into its X_64 and global rows (PointNet.py:270)
            w = np.concatenate([init[f"{name}/kernel_local"], init[f"{name}/kernel_global"]], 0)
        else:
            w = init[f"{name}/kernel"]
        assert np.array_equal(w, p[f"{name}/kernel"].reshape(cin, cout))
        assert (f"{name}/bn/scale" in init) == has_bn and (f"{name}/bias" in init) == (not has_bn)   # use_bias = not apply_bn (:540,630)
    for t, k in (("input_transform", 3), ("feature_transform", 64)):
        assert np.array_equal(init[f"{t}/w"], p[f"{t}/w"]) and np.array_equal(init[f"{t}/b"], p[f"{t}/b"].reshape(k, k))
    ops = {n[0] for n in model["nodes"]}
    assert ops <= set(onnx_eval._OPS) and "Dropout" not in ops
