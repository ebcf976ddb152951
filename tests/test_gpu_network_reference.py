"""GPU: the CUDA path, through the reference-interface mirror over the C-ABI, DIRECTLY against outputs of the reference's own
model code (tests/golden/network_reference.npz — the unmodified PointNet.py executed on tests/golden/keras_standin.py by
tests/golden/make_network_golden.py).  Tolerances: fp32 probabilities / rotation 1e-3 (north_star), bf16 2e-2, identical
class arg-max; gradients 1e-3 in the all-ReLUs-active regime, 6e-2 / 3e-2 with ordinary weights (tests/test_gpu_train.py)."""
import json
from pathlib import Path

import numpy as np
import pytest

from oracle import pointnet_numpy as onp

pytestmark = pytest.mark.gpu

G = np.load(Path(__file__).parent / "golden" / "network_reference.npz")
META = json.loads(str(G["meta"]))
C, P, SEED = META["C"], META["P"], META["seed"]


def err(a, ref):
    return float(np.abs(np.asarray(a, np.float64) - np.asarray(ref, np.float64)).max() / max(1.0, np.abs(ref).max()))


def model(vanilla=False, reg=False, precision="fp32", smooth=False):
    from pointcloudprocessing_b200.pointnet import PointNet
    p = onp.make_params(C, P, seed=SEED, vanilla=vanilla)
    if smooth:
        p = {k: (v + np.float32(6.0)).astype(np.float32) if k.endswith("/bn/beta") else v for k, v in p.items()}
    m = PointNet(C, P, 0.0, 3, vanilla=vanilla, regularize_input_transform=reg, regularize_feature_transform=reg, precision=precision)
    m.set_weights_dict(p)
    return m, p


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-3), ("bf16", 2e-2)])
@pytest.mark.parametrize("tag,vanilla", [("inf", False), ("van", True)])
def test_inference_matches_reference_graph(lib, tag, vanilla, precision, tol):
    m, _ = model(vanilla=vanilla, precision=precision)
    cls, seg, R = m(G["pc"])
    assert err(cls, G[f"{tag}_cls"]) <= tol and err(seg, G[f"{tag}_seg"]) <= tol and err(R, G[f"{tag}_R"]) <= tol
    assert np.array_equal(cls.argmax(-1), G[f"{tag}_cls"].argmax(-1))
    agree = float((seg.argmax(-1) == G[f"{tag}_seg"].argmax(-1)).mean())
    assert agree >= (0.999 if precision == "fp32" else 0.97), agree
    if vanilla:
        assert np.array_equal(R, np.tile(np.eye(3, dtype=np.float32), (G["pc"].shape[0], 1, 1)))


@pytest.mark.parametrize("tag", ["trn", "smo"])
def test_training_step_matches_reference_graph(lib, tag):
    from pointcloudprocessing_b200.pointnet import keras_compat as K
    smooth = tag == "smo"
    m, p = model(reg=True, smooth=smooth)
    m.compile(optimizer=K.Adam(learning_rate=1e-3),
              loss_weights={"classification_output": META["w_cls"], "segmentation_output": META["w_seg"], "se3": META["w_rot"]})
    b, n = G["pc"].shape[:2]
    y = {"classification_output": G["y_cls"].astype(np.int32), "segmentation_output": G["y_seg"].astype(np.int32), "se3": G["y_se3"]}
    met = m.train_on_batch(G["pc"], y, apply=False)                    # dropout_rate 0: no mask drawn
    want = G[tag + "_losses"]                                          # total, cls, seg, rot, sum(model.losses)
    assert abs(met["loss"] - want[0]) <= 1e-4 * want[0], (met, want)
    assert abs(met["classification_output_loss"] - want[1]) <= 1e-4 * max(1.0, want[1])
    assert abs(met["segmentation_output_loss"] - want[2]) <= 1e-4 * max(1.0, want[2])
    assert abs(met["se3_loss"] - want[3]) <= 1e-4 * max(1.0, want[3])
    assert err(m.train_output(0, b * C).reshape(b, C), G[tag + "_cls"]) <= 1e-3
    assert err(m.train_output(1, b * n * P).reshape(b, n, P), G[tag + "_seg"]) <= 1e-3
    assert err(m.train_output(2, b * 9).reshape(b, 3, 3), G[tag + "_R"]) <= 1e-3
    grads = m.gradients_dict()
    tol_norm, tol_elem = (1e-3, 1e-3) if smooth else (3e-2, 6e-2)
    gmax = G[tag + "_grad_norms"].max()
    for name, ref in zip(G[tag + "_grad_names"], G[tag + "_grad_norms"]):
        got = float(np.linalg.norm(grads[str(name)].astype(np.float64)))
        assert abs(got - ref) <= tol_norm * ref + 1e-5 * gmax, (str(name), got, ref)
    w = m.get_weights_dict()
    for key in G.files:
        if key.startswith(tag + "_grad:"):
            k = key.split(":", 1)[1]
            assert np.abs(grads[k].reshape(G[key].shape) - G[key]).max() <= tol_elem * np.abs(G[key]).max() + 1e-5 * gmax, k
        if key.startswith(tag + "_stat:"):
            k = key.split(":", 1)[1]
            assert err(w[k].reshape(G[key].shape), G[key]) <= 1e-4, k


def test_frozen_shared_network_matches_reference_graph(lib):
    from pointcloudprocessing_b200.pointnet import keras_compat as K
    m, p = model()
    m.freeze_shared_network()
    m.compile(optimizer=K.Adam(learning_rate=1e-3), loss_weights={"classification_output": 1.0, "segmentation_output": 1.0, "se3": 0.5})
    b, n = G["pc"].shape[:2]
    y = {"classification_output": G["y_cls"].astype(np.int32), "segmentation_output": G["y_seg"].astype(np.int32), "se3": G["y_se3"]}
    m.train_on_batch(G["pc"], y, apply=False)
    assert err(m.train_output(0, b * C).reshape(b, C), G["frz_cls"]) <= 1e-3
    assert err(m.train_output(1, b * n * P).reshape(b, n, P), G["frz_seg"]) <= 1e-3
    assert err(m.train_output(2, b * 9).reshape(b, 3, 3), G["frz_R"]) <= 1e-3
    w = m.get_weights_dict()
    k = "s1_l1_64_convolution_layer/bn/moving_mean"
    assert np.array_equal(w[k].ravel(), G["frz_stat:" + k].ravel())                       # frozen BN: moving statistics untouched
    k = "seg_l1_512_convolution_layer/bn/moving_mean"
    assert err(w[k].ravel(), G["frz_stat:" + k].ravel()) <= 1e-4


def test_freeze_thaw_bookkeeping_matches_reference(lib):
    """get_layer_trainability() after every step of a freeze/thaw sequence, as the reference's own methods report it
    (PointNet.py:294-349; names and order of _custom_layers included)."""
    m, _ = model()
    trace = json.loads(str(G["trainability_json"]))
    got = m.get_layer_trainability()
    assert list(got.items()) == list(trace[0][1].items())
    for call, want in trace[1:]:
        getattr(m, call)()
        got = m.get_layer_trainability()
        assert list(got.items()) == list(want.items()), call
