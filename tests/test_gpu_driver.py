"""GPU: the training driver mirror end to end (pointnet_train.py:63-257 of the reference): config JSON -> PointCloudSet per
profile (native frame parser, TFRecords) -> staged training (classification_pretrain then final, the second profile
continuing from the first one's checkpoint) -> per-profile artefacts."""
import json
import os

import numpy as np
import pytest

from tests.test_pointcloudset import CLASSES, PARTS, write_collect

pytestmark = pytest.mark.gpu


def test_train_pointnet_driver_two_profiles(lib, tmp_path):
    from pointcloudprocessing_b200 import pointnet_train
    from pointcloudprocessing_b200.pointnet import PointNet
    rng = np.random.default_rng(0)
    inp = tmp_path / "input"
    inp.mkdir()
    d, _ = write_collect(inp, rng, 40, pts=(60, 160))
    (tmp_path / "models").mkdir(); (tmp_path / "data").mkdir()
    prof = lambda tr, lw, mon: {"datasets": {"0": d.name}, "noise": {"x_stdev_m": 0.05, "y_stdev_m": 0.05, "z_stdev_m": 0.05},
                                "trainable": tr, "loss_weights": lw, "monitor": mon}
    cfg = {"info": {"name": "unit_lidar", "class_labels": {str(i): c for i, c in enumerate(CLASSES)},
                    "part_labels": {str(i): c for i, c in enumerate(PARTS)},
                    "training_profiles": {
                        "classification_pretrain": prof({"shared_network": True, "input_transform": True, "classification_head": True, "segmentation_head": False},
                                                        {"classification": 1.0, "segmentation": 0.0, "rotation": 0.0}, "val_classification_output_loss"),
                        "final": prof({"shared_network": False, "input_transform": False, "classification_head": False, "segmentation_head": True},
                                      {"classification": 0.0, "segmentation": 1.0, "rotation": 0.0}, "val_segmentation_output_loss")},
                    "continue_training_model": ""},
           "params": {"input_width": 128, "epochs": 2, "patience": 5, "batch_size": 4, "learning": {"rate": 1e-3, "decay_steps": 7000, "decay_rate": 0.7},
                      "random_seed": 42, "debugging": False, "regularize_input_transform": False, "regularize_feature_transform": False},   # no 'vanilla' key (q3)
           "file_system": {"model_path": str(tmp_path / "models") + "/", "input_path": str(inp) + "/", "data_path": str(tmp_path / "data") + "/"}}
    cf = tmp_path / "unit_lidar_config.json"
    cf.write_text(json.dumps(cfg))
    assert pointnet_train.train_pointnet(["prog", str(cf)])
    base = tmp_path / "models" / "unit_lidar"
    keys = ["loss", "classification_output_loss", "classification_output_sparse_categorical_accuracy", "segmentation_output_loss",
            "segmentation_output_sparse_categorical_accuracy", "se3_loss", "se3_root_mean_squared_error"]
    for p in ("classification_pretrain", "final"):
        h = json.loads((base / p / f"unit_lidar_{p}_history.json").read_text())
        assert sorted(h) == sorted(keys + ["val_" + k for k in keys]) and all(len(v) == 2 and np.isfinite(v).all() for v in h.values())
        assert (base / p / f"unit_lidar_{p}.npz").is_file() and (base / p / "unit_lidar_config.json").is_file()
        assert (tmp_path / "data" / f"unit_lidar_{p}" / "pc_set.joblib").is_file()
    assert any(f.startswith("log_") for f in os.listdir(base))
    # stage chaining: 'final' froze everything but the segmentation head, so its trunk equals the pretrain checkpoint's
    a = np.load(base / "classification_pretrain" / "unit_lidar_classification_pretrain.npz")
    b = np.load(base / "final" / "unit_lidar_final.npz")
    assert np.array_equal(a["s2_l3_1024_convolution_layer/kernel"], b["s2_l3_1024_convolution_layer/kernel"])
    assert not np.array_equal(a["seg_l1_512_convolution_layer/kernel"], b["seg_l1_512_convolution_layer/kernel"])
    m = PointNet.load(str(base / "final" / "unit_lidar_final.npz"))
    mk = PointNet.load(str(base / "final" / "unit_lidar_final.keras"))          # the reference's artefact format, same weights
    assert all(np.array_equal(v, mk.get_weights_dict()[k]) for k, v in m.get_weights_dict().items())
    cls, seg, R = m(np.zeros((2, 128, 3), np.float32) + rng.standard_normal((2, 128, 3)).astype(np.float32))
    assert cls.shape == (2, 3) and seg.shape == (2, 128, 4) and np.allclose(seg.sum(-1), 1, atol=1e-4)
    # the ONNX artefact (pointnet_train.py:237-247) holds the same network: executed by the independent reader it gives the model's outputs
    from oracle import onnx_eval
    xs = rng.standard_normal((2, 128, 3)).astype(np.float32)
    got = onnx_eval.run(onnx_eval.load((base / "final" / "unit_lidar_final.onnx").read_bytes()), {"pointnet_input": xs})
    cls, seg, R = m(xs)
    for a_, b_ in ((got["classification_output"], cls), (got["segmentation_output"], seg), (got["se3"], R)):
        assert np.abs(a_ - np.asarray(b_)).max() <= 2e-2 * max(1.0, np.abs(a_).max())
    # a second run reuses the existing PointCloudSets (pointnet_train.py:146-150)
    tp = pointnet_train.TrainProfile(str(cf))
    assert tp._training_profiles["final"]["pc"]._data_size["train"]["count"] == 30


def test_keras_archive_save_load_identical_outputs(lib, tmp_path):
    """PointNet.save('*.keras') / PointNet.load('*.keras') (pointnet_train.py:207-214, :280-295): the archive round trip
    reproduces every weight and therefore the outputs, bit for bit."""
    import torch
    from oracle import pointnet_numpy as onp
    from pointcloudprocessing_b200.pointnet import PointNet
    from tests import synth
    for vanilla in (False, True):
        m = PointNet(23, 12, 0.3, 42, vanilla=vanilla, precision="bf16")
        m.set_weights_dict(onp.make_params(23, 12, seed=7, vanilla=vanilla))
        path = str(tmp_path / f"model_{int(vanilla)}.keras")
        m.save(path, input_width=256)
        m2 = PointNet.load(path, precision="bf16")
        assert m2._vanilla == vanilla
        w1, w2 = m.get_weights_dict(), m2.get_weights_dict()
        assert set(w1) == set(w2) and all(np.array_equal(w1[k], w2[k]) for k in w1)
        x = torch.from_numpy(synth.gaussian_clouds(3, 4, 256)).cuda()
        for a, b in zip(m(x), m2(x)):
            assert torch.equal(a, b)
