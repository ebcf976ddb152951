"""The handful of Keras objects the reference's training driver passes to `compile` / `fit`
(point_cloud_analysis/pointnet_train.py:197-228, :310-351), as plain Python descriptors.

They carry hyper-parameters only: the arithmetic (losses, metrics, Adam, learning-rate schedule)
runs in libpcp_b200.so (`pcp_model_loss_backward`, `pcp_model_adam_step`, `pcp_model_evaluate`).
"""
from __future__ import annotations

import copy

import numpy as np


class ExponentialDecay:
    """keras.optimizers.schedules.ExponentialDecay (pointnet_train.py:310-315); staircase=False only."""

    def __init__(self, initial_learning_rate, decay_steps, decay_rate, staircase=False, name=None):
        if staircase:
            raise ValueError("the reference uses staircase=False; only that schedule runs on the device")
        self.initial_learning_rate, self.decay_steps, self.decay_rate = float(initial_learning_rate), int(decay_steps), float(decay_rate)

    def __call__(self, step):
        return self.initial_learning_rate * self.decay_rate ** (step / self.decay_steps)


class Adam:
    """keras.optimizers.Adam defaults (pointnet_train.py:317-319): beta_1 0.9, beta_2 0.999, epsilon 1e-7."""

    def __init__(self, learning_rate=1e-3, beta_1=0.9, beta_2=0.999, epsilon=1e-7, **_):
        self.learning_rate, self.beta_1, self.beta_2, self.epsilon = learning_rate, beta_1, beta_2, epsilon
        self.iterations = 0

    def schedule(self):
        lr = self.learning_rate
        if isinstance(lr, ExponentialDecay):
            return lr.initial_learning_rate, lr.decay_steps, lr.decay_rate
        return float(lr), 1, 1.0          # constant: rate^(step/1) with rate 1


class SparseCategoricalCrossentropy:
    """from_logits=False: clip probabilities to [1e-7, 1-1e-7], -log, mean (pointnet_train.py:337-338)."""

    def __init__(self, from_logits=False, **_):
        if from_logits:
            raise ValueError("the model outputs probabilities (PointNet.py:134,141); from_logits=True is not the reference contract")


class MeanSquaredError:
    pass


class SparseCategoricalAccuracy:
    name = "sparse_categorical_accuracy"


class RootMeanSquaredError:
    name = "root_mean_squared_error"


class Callback:
    def __init__(self):
        self.model = None

    def set_model(self, model):
        self.model = model

    def on_train_begin(self, logs=None):
        pass

    def on_epoch_end(self, epoch, logs=None):
        pass

    def on_train_end(self, logs=None):
        pass


class History(Callback):
    def __init__(self):
        super().__init__()
        self.history = {}
        self.epoch = []

    def on_epoch_end(self, epoch, logs=None):
        self.epoch.append(epoch)
        for k, v in (logs or {}).items():
            self.history.setdefault(k, []).append(float(v))


class EarlyStopping(Callback):
    """monitor / mode='min' / patience / restore_best_weights (pointnet_train.py:199-205)."""

    def __init__(self, monitor="val_loss", mode="min", patience=0, verbose=0, restore_best_weights=False, min_delta=0.0, **_):
        super().__init__()
        self.monitor, self.patience, self.verbose, self.restore_best_weights, self.min_delta = monitor, patience, verbose, restore_best_weights, min_delta
        self.sign = 1.0 if mode in ("min", "auto") else -1.0
        self.best, self.wait, self.best_weights, self.stopped_epoch = np.inf, 0, None, 0

    def on_train_begin(self, logs=None):
        self.best, self.wait, self.best_weights = np.inf, 0, None

    def on_epoch_end(self, epoch, logs=None):
        cur = (logs or {}).get(self.monitor)
        if cur is None:
            return
        if self.sign * cur < self.best - self.min_delta:
            self.best, self.wait = self.sign * cur, 0
            if self.restore_best_weights:
                self.best_weights = copy.deepcopy(self.model.get_weights_dict())
            return
        self.wait += 1
        if self.wait >= self.patience and epoch > 0:
            self.stopped_epoch = epoch
            self.model.stop_training = True

    def on_train_end(self, logs=None):
        # keras 3: the best weights are restored when training ends, early or not
        if self.restore_best_weights and self.best_weights is not None:
            if self.verbose:
                print("Restoring model weights from the end of the best epoch.")
            self.model.set_weights_dict(self.best_weights)


class ModelCheckpoint(Callback):
    """save_best_only on `monitor` (mode min), once per epoch (pointnet_train.py:207-214).  The container is this
    package's .npz (the reference's .keras zip needs h5py; same tensor names inside)."""

    def __init__(self, filepath, save_best_only=False, monitor="val_loss", mode="min", save_freq="epoch", verbose=0, **_):
        super().__init__()
        self.filepath, self.save_best_only, self.monitor, self.verbose = str(filepath), save_best_only, monitor, verbose
        self.sign = 1.0 if mode in ("min", "auto") else -1.0
        self.best = np.inf

    def on_epoch_end(self, epoch, logs=None):
        cur = (logs or {}).get(self.monitor)
        if self.save_best_only:
            if cur is None or not (self.sign * cur < self.best):
                return
            self.best = self.sign * cur
        if self.verbose:
            print(f"Epoch {epoch + 1}: saving model to {self.filepath}")
        self.model.save(self.filepath)
