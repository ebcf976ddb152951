"""Drop-in mirror of the reference's training driver `pointnet_train.py` (cited as pointnet_train.py:line under
/root/reference/point_cloud_analysis/): same JSON config schema, same per-profile staging (PointCloudSet per profile,
freeze/thaw of the four groups, Adam + ExponentialDecay, three losses with weights, EarlyStopping + ModelCheckpoint +
two-stage Ctrl-C), same artefacts per profile directory ({name}_{prof}_history.json with the 14 keras keys, the config
copy, the best checkpoint) — driving the sm_100a kernels of libpcp_b200.so instead of TensorFlow.

    python -m pointcloudprocessing_b200.pointnet_train kc46_lidar_config.json

Positions on the reference's quirks: q3 `params.vanilla` defaults to false when absent; q11 a named logger
and a portable log-file timestamp; checkpoints are this package's `.npz` container (the reference's `.keras` zip needs
h5py, absent here); the ONNX export (:237-247) is written by pointnet/onnx_export.py without onnx / tf2onnx.  Under torchrun every rank trains on its own shard of each batch and rank 0 writes the artefacts.
"""
from __future__ import annotations

import datetime
import json
import logging
import os
import shutil
import signal
import sys
from typing import Callable

import joblib
import numpy as np

from .pointcloud.PointCloudSet import PointCloudSet, get_dir_contents
from .pointnet.PointNet import PointNet
from .pointnet import keras_compat as keras
from .pointnet.keras_compat import Callback, EarlyStopping, ModelCheckpoint


class CtrlC_InterruptHandler(Callback):
    """First SIGINT: stop after the current epoch; second: exit (pointnet_train.py:42-61)."""

    def __init__(self, print_func: Callable[[str], None] = print):
        super().__init__()
        self._stop_requested = False
        self._print = print_func

    def stop_signaled(self, sig, frame):
        if not self._stop_requested:
            self._stop_requested = True
            self._print(">>> TRAINING INTERRUPT INITIATED BY USER <<<\nTraining will stop after the current epoch.\nPress Ctrl+C again to force quit.")
        else:
            self._print(">>> FORCE QUIT INITIATED BY USER <<<")
            sys.exit(0)

    def on_epoch_end(self, epoch, logs=None):
        if self._stop_requested:
            self._print("User stop received by the training loop.")
            self.model.stop_training = True


class TrainProfile:
    def __init__(self, config_file, device: str = "cuda", precision: str = "fp32", sync_bn: bool = True):
        with open(config_file, 'r') as cf:
            config = json.load(cf)
        self._config_file = config_file
        self._device, self._precision = device, precision
        self._sync_bn = sync_bn          # data parallel: BatchNorm statistics over the global batch, like the reference's single device
        info, params, fs = config['info'], config['params'], config['file_system']
        self._name: str = info['name']                                                       # pointnet_train.py:83-87
        self._class_labels: list = list(info['class_labels'].values())
        self._part_labels: list = list(info['part_labels'].values())
        self._training_profiles: dict = info['training_profiles']
        self._pretrained_model: str = info.get('continue_training_model', "")
        self._input_width: int = params['input_width']                                       # :90-101
        self._epochs: int = params['epochs']
        self._patience: int = params['patience']
        self._batch_size: int = params['batch_size']
        self._learning_rate: float = params['learning']['rate']
        self._learning_decay_steps: int = params['learning']['decay_steps']
        self._learning_decay_rate: float = params['learning']['decay_rate']
        self._random_seed: int = params['random_seed']
        self._debugging: bool = params.get('debugging', False)
        self._vanilla: bool = params.get('vanilla', False)                                   # q3
        self._reg_input_transform: bool = params.get('regularize_input_transform', False)
        self._reg_feature_transform: bool = params.get('regularize_feature_transform', False)
        self._model_path: str = fs['model_path']                                             # :104-106
        self._input_path: str = fs['input_path']
        self._data_path: str = fs['data_path']
        self._training_callbacks = []

        for p, what in ((self._model_path, "model_path"), (self._input_path, "input_path"), (self._data_path, "data_path")):
            if not os.path.isdir(p):
                raise FileNotFoundError(f"Error in TrainProfile:  {what} {p} does not exist")    # the reference only warns (:115-122)
        for prof in self._training_profiles:
            for ds in self._training_profiles[prof]['datasets'].values():
                if not os.path.isdir(f"{self._input_path}{ds}"):
                    raise FileNotFoundError(f"Error in TrainProfile:  {self._input_path}{ds} does not exist")
        if self._pretrained_model != "" and not os.path.isfile(f"{self._model_path}{self._pretrained_model}"):
            raise FileNotFoundError(f"Error in TrainProfile:  {self._model_path}{self._pretrained_model} does not exist")

        self._specific_model_path = f"{self._name}/"
        os.makedirs(f"{self._model_path}{self._specific_model_path}", exist_ok=True)

        dt = datetime.datetime.now()                                                         # q11: named logger, portable file name
        self._log = logging.getLogger(f"pointnet_train.{self._name}")
        self._log.setLevel(logging.DEBUG)
        self._log.propagate = False
        if not self._log.handlers:
            ch = logging.StreamHandler()
            fh = logging.FileHandler(f"{self._model_path}{self._specific_model_path}log_{dt.strftime('%Y%m%d_%H%M%S')}.log")
            ch.setFormatter(logging.Formatter('%(name)s - %(levelname)s - %(message)s'))
            fh.setFormatter(logging.Formatter('%(asctime)s - %(name)s - %(levelname)s - %(message)s'))
            self._log.addHandler(ch)
            self._log.addHandler(fh)

        # Data-parallel runs (torchrun): the PointCloudSet directory, pc_set.joblib and the TFRecord files are written by
        # rank 0 alone; the other ranks wait at a barrier and then load the finished set (all ranks writing the same files at
        # once could hand a half-written pc_set.joblib or a truncated TFRecord to a slower rank).
        from . import parallel
        rank, world = parallel.world()
        for prof in self._training_profiles:                                                 # :144-172
            P = self._training_profiles[prof]
            set_dir = f"{self._data_path}{self._name}_{prof}"
            if rank == 0:
                if os.path.isdir(set_dir):
                    self._log.info(f"Training profile {self._name}_{prof} already exists. Using existing profile...")
                    with open(f"{set_dir}/pc_set.joblib", "rb") as jl:
                        P['pc'] = joblib.load(jl)
                    P['pc']._print = self._log.info
                else:
                    noise = P['noise']
                    P['pc'] = PointCloudSet(name=f"{self._name}_{prof}", class_labels=self._class_labels,
                                                                part_labels=self._part_labels, network_input_width=self._input_width,
                                                                jitter_stdev_m=np.array([noise['x_stdev_m'], noise['y_stdev_m'], noise['z_stdev_m']]),
                                                                batch_size=self._batch_size, rand_seed=42, description=prof,
                                                                print_func=self._log.info, data_path=self._data_path)
                self._profile_datasets(prof)
            parallel.barrier()
            if rank != 0:
                with open(f"{set_dir}/pc_set.joblib", "rb") as jl:
                    P['pc'] = joblib.load(jl)
                P['pc']._print = self._log.info
            P['path'] = f"{self._specific_model_path}{prof}/"
            os.makedirs(f"{self._model_path}{P['path']}", exist_ok=True)

    def _profile_datasets(self, profile) -> None:                                            # :259-274
        P = self._training_profiles[profile]
        datasets = get_dir_contents(f"{self._data_path}{self._name}_{profile}", self._log.info)
        wanted = list(P['datasets'].values())
        if len(datasets) > 0:
            self._log.info(f"The following datasets were found in {self._data_path}{self._name}_{profile}:")
            for ds in datasets:
                self._log.info(f"\t-> {ds}")
        for k, set_name in enumerate(wanted):
            if set_name not in datasets:
                self._log.info(f"Adding data set {k + 1} of {len(wanted)}")
                P['pc'].add_from_aftr_output(dir_path=f"{self._input_path}{set_name}", shuffle_points=True)
        self._log.info('\nDatasets added successfully:\n')
        self._log.info(P['pc'].get_info())

    def _build_pointnet(self, profile: str):                                                 # :276-353
        self._training_callbacks = []
        P = self._training_profiles[profile]
        if self._pretrained_model != "":
            self._log.info(f"Continuing training on model {self._pretrained_model}")
            model = PointNet.load(f"{self._model_path}{self._pretrained_model}", precision=self._precision)
        else:
            model = PointNet(classification_output_width=len(self._class_labels),
                                            segmentation_output_width=len(self._part_labels), dropout_rate=0.3,
                                            random_seed=self._random_seed, debugging=self._debugging, vanilla=self._vanilla,
                                            regularize_input_transform=self._reg_input_transform,
                                            regularize_feature_transform=self._reg_feature_transform, precision=self._precision)
            model.build(input_shape=(None, self._input_width, 3))
        lr_schedule = keras.ExponentialDecay(self._learning_rate, decay_steps=self._learning_decay_steps,
                                             decay_rate=self._learning_decay_rate, staircase=False)
        optimizer = keras.Adam(learning_rate=lr_schedule)
        tr = P['trainable']
        (model.thaw_shared_network if tr['shared_network'] else model.freeze_shared_network)()
        (model.thaw_input_transform if tr['input_transform'] else model.freeze_input_transform)()
        (model.thaw_classification_head if tr['classification_head'] else model.freeze_classification_head)()
        (model.thaw_segmentation_head if tr['segmentation_head'] else model.freeze_segmentation_head)()
        lw = P['loss_weights']
        model.compile(optimizer=optimizer,
                      loss={'classification_output': keras.SparseCategoricalCrossentropy(),
                            'segmentation_output': keras.SparseCategoricalCrossentropy(), 'se3': keras.MeanSquaredError()},
                      loss_weights={'classification_output': lw['classification'], 'segmentation_output': lw['segmentation'], 'se3': lw['rotation']},
                      sync_bn=self._sync_bn,
                      metrics={'classification_output': [keras.SparseCategoricalAccuracy()],
                               'segmentation_output': [keras.SparseCategoricalAccuracy()], 'se3': [keras.RootMeanSquaredError()]})
        return model

    def train(self):                                                                         # :174-257
        from . import parallel
        rank, world = parallel.world()
        for prof in list(self._training_profiles.keys()):
            P = self._training_profiles[prof]
            model = self._build_pointnet(prof)
            self._log.info("PointNet Build")
            self._log.info("\tTrainable Layers")
            for l, v in model.get_layer_trainability().items():
                self._log.info(f"\t\t-> {l}: {v}")
            keyboard_interrupt = CtrlC_InterruptHandler(print_func=self._log.info)
            ckpt = f"{self._model_path}{P['path']}{self._name}_{prof}.npz"
            self._training_callbacks.append(EarlyStopping(monitor=P['monitor'], mode='min', patience=self._patience, verbose=1,
                                                          restore_best_weights=True))
            if rank == 0:
                self._training_callbacks.append(ModelCheckpoint(filepath=ckpt, save_best_only=True, monitor=P['monitor'], mode='min',
                                                                save_freq='epoch', verbose=1))
            self._training_callbacks.append(keyboard_interrupt)
            try:
                signal.signal(signal.SIGINT, keyboard_interrupt.stop_signaled)
            except ValueError:
                pass                                                                         # not the main thread
            pc = P['pc']
            history = model.fit(x=pc.get_train_set(device=self._device), validation_data=pc.get_val_set(device=self._device),
                                epochs=self._epochs, verbose=1, callbacks=self._training_callbacks,
                                steps_per_epoch=int(pc._data_size['train']['count'] / self._batch_size),
                                validation_steps=int(pc._data_size['val']['count'] / self._batch_size))
            if rank == 0:
                with open(f"{self._model_path}{P['path']}{self._name}_{prof}_history.json", 'w') as j:
                    json.dump(history.history, j)
                if not os.path.isfile(ckpt):
                    model.save(ckpt)
                # the reference's artefact name and format (ModelCheckpoint(...*.keras), pointnet_train.py:207-214): same weights as the .npz
                PointNet.load(ckpt, precision=self._precision).save(ckpt[:-4] + ".keras", input_width=self._input_width)
                onnx_path = f"{self._model_path}{P['path']}{self._name}_{prof}.onnx"                # :237-247, opset 13, (None, n, 3) f32
                model.export_onnx(onnx_path, self._input_width)
                self._log.info(f"ONNX model written to {onnx_path}")
                shutil.copy(self._config_file, f"{self._model_path}{P['path']}")
            # only rank 0 wrote the checkpoint / ONNX file: nobody loads it (or enters the next profile's collectives) before it is complete
            parallel.barrier()
            self._pretrained_model = f"{P['path']}{self._name}_{prof}.npz"                   # the next profile continues from this one (:257)


def train_pointnet(*args, **kwargs) -> bool:                                                 # :364-402
    argv = list(args[0]) if args else []
    if "-h" in argv or "--help" in argv:
        print_help()
        return False
    configs = [i for i in argv if i.split('.')[-1] == 'json']
    if len(configs) == 0:
        print_help()
        return False
    import torch
    if not torch.cuda.is_available():
        print("No GPUs available: this build has no CPU compute path.")
        return False
    print('GPUs Available: ', torch.cuda.device_count())
    from . import parallel
    parallel.init_from_env()
    for cf in configs:
        TrainProfile(cf, **kwargs).train()
    return True


def print_help():
    print('''PointNet Training Module (B200 build)

    usage: python -m pointcloudprocessing_b200.pointnet_train <name>_config.json [...]

    The configuration file follows the reference's schema (kc46_lidar_config.json):
      info:        name, class_labels{idx: str}, part_labels{idx: str}, continue_training_model,
                   training_profiles{<profile>: datasets{idx: dir}, noise{x,y,z_stdev_m},
                                     trainable{shared_network,input_transform,classification_head,segmentation_head},
                                     loss_weights{classification,segmentation,rotation}, monitor}
      params:      input_width, epochs, patience, batch_size, learning{rate,decay_steps,decay_rate}, random_seed,
                   debugging, vanilla, regularize_input_transform, regularize_feature_transform
      file_system: model_path, input_path, data_path (existing directories, trailing slash)''')


if __name__ == '__main__':
    if not any(i.split('_')[-1] == 'config.json' for i in sys.argv):
        print("No *_config.json file given.")
        print_help()
        sys.exit(2)
    print("Model training completed successfully." if train_pointnet(sys.argv) else "Model training failed.")
