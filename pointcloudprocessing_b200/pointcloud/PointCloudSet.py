"""Drop-in mirror of the reference's `pointcloud/PointCloudSet.py` (cited below as PointCloudSet.py:line, under
/root/reference/point_cloud_analysis/) without TensorFlow.

Same constructor, `add_from_aftr_output`, `add_data`, `get_train_set / get_val_set / get_test_set`, `get_info`,
`_adjust_to_input_width`, `_parse_state_info`, `_data_size[...]['count']` and the same on-disk artefacts: TFRecord files
`{data_path}{name}/{set}/{train|val|test}_{k}.tfrecord` of tf.train.Example records and `pc_set.joblib`.

What differs, on purpose:
  * frame text parsing, CRC32C, TFRecord scan and Example decode are native (libpcp_b200.so, csrc/host_io.cu);
  * the dataset iterators are plain Python generators (shuffle buffer 2048 -> repeat -> batch, like the tf.data graph of
    PointCloudSet.py:347-391); with `device='cuda'` the batch is copied H2D and the Gaussian jitter of `_parse_function`
    (PointCloudSet.py:335-337) runs on the GPU (`pcp_jitter_f32`, Philox); on the host the points are returned
    unjittered if the jitter is zero and it is an error otherwise (there is no CPU compute path);
  * q5: `rand_seed` really seeds the shuffles and pad draws (the reference creates and discards a Generator);
  * q10: `load_from_file` loads what `save` wrote (the reference appends '.pkl' to a '.joblib' name).
"""
from __future__ import annotations

import ctypes as C
import glob
import os
import sys
from collections.abc import Callable
from copy import deepcopy

import joblib
import numpy as np

from .. import _ffi
from . import tfrecord

SE3_ROWS, SE3_COLS, SE3_SIZE = 4, 4, 16           # utils/global_constants.py:1-4


class PointCloudSet:
    def __init__(self,
                 name: str,
                 class_labels: list,
                 part_labels: list,
                 network_input_width: int,
                 jitter_stdev_m: np.ndarray = np.array([0, 0, 0]),
                 val: float = 0.15,
                 test: float = 0.10,
                 batch_size: int = 32,
                 rand_seed=None,
                 description: str = '',
                 print_func: Callable[[str], None] = print,
                 data_path: str = ''):
        self._description = description
        self._batch_size = batch_size
        self._name = name
        self._class_labels = {label: i for i, label in enumerate(class_labels)}
        self._class_str = {i: label for i, label in enumerate(class_labels)}
        self._part_labels = {label: i for i, label in enumerate(part_labels)}
        self._part_str = {i: label for i, label in enumerate(part_labels)}
        self._network_input_width = int(network_input_width)
        self._jitter_stdev_m = np.asarray(jitter_stdev_m, np.float32)
        self._print = print_func
        self._data_path = data_path
        self._sets_added = 0
        self._data_size = {k: {'count': 0, 'class_count': {}, 'part_count': {}} for k in ('train', 'val', 'test')}   # PointCloudSet.py:66-82
        self._random_seed = rand_seed if type(rand_seed) == int else None
        self._rng = np.random.default_rng(self._random_seed)                                # q5
        if val < 1.0 and test < 1.0 and 1.0 - (val + test) < 1.0:                           # PointCloudSet.py:90-98
            self._train_amt, self._val_amt, self._test_amt = 1.0 - (val + test), val, test
        else:
            self._train_amt, self._val_amt, self._test_amt = 0.75, 0.15, 0.10
            self._print('PointCloudSet:  train_val_test_split incorrect format - set to default 75% / 15% / 10%')
        # schema of one record (PointCloudSet.py:100-105)
        self._feature_description = {'observations': (self._network_input_width * 3, 'float32'), 'class_label': (1, 'int64'),
                                     'part_labels': (self._network_input_width, 'int64'), 'se3': (9, 'float32')}
        os.makedirs(f"{self._data_path}{self._name}", exist_ok=True)
        self.save()

    # ------------------------------------------------------------------ persistence (PointCloudSet.py:111-114)
    def __getstate__(self):
        d = dict(self.__dict__)
        if not _picklable(d.get('_print')):
            d['_print'] = print
        return d

    def save(self):
        with open(f"{self._data_path}{self._name}/pc_set.joblib", "wb") as jl:
            joblib.dump(self, jl)

    # ------------------------------------------------------------------ ingest (PointCloudSet.py:116-218)
    def _parse_frame_file(self, path: str):
        """One frame file -> (obs (N,3) float64, class index, part indices (N,), skipped non-finite points); raises like
        the reference's per-frame try block does."""
        with open(path, 'rb') as f:
            text = f.read()
        cap = max(1, text.count(b'\n') + 1)
        xyz = np.empty((cap, 3), np.float64)
        part = np.empty((cap,), np.int32)
        cnt, bad, cls = C.c_size_t(), C.c_size_t(), C.c_int32()
        lib = _ffi.load()
        st = lib.pcp_parse_aftr_frame(text, len(text), "\n".join(self._class_labels).encode(), "\n".join(self._part_labels).encode(),
                                      _ffi.ptr(xyz), _ffi.ptr(part), cap, C.byref(cnt), C.byref(cls), C.byref(bad))
        if st != 0:
            raise Exception(lib.pcp_last_error().decode(errors="replace"))
        return xyz[:cnt.value], int(cls.value), part[:cnt.value].astype(np.int64), int(bad.value)

    def add_from_aftr_output(self, dir_path: str, shuffle_points: bool = True) -> bool:
        observations, class_labels, part_labels, se3 = [], [], [], []
        frames_searched = 0
        collect_contents = get_dir_contents(dir_path, self._print)
        lidar_contents = get_dir_contents(f'{dir_path}/Lidar', self._print)
        pose_log = [i for i in collect_contents if '_palindrome_state' in i]
        if len(pose_log) == 1:
            state_info = self._parse_state_info(f'{dir_path}/{pose_log[0]}')
        else:
            raise Exception(f"No state info found in {dir_path}")
        self._print(f"Parsing frames in {dir_path}...")
        # frames are parsed natively on a small thread pool (ctypes releases the GIL), 64 at a time, and consumed in frame
        # order, so the accepted frames, the messages and their order are those of the reference's sequential loop
        from concurrent.futures import ThreadPoolExecutor

        def _try_parse(i):
            try:
                return self._parse_frame_file(f'{dir_path}/Lidar/frame_{i}.txt')
            except Exception as e:                                                       # re-raised in order below
                return e
        n_frames = len(lidar_contents)
        pool = ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1))
        parsed = {}
        for i in range(n_frames):
            if i % 64 == 0:
                parsed = dict(zip(range(i, min(i + 64, n_frames)), pool.map(_try_parse, range(i, min(i + 64, n_frames)))))
            try:
                res = parsed.pop(i)
                if isinstance(res, Exception):
                    raise res
                obs, cl, pl, _bad = res
                if len(obs) != 0:
                    se = state_info[i]['tanker_in_sensor_frame'][:3, :3]
                    obs, pl = self._adjust_to_input_width(obs, pl)
                    if np.isfinite(obs).all():
                        observations.append(obs); class_labels.append(cl); part_labels.append(pl); se3.append(se)
                    else:
                        self._print(f'Per-line check failed - frame_{i} discarded after detecting non-finite value.')
            except Exception as e:
                if frames_searched == 0:
                    frames_searched = i
                self._print(f"Failed to add file {dir_path}/Lidar/frame_{i}.txt:\n\t{type(e).__name__} : {e}")
        pool.shutdown()
        self.add_data(dir_path.rstrip("/").split("/")[-1], np.array(observations), np.array(class_labels), np.array(part_labels), np.array(se3),
                      shuffle_points)
        return True

    def add_data(self, set_name: str, observations: np.ndarray, class_labels: np.ndarray, part_labels: np.ndarray, se3: np.ndarray,
                 shuffle_points: bool = True) -> None:
        """PointCloudSet.py:220-292: shuffle (aligned), split test = ceil(N*test), val = ceil(N*val), rest train, write
        the three TFRecord files of this set, account class / part occurrences, save."""
        N = observations.shape[0]
        if shuffle_points:
            indices = np.arange(N)
            self._rng.shuffle(indices)
            observations, class_labels, part_labels, se3 = observations[indices], class_labels[indices], part_labels[indices], se3[indices]
        n_test, n_val = int(np.ceil(N * self._test_amt)), int(np.ceil(N * self._val_amt))
        splits = {'test': (0, n_test), 'val': (n_test, n_test + n_val), 'train': (n_test + n_val, N)}
        os.makedirs(f"{self._data_path}{self._name}/{set_name}", exist_ok=True)
        for split in ('test', 'val', 'train'):
            lo, hi = splits[split]
            size = self._data_size[split]
            with tfrecord.TFRecordWriter(f"{self._data_path}{self._name}/{set_name}/{split}_{self._sets_added}.tfrecord") as writer:
                for i in range(lo, min(hi, N)):
                    cname = self._class_str[int(class_labels[i])]
                    size['class_count'][cname] = size['class_count'].get(cname, 0) + 1
                    counts = np.bincount(np.asarray(part_labels[i], np.int64), minlength=len(self._part_labels))
                    for lbl, idx in self._part_labels.items():
                        size['part_count'][lbl] = size['part_count'].get(lbl, 0) + int(counts[idx])
                    writer.write(self._serialize_sample(observations[i], class_labels[i], part_labels[i], se3[i]))
                    size['count'] += 1
        self._sets_added += 1
        self.save()

    def _serialize_sample(self, obs: np.ndarray, cls, seg: np.ndarray, se3: np.ndarray) -> bytes:
        return tfrecord.serialize_sample(obs, cls, seg, se3)                                # PointCloudSet.py:306-323

    # ------------------------------------------------------------------ dataset iterators (PointCloudSet.py:325-391)
    def _records(self, split: str):
        files = sorted(glob.glob(f"{self._data_path}{self._name}/*/{split}_*.tfrecord"))
        return [tfrecord.TFRecordFile(f) for f in files]

    def _record_stream(self, split: str, recs, rng, repeat: bool, shuffle_buffer: int):
        """(file index, record index) in the order the reference's pipeline hands records to batch() (PointCloudSet.py:341-391):
        list_files (file order reshuffled on every pass) -> train: interleave(cycle_length=2), one record from each of two
        open files in turn, the next file taking the place of an exhausted one; val / test: files read one after another
        -> shuffle(2048) -> repeat (the shuffle buffer drains at the end of a pass: every record exactly once per pass)."""
        def one_pass():
            files = [fi for fi in range(len(recs)) if len(recs[fi])]
            rng.shuffle(files)
            if split != 'train':
                for fi in files:
                    for ri in range(len(recs[fi])):
                        yield fi, ri
                return
            pending = list(files)
            slots = [[pending.pop(0), 0] for _ in range(min(2, len(pending)))]
            while slots:
                for sl in list(slots):
                    fi, ri = sl
                    yield fi, ri
                    sl[1] += 1
                    if sl[1] >= len(recs[fi]):
                        if pending:
                            sl[0], sl[1] = pending.pop(0), 0
                        else:
                            slots.remove(sl)
        while True:
            buf = []
            for item in one_pass():
                buf.append(item)
                if len(buf) >= shuffle_buffer:
                    j = int(rng.integers(len(buf)))
                    buf[j], buf[-1] = buf[-1], buf[j]
                    yield buf.pop()
            while buf:
                j = int(rng.integers(len(buf)))
                buf[j], buf[-1] = buf[-1], buf[j]
                yield buf.pop()
            if not repeat:
                return

    def _dataset(self, split: str, device=None, seed=None, jitter: bool = True, repeat: bool = True, shuffle_buffer: int = 2048,
                 prefetch: int = 3, decode_threads: int | None = None):
        """list_files -> [interleave] -> shuffle(2048) -> repeat -> map(parse + jitter, AUTOTUNE) -> batch -> prefetch(AUTOTUNE),
        as PointCloudSet.py:341-391.  Yields (x (b,n,3) float32, {'classification_output' (b,), 'segmentation_output' (b,n),
        'se3' (b,3,3)}).  Records are decoded a batch at a time by native host threads (pcp_example_decode_batch) out of the
        memory-mapped TFRecord files.  With a device, a producer thread decodes into a ring of `prefetch` PINNED host slots and
        issues the host-to-device copies on its own stream, so decoding and copying batch k + 1 overlap the training step on
        batch k; the Gaussian jitter of _parse_function (:335-337) runs on the device (pcp_jitter_f32)."""
        recs = self._records(split)
        if not any(len(r) for r in recs):
            return
        w, bs = self._network_input_width, self._batch_size
        rng = np.random.default_rng(self._random_seed if seed is None else seed)
        use_jitter = jitter and bool(np.any(self._jitter_stdev_m != 0))
        if use_jitter and device is None:
            raise RuntimeError("jitter runs on the GPU (pcp_jitter_f32): pass device='cuda' (there is no CPU compute path)")
        nthreads = decode_threads if decode_threads is not None else max(1, min(8, (os.cpu_count() or 2) // 2))
        lib = _ffi.load()
        stream_it = self._record_stream(split, recs, rng, repeat, shuffle_buffer)
        ptrs = np.empty(bs, np.uint64)
        lens = np.empty(bs, np.uint64)

        def fill(x, ycls, yseg, yse3) -> int:
            """next batch of records decoded into the given (bs, ...) host arrays; returns the number of records"""
            k = 0
            for fi, ri in stream_it:
                ptrs[k], lens[k] = recs[fi].record_ptr(ri)
                k += 1
                if k == bs:
                    break
            if k:
                _ffi.check(lib.pcp_example_decode_batch(_ffi.ptr(ptrs), _ffi.ptr(lens), k, w, _ffi.ptr(x), _ffi.ptr(ycls), _ffi.ptr(yseg), _ffi.ptr(yse3), nthreads))
            return k

        if device is None:
            while True:
                x = np.empty((bs, w, 3), np.float32)
                ycls = np.empty((bs,), np.int32)
                yseg = np.empty((bs, w), np.int32)
                yse3 = np.empty((bs, 3, 3), np.float32)
                k = fill(x, ycls, yseg, yse3)
                if k == 0:
                    return
                yield x[:k], {'classification_output': ycls[:k], 'segmentation_output': yseg[:k], 'se3': yse3[:k]}
                if k < bs:
                    return
        yield from self._device_batches(fill, device, use_jitter, max(2, int(prefetch)), split)

    def _device_batches(self, fill, device, use_jitter: bool, depth: int, split: str):
        """prefetch(AUTOTUNE) onto the GPU: see _dataset."""
        import queue
        import threading

        import torch
        from .. import ops
        w, bs = self._network_input_width, self._batch_size
        dev = torch.device(device)
        if dev.type == "cuda" and dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        copy_stream = torch.cuda.Stream(device=dev)

        class Slot:
            pass
        slots = []
        for _ in range(depth):
            sl = Slot()
            sl.h = (torch.empty((bs, w, 3), dtype=torch.float32).pin_memory(), torch.empty((bs,), dtype=torch.int32).pin_memory(),
                    torch.empty((bs, w), dtype=torch.int32).pin_memory(), torch.empty((bs, 3, 3), dtype=torch.float32).pin_memory())
            sl.d = tuple(torch.empty_like(t, device=dev) for t in sl.h)
            sl.copied = torch.cuda.Event()
            sl.released = None                     # event on the consumer's stream: the device buffers may be overwritten after it
            slots.append(sl)
        free_q, ready_q = queue.Queue(), queue.Queue(maxsize=depth)
        for sl in slots:
            free_q.put(sl)
        stop = threading.Event()

        def producer():
            try:
                torch.cuda.set_device(dev)
                while not stop.is_set():
                    sl = free_q.get()
                    if sl is None:
                        return
                    k = fill(*(t.numpy() for t in sl.h))
                    if k == 0:
                        ready_q.put((None, 0))
                        return
                    with torch.cuda.stream(copy_stream):
                        if sl.released is not None:
                            copy_stream.wait_event(sl.released)
                        for hd, dd in zip(sl.h, sl.d):
                            dd[:k].copy_(hd[:k], non_blocking=True)
                        sl.copied.record(copy_stream)
                    sl.copied.synchronize()        # the pinned slot is refilled next time round: its copy must have left the host
                    ready_q.put((sl, k))
                    if k < bs:
                        ready_q.put((None, 0))
                        return
            except BaseException as e:             # surfaced to the consumer
                ready_q.put((e, -1))

        th = threading.Thread(target=producer, name=f"pcs-prefetch-{split}", daemon=True)
        th.start()
        # jitter stream: keyed by the set's seed, the split and — for an unseeded set, where the reference draws fresh
        # tf.random.normal noise on every run — a nonce from the OS; the counter is the global point index of the iterator
        base = self._random_seed if isinstance(self._random_seed, int) else int.from_bytes(os.urandom(6), "little")
        seed_base = base + 1 + {"train": 0, "val": 1_000_003, "test": 2_000_003}.get(split, 0)
        step = 0
        prev = None
        try:
            while True:
                sl, k = ready_q.get()
                if prev is not None:               # the consumer has moved on: everything it enqueued on `prev` precedes this event
                    prev.released = torch.cuda.Event()
                    prev.released.record(torch.cuda.current_stream(dev))
                    free_q.put(prev)
                    prev = None
                if k < 0:
                    raise sl
                if sl is None:
                    return
                torch.cuda.current_stream(dev).wait_event(sl.copied)
                xd, ycls, yseg, yse3 = (t[:k] for t in sl.d)
                # the caller gets tensors of its own (like tf.data): the ring's device buffers are recycled behind its back
                xd = ops.jitter(xd, self._jitter_stdev_m, seed=seed_base, offset=step * bs * w) if use_jitter else xd.clone()
                labels = {'classification_output': ycls.clone(), 'segmentation_output': yseg.clone(), 'se3': yse3.clone()}
                prev = sl
                step += 1
                yield xd, labels
        finally:
            stop.set()
            free_q.put(None)

    def get_train_set(self, device=None, **kw):
        return self._dataset('train', device=device, **kw)

    def get_val_set(self, device=None, **kw):
        return self._dataset('val', device=device, **kw)        # q7: jitter applies to val / test too (PointCloudSet.py:373,387)

    def get_test_set(self, device=None, **kw):
        return self._dataset('test', device=device, **kw)

    # ------------------------------------------------------------------ info (PointCloudSet.py:393-441)
    def get_description(self):
        return self._description

    def get_info(self):
        out = f'{self._description}\n'
        out += f'Random seed: {self._random_seed}\n' if type(self._random_seed) == int else 'Is not seeded\n'
        out += f'Class labels: {self._class_labels.keys()}\n'
        out += f'Part labels: {self._part_labels.keys()}\n'
        total = sum(self._data_size[s]['count'] for s in ('train', 'val', 'test'))
        for title, key, amt in (('Train', 'train', self._train_amt), ('Validation', 'val', self._val_amt), ('Test', 'test', self._test_amt)):
            d = self._data_size[key]
            out += f'\n--- {title} Set ---\n'
            out += f'Specified proportion:  {amt}\n'
            out += f"Actual proportion: {d['count'] / total if total else float('nan')}\n"
            out += f"Total count: {d['count']}\n"
            out += 'Class count:\n'
            for label in self._class_labels:
                if label in d['class_count']:
                    out += f"\t{label}: {d['class_count'][label]}\n"
            out += 'Part count:\n'
            for label in self._part_labels:
                if label in d['part_count']:
                    out += f"\t{label}: {d['part_count'][label]}\n"
        return out

    # ------------------------------------------------------------------ resize (PointCloudSet.py:443-470)
    def _adjust_to_input_width(self, observations: np.ndarray, part_labels: np.ndarray) -> tuple:
        w = self._network_input_width
        if observations.shape[0] > w:
            return observations[:w], part_labels[:w]                                        # q6: truncate, no sampling
        repeated = self._rng.uniform(0, observations.shape[0], w - observations.shape[0]).astype(np.int_)
        observations = np.concatenate((observations, deepcopy(observations[repeated])), axis=0)
        part_labels = np.concatenate((part_labels, deepcopy(part_labels[repeated])), axis=0)
        assert observations.shape[0] == w and part_labels.shape[0] == w
        return observations, part_labels

    # ------------------------------------------------------------------ pose log (PointCloudSet.py:472-509)
    def _parse_state_info(self, filepath: str) -> dict:
        with open(filepath, 'r') as f:
            keys = [i for i in f.readline().strip().split("   ") if len(i) > 1]
            data = {}
            for line in f:
                d = line.strip().split(" ")
                if len(d) < 2:
                    continue
                fr = int(d[1])
                data[fr] = {keys[0]: d[0], keys[1]: d[1]}
                for i, key in enumerate(keys[2:]):
                    cols = [d[2 + i * SE3_SIZE + c * SE3_ROWS: 2 + i * SE3_SIZE + (c + 1) * SE3_ROWS] for c in range(SE3_COLS)]
                    data[fr][key] = np.array(cols, dtype=np.float64).T
                if 'Sensor Pose' in keys and 'Tanker Pose' in keys:
                    S, T = data[fr]['Sensor Pose'], data[fr]['Tanker Pose']
                    so3 = S[:3, :3].T @ T[:3, :3]
                    t_t_s = S[:3, :3].T @ (T[:3, 3:] - S[:3, 3:])
                    data[fr]['tanker_in_sensor_frame'] = np.concatenate([np.concatenate([so3, t_t_s], axis=1), np.array([[0, 0, 0, 1]])], axis=0)
        return data


def _picklable(obj) -> bool:
    import pickle
    try:
        pickle.dumps(obj)
        return True
    except Exception:
        return False


# ### FREE HELPER FUNCTIONS ### (PointCloudSet.py:511-539)
def load_from_file(joblib_file: str) -> PointCloudSet:
    """q10: loads what `save` writes (`.../pc_set.joblib`); a bare stem gets '.joblib' (then '.pkl') appended."""
    candidates = [joblib_file] if os.path.isfile(joblib_file) else [joblib_file + '.joblib', joblib_file + '.pkl']
    for c in candidates:
        if os.path.isfile(c):
            with open(c, 'rb') as pf:
                return joblib.load(pf)
    raise FileNotFoundError(joblib_file)


def get_dir_contents(dir_path: str, _print: Callable[[str], None] = print) -> list:
    try:
        return os.listdir(dir_path) or []
    except FileNotFoundError:
        _print(f"Error: The directory '{dir_path}' was not found.")
    except NotADirectoryError:
        _print(f"Error: The path '{dir_path}' is not a directory.")
    except PermissionError:
        _print(f"Error: Permission denied to read '{dir_path}'.")
    except Exception as e:
        _print(f"An error occurred: {e}")
    return []
