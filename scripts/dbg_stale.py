import sys, gc, ctypes, numpy as np, torch
sys.path.insert(0, ".")
from tests.test_gpu_train import build, make_batch, ALL
from pointcloudprocessing_b200.pointnet import keras_compat as K
rt = ctypes.CDLL("libcudart.so.12")
rt.cudaGetLastError.restype = ctypes.c_int
def probe(tag):
    e = rt.cudaGetLastError()
    print(tag, "last error =", e, flush=True)
torch.zeros(1).cuda()
probe("start")
C, P, b, n = 4, 3, 16, 128
m, _ = build(C, P, False, ALL, (1.0, 1.0, 0.0), lr=2e-3)
probe("build")
x, y, _ = make_batch(9, b, n, C, P)
def gen():
    while True:
        yield x, y
m.train_on_batch(x, y); probe("train_on_batch")
m.test_on_batch(x, y); probe("test_on_batch")
w = m.get_weights_dict(); probe("get_weights")
m.set_weights_dict(w); probe("set_weights")
hist = m.fit(x=gen(), validation_data=gen(), epochs=2, verbose=0, steps_per_epoch=3, validation_steps=1,
             callbacks=[K.EarlyStopping(monitor="val_loss", patience=10, restore_best_weights=True)])
probe("fit")
del m, hist; gc.collect(); probe("destroy")
