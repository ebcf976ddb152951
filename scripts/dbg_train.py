import sys, numpy as np, torch
sys.path.insert(0, '.')
from tests.test_gpu_train import *
C, P, b, n = 7, 5, 8, 64
trainable, weights, reg = PRETRAIN, (1.0, 0.0, 0.0), (False, False)
m, p = build(C, P, False, trainable, weights, reg, lr=1e-3, smooth=True)
tp = opt.to_torch(p, requires_grad=True)
keep = trainable_names(p, trainable, False)
am = {k: torch.zeros_like(tp[k]) for k in keep}
av = {k: torch.zeros_like(tp[k]) for k in keep}
for step in range(3):
    x, y, masks = make_batch(100 + step, b, n, C, P)
    met = m.train_on_batch(x, y, dropout_masks=masks)
    stats = {}
    out, loss, parts = oracle_step(x, y, masks, tp, False, trainable, weights, reg, stats)
    print(step, met["loss"], float(loss))
    grads = m.gradients_dict()
    gm = max(float(tp[k].grad.abs().max()) for k in keep)
    worst = sorted(((float(np.abs(grads[k].reshape(p[k].shape) - tp[k].grad.numpy()).max() / max(float(tp[k].grad.abs().max()), 1e-3*gm)), k) for k in keep), reverse=True)[:5]
    print("  grad err", worst)
    lr = onp.exp_decay_lr(1e-3, step, 7000, 0.7)
    opt.adam_step_(tp, {k: tp[k].grad for k in keep}, am, av, step + 1, lr)
    with torch.no_grad():
        for k, v in stats.items():
            tp[k].copy_(v)
    w = m.get_weights_dict()
    worst = sorted(((float(np.abs(w[k].reshape(p[k].shape) - tp[k].detach().numpy()).max()), k) for k in p), reverse=True)[:6]
    print("  weight abs diff", worst)
