"""bf16 (tensor-core) training step against the autograd oracle: per-tensor gradient errors."""
import sys, numpy as np, torch
sys.path.insert(0, ".")
from tests.test_gpu_train import build, make_batch, ALL, SEG_ONLY, oracle_step, trainable_names
from oracle import pointnet_torch as opt
C, P, b, n = 8, 8, 6, 128
for smooth in (True, False):
  for (van, tr, w, reg) in [(False, ALL, (1.0, 0.5, 0.25), (True, True)), (False, SEG_ONLY, (0.0, 1.0, 0.0), (False, False)), (True, ALL, (1.0, 1.0, 0.0), (False, False))]:
    m, p = build(C, P, van, tr, w, reg, precision="bf16", smooth=smooth)
    x, y, masks = make_batch(21, b, n, C, P)
    met = m.train_on_batch(x, y, dropout_masks=masks, apply=False)
    tp = opt.to_torch(p, requires_grad=True)
    with opt.bf16_gemms():
        out, loss, parts = oracle_step(x, y, masks, tp, van, tr, w, reg, {})
    print("smooth", smooth, "vanilla", van, "loss", met["loss"], float(loss))
    cls_d, seg_d = m.train_output(0, b * C).reshape(b, C), m.train_output(1, b * n * P).reshape(b, n, P)
    cls_r, seg_r = out["cls"].detach().numpy(), out["seg"].detach().numpy()
    print("  prob abs err", np.abs(cls_d - cls_r).max(), np.abs(seg_d - seg_r).max(), "argmax agree", (cls_d.argmax(-1) == cls_r.argmax(-1)).mean(), (seg_d.argmax(-1) == seg_r.argmax(-1)).mean())
    grads = m.gradients_dict()
    keep = trainable_names(p, tr, van)
    gmax = max(float(np.abs(tp[k].grad.numpy()).max()) for k in keep if tp[k].grad is not None)
    rows = []
    for k in keep:
        g_ref = tp[k].grad
        g_ref = np.zeros_like(p[k]) if g_ref is None else g_ref.numpy()
        d = grads[k].reshape(g_ref.shape).astype(np.float64) - g_ref
        rows.append((float(np.sqrt((d * d).sum()) / max(np.sqrt((g_ref.astype(np.float64) ** 2).sum()), 1e-2 * gmax)),
                     float(np.abs(d).max() / max(np.abs(g_ref).max(), 1e-2 * gmax)), k))
    rows.sort(reverse=True)
    for r in rows[:6]:
        print("   l2 %.4f max %.4f %s" % r)
