"""The implicitly handled max-pooled layer of the bf16 training step (csrc/train_pool.inl) against a float64 restatement of
conv 128 -> cout, BatchNormalization (batch statistics), ReLU, reduce_max over the points and its gradient
(pointnet/PointNet.py:425-429, :242-248) on the same bf16-rounded operands.  The device never forms the (rows x cout) tensor."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _bf16(a):
    return torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(torch.bfloat16).to(torch.float64).numpy()


def _reference(a, w, gamma, beta, dg, b, n, batch):
    ar, wr = _bf16(a), _bf16(w)
    rows, cout = b * n, w.shape[1]
    z = ar @ wr
    mean, var = (z.mean(0), z.var(0)) if batch else (np.zeros(cout), np.ones(cout))
    inv = 1.0 / np.sqrt(var + 1e-3)
    sc = gamma * inv
    zc = z.reshape(b, n, cout)
    arg = np.where(sc >= 0, zc.argmax(1), zc.argmin(1))                       # first extremum, like the device
    zsel = np.take_along_axis(zc, arg[:, None, :], 1)[:, 0, :]
    g = np.maximum((zsel - mean) * sc + beta, 0.0)
    dy = np.where(g > 0, dg, 0.0)
    dY = np.zeros((b, n, cout))
    np.put_along_axis(dY, arg[:, None, :], dy[:, None, :], 1)
    dY = dY.reshape(rows, cout)
    if batch:
        xhat = (z - mean) * inv
        s1, s2 = dY.sum(0), (dY * xhat).sum(0)
        dZ = sc * (dY - s1 / rows - xhat * s2 / rows)
    else:
        s1 = s2 = np.zeros(cout)
        dZ = sc * dY
    return g, mean, inv, ar.T @ dZ, s2, s1, dZ @ wr.T


@pytest.mark.parametrize("b,n,cout,batch,neg", [(4, 256, 1024, 1, False), (3, 200, 256, 1, True), (5, 64, 128, 1, False), (2, 1024, 1024, 1, True),
                                               (4, 256, 512, 0, True)])
def test_pooled_layer_implicit(lib, b, n, cout, batch, neg):
    rng = np.random.default_rng(b * 1000 + n + cout)
    rows = b * n
    a = np.maximum(rng.standard_normal((rows, 128)), 0).astype(np.float32)
    a[n // 2] = a[3]                                            # a duplicated point: tie between rows 3 and n/2 of cloud 0
    w = (0.1 * rng.standard_normal((128, cout))).astype(np.float32)
    gamma = (1.0 + 0.2 * rng.standard_normal(cout)).astype(np.float32)
    if neg:
        gamma[::7] *= -1.0
    beta = (0.3 * rng.standard_normal(cout)).astype(np.float32)
    dg = rng.standard_normal((b, cout)).astype(np.float32)
    out = {k: np.zeros(s, np.float32) for k, s in dict(g=(b, cout), mean=(cout,), inv=(cout,), dw=(128, cout), dgamma=(cout,), dbeta=(cout,),
                                                        da=(rows, 128)).items()}
    fp = lambda x: x.ctypes.data_as(C.POINTER(C.c_float))
    st = lib.pcp_selftest_pooled_layer(fp(a), fp(w), fp(gamma), fp(beta), fp(dg), b, n, cout, batch, fp(out["g"]), fp(out["mean"]), fp(out["inv"]),
                                       fp(out["dw"]), fp(out["dgamma"]), fp(out["dbeta"]), fp(out["da"]))
    assert st == 0, lib.pcp_last_error()
    g, mean, inv, dW, dgamma, dbeta, dA = _reference(a, w, gamma.astype(np.float64), beta.astype(np.float64), dg.astype(np.float64), b, n, batch)

    def err(x, ref):
        return float(np.abs(x - ref).max() / max(np.abs(ref).max(), 1e-30))
    assert err(out["mean"], mean) <= 1e-5 + 1e-5 * (not batch)
    assert err(out["inv"], inv) <= 1e-4
    assert err(out["g"], g) <= 1e-4
    if batch:
        assert err(out["dgamma"], dgamma) <= 1e-3
        assert err(out["dbeta"], dbeta) <= 1e-4
    assert err(out["dw"], dW) <= 1e-3
    # dA: the BN-correction term A . (W diag(c1) W^T) runs on tensor cores with the 128 x 128 matrix rounded to bf16
    assert err(out["da"], dA) <= 6e-3


def test_pooled_layer_data_parallel_protocol(lib):
    """sync-BN inside the implicit pooled layer, emulated on one device: two shards A, B of a batch are run separately, the
    all-reduce callback adds the OTHER shard's recorded sums (Gram matrix / column sums in the forward pass, BN backward sums
    in the backward pass).  Shard outputs must equal the corresponding rows of the single run on the concatenated batch and
    the weight gradients must add up to it."""
    from pointcloudprocessing_b200 import _ffi
    b, n, cout = 3, 256, 512                                   # per shard
    rng = np.random.default_rng(99)
    rows = b * n
    A = [np.maximum(rng.standard_normal((rows, 128)), 0).astype(np.float32) for _ in range(2)]
    w = (0.1 * rng.standard_normal((128, cout))).astype(np.float32)
    gamma = (1.0 + 0.2 * rng.standard_normal(cout)).astype(np.float32)
    gamma[::5] *= -1.0
    beta = (0.3 * rng.standard_normal(cout)).astype(np.float32)
    dg = [rng.standard_normal((b, cout)).astype(np.float32) for _ in range(2)]
    fp = lambda x: x.ctypes.data_as(C.POINTER(C.c_float))

    def run(a, d, bb):
        out = {k: np.zeros(s, np.float32) for k, s in dict(g=(bb, cout), mean=(cout,), inv=(cout,), dw=(128, cout), dgamma=(cout,), dbeta=(cout,),
                                                            da=(bb * n, 128)).items()}
        st = lib.pcp_selftest_pooled_layer(fp(a), fp(w), fp(gamma), fp(beta), fp(d), bb, n, cout, 1, fp(out["g"]), fp(out["mean"]), fp(out["inv"]),
                                           fp(out["dw"]), fp(out["dgamma"]), fp(out["dbeta"]), fp(out["da"]))
        assert st == 0, lib.pcp_last_error()
        return out

    rec = {}                                                   # per shard: the LOCAL buffers handed to each callback call
    state = {"shard": 0, "call": 0, "add_from": None}

    def cb(user, ptr, count, dtype, stream):
        class _Dev:
            __cuda_array_interface__ = {"shape": (int(count),), "typestr": "<f8" if dtype == 1 else "<f4", "data": (int(ptr), False), "version": 2}
        t = torch.as_tensor(_Dev(), device="cuda")
        torch.cuda.synchronize()
        rec.setdefault(state["shard"], {})[state["call"]] = t.clone()
        other = state["add_from"]
        if other is not None and state["call"] in rec.get(other, {}):
            t += rec[other][state["call"]]
        torch.cuda.synchronize()
        state["call"] += 1
        return 0
    thunk = _ffi.ALLREDUCE_FN(cb)
    try:
        lib.pcp_selftest_set_bn_allreduce(C.cast(thunk, C.c_void_p), None, 2.0)
        # pass 1: forward sums of both shards do not depend on the other shard — record them (backward sums recorded here are stale)
        for s in (0, 1):
            state.update(shard=s, call=0, add_from=None)
            run(A[s], dg[s], b)
        fwd = {s: {c: rec[s][c] for c in (0, 1)} for s in (0, 1)}
        # pass 2: shard B under the global statistics -> its true backward sums; then shard A complete; then shard B complete
        rec.clear(); rec.update({0: dict(fwd[0]), 1: dict(fwd[1])})
        state.update(shard=1, call=0, add_from=0); run(A[1], dg[1], b)
        keep_b = dict(rec[1])
        state.update(shard=0, call=0, add_from=1); out_a = run(A[0], dg[0], b)
        rec[1] = keep_b
        rec[0] = {c: v for c, v in rec[0].items()}
        state.update(shard=1, call=0, add_from=0); out_b = run(A[1], dg[1], b)
    finally:
        lib.pcp_selftest_set_bn_allreduce(None, None, 1.0)
    ref = run(np.concatenate(A), np.concatenate(dg), 2 * b)

    def err(x, r):
        return float(np.abs(x - r).max() / max(np.abs(r).max(), 1e-30))
    assert err(out_a["mean"], ref["mean"]) <= 1e-5 and err(out_a["inv"], ref["inv"]) <= 1e-5
    assert err(np.concatenate([out_a["g"], out_b["g"]]), ref["g"]) <= 1e-5
    assert err(np.concatenate([out_a["da"], out_b["da"]]), ref["da"]) <= 1e-3
    assert err(out_a["dw"] + out_b["dw"], ref["dw"]) <= 1e-3
    assert err(out_a["dgamma"] + out_b["dgamma"], ref["dgamma"]) <= 1e-4
    assert err(out_a["dbeta"] + out_b["dbeta"], ref["dbeta"]) <= 1e-4
