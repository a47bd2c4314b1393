"""GPU parity of the tcgen05 trunk (gz_net_tc.cu), layer by layer, against a plain PyTorch
reference of the same ops (policy_value_net_pytorch.py:43-57) evaluated in float64 and float32.

Tolerance: 1e-4 absolute on probabilities and value (BASELINE.json north_star); the per-layer
activations are held to 2e-5 relative to the layer's largest activation, which is what the
hi/lo fp16 split (22 mantissa bits, fp32 accumulation) has to deliver."""
import os

import numpy as np
import pytest

from helpers import GOLDEN, load_golden_weights

pytestmark = pytest.mark.gpu
NET_TOL = 1e-4


def _layers_ref(arrs, planes, dtype):
    import torch
    import torch.nn.functional as F
    p = [torch.from_numpy(np.ascontiguousarray(a)).to(dtype) for a in arrs]
    x = torch.from_numpy(planes).to(dtype)
    out = {}
    with torch.no_grad():
        x = F.relu(F.conv2d(x, p[0], p[1], padding=1)); out["act1"] = x
        x = F.relu(F.conv2d(x, p[2], p[3], padding=1)); out["act2"] = x
        x = F.relu(F.conv2d(x, p[4], p[5], padding=1)); out["act3"] = x
        a = F.relu(F.conv2d(x, p[6], p[7]))
        v = F.relu(F.conv2d(x, p[10], p[11]))
        out["head"] = torch.cat([a, v], dim=1).flatten(2)
        pr = F.softmax(F.linear(a.flatten(1), p[8], p[9]), dim=1)
        val = torch.tanh(F.linear(F.relu(F.linear(v.flatten(1), p[12], p[13])), p[14], p[15]))
    res = {k: t.double().numpy() for k, t in out.items()}
    for k in ("act1", "act2", "act3"):  # [B,C,H,W] -> [B,S,C]
        B, C = res[k].shape[:2]
        res[k] = res[k].reshape(B, C, -1).transpose(0, 2, 1)
    res["probs"] = pr.double().numpy()
    res["values"] = val.double().numpy().reshape(-1)
    return res


def _random_weights(rng, w, scale=1.0, trunk=(32, 64, 128)):
    from alphazero_gomoku_b200.engine import weight_shapes
    arrs = []
    for s in weight_shapes(w, w, trunk):
        fan = int(np.prod(s[1:])) if len(s) > 1 else 1
        arrs.append((rng.randn(*s) * (scale * np.sqrt(2.0 / fan) if len(s) > 1 else 0.1)).astype(np.float32))
    return arrs


def _report(tag, got, ref):
    worst = {}
    for k in ("act1", "act2", "act3", "head", "probs", "values"):
        d = np.abs(got[k].astype(np.float64) - ref[k])
        scale = max(np.abs(ref[k]).max(), 1e-30)
        worst[k] = (d.max(), d.max() / scale)
        extra = ""
        if k.startswith("act") and d.max() / scale > 1e-3:
            # where is it wrong?  per board / per cell / per channel maxima help to tell a layout bug
            extra = " | bad boards %s cells %s chans %s" % (
                np.nonzero(d.max(axis=(1, 2)) / scale > 1e-3)[0][:12].tolist(),
                np.nonzero(d.max(axis=(0, 2)) / scale > 1e-3)[0][:24].tolist(),
                np.nonzero(d.max(axis=(0, 1)) / scale > 1e-3)[0][:24].tolist())
        print("tc %-10s %-6s max|d| %.3e  rel %.3e  (ref max %.3e)%s" % (tag, k, d.max(), d.max() / scale, scale, extra))
    return worst


@pytest.mark.parametrize("w,B,binary", [(8, 7, True), (8, 300, True), (6, 41, True), (9, 33, True), (15, 10, True),
                                        (3, 5, True), (5, 64, False), (8, 16, False), (11, 29, True)])
def test_tc_layers_vs_torch(w, B, binary):
    from alphazero_gomoku_b200 import Engine
    rng = np.random.RandomState(1000 * w + B)
    arrs = _random_weights(rng, w, scale=1.3)
    planes = (rng.rand(B, 4, w, w) < 0.35).astype(np.float32) if binary else rng.rand(B, 4, w, w).astype(np.float32)
    planes[0] = 0.0
    e = Engine(width=w, height=w, n_in_row=3, net_impl=2)
    e.set_weights(arrs)
    got = e.debug_net_layers(planes)
    ref = _layers_ref(arrs, planes, __import__("torch").float64)
    worst = _report("%dx%d B=%d" % (w, w, B), got, ref)
    for k in ("act1", "act2", "act3", "head"):
        assert worst[k][1] < 2e-5, (k, worst[k])
    assert worst["probs"][0] < NET_TOL and worst["values"][0] < NET_TOL
    # the public entry point runs the same path and agrees with the fp32 CUDA-core path
    probs, vals = e.net_forward(planes)
    assert np.array_equal(probs, got["probs"]) and np.array_equal(vals, got["values"])
    e2 = Engine(width=w, height=w, n_in_row=3, net_impl=1)
    e2.set_weights(arrs)
    p1, v1 = e2.net_forward(planes)
    assert np.abs(p1 - probs).max() < NET_TOL and np.abs(v1 - vals).max() < NET_TOL
    e.close(); e2.close()


@pytest.mark.parametrize("tag,w,n", [("6_6_4", 6, 4), ("8_8_5", 8, 5)])
def test_tc_golden_shipped_weights(tag, w, n):
    """tcgen05 path vs the reference PyTorch fp32 Net outputs on the shipped weights."""
    from alphazero_gomoku_b200 import Engine
    z = np.load(os.path.join(GOLDEN, "net.npz"))
    arrs = load_golden_weights("weights_" + tag)
    e = Engine(width=w, height=w, n_in_row=n, net_impl=2)
    e.set_weights(arrs)
    probs, vals = e.net_forward(z[tag + "_planes"])
    dp, dv = np.abs(probs - z[tag + "_probs"]).max(), np.abs(vals - z[tag + "_values"]).max()
    print("tc golden", tag, "max|dp| %.3e max|dv| %.3e" % (dp, dv))
    assert dp < NET_TOL and dv < NET_TOL
    e.close()


def test_tc_large_batch_matches_small_batches():
    """Persistent grid, ragged last group: a 16384-board batch equals the same boards in pieces."""
    from alphazero_gomoku_b200 import Engine
    rng = np.random.RandomState(5)
    arrs = load_golden_weights("weights_8_8_5")
    planes = (rng.rand(4099, 4, 8, 8) < 0.3).astype(np.float32)
    e = Engine(width=8, height=8, n_in_row=5, net_impl=2)
    e.set_weights(arrs)
    p, v = e.net_forward(planes)
    for lo, hi in [(0, 1), (1, 3), (3, 450), (450, 4099)]:
        pp, vv = e.net_forward(planes[lo:hi])
        assert np.array_equal(pp, p[lo:hi]) and np.array_equal(vv, v[lo:hi])
    e.close()


def test_tc_repeated_launches_are_identical():
    """The fused trunk hands tiles between its warp sets through shared memory, TMEM and a dozen barriers; a missing
    ordering shows up as an occasional different bit.  150 launches over the same 16384 boards (28 groups per CTA,
    every pipeline phase overlapped with its neighbours) and 40 over ragged sizes must reproduce the first result."""
    import torch
    from alphazero_gomoku_b200 import Engine
    rng = np.random.RandomState(17)
    e = Engine(width=8, height=8, n_in_row=5, net_impl=2)
    e.set_weights(load_golden_weights("weights_8_8_5"))
    planes = torch.from_numpy((rng.rand(16384, 4, 8, 8) < 0.3).astype(np.float32)).cuda()
    p0, v0 = e.net_forward(planes)
    e.sync()
    p0, v0 = p0.clone(), v0.clone()
    for rep in range(150):
        p, v = e.net_forward(planes)
        e.sync()
        assert torch.equal(p, p0) and torch.equal(v, v0), rep
    for rep in range(40):
        B = int(rng.randint(1, 16384))
        p, v = e.net_forward(planes[:B])
        e.sync()
        assert torch.equal(p, p0[:B]) and torch.equal(v, v0[:B]), (rep, B)
    e.close()


_PAIR_SCRIPT = r"""
import sys, numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[1] + "/tests")
from alphazero_gomoku_b200 import Engine
from helpers import load_golden_weights
out = {}
for tag, w, n, B in (("8", 8, 5, 4099), ("6", 6, 4, 1), ("9", 9, 5, 777), ("15", 15, 5, 130)):
    rng = np.random.RandomState(11 + w)
    if w in (6, 8):
        arrs = load_golden_weights("weights_%d_%d_%d" % (w, w, n))
    else:
        from test_gpu_net_tc import _random_weights
        arrs = _random_weights(rng, w)
    planes = (rng.rand(B, 4, w, w) < 0.3).astype(np.float32)
    e = Engine(width=w, height=w, n_in_row=n, net_impl=2)
    e.set_weights(arrs)
    p, v = e.net_forward(planes)
    out["p" + tag], out["v" + tag] = p, v
    e.close()
np.savez(sys.argv[2], **out)
"""


def test_tc_cta_pair_variant_matches_single_cta(tmp_path):
    """GZ_TC_PAIR=1 runs the trunk as clusters of two CTAs issuing cta_group::2 MMAs (M = 256 over the pair, each CTA
    holding half of every B operand; DESIGN 4.5: built, measured slower, kept as an A/B switch).  Same products in the
    same order, so it must reproduce the shipped kernel bit for bit - dense 8x8 and padded layouts, odd group counts
    (a peer CTA without a group), a single board."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for pair in (False, True):
        env = dict(os.environ)
        env.pop("GZ_TC_PAIR", None)
        if pair:
            env["GZ_TC_PAIR"] = "1"
        f = str(tmp_path / ("pair%d.npz" % pair))
        subprocess.run([sys.executable, "-c", _PAIR_SCRIPT, root, f], check=True, env=env, timeout=240)
        outs.append(np.load(f))
    for k in outs[0].files:
        assert np.array_equal(outs[0][k], outs[1][k]), k


@pytest.mark.parametrize("w,B,trunk", [(15, 5, (64, 128, 256)), (8, 700, (64, 128, 256)), (6, 41, (96, 64, 64)),
                                       (9, 333, (64, 128, 128)), (15, 160, (32, 64, 256)), (3, 1, (128, 128, 192))])
def test_tc_wide_trunk_layers_vs_torch(w, B, trunk):
    """Wider trunks (BASELINE config C4) take the layer-wise tcgen05 kernels of gz_net_tcw.cu: same
    per-layer and output tolerances as the fused kernel, against float64 torch; the fp32 CUDA-core
    path on the same weights is the second witness."""
    from alphazero_gomoku_b200 import Engine
    rng = np.random.RandomState(7000 * w + B)
    arrs = _random_weights(rng, w, scale=1.3, trunk=trunk)
    planes = (rng.rand(B, 4, w, w) < 0.35).astype(np.float32)
    planes[0] = 0.0
    e = Engine(width=w, height=w, n_in_row=3, net_impl=2, trunk=trunk)
    e.set_weights(arrs)
    got = e.debug_net_layers(planes)
    ref = _layers_ref(arrs, planes, __import__("torch").float64)
    worst = _report("wide %dx%d B=%d %s" % (w, w, B, trunk), got, ref)
    for k in ("act1", "act2", "act3", "head"):
        assert worst[k][1] < 2e-5, (k, worst[k])
    assert worst["probs"][0] < NET_TOL and worst["values"][0] < NET_TOL
    probs, vals = e.net_forward(planes)
    assert np.array_equal(probs, got["probs"]) and np.array_equal(vals, got["values"])
    # a smaller batch afterwards (stale groups in the activation images must not leak)
    k = max(1, B // 3)
    p2, v2 = e.net_forward(planes[:k])
    assert np.array_equal(p2, probs[:k]) and np.array_equal(v2, vals[:k])
    e2 = Engine(width=w, height=w, n_in_row=3, net_impl=1, trunk=trunk)
    e2.set_weights(arrs)
    p1, v1 = e2.net_forward(planes)
    assert np.abs(p1 - probs).max() < NET_TOL and np.abs(v1 - vals).max() < NET_TOL
    e.close(); e2.close()


def test_tc_rejects_other_trunks():
    from alphazero_gomoku_b200 import Engine
    e = Engine(width=6, height=6, n_in_row=4, net_impl=2, trunk=(64, 192, 256))
    rng = np.random.RandomState(0)
    from alphazero_gomoku_b200.engine import weight_shapes
    e.set_weights([rng.randn(*s).astype(np.float32) * 0.05 for s in weight_shapes(6, 6, (64, 192, 256))])
    with pytest.raises(Exception):
        e.net_forward(np.zeros((1, 4, 6, 6), np.float32))
    e.close()
