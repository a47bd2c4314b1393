"""gz_net_forward_noalloc (no counterpart in the reference): the net
forward with its inter-kernel scratch in a static 8 MiB device array instead of a cudaMalloc'ed buffer.
Checked through the C ABI against gz_net_forward, the existing entry point, on the same device inputs: the same two
kernels run on the same data, so the outputs must agree far inside the 1e-4 tolerance of the net tier (they are
expected to be bit-identical), on two calls in a row (the second with another batch size, so that it meets the
first call's leftovers in the shared array), across two engines of different geometry, and the call must reject
what it cannot serve.  This file is collected last on purpose."""
import numpy as np
import pytest

from helpers import load_golden_weights

pytestmark = pytest.mark.gpu

TOL = 1e-4  # the net tier's tolerance (north_star); observed difference: 0


def _random_planes(rng, B, w):
    import torch
    x = np.zeros((B, 4, w, w), np.float32)
    x[:, 0] = rng.rand(B, w, w) < 0.2
    x[:, 1] = (rng.rand(B, w, w) < 0.2) & (x[:, 0] == 0)
    x[:, 3] = rng.randint(0, 2, size=(B, 1, 1))
    return torch.from_numpy(x).cuda()


def _both(e, planes):
    import torch
    B = planes.shape[0]
    ref_p, ref_v = e.net_forward(planes)  # existing entry point, device pointers in and out
    torch.cuda.synchronize()
    p = torch.full((B, e.S), -7.0, dtype=torch.float32, device=planes.device)
    v = torch.full((B,), -7.0, dtype=torch.float32, device=planes.device)
    torch.cuda.synchronize()
    e.net_forward_noalloc(planes, p, v)
    e.sync()
    return ref_p.cpu().numpy(), ref_v.cpu().numpy(), p.cpu().numpy(), v.cpu().numpy()


def test_noalloc_forward_matches_net_forward_on_consecutive_calls():
    from alphazero_gomoku_b200 import Engine
    rng = np.random.RandomState(5)
    e8 = Engine(width=8, height=8, n_in_row=5, n_games=1, n_playout=1, nodes_per_game=4, net_impl=2)
    e8.set_weights(load_golden_weights("weights_8_8_5"))
    e6 = Engine(width=6, height=6, n_in_row=4, n_games=1, n_playout=1, nodes_per_game=4, net_impl=2)
    e6.set_weights(load_golden_weights("weights_6_6_4"))
    # two calls in a row on one engine (a full last tile, then a ragged smaller batch over its leftovers), then
    # another geometry over the same static array, then the first engine again
    for e, w, B in ((e8, 8, 384), (e8, 8, 77), (e6, 6, 301), (e8, 8, 1000), (e6, 6, 5)):
        planes = _random_planes(rng, B, w)
        rp, rv, p, v = _both(e, planes)
        assert np.isfinite(p).all() and np.isfinite(v).all()
        assert np.abs(p.sum(axis=1) - 1.0).max() < 1e-4
        assert np.abs(p - rp).max() <= TOL and np.abs(v - rv).max() <= TOL, (w, B, np.abs(p - rp).max())
        assert np.array_equal(p, rp) and np.array_equal(v, rv)  # same kernels, same inputs
    e8.close()
    e6.close()


def test_noalloc_forward_rejects_what_it_cannot_serve():
    import torch
    from alphazero_gomoku_b200 import Engine
    e = Engine(width=8, height=8, n_in_row=5, n_games=1, n_playout=1, nodes_per_game=4, net_impl=2)
    e.set_weights(load_golden_weights("weights_8_8_5"))
    rng = np.random.RandomState(6)
    big = 5376 + 1  # one board past the 8 MiB image at 8x8 (42 tiles of 128 boards)
    planes = _random_planes(rng, big, 8)
    p = torch.empty((big, 64), dtype=torch.float32, device="cuda")
    v = torch.empty((big,), dtype=torch.float32, device="cuda")
    with pytest.raises(Exception, match="at most 5376 boards"):
        e.net_forward_noalloc(planes, p, v)
    e.net_forward_noalloc(planes[:5376], p[:5376], v[:5376])  # exactly the limit is served
    e.sync()
    rp, rv = e.net_forward(planes[:5376])
    torch.cuda.synchronize()
    assert torch.equal(rp, p[:5376]) and torch.equal(rv, v[:5376])
    # host pointers would need a staging buffer: rejected through the raw C ABI
    hp = np.zeros((4, 4, 8, 8), np.float32)
    ho, hv = np.zeros((4, 64), np.float32), np.zeros(4, np.float32)
    rc = e.L.gz_net_forward_noalloc(e._h, hp.ctypes.data, 4, ho.ctypes.data, hv.ctypes.data)
    assert rc != 0 and b"device pointers" in e.L.gz_last_error(e._h)
    e.close()
    # a wider trunk has no fused kernel, hence no no-allocation variant
    from alphazero_gomoku_b200.engine import weight_shapes
    wide = Engine(width=9, height=9, n_in_row=5, n_games=1, n_playout=1, nodes_per_game=4, trunk=(64, 128, 256))
    wide.set_weights([np.zeros(s, np.float32) for s in weight_shapes(9, 9, (64, 128, 256))])
    x = _random_planes(rng, 2, 9)
    with pytest.raises(Exception, match="no-allocation variant"):
        wide.net_forward_noalloc(x, torch.empty((2, 81), device="cuda"), torch.empty(2, device="cuda"))
    wide.close()
