"""Device train step (gz_train_step) timing: ms/step at the reference's batch size
(train.py: batch_size 512) with inputs resident on the device and from host memory, next to
torch autograd + torch.optim.Adam on the same GPU (cuDNN/cuBLAS, the library baseline) and on the
host cores (what the reference's train_step runs with use_gpu=False).
    python tools/bench_train.py [--width 8] [--batch 512] [--steps 50]"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from alphazero_gomoku_b200.engine import Engine, weight_shapes  # noqa: E402


def torch_step(p, opt, x, pi, z):
    opt.zero_grad()
    h = F.relu(F.conv2d(x, p[0], p[1], padding=1))
    h = F.relu(F.conv2d(h, p[2], p[3], padding=1))
    h = F.relu(F.conv2d(h, p[4], p[5], padding=1))
    a = F.relu(F.conv2d(h, p[6], p[7])).flatten(1)
    logp = F.log_softmax(F.linear(a, p[8], p[9]), dim=1)
    v = F.relu(F.conv2d(h, p[10], p[11])).flatten(1)
    v = torch.tanh(F.linear(F.relu(F.linear(v, p[12], p[13])), p[14], p[15]))
    loss = F.mse_loss(v.view(-1), z) - torch.mean(torch.sum(pi * logp, 1))
    loss.backward()
    opt.step()
    ent = -torch.mean(torch.sum(torch.exp(logp) * logp, 1))
    return loss.item(), ent.item()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--width", type=int, default=8)
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--cpu-steps", type=int, default=5)
    a = ap.parse_args()
    w, B, S = a.width, a.batch, a.width * a.width
    rng = np.random.RandomState(0)
    arrs = [(rng.uniform(-1, 1, size=s) / np.sqrt(max(int(np.prod(s[1:])), 16))).astype(np.float32)
            for s in weight_shapes(w, w)]
    st = (rng.rand(B, 4, w, w) < 0.3).astype(np.float32)
    pi = rng.dirichlet(0.3 * np.ones(S), size=B).astype(np.float32)
    z = rng.choice([-1.0, 0.0, 1.0], size=B).astype(np.float32)
    e = Engine(width=w, height=w, n_in_row=min(w, 5), n_games=1, n_playout=1, nodes_per_game=4)
    e.set_weights(arrs)
    dev = torch.device("cuda", 0)
    dst, dpi, dz = (torch.from_numpy(x).to(dev) for x in (st, pi, z))
    out = {"workload": "train_step %dx%d batch %d (policy_value_net_pytorch.py:117-148)" % (w, w, B)}

    def timed(fn, n):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / n * 1e3

    out["device_inputs_ms"] = timed(lambda: e.train_step(dst, dpi, dz, 2e-3), a.steps)
    out["host_inputs_ms"] = timed(lambda: e.train_step(st, pi, z, 2e-3), a.steps)
    # kernels only (no loss read-back, no host sync per step): backward + apply issued back to back
    L, h = e.L, e._h
    import ctypes as C
    def raw():
        L.gz_train_step(h, dst.data_ptr(), dpi.data_ptr(), dz.data_ptr(), B, C.c_float(2e-3), None, None)
    out["kernels_only_ms"] = timed(raw, a.steps)
    flops = 3 * 2 * B * S * (9 * 4 * 32 + 9 * 32 * 64 + 9 * 64 * 128 + 128 * 6) + 3 * 2 * B * (4 * S * S + 2 * S * 64 + 64)
    out["algorithmic_gflop_per_step"] = flops / 1e9
    out["kernels_only_tflops"] = flops / (out["kernels_only_ms"] * 1e-3) / 1e12

    for tf32 in (True, False):
        torch.backends.cudnn.allow_tf32 = tf32
        torch.backends.cuda.matmul.allow_tf32 = tf32
        p = [torch.tensor(x, device=dev, requires_grad=True) for x in arrs]
        opt = torch.optim.Adam(p, weight_decay=1e-4)
        out["torch_gpu_%s_ms" % ("tf32" if tf32 else "fp32")] = timed(lambda: torch_step(p, opt, dst, dpi, dz), a.steps)
    p = [torch.tensor(x, requires_grad=True) for x in arrs]
    opt = torch.optim.Adam(p, weight_decay=1e-4)
    cst, cpi, cz = torch.from_numpy(st), torch.from_numpy(pi), torch.from_numpy(z)
    torch_step(p, opt, cst, cpi, cz)
    t0 = time.perf_counter()
    for _ in range(a.cpu_steps):
        torch_step(p, opt, cst, cpi, cz)
    out["torch_cpu_ms"] = (time.perf_counter() - t0) / a.cpu_steps * 1e3
    out["torch_cpu_threads"] = torch.get_num_threads()
    out["samples_per_sec_device"] = B / (out["device_inputs_ms"] * 1e-3)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
