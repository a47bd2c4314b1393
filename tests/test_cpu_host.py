"""CPU tests (no GPU): the C-ABI library loads and exports every declared symbol, the package
fails loudly without a GPU (no fallback), host-side logic (game sharding, sample packing,
dihedral augmentation) and the two multi-GPU collectives over gloo with world_size 2."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _has_gpu():
    from alphazero_gomoku_b200 import _lib
    return _lib.lib().gz_device_count() > 0


def test_abi_exports_every_declared_symbol():
    from alphazero_gomoku_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "gomoku_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = sorted(set(re.findall(r"\b(gz_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 35
    L = _lib.lib()  # raises if the .so is missing or lacks a bound symbol
    for name in declared:
        assert hasattr(L, name), name
    assert sorted(_lib.exported_symbols()) == declared  # the ctypes table covers the whole header
    assert b"sm_100a" in L.gz_version()
    # no torch / C++ types cross the boundary: every exported gz_ symbol is unmangled C
    out = subprocess.check_output(["nm", "-D", "--defined-only", _lib.LIB_PATH], text=True)
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    assert set(declared) <= exported


def test_library_has_blackwell_kernels():
    """The shipped .so carries sm_100a SASS with tcgen05 MMAs, TMEM loads and bulk copies."""
    from alphazero_gomoku_b200 import _lib
    try:
        sass = subprocess.check_output(["cuobjdump", "-sass", _lib.LIB_PATH], text=True, stderr=subprocess.DEVNULL)
    except (OSError, subprocess.CalledProcessError):
        pytest.skip("cuobjdump not available")
    assert "sm_100a" in sass
    for mnemonic in ("UTCHMMA", "LDTM", "UBLKCP"):
        assert mnemonic in sass, mnemonic


def test_no_cpu_fallback():
    if _has_gpu():
        pytest.skip("a GPU is visible")
    from alphazero_gomoku_b200 import Board, Engine, PolicyValueNet
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        Engine(width=6, height=6, n_in_row=4)
    with pytest.raises(RuntimeError):
        PolicyValueNet(6, 6)
    with pytest.raises(RuntimeError):
        PolicyValueNet(6, 6, use_gpu=False)
    b = Board(width=6, height=6, n_in_row=4)
    b.init_board()
    b.do_move(3)  # host bookkeeping works ...
    assert b.states == {3: 1} and b.current_player == 2 and 3 not in b.availables
    with pytest.raises(ValueError):
        b.do_move(3)  # ... with the reference's error for an occupied cell (game.py:122)
    with pytest.raises(RuntimeError):
        b.game_end()  # ... but rule evaluation needs the device
    with pytest.raises(Exception):
        Board(width=3, height=3, n_in_row=5).init_board()  # game.py:61-62


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "alphazero_gomoku_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                for bad in (r"^\s*(from|import)\s+oracle", r'#include\s+"[^"]*oracle', r"liborc", r"\borc_\w+\s*\("):
                    assert not re.search(bad, src, flags=re.M), (os.path.join(dp, f), bad)


def test_shard_games():
    from alphazero_gomoku_b200.train import shard_games
    for n in (0, 1, 7, 8, 16384, 100003):
        for world in (1, 2, 3, 8):
            spans = [shard_games(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_pack_unpack_samples_roundtrip():
    from alphazero_gomoku_b200.train import pack_samples, unpack_samples
    rng = np.random.RandomState(0)
    n, w = 37, 8
    states = (rng.rand(n, 4, w, w) < 0.3).astype(np.float32)
    pis = rng.dirichlet(np.ones(w * w), size=n).astype(np.float32)
    zs = rng.choice([-1.0, 0.0, 1.0], size=n).astype(np.float32)
    gids = (np.int64(5) << 40) + rng.randint(0, 16384, size=n).astype(np.int64)
    plies = rng.randint(0, 64, size=n).astype(np.int32)
    s2, p2, z2, g2, l2 = unpack_samples(pack_samples(states, pis, zs, gids, plies), w, w)
    assert np.array_equal(s2, states) and np.array_equal(p2, pis) and np.array_equal(z2, zs)
    assert np.array_equal(g2, gids) and np.array_equal(l2, plies)
    assert pack_samples(states[:0], pis[:0], zs[:0], gids[:0], plies[:0]).shape[0] == 0


def test_equi_data_is_the_dihedral_orbit():
    """train.py:116-135: 8 images; a one-hot pi keeps pointing at the image of the same stone."""
    from helpers import dihedral_images
    w = 6
    rng = np.random.RandomState(1)
    move = 2 * w + 5                      # cell (h=2, w=5)
    state = np.zeros((4, w, w))
    state[0][w - 1 - move // w, move % w] = 1.0   # stored row-reversed (game.py:116)
    state[1] = (rng.rand(w, w) < 0.3)
    pi = np.zeros(w * w); pi[move] = 1.0
    out = dihedral_images(state, pi, 1.0, w, w)
    assert len(out) == 8
    seen = set()
    for s, p, z in out:
        assert z == 1.0 and s.shape == (4, w, w) and abs(p.sum() - 1.0) < 1e-12
        m = int(np.argmax(p))
        assert s[0][w - 1 - m // w, m % w] == 1.0 and s[0].sum() == 1.0
        assert s[1].sum() == state[1].sum()
        seen.add(s.tobytes())
    assert len(seen) == 8


_WORKER = r"""
import os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, %(root)r)
from alphazero_gomoku_b200.train import (all_gather_samples, all_gather_samples_device, allreduce_mean, broadcast_weights,
                                         pack_samples, shard_games)
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
# every rank "plays" its shard of 11 games; game g yields (g %% 4) + 1 samples tagged with g
lo, hi = shard_games(11, rank, world)
rows = [(g, k) for g in range(lo, hi) for k in range(g %% 4 + 1)]
S = 9
states = np.zeros((len(rows), 4, 3, 3), np.float32)
pis = np.zeros((len(rows), S), np.float32)
for i, (g, k) in enumerate(rows):
    states[i, 0].flat[g %% S] = 1.0
    pis[i, k %% S] = 1.0
rec = pack_samples(states, pis, np.full(len(rows), float(rank), np.float32),
                   np.array([g for g, _ in rows], np.int64), np.array([k for _, k in rows], np.int32))
allrec = all_gather_samples(rec)
want = [(g, k) for g in range(11) for k in range(g %% 4 + 1)]
got = [(int(r[1]), int(r[2])) for r in allrec]
assert got == want, (got, want)            # union ordered (rank, game, ply) == data_buffer.extend order
# the tensor-resident variant (NCCL on the GPU box, gloo here): same union, same order, ragged counts incl. zero
zs = torch.tensor([100.0 * g + k for g, k in rows])
st2, pi2, z2 = all_gather_samples_device(torch.from_numpy(states), torch.from_numpy(pis), zs)
assert z2.tolist() == [100.0 * g + k for g, k in want]
assert st2.shape == (len(want), 4, 3, 3) and [int(x.flatten().argmax()) for x in st2[:, 0]] == [g %% S for g, _ in want]
assert [int(x.argmax()) for x in pi2] == [k %% S for _, k in want]
empty = rank == 1
e_st, e_pi, e_z = all_gather_samples_device(torch.from_numpy(states[:0] if empty else states),
                                           torch.from_numpy(pis[:0] if empty else pis), zs[:0] if empty else zs)
assert e_z.shape[0] == len([1 for g, k in want if g < shard_games(11, 0, world)[1]])
ws = [np.full((3, 2), 7.0 + i, np.float32) if rank == 0 else np.zeros((3, 2), np.float32) for i in range(16)]
out = broadcast_weights(ws, src=0)
assert all(float(o[0, 0]) == 7.0 + i for i, o in enumerate(out))
# gradient exchange of the data-parallel update: in-place mean over ranks, numpy and torch inputs
g = np.arange(5, dtype=np.float32) * (rank + 1)
assert np.allclose(allreduce_mean(g), np.arange(5) * 1.5) and np.allclose(g, np.arange(5) * 1.5)
t = torch.full((4,), float(rank))
assert torch.allclose(allreduce_mean(t), torch.full((4,), 0.5))
dist.barrier()
dist.destroy_process_group()
print("rank", rank, "ok")
"""


def test_collectives_gloo_world2(tmp_path):
    """sample all-gather (ragged), weight broadcast and gradient mean on 2 ranks over gloo."""
    script = tmp_path / "worker.py"
    script.write_text(_WORKER % {"root": ROOT})
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29731", WORLD_SIZE="2")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=300)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, o
        assert "rank %d ok" % r in o


def test_bench_reference_arm_contract():
    """bench.py --impl reference prints one JSON line with the keys the driver reads."""
    import json
    out = subprocess.check_output([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                                   "--warmup", "0", "--playouts", "20", "--board", "6", "--n-in-row", "4"],
                                  text=True, timeout=600)
    line = json.loads(out.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "mcts_playouts_per_sec" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] in ("port", "reference") and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0


def test_equi_data_host_vs_reference_golden():
    """The numpy checker tests/helpers.py:dihedral_images against fixtures generated by the live reference's
    TrainPipeline.get_equi_data (tests/golden/make_golden.py equi); the product path is gz_equi_data (GPU tests)."""
    from helpers import dihedral_images
    z = np.load(os.path.join(ROOT, "tests", "golden", "equi.npz"))
    for w in (6, 8, 9, 15):
        st, pi, zs = z["w%d_states" % w], z["w%d_pis" % w], z["w%d_zs" % w]
        out = [o for i in range(len(zs)) for o in dihedral_images(st[i], pi[i], zs[i], w, w)]
        assert np.array_equal(np.array([o[0] for o in out]), z["w%d_out_states" % w])
        assert np.array_equal(np.array([o[1] for o in out]), z["w%d_out_pis" % w])
        assert np.array_equal(np.array([o[2] for o in out]), z["w%d_out_zs" % w])


def test_device_replay_buffer_is_deque_plus_random_sample():
    """DeviceReplayBuffer (ring of tensors, here on the CPU device) == deque(maxlen) + random.sample
    (train.py:82,167) under the same host RNG state, through several wrap-arounds."""
    import random
    from collections import deque
    import torch
    from alphazero_gomoku_b200.train import DeviceReplayBuffer
    rng = np.random.RandomState(1)
    buf, ref = DeviceReplayBuffer(50, 3, 3, "cpu"), deque(maxlen=50)
    tag = 0
    for n in (7, 20, 30, 3, 50, 64, 1, 49, 120, 5):
        st = np.zeros((n, 4, 3, 3), np.float32)
        st[:, 0, 0, 0] = np.arange(tag, tag + n)
        pi = rng.rand(n, 9).astype(np.float32)
        z = np.arange(tag, tag + n).astype(np.float32)
        tag += n
        buf.extend(torch.from_numpy(st), torch.from_numpy(pi), torch.from_numpy(z))
        ref.extend((st[i], pi[i], z[i]) for i in range(n))
        assert len(buf) == len(ref)
        k = min(len(ref), 6)
        random.seed(tag)
        want = random.sample(ref, k)
        random.seed(tag)
        gs, gp, gz = buf.sample(k)
        assert np.array_equal(gz.numpy(), np.array([w[2] for w in want]))
        assert np.array_equal(gs.numpy(), np.array([w[0] for w in want]))
        assert np.array_equal(gp.numpy(), np.array([w[1] for w in want]))


def test_games_from_steps_log(tmp_path):
    """train.py:446-466: steps-log lines -> unique decisive games, winner from the parity of the length."""
    from alphazero_gomoku_b200.train import games_from_steps_log
    path = tmp_path / "steps.log"
    path.write_text("{'steps': [112, 97, 125], 'winner': 1, 'start_player': 1}\n"
                    "{'steps': [1, 2, 3, 4], 'winner': 2, 'start_player': 2}\n"
                    "{'steps': [5, 6], 'winner': -1, 'start_player': 1}\n"
                    "{'steps': [112, 97, 125], 'winner': 1, 'start_player': 1}\n"
                    "\n{'steps': [9], 'winner': 1}\n")
    assert games_from_steps_log(str(path)) == [[2, [1, 2, 3, 4]], [1, [112, 97, 125]]]


def test_one_word_win_test_matches_the_wide_detector(tmp_path):
    """bb8_has_line_dir (gz_common.cuh), the win test of the fused tree kernels on boards up to 8 x 8, is host-callable:
    tests/csrc/bb8_check.cu compares it on the CPU, direction by direction, with the 256-bit detector it replaces and
    with a plain scan written from game.py:131-160 - 288 000 (position, direction) pairs over board sides 3..8 and
    run lengths 3..6."""
    import shutil
    nvcc = os.environ.get("NVCC") or shutil.which("nvcc")
    if not nvcc:
        pytest.skip("nvcc not found")
    exe = str(tmp_path / "bb8_check")
    src = os.path.join(ROOT, "tests", "csrc", "bb8_check.cu")
    subprocess.check_call([nvcc, "-O2", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-o", exe, src],
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=300)
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and out.stdout.startswith("ok "), out.stdout + out.stderr
    assert int(out.stdout.split()[1]) >= 200000
