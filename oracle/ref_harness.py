"""ORACLE — TEST INFRASTRUCTURE ONLY.

Imports the UNMODIFIED Python reference from /root/reference (present in the build
container only, never on the GPU box) with the shims SURVEY.md 8c lists, so that the C
oracle can be pinned against it and golden fixtures can be generated
(tests/golden/make_golden.py).  Nothing here is copied from the reference.
"""
import contextlib
import importlib
import io
import os
import pickle
import sys

import numpy as np

REF_DIR = os.environ.get("GOMOKU_REFERENCE_DIR", "/root/reference")


def available():
    return os.path.isfile(os.path.join(REF_DIR, "game.py"))


def _import(name):
    if REF_DIR not in sys.path:
        sys.path.append(REF_DIR)
    # The product package has same-named modules *inside* alphazero_gomoku_b200/, never
    # top-level, so a top-level import of `game` etc. resolves to the reference.
    return importlib.import_module(name)


def game():
    return _import("game")


def mcts_alphazero():
    return _import("mcts_alphaZero")


def mcts_pure():
    return _import("mcts_pure")


def net_pytorch():
    import warnings
    warnings.filterwarnings("ignore")
    return _import("policy_value_net_pytorch")


def theano_to_torch_arrays(params):
    """Theano/Lasagne pickle (true convolution, FC [in,out]) ->
    torch state_dict layout (cross-correlation OIHW, FC [out,in])."""
    out = []
    for i, a in enumerate(params):
        a = np.asarray(a, dtype=np.float32)
        if a.ndim == 4:
            a = a[:, :, ::-1, ::-1]
        elif a.ndim == 2:
            a = a.T
        out.append(np.array(a, dtype=np.float32, order="C", copy=True))
    return out


def load_theano_weights(name):
    """name: e.g. 'best_policy_8_8_5.model' under reference mode/."""
    with open(os.path.join(REF_DIR, "mode", name), "rb") as f:
        params = pickle.load(f, encoding="bytes")
    return theano_to_torch_arrays(params)


def make_ref_net(width, height, arrays=None, seed=0):
    """Reference PolicyValueNet (CPU) optionally loaded with state_dict-layout arrays."""
    import torch
    pv = net_pytorch()
    torch.manual_seed(seed)
    net = pv.PolicyValueNet(width, height)
    if arrays is not None:
        sd = net.policy_value_net.state_dict()
        for k, a in zip(sd.keys(), arrays):
            assert tuple(sd[k].shape) == tuple(a.shape), (k, sd[k].shape, a.shape)
            sd[k] = torch.from_numpy(np.array(a, dtype=np.float32, order="C", copy=True))
        net.policy_value_net.load_state_dict(sd)
    return net


def net_arrays(ref_net):
    return [v.detach().cpu().numpy().astype(np.float32).copy()
            for v in ref_net.policy_value_net.state_dict().values()]


def float_value_fn(ref_net):
    """policy_value_fn with the value as a Python float (SURVEY 8c-3 shim), so the
    reference search runs in float64 like its Theano/NumPy back ends did."""
    def fn(board):
        with contextlib.redirect_stderr(io.StringIO()):
            probs, v = ref_net.policy_value_fn(board)
        return probs, float(v)
    return fn


@contextlib.contextmanager
def quiet():
    """mcts_pure prints the tree and board.states on every move."""
    with contextlib.redirect_stdout(io.StringIO()):
        yield


def random_position(Board, width, height, n_in_row, rng, max_plies=None, start_player=0):
    """Seeded sampler: k ~ U[0, max_plies] uniformly random legal plies from the empty
    board, stopping at the first terminal state (SURVEY 8d parity set).  Returns
    (board, moves)."""
    S = width * height
    if max_plies is None:
        max_plies = S // 2
    b = Board(width=width, height=height, n_in_row=n_in_row)
    b.init_board(start_player)
    k = int(rng.randint(0, max_plies + 1))
    moves = []
    for _ in range(k):
        av = b.availables
        mv = int(av[rng.randint(0, len(av))])
        b.do_move(mv)
        moves.append(mv)
        if b.game_end()[0]:
            break
    return b, moves


def train_module():
    """The reference's train.py imported with the two shims of SURVEY 8c (PyYAML 6 safe_load, the
    absent keras back end aliased to the pytorch one), from a scratch cwd holding conf/ and logs/."""
    import shutil
    import tempfile
    if "train" in sys.modules and getattr(sys.modules["train"], "__file__", "").startswith(REF_DIR):
        return sys.modules["train"]
    import yaml
    scratch = tempfile.mkdtemp(prefix="gz_ref_train_")
    shutil.copytree(os.path.join(REF_DIR, "conf"), os.path.join(scratch, "conf"))
    os.makedirs(os.path.join(scratch, "logs"), exist_ok=True)
    mytoolkit = _import("mytoolkit")
    mytoolkit.load_config = lambda *a, **k: yaml.safe_load(open(os.path.join(scratch, "conf", "train_config.yaml")))
    sys.modules["policy_value_net_keras"] = _import("policy_value_net_pytorch")
    cwd = os.getcwd()
    os.chdir(scratch)
    try:
        return _import("train")
    finally:
        os.chdir(cwd)


def equi_data(play_data, width, height):
    """TrainPipeline.get_equi_data (train.py:116-135) of the live reference."""
    class _Self(object):
        board_width, board_height = width, height
    return train_module().TrainPipeline.get_equi_data(_Self(), play_data)



@contextlib.contextmanager
def item_shim():
    """the reference's train_step ends with `loss.data[0]`, an IndexError on a 0-dim
    tensor under torch >= 0.5.  While active, indexing a 0-dim tensor with 0 returns `.item()`; the
    reference source runs unmodified."""
    import torch
    orig = torch.Tensor.__getitem__

    def getitem(self, idx):
        if self.dim() == 0 and isinstance(idx, int) and idx == 0:
            return self.item()
        return orig(self, idx)
    torch.Tensor.__getitem__ = getitem
    try:
        yield
    finally:
        torch.Tensor.__getitem__ = orig


def train_steps(width, height, arrays, batches, lrs):
    """The live reference's PolicyValueNet.train_step (policy_value_net_pytorch.py:117-148, CPU fp32)
    run on `batches` = [(states, pis, zs), ...] starting from state_dict-layout `arrays`.  Returns
    ([(loss, entropy), ...], final arrays)."""
    net = make_ref_net(width, height, arrays)
    out = []
    with item_shim(), contextlib.redirect_stderr(io.StringIO()):
        for (st, pi, z), lr in zip(batches, lrs):
            loss, ent = net.train_step(st, pi, z, lr)
            out.append((float(loss), float(ent)))
    return out, net_arrays(net)
