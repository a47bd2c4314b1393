"""PolicyValueNet: drop-in for the reference's policy_value_net_pytorch.py (:60-157).

Inference (`policy_value_fn`, `policy_value`) runs through the C ABI (gz_net_forward: the
tcgen05 fused trunk + FC heads).  `train_step` (reference :117-148) is gz_train_step:
hand-written forward/backward/Adam kernels over a flat master weight vector on the device; engines
that search with this net pick the new weights up through `version` / `get_policy_param()`.
"""
import pickle

import numpy as np

from .engine import Engine, STATE_DICT_KEYS, weight_shapes


def set_learning_rate(optimizer, lr):
    for param_group in optimizer.param_groups:
        param_group['lr'] = lr


def _default_init(width, height, seed=None):
    """torch.nn default init of the reference Net (kaiming-uniform a=sqrt(5) == U(+-1/sqrt(fan_in)))."""
    rng = np.random.RandomState(seed)
    arrs = []
    shapes = weight_shapes(width, height)
    for i in range(0, 16, 2):
        fan_in = int(np.prod(shapes[i][1:]))
        bound = 1.0 / np.sqrt(fan_in)
        arrs.append(rng.uniform(-bound, bound, size=shapes[i]).astype(np.float32))
        arrs.append(rng.uniform(-bound, bound, size=shapes[i + 1]).astype(np.float32))
    return arrs


def load_model_file(path, width, height):
    """Reads a torch state_dict (policy_value_net_pytorch.py:76-78) or a Theano/Lasagne pickle of
    16 arrays (policy_value_net.py:118-121); returns (arrays, layout)."""
    from . import _lib
    with open(path, "rb") as f:
        head = f.read(8)
    # torch.save writes a zip archive (PK\x03\x04) or, before torch 1.6, a pickle of the magic number
    # 0x1950a86a20f9469cfc6c; anything else is treated as a Theano/Lasagne pickle.  A torch archive that
    # fails to load raises its own error instead of falling through to an unrelated unpickling error.
    is_torch = head[:4] == b"PK\x03\x04"
    if not is_torch and head[:2] == b"\x80\x02":
        with open(path, "rb") as f:
            is_torch = b"\x8a\x0a\x6c\xfc\x9c\x46\xf9\x20\x6a\xa8\x50\x19" in f.read(32)
    if is_torch:
        import torch
        sd = torch.load(path, map_location="cpu", weights_only=True)
        if not (isinstance(sd, dict) and all(k in sd for k in STATE_DICT_KEYS)):
            raise Exception("model file %s: not a state_dict of the reference Net (keys %s...)"
                            % (path, list(sd)[:3] if isinstance(sd, dict) else type(sd)))
        return [sd[k].detach().cpu().float().numpy() for k in STATE_DICT_KEYS], _lib.LAYOUT_TORCH
    with open(path, "rb") as f:
        params = pickle.load(f, encoding="bytes")
    if len(params) != 16:
        raise Exception("model file %s: expected 16 parameter arrays, got %d" % (path, len(params)))
    return [np.asarray(p, dtype=np.float32) for p in params], _lib.LAYOUT_THEANO


class PolicyValueNet():
    """policy-value network"""

    def __init__(self, board_width, board_height, model_file=None, use_gpu=True, seed=None, device=0):
        if not use_gpu:
            raise RuntimeError("alphazero_gomoku_b200 has no CPU path: use_gpu must be True")
        self.use_gpu = True
        self.board_width = board_width
        self.board_height = board_height
        self.l2_const = 1e-4
        self.version = 0  # bumped whenever the weights change; MCTS engines re-upload lazily
        n = min(board_width, board_height, 5)
        self._engine = Engine(width=board_width, height=board_height, n_in_row=n, n_games=1, n_playout=1,
                              nodes_per_game=4, device=device)
        from . import _lib
        if model_file:
            arrs, layout = load_model_file(model_file, board_width, board_height)
        else:
            arrs, layout = _default_init(board_width, board_height, seed), _lib.LAYOUT_TORCH
        self._engine.set_weights(arrs, layout=layout)
        self._params = self._engine.get_weights()  # torch layout, host copies (None = stale after training)

    # ---- inference -----------------------------------------------------------------------------
    def policy_value(self, state_batch):
        """:80-94: batch of states -> (act_probs [B,S], values [B,1]) as numpy float32."""
        x = np.ascontiguousarray(np.asarray(state_batch, dtype=np.float32).reshape(
            -1, 4, self.board_width, self.board_height))
        probs, values = self._engine.net_forward(x)
        return probs, values.reshape(-1, 1)

    def policy_value_fn(self, board):
        """:96-115: board -> (zip(legal moves, probs), value of the side to move)."""
        legal_positions = board.availables
        current_state = np.ascontiguousarray(board.current_state().reshape(
            -1, 4, self.board_width, self.board_height), dtype=np.float32)
        probs, values = self._engine.net_forward(current_state)
        act_probs = zip(legal_positions, probs[0][legal_positions])
        return act_probs, float(values[0])

    # ---- weights -------------------------------------------------------------------------------
    def _host_params(self):
        if self._params is None:
            self._params = self._engine.get_weights()
        return self._params

    def get_policy_param(self):
        """16 float32 arrays in state_dict order and layout (the reference returns the state_dict)."""
        return [p.copy() for p in self._host_params()]

    def state_dict(self):
        import torch
        return {k: torch.from_numpy(p.copy()) for k, p in zip(STATE_DICT_KEYS, self._host_params())}

    def set_policy_param(self, arrays, layout=0):
        self._engine.set_weights(arrays, layout=layout)
        self._params = self._engine.get_weights()
        self.version += 1

    def save_model(self, model_file):
        """:154-157: torch.save(state_dict) -- loadable by the reference's PolicyValueNet."""
        import torch
        torch.save(self.state_dict(), model_file)

    # ---- training ----------------------------------------------------------------
    def _batch(self, state_batch, mcts_probs=None, winner_batch=None):
        """Lists / numpy arrays (the reference's inputs) -> float32 arrays; torch cuda tensors (a batch
        drawn from the device replay buffer) pass through untouched."""
        if hasattr(state_batch, "data_ptr"):
            return state_batch, mcts_probs, winner_batch
        st = np.ascontiguousarray(np.asarray(state_batch, dtype=np.float32).reshape(
            -1, 4, self.board_width, self.board_height))
        pi = None if mcts_probs is None else np.asarray(mcts_probs, dtype=np.float32)
        z = None if winner_batch is None else np.asarray(winner_batch, dtype=np.float32)
        return st, pi, z

    def train_step(self, state_batch, mcts_probs, winner_batch, lr):
        """:117-148: loss = MSE(v, z) - mean(sum(pi * log p)), Adam with weight decay 1e-4; forward,
        backward and the optimiser step are CUDA kernels behind gz_train_step (the Adam moments live
        in the engine for the life of this object, like the reference's self.optimizer)."""
        st, pi, z = self._batch(state_batch, mcts_probs, winner_batch)
        loss, entropy = self._engine.train_step(st, pi, z, lr)
        self._params = None  # pulled from the device on demand
        self._trained = True
        self.version += 1
        return loss, entropy

    def train_backward(self, state_batch, mcts_probs, winner_batch):
        """Forward + backward of train_step without the optimiser step -> (loss, entropy); the flat
        gradient stays on the device (`train_grads`) so that ranks can average it first."""
        return self._engine.train_backward(*self._batch(state_batch, mcts_probs, winner_batch))

    def train_grads(self, out=None):
        return self._engine.train_grads(out)

    def train_apply(self, lr, grads=None):
        """The Adam step of train_step, with `grads` (flat, host or device) or the local gradient."""
        self._engine.train_apply(lr, grads)
        self._params = None
        self.version += 1

    def policy_value_train(self, state_batch):
        """policy_value (:80-94) with the weights as trained so far, without refreshing the tcgen05
        inference images: what policy_update's KL check calls between epochs (train.py:178-193)."""
        probs, values = self._engine.train_forward(self._batch(state_batch)[0])
        return probs, values.reshape(-1, 1)
