"""TrainPipeline: drop-in for the self-play side of the reference's train.py (:63-214).

`collect_selfplay_data(n_games)` is the hot-path entry point: the n_games are played
CONCURRENTLY on the device (one slot each; gz_selfplay_*), not one after another, and the
finished games come back as packed (state, pi, z) records in game order.  With
torch.distributed initialised, every rank plays its own games and `sync_selfplay_data` /
`broadcast_weights` are the only two collectives: an all-gather of the samples and a
broadcast of the 16 weight tensors after an update.
"""
import logging
import random
from collections import deque

import numpy as np

from .engine import Engine
from .game import Board, Game
from .mcts_alphaZero import MCTSPlayer
from .policy_value_net import PolicyValueNet

_logger = logging.getLogger(__name__)

# conf/train_config.yaml:11-37 of the reference
DEFAULT_CONF = {"board_width": 15, "board_height": 15, "n_in_row": 5, "learn_rate": 0.002, "lr_multiplier": 0.2,
                "temp": 1.0, "n_playout": 600, "c_puct": 5, "buffer_size": 3000, "batch_size": 512, "epochs": 6,
                "play_batch_size": 1, "kl_targ": 0.02, "check_freq": 20, "pure_mcts_playout_num": 3000,
                "game_batch_num": 5000}


def shard_games(n_games, rank, world):
    """Games [0, n_games) split over ranks: contiguous, sizes differ by at most one."""
    base, rem = divmod(int(n_games), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def pack_samples(states, pis, zs, game_ids, plies):
    """One float32 record per sample: [game_id_hi, game_id_lo, ply, z, pi[S], state[4S]]."""
    n = len(zs)
    S = pis.shape[1] if n else 0
    rec = np.zeros((n, 4 + 5 * S), np.float32)
    if n:
        gid = np.asarray(game_ids, np.int64)
        rec[:, 0] = (gid >> 20).astype(np.float32)
        rec[:, 1] = (gid & ((1 << 20) - 1)).astype(np.float32)
        rec[:, 2] = plies
        rec[:, 3] = zs
        rec[:, 4:4 + S] = pis
        rec[:, 4 + S:] = states.reshape(n, -1)
    return rec


def unpack_samples(rec, width, height):
    S = width * height
    gid = (rec[:, 0].astype(np.int64) << 20) | rec[:, 1].astype(np.int64)
    return (rec[:, 4 + S:].reshape(-1, 4, width, height), rec[:, 4:4 + S], rec[:, 3], gid, rec[:, 2].astype(np.int32))


def all_gather_samples(rec, group=None):
    """Variable-length all-gather of packed sample records (counts first, then padded rows).
    Works on NCCL (device tensors) and gloo (host tensors); returns the union ordered by rank."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return rec
    world = dist.get_world_size(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    n = torch.tensor([rec.shape[0]], dtype=torch.int64, device=dev)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n, group=group)
    counts = [int(c.item()) for c in counts]
    m = max(max(counts), 1)
    pad = torch.zeros((m, rec.shape[1]), dtype=torch.float32, device=dev)
    if rec.shape[0]:
        pad[:rec.shape[0]] = torch.from_numpy(rec).to(dev)
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad, group=group)
    return np.concatenate([o[:c].cpu().numpy() for o, c in zip(out, counts)], axis=0)


def all_gather_samples_device(states, pis, zs, group=None):
    """all_gather_samples for device-resident samples (torch CUDA tensors in, union out, ordered by rank): the
    counts travel first, then one padded all-gather of [z | pi | state] rows over NCCL; nothing touches the host
    except the counts."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return states, pis, zs
    world = dist.get_world_size(group)
    n, S = int(zs.shape[0]), int(pis.shape[1])
    dev = zs.device
    cnt = torch.tensor([n], dtype=torch.int64, device=dev)
    cnts = torch.zeros(world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(cnts, cnt, group=group)
    counts = [int(c) for c in cnts.tolist()]
    m = max(max(counts), 1)
    rec = torch.zeros((m, 1 + 5 * S), dtype=torch.float32, device=dev)
    if n:
        rec[:n, 0] = zs
        rec[:n, 1:1 + S] = pis
        rec[:n, 1 + S:] = states.reshape(n, -1)
    out = torch.empty((world * m, 1 + 5 * S), dtype=torch.float32, device=dev)  # rank r's rows at [r * m, (r + 1) * m)
    dist.all_gather_into_tensor(out, rec, group=group)
    rows = torch.cat([out[r * m:r * m + c] for r, c in enumerate(counts)], dim=0)
    return (rows[:, 1 + S:].reshape((-1,) + tuple(states.shape[1:])).contiguous(), rows[:, 1:1 + S].contiguous(),
            rows[:, 0].contiguous())


def broadcast_weights(arrays, src=0, group=None):
    """One flat broadcast of the 16 weight tensors (NVLink/NCCL when available)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return arrays
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    flat = torch.from_numpy(np.concatenate([np.asarray(a, np.float32).reshape(-1) for a in arrays])).to(dev)
    dist.broadcast(flat, src=src, group=group)
    flat = flat.cpu().numpy()
    out, o = [], 0
    for a in arrays:
        out.append(flat[o:o + a.size].reshape(a.shape).copy())
        o += a.size
    return out


def samples_from_train_data(engine, games):
    """Game.start_from_train_data (game.py:292-358) for a list of recorded games `[winner, moves]` at
    once: every prefix position of every game goes through ONE gz_rules_eval call (planes and move
    legality on the device).  Returns per-game lists of (state float64 [4,W,H], pi float64 [S], z)."""
    prefixes, owner = [], []
    for gi, (winner, moves) in enumerate(games):
        moves = [int(m) for m in moves]
        for k in range(len(moves)):
            prefixes.append(moves[:k])
            owner.append(gi)
        prefixes.append(moves)  # the final position only checks the last move's legality
        owner.append(gi)
    out = engine.rules_eval(prefixes, [0] * len(prefixes), want=("planes", "status"))
    bad = np.nonzero(out["status"] != 0)[0]
    if len(bad):
        raise ValueError("recorded game %d contains an illegal move" % owner[int(bad[0])])
    res, i = [], 0
    S = engine.S
    for winner, moves in games:
        plays = []
        for k, mv in enumerate(moves):
            pi = np.full(S, 0.000001)
            pi[int(mv)] = 0.99999
            mover = 1 + (k % 2)
            z = 0.0 if winner == -1 else (1.0 if mover == winner else -1.0)
            plays.append((out["planes"][i + k].astype(np.float64), pi, z))
        i += len(moves) + 1
        res.append(plays)
    return res


def games_from_steps_log(path):
    """train.py:446-466 (gen_moves_from_yixin): `{'steps': [...], 'winner': w, ...}` lines -> unique
    decisive games as [winner, moves] with winner = 2 - len(moves) % 2 (player 1 moves first)."""
    import ast
    seen, res = set(), []
    with open(path, "r") as f:
        for line in f:
            line = line.strip()
            if not line:
                break
            rec = ast.literal_eval(line)
            if rec["winner"] == -1 or tuple(rec["steps"]) in seen:
                continue
            seen.add(tuple(rec["steps"]))
    for steps in sorted(seen):
        res.append([2 - len(steps) % 2, list(steps)])
    return res


class DeviceReplayBuffer(object):
    """The reference's `deque(maxlen=buffer_size)` + `random.sample` (train.py:82,167) with the samples
    resident on the device: a ring of float32 tensors; `extend` keeps the newest maxlen
    samples, `sample(k)` draws k distinct positions with the host `random` module exactly as
    random.sample does on a sequence of the same length, and gathers them on the device."""

    def __init__(self, maxlen, width, height, device):
        import torch
        self.maxlen, self.count, self.head = int(maxlen), 0, 0
        dev = device if isinstance(device, (str, torch.device)) else torch.device("cuda", device)
        self.states = torch.zeros((self.maxlen, 4, width, height), dtype=torch.float32, device=dev)
        self.pis = torch.zeros((self.maxlen, width * height), dtype=torch.float32, device=dev)
        self.zs = torch.zeros((self.maxlen,), dtype=torch.float32, device=dev)

    def __len__(self):
        return self.count

    def extend(self, states, pis, zs):
        n = int(zs.shape[0])
        if n >= self.maxlen:
            states, pis, zs, n = states[-self.maxlen:], pis[-self.maxlen:], zs[-self.maxlen:], self.maxlen
            self.head = 0
        first = min(n, self.maxlen - self.head)
        for dst, src in ((self.states, states), (self.pis, pis), (self.zs, zs)):
            dst[self.head:self.head + first] = src[:first]
            if n > first:
                dst[:n - first] = src[first:]
        self.head = (self.head + n) % self.maxlen
        self.count = min(self.maxlen, self.count + n)

    def physical(self, logical):
        """deque position (0 = oldest) -> ring slot"""
        return (self.head - self.count + np.asarray(logical, np.int64)) % self.maxlen

    def sample(self, k):
        import torch
        idx = torch.from_numpy(self.physical(random.sample(range(self.count), k))).to(self.zs.device)
        return self.states.index_select(0, idx), self.pis.index_select(0, idx), self.zs.index_select(0, idx)


def allreduce_mean(x, group=None):
    """In-place mean over ranks of a flat float32 vector (torch tensor on the backend's device, or a
    numpy array that is staged through one); the gradient exchange of the data-parallel update."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return x
    world = dist.get_world_size(group)
    if isinstance(x, np.ndarray):
        dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
        t = torch.from_numpy(x).to(dev)
        dist.all_reduce(t, group=group)
        x[...] = (t / world).cpu().numpy()
        return x
    dist.all_reduce(x, group=group)
    x /= world
    return x


class TrainPipeline():
    def __init__(self, conf=None, init_model=None, best_model="best_policy.model", seed=0, device_buffer=False,
                 device=None, continuous=False, selfplay_slots=None):
        """device_buffer=True keeps the replay buffer on the device (DeviceReplayBuffer): samples go
        self-play ring -> augmentation -> buffer -> train step without touching the host.
        device: CUDA ordinal of this process (default: LOCAL_RANK, one process per GPU).
        continuous=True: collect_selfplay_data keeps `selfplay_slots` games in flight ACROSS calls (slots
        recycle, weight updates are picked up by the games in progress) instead of draining every batch to its
        last game; see _collect_continuous.  Default: the reference's shape, n_games from empty boards per call."""
        import os
        self.device = int(os.environ.get("LOCAL_RANK", "0")) if device is None else int(device)
        c = dict(DEFAULT_CONF)
        c.update(conf or {})
        self.board_width, self.board_height, self.n_in_row = c['board_width'], c['board_height'], c['n_in_row']
        self.board = Board(width=self.board_width, height=self.board_height, n_in_row=self.n_in_row)
        self.game = Game(self.board)
        self.learn_rate, self.lr_multiplier, self.temp = c['learn_rate'], c['lr_multiplier'], c['temp']
        self.n_playout, self.c_puct = c['n_playout'], c['c_puct']
        self.buffer_size, self.batch_size = c['buffer_size'], c['batch_size']
        self.data_buffer = deque(maxlen=self.buffer_size)
        self.device_buffer = bool(device_buffer)
        self.play_batch_size, self.epochs, self.kl_targ = c['play_batch_size'], c['epochs'], c['kl_targ']
        self.check_freq, self.game_batch_num = c['check_freq'], c['game_batch_num']
        self.best_win_ratio = 0.2
        self.pure_mcts_playout_num = c['pure_mcts_playout_num']
        self.seed = seed
        self.policy_value_net = PolicyValueNet(self.board_width, self.board_height, model_file=init_model,
                                               device=self.device)
        self.mcts_player = MCTSPlayer(self.policy_value_net.policy_value_fn, c_puct=self.c_puct,
                                      n_playout=self.n_playout, is_selfplay=1)
        self._engine = None
        self._arena = None
        self._arena_rounds = 0
        self.best_model = best_model
        self._engine_version = -1
        self._games_played = 0
        self.last_games = []
        self.continuous = bool(continuous)
        self.selfplay_slots = selfplay_slots
        # continuous mode on several ranks: agree after every ply whether ALL ranks hold their quota and leave the
        # collect loop together, so that no GPU idles at the next collective waiting for a slower rank (the extra
        # games a fast rank plays meanwhile are carried to its next call, not wasted)
        self.sync_plies = False
        self._cont = None  # continuous mode: {"next": first unreleased game serial, "games": {...}, "chunks": [...]}

    def get_equi_data(self, play_data):
        """train.py:116-135: [(state, pi, z), ...] -> the 8 dihedral images of every sample, in the reference's
        order, produced by the device kernel (gz_equi_data) and handed back as float64 arrays."""
        play_data = list(play_data)
        if not play_data:
            return []
        st, pi, z = self.policy_value_net._engine.equi_data(
            np.array([d[0] for d in play_data], np.float32), np.array([d[1] for d in play_data], np.float32),
            np.array([d[2] for d in play_data], np.float32))
        st, pi = st.astype(np.float64), pi.astype(np.float64)
        return [(st[i], pi[i], float(z[i])) for i in range(len(z))]

    def _selfplay_engine(self, n_slots):
        if self._engine is None or self._engine.G != n_slots:
            if self._engine is not None:
                self._engine.close()
            rank = _rank()
            self._engine = Engine(width=self.board_width, height=self.board_height, n_in_row=self.n_in_row,
                                  n_games=n_slots, n_playout=self.n_playout, c_puct=float(self.c_puct),
                                  seed=self.seed + rank, game_id_base=(rank << 40), device=self.device)
            self._engine_version = -1
        if self._engine_version != self.policy_value_net.version:
            self._engine.set_weights(self.policy_value_net.get_policy_param())
            self._engine_version = self.policy_value_net.version
        return self._engine

    def _play(self, n_games):
        """n_games concurrent device games; returns the drained records ordered by (game, ply)."""
        e = self._selfplay_engine(int(n_games))
        e.selfplay_begin(temp=float(self.temp), recycle=False)
        for _ in range(self.board_width * self.board_height):
            e.selfplay_step(1)
            if e.selfplay_pending()[2] == 0:
                break
        d = e.selfplay_drain()
        self.last_games = e.selfplay_drain_games()
        order = np.lexsort((d["plies"], d["game_ids"]))
        for gid, winner, moves in sorted(self.last_games):
            _logger.warning("{'steps':[%s],'winner':%d}" % (",".join(map(str, moves)), winner))  # game.py:411
        self._games_played += len(self.last_games)
        return e, {k: v[order] for k, v in d.items()}

    def selfplay_games(self, n_games):
        """Play n_games concurrently on the device; returns per-game [(state, pi, z), ...] lists
        (float64 arrays, the reference's dtypes) in slot order."""
        _, d = self._play(n_games)
        games, last = [], None
        for i in range(len(d["zs"])):
            if d["game_ids"][i] != last:
                games.append([])
                last = d["game_ids"][i]
            games[-1].append((d["states"][i].astype(np.float64), d["pis"][i].astype(np.float64), float(d["zs"][i])))
        return games

    def collect_selfplay_data(self, n_games=1, train_data=[]):
        """train.py:137-159 (self-play branch): n_games games -> x8 augmentation -> data_buffer.  The
        games run concurrently and the dihedral images are produced on the device (gz_equi_data);
        the buffer receives the reference's items in the reference's order (game, ply, image)."""
        if len(train_data) != 0:
            return self._collect_from_train_data(n_games, train_data)
        st, pi, z = self._raw_samples(n_games)
        self.episode_len = int(z.shape[0])
        self._extend_buffer(st, pi, z)
        return

    def _raw_samples(self, n_games):
        """The (state, pi, z) records of n_games finished games in (game, ply) order, before augmentation: numpy
        arrays, or torch CUDA tensors when device_buffer; batch or continuous mode."""
        if self.continuous:
            return self._collect_continuous(n_games)
        if self.device_buffer:
            return self._collect_on_device(n_games)
        e, d = self._play(n_games)
        return d["states"], d["pis"], d["zs"]

    def _extend_buffer(self, st, pi, z):
        """x8 dihedral images on the device (gz_equi_data), then data_buffer.extend in the reference's order"""
        if int(z.shape[0]) == 0:
            return
        e = self._engine if self._engine is not None else self.policy_value_net._engine
        st, pi, z = e.equi_data(st, pi, z)
        if hasattr(st, "data_ptr"):
            if isinstance(self.data_buffer, deque):
                self.data_buffer = DeviceReplayBuffer(self.buffer_size, self.board_width, self.board_height, e.device)
            self.data_buffer.extend(st, pi, z)
        else:
            st, pi = st.astype(np.float64), pi.astype(np.float64)
            self.data_buffer.extend((st[i], pi[i], float(z[i])) for i in range(len(z)))

    def collect_selfplay_data_sharded(self, n_games, group=None):
        """collect_selfplay_data over several GPUs the way SURVEY 8e lays it out: every rank plays its shard of the
        n_games (train.shard_games), the raw samples are all-gathered over NCCL on the device (counts, then padded
        records) and the trainer rank (0) augments the union and extends ITS buffer in (rank, game, ply) order - the
        order train.py:155 would have produced playing the games one after another.  Needs device_buffer=True."""
        import torch.distributed as dist
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        rank = dist.get_rank(group) if dist.is_initialized() else 0
        lo, hi = shard_games(n_games, rank, world)
        st, pi, z = self._raw_samples(hi - lo)
        import time
        import torch
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(group)  # ranks finish their shards at different times: that wait is self-play, not all-gather
        t0 = time.perf_counter()
        local_rows = int(z.shape[0])
        st, pi, z = all_gather_samples_device(st, pi, z, group)
        torch.cuda.synchronize()
        self.last_allgather = {"seconds": time.perf_counter() - t0, "local_samples": local_rows,
                               "union_samples": int(z.shape[0]), "bytes_per_sample": 4 * (1 + 5 * int(pi.shape[1]))}
        self.episode_len = int(z.shape[0])
        if rank == 0:
            self._extend_buffer(st, pi, z)

    def _collect_from_train_data(self, n_games, train_data):
        """train.py:137-159, recorded-game branch: the first n_games entries of train_data ([winner,
        moves] each) become samples through one batched rules call, then the usual x8 augmentation."""
        e = self.policy_value_net._engine  # stateless rules call: any engine of this board size
        games = samples_from_train_data(e, train_data[:int(n_games)])
        flat = [s for g in games for s in g]
        self.episode_len = len(flat)
        if not flat:
            return
        st, pi, z = e.equi_data(np.array([s[0] for s in flat], np.float32), np.array([s[1] for s in flat], np.float32),
                                np.array([s[2] for s in flat], np.float32))
        if self.device_buffer:
            import torch
            if isinstance(self.data_buffer, deque):
                self.data_buffer = DeviceReplayBuffer(self.buffer_size, self.board_width, self.board_height, e.device)
            dev = self.data_buffer.zs.device
            self.data_buffer.extend(torch.from_numpy(st).to(dev), torch.from_numpy(pi).to(dev), torch.from_numpy(z).to(dev))
        else:
            st, pi = st.astype(np.float64), pi.astype(np.float64)
            self.data_buffer.extend((st[i], pi[i], float(z[i])) for i in range(len(z)))

    def _collect_on_device(self, n_games):
        """Batch mode with every array on the device: the finished games' samples are drained into torch tensors
        in the engine's output order (games in finishing order, plies ascending inside a game)."""
        import torch
        e = self._selfplay_engine(int(n_games))
        e.selfplay_begin(temp=float(self.temp), recycle=False)
        for _ in range(self.board_width * self.board_height):
            e.selfplay_step(1)
            if e.selfplay_pending()[2] == 0:
                break
        n = int(e.selfplay_pending()[0])
        dev = torch.device("cuda", e.device)
        bufs = {"states": torch.empty((max(n, 1), 4, self.board_width, self.board_height), dtype=torch.float32, device=dev),
                "pis": torch.empty((max(n, 1), self.board_width * self.board_height), dtype=torch.float32, device=dev),
                "zs": torch.empty((max(n, 1),), dtype=torch.float32, device=dev),
                "game_ids": torch.empty((max(n, 1),), dtype=torch.int64, device=dev),
                "plies": torch.empty((max(n, 1),), dtype=torch.int32, device=dev)}
        d = e.selfplay_drain(n, bufs)
        self.last_games = e.selfplay_drain_games()
        self._games_played += len(self.last_games)
        return d["states"], d["pis"], d["zs"]

    def _collect_continuous(self, n_games):
        """collect_selfplay_data without the tail: train.py:137-159 plays its n games one after another, the batch
        mode here plays them side by side and waits for the longest - at the end of every batch most slots idle.
        Here the engine's slots never idle (gz_selfplay_begin(recycle=2) once; a slot that finishes starts the next
        game at once, with the weights current at that moment) and games in flight survive across calls and across
        weight updates.  Every game carries a START serial; a call hands out the n games with the next n serials,
        in that order, as soon as all of them have finished - every game exactly once, in the order the reference's
        loop would have started them, whatever its length: taking the first n games to FINISH instead would
        over-represent short games.  Finished games with later serials wait (on the device when device_buffer) for
        a later call.  Returns the released games' raw samples in (serial, ply) order."""
        n_games = int(n_games)
        slots = int(self.selfplay_slots or n_games)
        e = self._selfplay_engine(slots)
        S = self.board_width * self.board_height
        if self._cont is None or self._cont["engine"] is not e:
            e.selfplay_begin(temp=float(self.temp), recycle=2)
            first = e.selfplay_first_serial()
            self._cont = {"engine": e, "next": first, "done": {}, "chunks": [], "plies": 0}
        c = self._cont
        lo, hi = c["next"], c["next"] + n_games
        on_device = self.device_buffer
        if on_device:
            import torch
            dev = torch.device("cuda", e.device)

        def have():
            return sum(1 for g in range(lo, hi) if g in c["done"])

        def all_ready():
            ready = have() >= n_games
            if not self.sync_plies:
                return ready
            import torch
            import torch.distributed as dist
            if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
                return ready
            flag = torch.tensor([1 if ready else 0], dtype=torch.int32,
                                device=torch.device("cuda", e.device) if dist.get_backend() == "nccl" else "cpu")
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            return bool(flag.item())
        while not all_ready():
            e.selfplay_step(1)
            c["plies"] += 1
            n = int(e.selfplay_pending()[0])
            if n == 0:
                continue
            if on_device:
                bufs = {"states": torch.empty((n, 4, self.board_width, self.board_height), dtype=torch.float32, device=dev),
                        "pis": torch.empty((n, S), dtype=torch.float32, device=dev),
                        "zs": torch.empty((n,), dtype=torch.float32, device=dev),
                        "game_ids": torch.empty((n,), dtype=torch.int64, device=dev),
                        "plies": torch.empty((n,), dtype=torch.int32, device=dev)}
                d = e.selfplay_drain(n, bufs)
            else:
                d = e.selfplay_drain(n)
            c["chunks"].append(d)
            for gid, winner, moves in e.selfplay_drain_games():
                c["done"][gid - e.game_id_base] = (gid, winner, moves)
            if e.selfplay_stats()["pool_overflows"] or e.selfplay_dropped():
                raise MemoryError("continuous self-play: device sample ring or tree pool exhausted")
        # release serials [lo, hi) in order; later games stay pending
        self.last_games = [c["done"].pop(g) for g in range(lo, hi)]
        c["next"] = hi
        for gid, winner, moves in self.last_games:
            _logger.warning("{'steps':[%s],'winner':%d}" % (",".join(map(str, moves)), winner))  # game.py:411
        self._games_played += n_games
        lo_id, hi_id = lo + e.game_id_base, hi + e.game_id_base
        if on_device:
            cat = {k: torch.cat([d[k] for d in c["chunks"]]) for k in c["chunks"][0]}
            sel = (cat["game_ids"] >= lo_id) & (cat["game_ids"] < hi_id)
            keep = ~sel
            c["chunks"] = [{k: v[keep] for k, v in cat.items()}] if bool(keep.any()) else []
            rel = {k: v[sel] for k, v in cat.items()}
            order = torch.argsort(rel["game_ids"] * 1024 + rel["plies"].to(torch.int64))
            return rel["states"][order].contiguous(), rel["pis"][order].contiguous(), rel["zs"][order].contiguous()
        else:
            cat = {k: np.concatenate([d[k] for d in c["chunks"]]) for k in c["chunks"][0]}
            sel = (cat["game_ids"] >= lo_id) & (cat["game_ids"] < hi_id)
            c["chunks"] = [{k: v[~sel] for k, v in cat.items()}] if (~sel).any() else []
            rel = {k: v[sel] for k, v in cat.items()}
            order = np.lexsort((rel["plies"], rel["game_ids"]))
            return rel["states"][order], rel["pis"][order], rel["zs"][order]

    def _minibatch(self, k):
        """(states, pis, zs, zs as numpy): random.sample of the buffer (train.py:167)."""
        if isinstance(self.data_buffer, DeviceReplayBuffer):
            st, pi, z = self.data_buffer.sample(k)
            return st, pi, z, z.cpu().numpy().astype(np.float64)
        mini_batch = random.sample(self.data_buffer, k)
        z = [data[2] for data in mini_batch]
        return [data[0] for data in mini_batch], [data[1] for data in mini_batch], z, np.array(z)

    def broadcast_initial_weights(self, src=0, group=None):
        """Every rank starts from rank `src`'s weights (one flat broadcast); after that the data-parallel
        updates keep them bit-identical without further weight traffic."""
        self.policy_value_net.set_policy_param(broadcast_weights(self.policy_value_net.get_policy_param(), src, group))

    def policy_update_dp(self, group=None):
        """train.py:161-214 as a data-parallel update: every rank draws batch_size / world
        samples from ITS OWN buffer, runs forward + backward on the device, the flat gradients are
        averaged with one all-reduce (NCCL over NVLink) and every rank applies the same Adam step, so
        the weights stay identical without a broadcast and samples never leave their GPU.  The KL used
        for the early stop and the lr schedule is the mean over ranks."""
        import torch
        import torch.distributed as dist
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        local_bs = max(1, self.batch_size // world)
        state_batch, mcts_probs_batch, winner_batch, _ = self._minibatch(local_bs)
        net = self.policy_value_net
        old_probs, old_v = net.policy_value_train(state_batch)
        nccl = dist.is_initialized() and dist.get_backend(group) == "nccl"
        grads = torch.empty(net._engine.train_param_count(), dtype=torch.float32,
                            device=torch.device("cuda", net._engine.device)) if nccl else None
        for i in range(self.epochs):
            loss, entropy = net.train_backward(state_batch, mcts_probs_batch, winner_batch)
            grads = allreduce_mean(net.train_grads(grads), group)
            net.train_apply(self.learn_rate * self.lr_multiplier, grads)
            new_probs, new_v = net.policy_value_train(state_batch)
            kl = np.mean(np.sum(old_probs * (np.log(old_probs + 1e-10) - np.log(new_probs + 1e-10)), axis=1))
            stats = allreduce_mean(np.array([kl, loss, entropy], np.float32), group)
            kl, loss, entropy = float(stats[0]), float(stats[1]), float(stats[2])
            if kl > self.kl_targ * 4:
                break
        if kl > self.kl_targ * 2 and self.lr_multiplier > 0.05:
            self.lr_multiplier /= 1.5
        elif kl < self.kl_targ / 2 and self.lr_multiplier < 10:
            self.lr_multiplier *= 1.5
        _logger.info("kl=%4.3f, lr_multiplier=%4.3f, loss=%4.3f, entropy=%4.3f (data parallel, %d ranks)"
                     % (kl, self.lr_multiplier, loss, entropy, world))
        return loss, entropy

    def policy_update(self):
        """train.py:161-214."""
        state_batch, mcts_probs_batch, winner_batch, z_host = self._minibatch(self.batch_size)
        # both sides of the KL come from the fp32 training path (gz_train_forward): the tcgen05
        # inference images are refreshed once, when self-play next needs the weights
        old_probs, old_v = self.policy_value_net.policy_value_train(state_batch)
        for i in range(self.epochs):
            loss, entropy = self.policy_value_net.train_step(state_batch, mcts_probs_batch, winner_batch,
                                                             self.learn_rate * self.lr_multiplier)
            new_probs, new_v = self.policy_value_net.policy_value_train(state_batch)
            kl = np.mean(np.sum(old_probs * (np.log(old_probs + 1e-10) - np.log(new_probs + 1e-10)), axis=1))
            if kl > self.kl_targ * 4:
                break
        if kl > self.kl_targ * 2 and self.lr_multiplier > 0.05:
            self.lr_multiplier /= 1.5
        elif kl < self.kl_targ / 2 and self.lr_multiplier < 10:
            self.lr_multiplier *= 1.5
        explained_var_old = 1 - np.var(z_host - old_v.flatten()) / np.var(z_host)
        explained_var_new = 1 - np.var(z_host - new_v.flatten()) / np.var(z_host)
        _logger.info("kl=%4.3f, lr_multiplier=%4.3f, loss=%4.3f, entropy=%4.3f, explained_var_old=%4.3f, "
                     "explained_var_new=%4.3f" % (kl, self.lr_multiplier, loss, entropy, explained_var_old,
                                                  explained_var_new))
        return loss, entropy

    def policy_evaluate(self, n_games=10, player=0):
        """train.py:216-276: the current policy (player 1) against pure MCTS with
        pure_mcts_playout_num rollouts (player 2), first move alternating; all n_games are played
        concurrently on the device.  Returns (wins + 0.5 ties) / n_games."""
        if player != 0:
            raise NotImplementedError("the Yixin engine bridge is outside the hot path (SURVEY 2)")
        n_games = int(n_games)
        if self._arena is None or self._arena.G != n_games:
            if self._arena is not None:
                self._arena.close()
            self._arena = Engine(width=self.board_width, height=self.board_height, n_in_row=self.n_in_row,
                                 n_games=n_games, n_playout=max(self.n_playout, self.pure_mcts_playout_num),
                                 c_puct=float(self.c_puct), seed=self.seed + 7919 + _rank(), device=self.device)
        self._arena.set_weights(self.policy_value_net.get_policy_param())
        self._arena_rounds += 1
        winners = arena_play(self._arena, self.n_playout, self.pure_mcts_playout_num,
                             start_players=np.arange(n_games) % 2, search_salt=1000 * self._arena_rounds)
        win_cnt = {w: int((winners == w).sum()) for w in (1, 2, -1)}
        for i, w in enumerate(winners):
            _logger.info("%d-%d : winner is player%d (start from player%d)" % (n_games, i + 1, w, i % 2 + 1))
        win_ratio = 1.0 * (win_cnt[1] + 0.5 * win_cnt[-1]) / n_games
        _logger.info("Win: %d, Lose: %d, Tie: %d" % (win_cnt[1], win_cnt[2], win_cnt[-1]))
        return win_ratio

    def run(self, eval_games=10):
        """train.py:278-350 (the self-play branch): collect -> update -> every check_freq batches
        save, evaluate against pure MCTS, keep the best model and raise the opponent's playouts."""
        for i in range(self.game_batch_num):
            self.collect_selfplay_data(n_games=self.play_batch_size, train_data=[])
            _logger.info("batch %d-%d, episode_len=%02d: " % (self.game_batch_num, i + 1, self.episode_len))
            if len(self.data_buffer) > self.batch_size:
                self.policy_update()
            if (i + 1) % self.check_freq == 0:
                _logger.info("current self-play batch: %d" % (i + 1))
                self.policy_value_net.save_model('./current_policy.model')
                win_ratio = self.policy_evaluate(n_games=eval_games)
                if win_ratio > self.best_win_ratio:
                    self.policy_value_net.save_model(self.best_model)
                    _logger.info("New best policy(%3.2f>%3.2f), save to %s" % (win_ratio, self.best_win_ratio, self.best_model))
                    self.best_win_ratio = win_ratio
                    if self.best_win_ratio == 1.0 and self.pure_mcts_playout_num < 5000:
                        self.pure_mcts_playout_num += 1000
                        self.best_win_ratio = 0.0
                        _logger.info("pure_mcts_playout increase to %d" % self.pure_mcts_playout_num)
        self.policy_value_net.save_model('./current_policy.model')


def arena_play(engine, n_playout, pure_playouts, start_players, temp=1e-3, search_salt=0, leaves_in_flight=1):
    """Batched Game.start_play (game.py:211-290) of MCTSPlayer (player 1, net-guided, is_selfplay=0)
    against mcts_pure.MCTSPlayer (player 2, rollouts) in every slot of `engine` at once.

    Both reference players search from a FRESH tree every move (update_with_move(-1),
    mcts_alphaZero.py:255, mcts_pure.py:268), so one engine serves both: on every ply the slots
    where player 1 is to move run the net search and the others the rollout search.  The AZ move
    is sampled from softmax(log(visits)/temp) (np.random.choice(acts, p=probs), :247; the 15x15
    first-move override :251-253 is applied), the pure move is the most visited child (:230).
    Returns winners int8[G] (1, 2 or -1 = tie)."""
    G, S = engine.G, engine.S
    engine.set_active(None)
    engine.boards_reset(np.asarray(start_players, np.int8))
    winners = np.zeros(G, np.int8)
    done = np.zeros(G, bool)
    for ply in range(S + 1):
        st = engine.boards_status()
        ended = st["ended"].astype(bool)
        winners[ended & ~done] = st["winner"][ended & ~done]
        done |= ended
        if done.all():
            break
        moves = np.full(G, -2, np.int32)
        for who, mask in ((1, ~done & (st["cur_player"] == 1)), (2, ~done & (st["cur_player"] == 2))):
            if not mask.any():
                continue
            engine.set_active(mask)
            engine.update_with_move(np.where(mask, -1, -2).astype(np.int32))  # fresh trees for the movers
            if who == 1:
                engine.search(n_playout, leaves_in_flight=leaves_in_flight)
                mv, _ = engine.root_select_move(temp=temp, selfplay=False)
                if engine.width == 15 and engine.height == 15:  # mcts_alphaZero.py:251-253
                    first = mask & (st["n_stones"] == 0)
                    mv = np.where(first, np.random.choice([96, 97, 98, 111, 112, 113, 126, 127, 128], size=G), mv)
            else:
                engine.pure_search(pure_playouts, search_id=search_salt + ply)
                mv = engine.pure_best_move()
            moves[mask] = mv[mask]
        engine.set_active(None)
        engine.boards_do_move(moves)
    engine.update_with_move(np.full(G, -1, np.int32))
    return winners


def _rank():
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank()
    except Exception:
        pass
    return 0
