#!/usr/bin/env python
"""bench.py — MCTS playouts/s of the self-play hot path (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            our engine (one rank per GPU)
  python bench.py --impl reference --gpus N --steps K ...  CPU arm: the oracle port of the
                                                           reference's CPU path, all host cores

A "step" is one self-play ply in every resident game: n_playout waves of
(select -> net on the gathered leaves -> expand + backup), then move choice, record, re-root.
Workload at N=1 = BASELINE.json configs[2] per-GPU share: AlphaZero MCTS 8x8 five-in-a-row,
400 playouts/move, 16384 concurrent games per GPU, shipped 8x8 weights (tests/golden).
Prints ONE JSON line (see the task contract for the keys).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

_JSON_LINE = []  # the one line main() prints to the real stdout at the end
METRIC = "mcts_playouts_per_sec"
UNIT = "playouts/s"


def net_flops(S, trunk=(32, 64, 128)):
    """Algorithmic FLOPs (2*MAC, dense taps) per evaluated leaf, SURVEY Appendix D."""
    c1, c2, c3 = trunk
    conv = {"conv1": 2 * 9 * 4 * c1 * S, "conv2": 2 * 9 * c1 * c2 * S, "conv3": 2 * 9 * c2 * c3 * S,
            "heads": 2 * (c3 * 6 * S + 4 * S * S + 2 * S * 64 + 64)}
    conv["total"] = sum(conv.values())
    # fused tensor-core trunk = conv1..3 + the two 1x1 head convs; heads_fc = the three FCs
    conv["trunk_tc"] = conv["conv1"] + conv["conv2"] + conv["conv3"] + 2 * c3 * 6 * S
    conv["heads_fc"] = 2 * (4 * S * S + 2 * S * 64 + 64)
    return conv


def load_weights(width, trunk=(32, 64, 128)):
    if tuple(trunk) != (32, 64, 128):  # no shipped weights for wider trunks: seeded He init
        from alphazero_gomoku_b200.engine import weight_shapes
        rng = np.random.RandomState(0)
        arrs = [(rng.randn(*s) * (np.sqrt(2.0 / int(np.prod(s[1:]))) if len(s) > 1 else 0.1)).astype(np.float32)
                for s in weight_shapes(width, width, tuple(trunk))]
        return arrs, "seeded random init (no shipped weights for trunk %s)" % (tuple(trunk),)
    path = os.path.join(ROOT, "tests", "golden", "weights_%d_%d_%d.npz" % (width, width, 5 if width >= 8 else 4))
    if os.path.exists(path):
        z = np.load(path)
        return [z["p%02d" % i] for i in range(16)], os.path.basename(path)
    from alphazero_gomoku_b200.engine import weight_shapes
    rng = np.random.RandomState(0)
    arrs = []
    for s in weight_shapes(width, width):
        fan = int(np.prod(s[1:])) if len(s) > 1 else 1
        arrs.append((rng.randn(*s) * (np.sqrt(2.0 / fan) if len(s) > 1 else 0.1)).astype(np.float32))
    return arrs, "random-init (seed 0)"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d, "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.nvml = None

    def _nvml_start(self):
        """NVML polled every 10 ms from a thread (the main thread sits in cudaStreamSynchronize with the GIL
        released): a short timed region still gets samples, which nvidia-smi's 200 ms loop does not."""
        import pynvml as nv
        import torch
        nv.nvmlInit()
        try:
            uuid = str(torch.cuda.get_device_properties(self.index).uuid)
            h = nv.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        except Exception:
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
        reasons_fn = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            nv.nvmlDeviceGetCurrentClocksThrottleReasons
        mx = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
        st = {"sm": [], "mask": 0, "stop": False, "max": mx}

        def loop():
            while not st["stop"]:
                try:
                    st["sm"].append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                    st["mask"] |= int(reasons_fn(h))
                except Exception:
                    pass
                time.sleep(0.01)
        st["thread"] = threading.Thread(target=loop, daemon=True)
        st["thread"].start()
        self.nvml = st

    def _nvml_stop(self):
        st = self.nvml
        st["stop"] = True
        st["thread"].join(timeout=2)
        # nvml.h: SwPowerCap 0x4, HwSlowdown 0x8, SwThermalSlowdown 0x20, HwThermalSlowdown 0x40
        bits = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}
        return {"sm_mhz": float(np.median(st["sm"])) if st["sm"] else None, "sm_max_mhz": st["max"],
                "samples": len(st["sm"]), "reasons": sorted(k for k, b in bits.items() if st["mask"] & b),
                "source": "nvml, 10 ms period"}

    def start(self):
        try:
            self._nvml_start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.nvml is not None:
            return self._nvml_stop()
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---- CPU legs (oracle port of the reference's CPU path) ----------------------------------------
class CpuSelfplayWorker:
    """One single-threaded self-play game on the oracle port that LIVES ACROSS STEPS: board, tree (subtree reuse,
    mcts_alphaZero.py:183-192) and rng persist, a finished game is followed by a new one (game.py:360-416), so a
    step of the CPU arm is what a step of the GPU arm is - one more ply of games that are already under way."""

    def __init__(self, arrs, width, n_in_row, n_playout, seed):
        from oracle import orc
        self.orc = orc
        self.net = orc.Net(width, width, arrs)
        self.b = orc.Board(width, width, n_in_row)
        self.b.init_board(0)
        self.n_playout = n_playout
        self.m = orc.MCTS(5, n_playout, orc.EVAL_NET, net=self.net)
        self.rng = np.random.RandomState(seed)
        self.playouts = 0
        self.plies = 0

    def ply(self, n_playout=None):
        n = self.n_playout if n_playout is None else n_playout
        acts, vis, _, _ = self.m.search(self.b, n)  # the C call releases the GIL
        self.playouts += n
        self.plies += 1
        p = vis / vis.sum()
        mv = int(acts[self.rng.choice(len(acts), p=p)])
        self.m.update_with_move(mv)
        self.b.do_move(mv)
        if self.b.game_end()[0]:
            self.b.init_board(0)
            self.m.update_with_move(-1)

    def prefill(self, n_plies):
        """untimed: bring the game to a mixed phase with cheap searches (the GPU arm's --prefill-plies)"""
        for _ in range(n_plies):
            self.ply(16)
        self.playouts = 0
        self.plies = 0


def run_workers(workers, n_plies=1, budget_s=None):
    """every worker plays n_plies on its own thread; with budget_s a worker stops after the ply that passes it"""
    def work(w):
        t0 = time.perf_counter()
        for k in range(n_plies):
            if budget_s is not None and k >= 2 and time.perf_counter() - t0 >= budget_s:
                break
            w.ply()
    ts = [threading.Thread(target=work, args=(w,)) for w in workers]
    for t in ts: t.start()
    for t in ts: t.join()


def cpu_baseline(arrs, width, n_in_row, n_playout, n_plies=64, budget_s=12.0):
    """Single-thread oracle port (reference = single-threaded Python) on a bounded sample: self-play plies of one
    game that is already under way (6 cheap prefill plies), until budget_s seconds of CPU work are done."""
    w = CpuSelfplayWorker(arrs, width, n_in_row, n_playout, 0)
    w.prefill(6)
    t0 = time.perf_counter()
    run_workers([w], n_plies, budget_s)
    dt = time.perf_counter() - t0
    return {"value": w.playouts / dt, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": "oracle/ C port of mcts_alphaZero.MCTS + Net forward, 1 thread, single-game self-play on the "
                      "%dx%d board continuing a game in progress (tree reuse, a new game when one ends), %d plies x %d "
                      "playouts, %.1f s" % (width, width, w.plies, n_playout, dt)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import orc
    orc.build()
    arrs, wname = load_weights(args.board)
    threads = os.cpu_count() or 1
    workers = [CpuSelfplayWorker(arrs, args.board, args.n_in_row, args.playouts, i) for i in range(threads)]
    # untimed: spread the games over the phases of a game, as the GPU arm's prefill does
    ts = [threading.Thread(target=w.prefill, args=((5 * i) % 17,)) for i, w in enumerate(workers)]
    for t in ts: t.start()
    for t in ts: t.join()
    for _ in range(args.warmup):
        run_workers(workers, 1)
    for w in workers:
        w.playouts = 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        run_workers(workers, 1)
    dt = time.perf_counter() - t0
    total = sum(w.playouts for w in workers)
    val = total / dt
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * dt / max(args.steps, 1),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, wname, games=args.games),
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": "bounded sample of the same workload: %d host threads, one game each, carried from "
                                       "step to step (board + tree with subtree reuse; games spread over 0-16 plies by an "
                                       "untimed prefill); each step = 1 ply x %d playouts per game (oracle/ C port of the "
                                       "reference's CPU path; the Python reference cannot travel to this box)"
                                       % (threads, args.playouts)},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    _JSON_LINE.append(json.dumps(line))


def workload_config(args, wname, games):
    return {"workload": "AlphaZero MCTS self-play %dx%d %d-in-a-row, %d playouts/move, %d concurrent games per GPU "
                        "(BASELINE.json configs[2] per-GPU share)" % (args.board, args.board, args.n_in_row,
                                                                      args.playouts, games),
            "board": args.board, "n_in_row": args.n_in_row, "n_playout": args.playouts, "games_per_gpu": games,
            "c_puct": 5, "temp": 1.0, "dirichlet": [0.1, 0.25], "weights": wname,
            "step": "one self-play ply in every game (n_playout waves + move choice + record + re-root)",
            "phase": "steady state: %d untimed prefill plies before the warm-up, so the recycled slots sit at mixed "
                     "game phases (SURVEY 8d C3)" % args.prefill_plies,
            "l2": "inputs larger than L2 (tree pool >> 126 MB)"}


# ---- GPU arm -------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from alphazero_gomoku_b200 import Engine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.Stream(device=dev)
    arrs, wname = load_weights(args.board, args.trunk)
    # weights travel the way an update would: rank 0 owns them, NCCL broadcast (NVLink)
    flat = torch.from_numpy(np.concatenate([a.reshape(-1) for a in arrs])).to(dev)
    if world > 1:
        if rank != 0:
            flat.zero_()
        dist.broadcast(flat, src=0)
    sizes = [a.size for a in arrs]
    dev_arrs = list(torch.split(flat, sizes))
    eng = Engine(width=args.board, height=args.board, n_in_row=args.n_in_row, n_games=args.games,
                 n_playout=args.playouts, seed=args.seed + rank, game_id_base=rank << 40, device=local,
                 stream=stream.cuda_stream, nodes_per_game=args.nodes_per_game, trunk=tuple(args.trunk))
    eng.set_weights(dev_arrs)
    eng.selfplay_begin(temp=1.0, recycle=True)

    def barrier():
        if world > 1:
            dist.barrier()

    # Setup (untimed): bring the slots to the steady state BASELINE's C3 measurement asks for (SURVEY 8d:
    # "every slot finishes >= 1 game ... steady state").  Games that all start together are in
    # lock-step for the first plies; after a prefill of about one and a half mean game lengths the
    # recycled slots are spread over all game phases, which is what a long self-play run looks like.
    done_prefill = 0
    while done_prefill < args.prefill_plies:
        k = min(8, args.prefill_plies - done_prefill)
        eng.selfplay_step(k)
        eng.selfplay_drain()
        eng.selfplay_drain_games()
        done_prefill += k
    for _ in range(args.warmup):
        eng.selfplay_step(1)
    eng.sync()
    eng.selfplay_drain()
    eng.selfplay_drain_games()
    eng.selfplay_stats(reset=True)
    eng.counters(reset=True)
    eng.profile_read(reset=True)
    eng.profile_enable(True, stride=args.prof_stride)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(args.steps):
        eng.selfplay_step(1)
    ev1.record(stream)
    torch.cuda.synchronize()
    barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    eng.profile_enable(False)
    prof = eng.profile_read(reset=True)
    stats = eng.selfplay_stats(reset=False)  # same device counters as counters(): read before the reset
    ctr = eng.counters(reset=True)

    # ---- end to end: host weights in, host results out, every step -------------------------------
    # N > 1: the step does what a training update does around self-play (reference train.py:137-159 + :306-320 on
    # several GPUs): rank 0's new weights go host -> device -> NCCL broadcast -> gz_set_weights on every rank, the
    # step runs, and the step's finished samples are all-gathered over NCCL (counts, then padded packed records),
    # so that every rank - the trainer among them - holds the union in (rank, game, ply) order.
    host_w = [torch.from_numpy(a.copy()).pin_memory() for a in arrs]
    flat_host = torch.from_numpy(np.concatenate([a.reshape(-1) for a in arrs])).pin_memory()
    mv = torch.zeros(args.games, dtype=torch.int32).pin_memory()
    pr = torch.zeros((args.games, args.board * args.board), dtype=torch.float32).pin_memory()
    h2d = sum(a.numel() * 4 for a in host_w)
    d2h_total = 0
    # pinned host buffers for the finished-game samples of a step (a step finishes about G plies' worth of
    # samples on average; whatever does not fit stays in the device ring for the next drain)
    cap = 6 * args.games
    S_ = args.board * args.board
    pinned = {"states": torch.zeros((cap, 4, args.board, args.board), dtype=torch.float32).pin_memory(),
              "pis": torch.zeros((cap, S_), dtype=torch.float32).pin_memory(),
              "zs": torch.zeros(cap, dtype=torch.float32).pin_memory(),
              "game_ids": torch.zeros(cap, dtype=torch.int64).pin_memory(),
              "plies": torch.zeros(cap, dtype=torch.int32).pin_memory()}
    coll = None
    if world == 1:
        # the finished-game samples of step i are drained into one of two DEVICE buffer sets and copied to the pinned
        # host buffers on a side stream underneath step i + 1 (their size is known once the step is over; only the
        # bulk copy is deferred - it is complete, and counted, before the timed region ends)
        def _dev_set():
            return {k: torch.zeros(v.shape, dtype=v.dtype, device=dev) for k, v in pinned.items()}
        dev_bufs1 = [_dev_set() for _ in range(2)]
        pinned2 = [pinned, {k: torch.zeros_like(v).pin_memory() for k, v in pinned.items()}]
        copy_stream = torch.cuda.Stream(device=dev)
        ev_dr1 = [torch.cuda.Event() for _ in range(2)]
        ev_cp1 = [torch.cuda.Event() for _ in range(2)]
    if world > 1:
        rec_w = 4 + 5 * S_  # packed record: game id (2 floats), ply, z, pi[S], state[4S]  (train.pack_samples)
        # two sets of device buffers: the all-gather of step i runs on a side stream underneath step i + 1
        # (SURVEY 8e), fixed-size (cap rows per rank, the count in row 0 of a header) so that no rank reads a
        # count on the host - a host read would make every step wait for the slowest rank twice
        dev_bufs = [{"states": torch.zeros((cap, 4, args.board, args.board), dtype=torch.float32, device=dev),
                     "pis": torch.zeros((cap, S_), dtype=torch.float32, device=dev),
                     "zs": torch.zeros(cap, dtype=torch.float32, device=dev),
                     "game_ids": torch.zeros(cap, dtype=torch.int64, device=dev),
                     "plies": torch.zeros(cap, dtype=torch.int32, device=dev)} for _ in range(2)]
        rec = [torch.zeros((cap + 1, rec_w), dtype=torch.float32, device=dev) for _ in range(2)]
        gathered = [torch.zeros((world * (cap + 1), rec_w), dtype=torch.float32, device=dev) for _ in range(2)]
        coll_stream = torch.cuda.Stream(device=dev)
        ev_b = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(args.steps + 1)]
        ev_g = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(args.steps + 1)]
        ev_drained = [torch.cuda.Event() for _ in range(2)]
        ev_gathered = [torch.cuda.Event() for _ in range(2)]
        coll = {"bytes_broadcast": int(flat.numel() * 4), "bytes_allgather": int(world * (cap + 1) * rec_w * 4)}
        with torch.cuda.stream(coll_stream):  # untimed: NCCL sets its channels up on the first collective of each kind and size
            for _ in range(2):
                dist.broadcast(flat, src=0)
                dist.all_gather_into_tensor(gathered[0], rec[0])
        coll_stream.synchronize()

    def e2e_step(it):
        nonlocal d2h_total
        if world == 1:
            eng.set_weights([a.numpy() for a in host_w])          # H2D: the step's inputs (weights)
        else:
            with torch.cuda.stream(stream):
                ev_b[it][0].record(stream)
                if rank == 0:
                    flat.copy_(flat_host, non_blocking=True)      # H2D on the trainer rank
                dist.broadcast(flat, src=0)                       # NCCL over NVLink
                ev_b[it][1].record(stream)
            stream.synchronize()
            eng.set_weights(list(torch.split(flat, sizes)))
        eng.selfplay_step(1)
        eng.selfplay_last(mv.numpy(), pr.numpy())             # D2H: (move, move_probs) of the step
        if world == 1:
            k = it & 1
            ev_cp1[k].synchronize()                               # the copy that last used this buffer set is over
            d = eng.selfplay_drain(max_samples=cap, bufs=dev_bufs1[k])  # samples of games that finished, on the device
            n = int(d["zs"].shape[0])
            ev_dr1[k].record(stream)
            with torch.cuda.stream(copy_stream):                  # D2H underneath the next step
                copy_stream.wait_event(ev_dr1[k])
                for name, t in d.items():
                    pinned2[k][name][:n].copy_(t, non_blocking=True)
                ev_cp1[k].record(copy_stream)
            d2h_total += mv.numel() * 4 + pr.numel() * 4 + sum(t.numel() * t.element_size() for t in d.values())
        else:
            k = it & 1
            ev_gathered[k].synchronize()                          # the all-gather that last used this buffer set is over
            d = eng.selfplay_drain(max_samples=cap, bufs=dev_bufs[k])  # stays on the device for the all-gather
            n = int(d["zs"].shape[0])
            ev_drained[k].record(stream)
            with torch.cuda.stream(coll_stream):
                coll_stream.wait_event(ev_drained[k])
                ev_g[it][0].record(coll_stream)
                r = rec[k]
                r[0, 0] = float(n)                                # header row: this rank's sample count
                torch.bitwise_right_shift(d["game_ids"], 20, out=tmp_i64[:n])
                r[1:n + 1, 0] = tmp_i64[:n]
                torch.bitwise_and(d["game_ids"], (1 << 20) - 1, out=tmp_i64[:n])
                r[1:n + 1, 1] = tmp_i64[:n]
                r[1:n + 1, 2] = d["plies"]
                r[1:n + 1, 3] = d["zs"]
                r[1:n + 1, 4:4 + S_] = d["pis"]
                r[1:n + 1, 4 + S_:] = d["states"].reshape(n, -1)
                dist.all_gather_into_tensor(gathered[k], r)       # rank q's rows at [q * (cap + 1), ...)
                ev_g[it][1].record(coll_stream)
                ev_gathered[k].record(coll_stream)
            # D2H of the step's result besides the moves: this rank's own finished samples' targets
            d2h_total += mv.numel() * 4 + pr.numel() * 4 + n * 4
            zs_pinned[:n].copy_(d["zs"], non_blocking=True)

    if world > 1:
        tmp_i64 = torch.zeros(cap, dtype=torch.int64, device=dev)
        zs_pinned = torch.zeros(cap, dtype=torch.float32).pin_memory()
        e2e_step(args.steps)  # one untimed e2e step: allocator pools and NCCL buffers of this code path are warm afterwards
        torch.cuda.synchronize()
        d2h_total = 0
    barrier()
    torch.cuda.synchronize()
    eng.counters(reset=True)
    t0 = time.perf_counter()
    for it in range(args.steps):
        e2e_step(it)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    barrier()
    if coll is not None:
        coll["broadcast_ms_per_step"] = sum(ev_b[i][0].elapsed_time(ev_b[i][1]) for i in range(args.steps)) / max(args.steps, 1)
        coll["allgather_ms_per_step"] = sum(ev_g[i][0].elapsed_time(ev_g[i][1]) for i in range(args.steps)) / max(args.steps, 1)
        coll["what"] = ("per step inside the e2e region: weights host -> rank 0 device -> dist.broadcast (engine stream; includes the wait "
                        "for the slowest rank of the previous step) -> gz_set_weights on every rank; after the step the finished samples "
                        "are packed and dist.all_gather_into_tensor'ed on a side stream underneath the next step (fixed-size: %d rows "
                        "per rank, count in a header row); times from CUDA events on those streams" % (cap + 1))
    stats2 = eng.selfplay_stats(reset=False)
    ctr2 = eng.counters(reset=True)

    # ---- one generation: keep playing until every slot has finished a game (bounded) -------------
    # The other half of BASELINE.json's metric (self-play games/hour) needs whole games; this
    # region is timed on its own and does not enter `value` / `e2e`.
    gen = None
    if args.generation_plies > 0:
        barrier()
        torch.cuda.synchronize()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record(stream)
        n_gen = 0
        for _ in range(args.generation_plies):
            eng.selfplay_step(1)
            n_gen += 1
            if n_gen % 8 == 0:
                eng.selfplay_drain()      # keep the device sample ring from filling up
                if eng.selfplay_stats(reset=False)["n_games"] >= args.games:
                    break
        g1.record(stream)
        torch.cuda.synchronize()
        eng.selfplay_drain()
        st3 = eng.selfplay_stats(reset=False)
        ctr3 = eng.counters(reset=True)
        gen = [g0.elapsed_time(g1), float(n_gen), float(st3["n_games"]), float(st3["n_samples"]), float(ctr3["playouts"]),
               float(ctr3["leaf_evals"]), float(st3["n_plies"]), float(st3["pool_overflows"])]

    playouts = float(ctr["playouts"])
    t = torch.tensor([ms, e2e_s * 1000.0], dtype=torch.float64, device=dev)
    s = torch.tensor([playouts, float(ctr2["playouts"]), float(ctr["leaf_evals"]), float(stats["n_plies"]),
                      float(stats["n_games"] + stats2["n_games"]), float(stats["n_samples"] + stats2["n_samples"]),
                      float(ctr["kernel_launches"]), float(stats["pool_overflows"] + stats2["pool_overflows"])],
                     dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
    ms_max, e2e_ms_max = t.tolist()
    gen_line = None
    if gen is not None:
        gt = torch.tensor(gen[:1], dtype=torch.float64, device=dev)
        gs = torch.tensor(gen[1:], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(gt, op=dist.ReduceOp.MAX)
            dist.all_reduce(gs, op=dist.ReduceOp.SUM)
        g_ms = gt.item()
        g_steps, g_games, g_samples, g_playouts, g_leaf, g_plies, g_over = gs.tolist()
        g_len = (g_samples / g_games) if g_games > 0 else None
        gen_line = {"plies_played_per_slot": int(round(g_steps / world)), "seconds": g_ms * 1e-3,
                    "games_finished": int(g_games), "mean_game_len": g_len,
                    "games_per_hour_counted": g_games / (g_ms * 1e-3) * 3600.0,
                    "games_per_hour_steady": ((g_plies / (g_ms * 1e-3)) / g_len * 3600.0) if g_len else None,
                    "playouts_per_sec": g_playouts / (g_ms * 1e-3),
                    "net_evals_per_playout": g_leaf / max(g_playouts, 1.0),
                    "pool_overflows": int(g_over), "max_tree_nodes": int(st3["max_tree_nodes"]),
                    "nodes_per_game": eng.nodes_per_game,
                    "what": "untimed for value/e2e: the same engine keeps playing (slots recycle) until every slot "
                            "has finished a game or --generation-plies; counted = games finished in this window / "
                            "time, steady = plies per second / mean game length"}
    tot_playouts, tot_playouts_e2e, tot_leaf, tot_plies, tot_games, tot_samples, tot_launch, tot_over = s.tolist()

    if rank == 0:
        S = args.board * args.board
        fl = net_flops(S, tuple(args.trunk))
        pk, pk_src = peaks()
        # dominant kernel = the class with the largest share of device time
        dom = max(prof, key=lambda k: prof[k][0])
        dom_ms, dom_n = prof[dom]
        total_prof_ms = sum(v[0] for v in prof.values())
        roof = {"bound": "tensor", "kernel": dom, "unit": "TFLOP/s", "peak": pk["bf16_tflops_sustained"],
                "peak_source": pk_src + ", sustained bf16 (kernel timed inside a long step)",
                "traffic": None}
        # DRAM bytes per launch of the dominant kernel: read from the committed ncu --set full capture
        # (profiles/r2_ncu_traffic.json, written by tools/ncu_traffic.py from a launch over a KNOWN number of leaves)
        # and scaled to the mean number of leaves a launch of this run evaluated
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")))
            cap = [v[0] for k, v in tj["kernels"].items() if k.startswith(dom)]
            if cap and args.board == 8 and tuple(args.trunk) == (32, 64, 128) and dom_n > 0:
                per_unit = (cap[0]["dram_read_bytes"] + cap[0]["dram_write_bytes"]) / tj["units_per_launch"]
                roof["traffic"] = per_unit * ctr["leaf_evals"] / dom_n
                roof["traffic_source"] = ("profiles/r2_ncu_traffic.json: %.0f B per leaf measured by ncu (dram__bytes_read + write, one launch "
                                          "over %d leaves of the search path) x the mean leaves per launch of this run; algorithmic: 4*S "
                                          "plane bits = 32 B per leaf (the kernel reads the leaf's 64 B bitboards + 16 B meta + 4 B slot map; "
                                          "the rest is the 0.4 MB weight image per launch)"
                                          % (per_unit, tj["units_per_launch"]))
        except Exception as ex:
            roof["traffic_source"] = "profiles/r2_ncu_traffic.json unreadable: %r" % (ex,)
        if dom in fl and dom_ms > 0:
            ach = ctr["leaf_evals"] * fl[dom] / (dom_ms * 1e-3) / 1e12
            roof.update({"achieved": ach, "frac": ach / pk["bf16_tflops_sustained"],
                         "issued_tflops": ach * (3.0 if dom == "trunk_tc" else 1.0),
                         "note": ("algorithmic FLOPs (2*MAC of the fp32 net, dense taps); the kernel issues 3 fp16 MMA passes "
                                  "per product for fp32-grade accuracy; at 8x8 every M row of an MMA is a board cell (dense line space, 4 boards per 256-row group)"
                                  if dom == "trunk_tc" else ""),
                         "flops_per_leaf": fl[dom], "avg_launch_ms": dom_ms / max(dom_n, 1), "launches": dom_n,
                         "share_of_device_time": dom_ms / max(total_prof_ms, 1e-9)})
        else:  # a tree/rule kernel dominates: HBM roofline with the algorithmic byte formula
            lv, ch, ex = ctr["select_levels"], ctr["select_children"], ctr["expanded_children"]
            byts = 16.0 * ch + 16.0 * (lv + ctr["playouts"]) + 20.0 * ex + ctr["leaf_evals"] * (16.0 * S + 4.0 * S + 4.0) + 128.0 * ctr["playouts"]
            ach = byts / (max(dom_ms, 1e-9) * 1e-3) / 1e9
            roof.update({"bound": "hbm", "unit": "GB/s", "peak": pk["hbm_gbs"], "achieved": ach,
                         "frac": ach / pk["hbm_gbs"], "avg_launch_ms": dom_ms / max(dom_n, 1), "launches": dom_n})
        net_ms = sum(prof[k][0] for k in ("conv1", "conv2", "conv3", "heads", "trunk_tc", "heads_fc"))
        tc = prof["trunk_tc"][1] > 0
        mean_len = (tot_samples / tot_games) if tot_games > 0 else None
        value = tot_playouts / (ms_max * 1e-3)
        plies_per_s = tot_plies / (ms_max * 1e-3)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_max / max(args.steps, 1), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32 (fp16x2-split tensor-core operands, fp32 accumulate)" if tc else "f32",
            "data": "synthetic",
            "config": workload_config(args, wname, args.games),
            "e2e": {"value": tot_playouts_e2e / (e2e_ms_max * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h_total // max(args.steps, 1),
                    "ms_per_step": e2e_ms_max / max(args.steps, 1),
                    "net_evals_per_playout": float(ctr2["leaf_evals"]) / max(float(ctr2["playouts"]), 1.0),
                    "what": ("per step: gz_set_weights from pinned host arrays, gz_selfplay_step(1), "
                             "gz_selfplay_last into pinned host buffers (synchronous), gz_selfplay_drain of the finished samples "
                             "into a device buffer whose copy to pinned host memory runs on a side stream underneath the next "
                             "step (all copies are complete before the timed region ends); " if world == 1 else
                             "per step: weights pinned host -> rank 0 -> NCCL broadcast -> gz_set_weights, gz_selfplay_step(1), "
                             "gz_selfplay_last into pinned host buffers, finished samples drained on the device and NCCL "
                             "all-gathered (see collectives); ") + "these steps follow the "
                            "device-timed ones in the same games (net_evals_per_playout says how many playouts "
                            "needed the net: terminal leaves do not)"},
            "gpu_launches": int(tot_launch),
            "collectives": coll,
            "clocks": clocks,
            "roofline": roof,
            "net": {"impl": ("tcgen05 fused trunk (fp16 hi/lo split, 3 MMA passes, fp32 accumulate in TMEM) + tcgen05 FC heads (same split)"
                             if tc else "fp32 CUDA-core (FFMA) implicit GEMM"), "net_evals_per_playout": float(ctr["leaf_evals"]) / max(float(ctr["playouts"]), 1.0), "leaf_evals": int(tot_leaf),
                    "leaf_evals_per_s": tot_leaf / (ms_max * 1e-3),
                    "algorithmic_tflops": ctr["leaf_evals"] * fl["total"] / max(net_ms * 1e-3, 1e-12) / 1e12,
                    "flops_per_leaf": fl["total"]},
            "tree": {"mean_depth": ctr["select_levels"] / max(ctr["playouts"], 1),
                     "mean_children_per_level": ctr["select_children"] / max(ctr["select_levels"], 1),
                     "pool_overflows": int(tot_over)},
            "kernel_ms": {k: round(v[0], 3) for k, v in prof.items() if v[1]},
            "selfplay": {"plies_per_s": plies_per_s, "games_finished": int(tot_games),
                         "mean_game_len": mean_len,
                         "games_per_hour": (plies_per_s / mean_len * 3600.0) if mean_len else None},
        }
        if gen_line is not None:
            line["generation"] = gen_line
        if world == 1 and not args.no_cpu_baseline:
            from oracle import orc
            orc.build()
            line["cpu_baseline"] = cpu_baseline(arrs, args.board, args.n_in_row, args.playouts, args.cpu_plies, args.cpu_seconds)
        if tot_over:
            line["invalid"] = "tree pool overflowed: raise nodes_per_game"
        _JSON_LINE.append(json.dumps(line))
    eng.close()
    if world > 1:
        dist.destroy_process_group()


# ---- BASELINE config C5: the train.py loop on N GPUs -------------------------------------------------
def run_c5(args):
    """collect self-play data -> policy update, 9x9 five-in-a-row, n_playout 600 (conf/train_config.yaml:20), batch 512,
    <= 5 epochs, one process per GPU (reference train.py:137-159, :161-214, :306-320).
      --c5-mode gather : every rank plays its shard, NCCL all-gather of the samples, rank 0 runs policy_update on the
                         union, NCCL broadcast of the new weights (the reference's single trainer)
      --c5-mode dp     : every rank keeps its samples, policy_update_dp (gradient all-reduce), no weight traffic
    Self-play is the continuous mode (slots never idle; --c5-batch = the reference-shaped batches)."""
    import torch
    import torch.distributed as dist
    from alphazero_gomoku_b200.train import TrainPipeline, broadcast_weights
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    board, playouts = (9, 600) if args.board == 8 and args.playouts == 400 else (args.board, args.playouts)
    games = args.games if args.games != 16384 else 4096
    conf = {"board_width": board, "board_height": board, "n_in_row": 5, "n_playout": playouts,
            "buffer_size": 1 << 20, "batch_size": 512, "epochs": 5}
    tp = TrainPipeline(conf, seed=args.seed, device_buffer=True, device=local, continuous=not args.c5_batch,
                       selfplay_slots=games)
    tp.sync_plies = world > 1
    tp.broadcast_initial_weights()
    t_play, t_upd, t_bc, t_ag, n_games, n_plies, ag_bytes = [], [], [], [], 0, 0, 0
    loss = ent = float("nan")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    t_wall0 = None
    for it in range(args.steps + args.warmup):
        barrier()
        t0 = time.perf_counter()
        if it == args.warmup:
            t_wall0 = t0
        if args.c5_mode == "gather":
            tp.collect_selfplay_data_sharded(games * world)
        else:
            tp.collect_selfplay_data(games)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        barrier()  # rank skew of the self-play phase is not update time
        t1b = time.perf_counter()
        bc = 0.0
        if args.c5_mode == "gather":
            if rank == 0 and len(tp.data_buffer) > tp.batch_size:
                loss, ent = tp.policy_update()
            w = tp.policy_value_net.get_policy_param()
            torch.cuda.synchronize()
            tb = time.perf_counter()
            w = broadcast_weights(w, src=0)
            if rank != 0:
                tp.policy_value_net.set_policy_param(w)
            torch.cuda.synchronize()
            bc = time.perf_counter() - tb
        else:
            loss, ent = tp.policy_update_dp() if world > 1 else tp.policy_update()
            tp.policy_value_net.get_policy_param()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        if it < args.warmup:
            continue
        t_play.append(t1 - t0); t_upd.append(t2 - t1b); t_bc.append(bc)
        if args.c5_mode == "gather":
            t_ag.append(tp.last_allgather["seconds"])
            ag_bytes += tp.last_allgather["union_samples"] * tp.last_allgather["bytes_per_sample"]
        n_games += len(tp.last_games)
        n_plies += sum(len(m) for _, _, m in tp.last_games)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    stat = torch.tensor([sum(t_play), sum(t_upd), float(n_games), float(n_plies), sum(t_bc), sum(t_ag), t_wall],
                        dtype=torch.float64, device=dev)
    if world > 1:
        mx = stat.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(stat, op=dist.ReduceOp.SUM)
        t_p, t_u, t_b, t_a, t_wall = float(mx[0]), float(mx[1]), float(mx[4]), float(mx[5]), float(mx[6])
        flat = torch.from_numpy(np.concatenate([x.ravel() for x in tp.policy_value_net.get_policy_param()])).to(dev)
        ref = flat.clone()
        dist.broadcast(ref, 0)
        same = torch.tensor([float(torch.equal(flat, ref))], device=dev)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        same = bool(same.item() == 1.0)
    else:
        t_p, t_u, t_b, t_a, same = float(stat[0]), float(stat[1]), float(stat[4]), float(stat[5]), True
    if rank == 0:
        g, pl = float(stat[2]), float(stat[3])
        k = max(args.steps, 1)
        line = {"metric": "selfplay_games_per_hour", "value": g / t_wall * 3600.0, "unit": "games/h", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * t_wall / k,
                "timing": "wall clock over the timed iterations between barriers (max over ranks): rank skew included",
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32 (fp16x2-split tensor-core operands in self-play, fp32 train step)", "data": "synthetic",
                "config": {"workload": "BASELINE config C5: train.py loop %dx%d five-in-a-row, n_playout %d, %d games per GPU "
                                       "per iteration, batch 512, <= 5 epochs; a step = collect + update" % (board, board, playouts, games),
                           "mode": args.c5_mode, "selfplay": "batch" if args.c5_batch else "continuous",
                           "weights": "torch default init, seed %d" % args.seed},
                "games_per_iteration": g / k, "mean_game_len": pl / max(g, 1),
                "selfplay_s_per_iter": t_p / k, "update_s_per_iter": t_u / k,
                "games_per_hour_selfplay_only": g / t_p * 3600.0,
                "playouts_per_sec_selfplay": pl * playouts / t_p,
                "collectives": {"allgather_s_per_iter": t_a / k, "allgather_bytes_per_iter": ag_bytes // k if ag_bytes else 0,
                                "broadcast_s_per_iter": t_b / k,
                                "what": "gather: samples all-gathered on the device inside selfplay_s, weights broadcast inside "
                                        "update_s (incl. the host staging of the 16 arrays); dp: one gradient all-reduce per epoch inside update_s"},
                "loss": loss, "entropy": ent, "weights_identical_across_ranks": same, "lr_multiplier": tp.lr_multiplier}
        _JSON_LINE.append(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--games", type=int, default=16384, help="concurrent games per GPU")
    ap.add_argument("--board", type=int, default=8)
    ap.add_argument("--n-in-row", type=int, default=5)
    ap.add_argument("--playouts", type=int, default=400)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--trunk", type=int, nargs=3, default=[32, 64, 128],
                    help="conv trunk widths (BASELINE config C4 'wider trunk': 64 128 256; seeded random weights)")
    ap.add_argument("--cpu-plies", type=int, default=64, help="cpu_baseline: at most this many plies")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="cpu_baseline: stop after this much CPU work")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--prefill-plies", type=int, default=24,
                    help="untimed setup plies before the warm-up that bring the slots to mixed game phases (0 = time the "
                         "opening plies of synchronised games)")
    ap.add_argument("--nodes-per-game", type=int, default=0, help="tree pool blocks per game (0 = engine default)")
    ap.add_argument("--generation-plies", type=int, default=48,
                    help="after the timed regions, play up to this many more plies (until every slot finished a game) "
                         "and report games/hour; 0 disables")
    ap.add_argument("--prof-stride", type=int, default=16,
                    help="CUDA-event pairs around every N-th launch of each kernel class inside the timed region")
    ap.add_argument("--config", default="c3", choices=["c3", "c5"],
                    help="c3 (default): the self-play step of BASELINE configs[2]; c5: the train.py loop (collect + update) at 9x9")
    ap.add_argument("--c5-mode", default="gather", choices=["gather", "dp"])
    ap.add_argument("--c5-batch", action="store_true", help="c5: reference-shaped self-play batches instead of the continuous mode")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line (the driver parses it): libraries that print to fd 1 (NCCL's
    # "NCCL version ..." banner, for one) are sent to stderr for the duration of the run
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        if args.impl == "reference":
            run_reference(args)
        elif args.config == "c5":
            run_c5(args)
        else:
            run_ours(args)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    if _JSON_LINE:
        print(_JSON_LINE[-1], flush=True)


if __name__ == "__main__":
    main()
