"""BASELINE config C5: the train.py loop (collect self-play data -> policy update) at 9x9 / 5-in-row,
n_playout 600 (conf/train_config.yaml:20), one process per GPU.  Per iteration every rank plays
--games concurrent games into its device replay buffer, then all ranks run ONE data-parallel update
(batch 512 split over ranks, up to `epochs` Adam steps, gradients averaged with one NCCL all-reduce per
step).  Prints one JSON line: games/hour (whole job), playouts/s, seconds per update.
    python tools/bench_c5.py [--games 4096] [--iters 2]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/bench_c5.py"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from alphazero_gomoku_b200.train import TrainPipeline  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--games", type=int, default=4096, help="concurrent games per GPU per iteration")
    ap.add_argument("--iters", type=int, default=2)
    ap.add_argument("--board", type=int, default=9)
    ap.add_argument("--playouts", type=int, default=600)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    conf = {"board_width": a.board, "board_height": a.board, "n_in_row": 5, "n_playout": a.playouts,
            "buffer_size": 10000, "batch_size": 512, "epochs": 5}
    tp = TrainPipeline(conf, seed=0, device_buffer=True, device=local)
    tp.broadcast_initial_weights()
    t_play, t_upd, games, plies, steps = [], [], 0, 0, 0
    for it in range(a.iters + 1):  # iteration 0 is the warm-up (allocations, first-touch)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        tp.collect_selfplay_data(a.games)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        if world > 1:
            dist.barrier()  # rank skew of the self-play phase is not update time
        torch.cuda.synchronize()
        t1b = time.perf_counter()
        loss, ent = tp.policy_update_dp() if world > 1 else tp.policy_update()
        w = tp.policy_value_net.get_policy_param()  # refresh point: self-play engines pick these up
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        if it == 0:
            continue
        t_play.append(t1 - t0)
        t_upd.append(t2 - t1b)
        games += len(tp.last_games)
        plies += tp.episode_len
    stat = torch.tensor([sum(t_play), sum(t_upd), float(games), float(plies)], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = stat.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(stat, op=dist.ReduceOp.SUM)
        t_p, t_u = float(mx[0]), float(mx[1])
        # identical weights on every rank after the data-parallel updates
        flat = torch.from_numpy(np.concatenate([x.ravel() for x in w])).cuda()
        ref = flat.clone()
        dist.broadcast(ref, 0)
        same = torch.tensor([float(torch.equal(flat, ref))], device="cuda")
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
    else:
        t_p, t_u, same = float(stat[0]), float(stat[1]), torch.ones(1)
    if rank == 0:
        g, p = float(stat[2]), float(stat[3])
        print(json.dumps({
            "config": "C5: train loop %dx%d/5, n_playout %d, %d games/GPU/iteration, batch 512, epochs<=5" % (a.board, a.board, a.playouts, a.games),
            "n_gpus": world, "iters": a.iters, "games": g, "mean_game_len": p / max(g, 1),
            "selfplay_s_per_iter": t_p / a.iters, "update_s_per_iter": t_u / a.iters,
            "games_per_hour": g / (t_p + t_u) * 3600.0, "games_per_hour_selfplay_only": g / t_p * 3600.0,
            "playouts_per_sec_selfplay": p * a.playouts / t_p, "loss": loss, "entropy": ent,
            "weights_identical_across_ranks": bool(same.item() == 1.0), "lr_multiplier": tp.lr_multiplier}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
