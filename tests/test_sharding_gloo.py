"""The N>1 path on CPU: world_size-2 gloo run of the row sharding (sharding.shard_rows / owner: what the pipeline stage
and tools/stream_config5.py deal rows with) and of the timing aggregation bench.py computes `value` with
(sharding.job_throughput)."""
import os
import socket
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "oracle")):          # spawned ranks import this module without conftest.py
    if _p not in sys.path:
        sys.path.insert(0, _p)
import pgen_pkg  # noqa: E402

pgen_pkg.load()

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import pgen_oracle as O
from textsummarization_on_flink_b200 import sharding, synth


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    V, H, E = 120, 32, 16
    W = synth.make_weights("W1", vsize=V, hidden=H, emb=E, seed=3)
    hps = O.parse_hyperparams("p --mode=decode --beam_size=2 --max_dec_steps=6 --min_dec_steps=2 --coverage=1")
    arts = synth.make_article_ids(7, vsize=V, seed=9, max_len=10, oov_rate=0.2)     # 7 rows over 2 ranks: 4 + 3
    mine = sharding.shard_rows(arts, rank, world)
    toks = [O.decode_ids(W, hps, a, e, k, mode="dedup")[0] for a, e, k in mine]
    # every rank pretends to have needed (rank + 1) seconds: the job rate uses the slowest rank
    rate = sharding.job_throughput(len(mine), float(rank + 1), dist)
    gathered = [None] * world
    dist.all_gather_object(gathered, toks)
    if rank == 0:
        np.save(os.path.join(out_dir, "rate.npy"), np.array([rate]))
        merged = sharding.unshard_rows(gathered)
        ref = [O.decode_ids(W, hps, a, e, k, mode="dedup")[0] for a, e, k in arts]
        assert merged == ref
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_matches_single_process(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    rate = float(np.load(tmp_path / "rate.npy")[0])
    assert abs(rate - 7 / 2.0) < 1e-9            # (4 + 3) rows / max(1 s, 2 s)


def test_shard_round_trip():
    rows = list(range(11))
    for world in (1, 2, 3, 8):
        shards = [sharding.shard_rows(rows, r, world) for r in range(world)]
        assert sorted(sum(shards, [])) == rows
        assert sharding.unshard_rows(shards) == rows
        assert max(len(s) for s in shards) - min(len(s) for s in shards) <= 1
    assert sharding.job_throughput(10, 2.0) == 5.0


def _wait_worker(rank, world, port, out_dir):
    import time
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    t0 = time.time()
    if rank == 0:
        sharding.wait_for_rank0(dist, "leg")            # not done yet: returns at once on rank 0
        time.sleep(1.0)                                 # "the CPU leg"
        sharding.wait_for_rank0(dist, "leg", done=True)
    else:
        sharding.wait_for_rank0(dist, "leg")            # sleeps on the store's socket until rank 0 is done
        np.save(os.path.join(out_dir, "waited.npy"), np.array([time.time() - t0]))
    dist.barrier()
    dist.destroy_process_group()


def test_other_ranks_sleep_through_rank0s_cpu_leg(tmp_path):
    """bench.py: ranks != 0 wait for rank 0's CPU-baseline leg on the store (no collective, no busy core)."""
    mp.spawn(_wait_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    assert float(np.load(tmp_path / "waited.npy")[0]) >= 0.9
