"""Multi-process path of the sequence sharding (world_size 2, gloo, CPU): every sequence is processed
exactly once, the final gather returns the same merged result on every rank, timings reduce by max."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n_seq, out_dir):
    from rmem_ocu_b200 import shard
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = shard.assign_sequences(n_seq, world, rank)
    # stand-in for a clip's result: a checksum that depends only on the sequence id
    local = {s: {"frames": 10 + s, "checksum": (s * 2654435761) % 1000003, "rank": rank} for s in mine}
    merged = shard.gather_results(local)
    t = shard.max_over_ranks(1.0 + rank)
    torch.save({"merged": merged, "t": t, "mine": mine}, os.path.join(out_dir, "r%d.pt" % rank))
    dist.destroy_process_group()


def test_two_rank_sharding_and_gather(tmp_path):
    world, n_seq = 2, 7
    mp.spawn(_worker, args=(world, _free_port(), n_seq, str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(os.path.join(str(tmp_path), "r%d.pt" % r)) for r in range(world)]
    assert res[0]["merged"] == res[1]["merged"]
    assert sorted(res[0]["merged"]) == list(range(n_seq))
    assert sorted(res[0]["mine"] + res[1]["mine"]) == list(range(n_seq))
    assert all(res[0]["merged"][s]["rank"] == s % world for s in range(n_seq))
    assert res[0]["t"] == res[1]["t"] == 2.0


def test_assign_sequences_partition():
    from rmem_ocu_b200 import shard
    for world in (1, 2, 4, 8):
        seen = sorted(s for r in range(world) for s in shard.assign_sequences(16, world, r))
        assert seen == list(range(16))
    assert shard.gather_results({3: "x"}) == {3: "x"}     # no process group: identity
