"""CPU, world_size 2 and 3 over gloo: the multi-GPU host logic (receiver shards, tx-record exchange plan,
result merge).  Each rank evaluates ITS receivers with the oracle standing in for the GPU shard
(this is a test: the oracle is the checker of the merge logic), all-gathers its list, merges, and
must reproduce the unsharded tick exactly."""
import os
import sys
import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle_py as orc
    from radio_sim_b200.sharding import shard_bounds, TickPlan, merge_results
    from helpers import random_deployment, pick_sources
    n, t = 3000, 90
    x, y, z, ch = random_deployment(31, n, 500.0, channels=2)
    src = pick_sources(31, n, t)
    power = np.random.default_rng(31).choice([-7.0, 0.0], size=t)
    bounds = shard_bounds(n, world)
    plan = TickPlan(src, bounds)
    lo, hi = bounds[rank]
    # "all-gather of tx records": every rank contributes (src, power, channel) of the packets it owns
    mine = np.stack([plan.local_src[rank].astype(np.float64), plan.local_fields(rank, power, 0.0),
                     plan.local_fields(rank, ch[src], -1).astype(np.float64)])
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    rec = np.concatenate(gathered, axis=1)
    g_src, g_pow, g_ch = rec[0].astype(np.int32), rec[1], rec[2].astype(np.int32)
    valid = np.flatnonzero(g_src >= 0)
    # the shard: all nodes are needed as potential sources, but only receivers [lo,hi) are enabled here
    en = np.zeros(n, dtype=np.uint8); en[lo:hi] = 1
    ref = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x, y, z, channel=ch, enabled=en),
                   orc.make_packets(g_src[valid], g_pow[valid], g_ch[valid]))
    p = ref["pairs"]
    part = dict(tx=valid[p["tx"]].astype(np.int32), rx=p["rx"].copy(), outcome=p["outcome"].copy(), rssi=p["rssi_dbm"].copy())
    parts = [None] * world
    dist.all_gather_object(parts, part)
    merged = merge_results(plan, parts)
    if rank == 0:
        full = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x, y, z, channel=ch), orc.make_packets(src, power, ch[src]))["pairs"]
        # sharded interference sums see the same transmitters, so everything is identical
        ok = (np.array_equal(merged["tx"], full["tx"]) and np.array_equal(merged["rx"], full["rx"]) and
              np.array_equal(merged["outcome"], full["outcome"]) and np.array_equal(merged["rssi"], full["rssi_dbm"]))
        q.put((ok, len(full), plan.counts, plan.tpad))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_tick_merges_to_the_global_order(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() + 7 * world) % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    ok, npairs, counts, tpad = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok and npairs > 50 and sum(counts) == 90 and tpad == max(counts)


def test_plan_padding_and_mapping():
    from radio_sim_b200.sharding import shard_bounds, TickPlan
    b = shard_bounds(10, 3)
    assert b == [(0, 3), (3, 6), (6, 10)]
    plan = TickPlan(np.array([9, 0, 4, 1, 8, 2]), b)
    assert plan.counts == [3, 1, 2] and plan.tpad == 3 and plan.n_records == 9
    assert plan.local_src[0].tolist() == [0, 1, 2] and plan.local_src[1].tolist() == [4, -1, -1]
    assert plan.packet_of.tolist() == [1, 3, 5, 2, -1, -1, 0, 4, -1]
    with pytest.raises(ValueError):
        TickPlan(np.array([10]), b)
