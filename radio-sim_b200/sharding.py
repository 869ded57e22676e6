"""Receiver sharding across GPUs (one process per GPU): pure host-side index arithmetic.

Receivers are block-partitioned by node index -- rank g owns [g*N/G, (g+1)*N/G) -- so that the
per-rank result lists, each in (packet, node index) order, concatenate into the global Java
call order of Simulator.generateReceptionEvents with a stable merge by packet.  The only
exchange per tick is the all-gather of the tx records: every rank contributes the packets whose
source node it owns, padded to a common length; a record's "arrival index" on the device is its
position in the gathered buffer, mapped back to the tick's packet order here.
"""
import numpy as np


def shard_bounds(n, world):
    return [(r * n // world, (r + 1) * n // world) for r in range(world)]


class TickPlan:
    """Per-tick exchange plan for packets `src` (global node indices, arrival order)."""

    def __init__(self, src, bounds):
        src = np.asarray(src, dtype=np.int32)
        self.world = len(bounds)
        owners = np.full(src.shape, -1, dtype=np.int64)
        for r, (lo, hi) in enumerate(bounds):
            owners[(src >= lo) & (src < hi)] = r
        if (owners < 0).any():
            raise ValueError("source index outside every shard")
        self.counts = [int((owners == r).sum()) for r in range(self.world)]
        self.tpad = max(self.counts) if src.size else 0
        self.n_records = self.world * self.tpad
        # gathered position -> original packet index (-1 = padding)
        self.packet_of = np.full(self.n_records, -1, dtype=np.int32)
        self.local_src, self.local_packet = [], []
        for r in range(self.world):
            mine = np.flatnonzero(owners == r).astype(np.int32)
            padded = np.full(self.tpad, -1, dtype=np.int32)
            padded[:mine.size] = src[mine]
            self.local_src.append(padded)
            self.local_packet.append(mine)
            self.packet_of[r * self.tpad: r * self.tpad + mine.size] = mine

        # original packet k -> its position in the gathered list (the order the reference's draws follow)
        self.order = np.zeros(src.size, dtype=np.int32)
        valid = np.flatnonzero(self.packet_of >= 0)
        self.order[self.packet_of[valid]] = valid

    def local_fields(self, rank, values, fill):
        """Per-packet values (power, channel) re-ordered for rank's padded local list."""
        out = np.full(self.tpad, fill, dtype=np.asarray(values).dtype)
        out[:self.local_packet[rank].size] = np.asarray(values)[self.local_packet[rank]]
        return out


def merge_results(plan, per_rank):
    """per_rank[r] = dict(tx, rx, ...) of rank r with tx = gathered record position and rx = global
    node index.  Returns the same columns in the global order: packets in arrival order, receivers in
    node order (ranks own increasing index blocks, so rank order == node order within a packet)."""
    cols = {k: np.concatenate([np.asarray(p[k]) for p in per_rank]) for k in per_rank[0]}
    tx = plan.packet_of[cols["tx"]] if cols["tx"].size else cols["tx"].astype(np.int32)
    order = np.lexsort((cols["rx"], tx))
    out = {k: v[order] for k, v in cols.items()}
    out["tx"] = tx[order]
    return out
