"""Generates tests/golden/fullsize_hashes.json: SHA-256 of the heard-pair lists (tx, rx, outcome) the ORACLE
produces for BASELINE's full-size configurations (tools/workloads.py: C2, C3, C4 random, C4 grid, C5 ticks 1-3).

    python tests/golden/make_fullsize_hashes.py [threads]

The lists are too large to commit (C4: 1.4 M pairs), a digest is not.  Digest = sha256 over
int32 tx[] | int32 rx[] | uint8 outcome[] (little endian, list order = packets in arrival order, receivers in
node-array order).  tests/test_gpu_fullsize.py and `bench.py --verify` recompute it from the GPU's lists; the
deployments come from java.util.Random, so a JVM harness (java/tools/GenGolden.java) can produce the same
digests from the reference's side once a SINR medium exists there."""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tools")]
import oracle_py as orc          # noqa: E402
import workloads as W            # noqa: E402


def digest(tx, rx, outcome):
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(tx, dtype="<i4").tobytes())
    h.update(np.ascontiguousarray(rx, dtype="<i4").tobytes())
    h.update(np.ascontiguousarray(outcome, dtype="u1").tobytes())
    return h.hexdigest()


def packets_for(dep, src):
    """C5: a packet goes out on its source's channel (RadioPacket.java:49-50 snapshot)."""
    return orc.make_packets(src, 0.0, dep["channel"][src])


def entry(ref):
    p = ref["pairs"]
    assert ref["status"] == 0
    return dict(n_pairs=int(ref["n_pairs"]), n_decisions=int(ref["n_decisions"]),
                n_delivered=int((p["outcome"] == 2).sum()), sha256=digest(p["tx"], p["rx"], p["outcome"]))


def main():
    threads = int(sys.argv[1]) if len(sys.argv) > 1 else os.cpu_count()
    out = {}
    for name, grid in (("c2", False), ("c3", False), ("c4", False), ("c4", True)):
        dep = W.deployment(name, grid=grid)
        nodes = orc.NodeSet(dep["x"], dep["y"], channel=dep["channel"])
        ref = orc.tick(orc.params(orc.MODEL_SINR), nodes, packets_for(dep, dep["src"]), threads=threads, cap=6_000_000)
        out[name + ("-grid" if grid else "")] = entry(ref)
        print(name, grid, out[name + ("-grid" if grid else "")], flush=True)
    dep = W.deployment("c5")
    for k in (1, 2, 3):
        src = W.c5_tick(dep, k)
        nodes = orc.NodeSet(dep["x"], dep["y"], channel=dep["channel"])
        ref = orc.tick(orc.params(orc.MODEL_SINR), nodes, packets_for(dep, src), threads=threads, cap=6_000_000)
        out[f"c5-tick{k}"] = entry(ref)
        print("c5", k, out[f"c5-tick{k}"], flush=True)
    with open(os.path.join(HERE, "fullsize_hashes.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
        f.write("\n")


if __name__ == "__main__":
    main()
