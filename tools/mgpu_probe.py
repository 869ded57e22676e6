"""Where does a sharded tick spend its time? (development aid; run under torchrun)
Per rank: device time of exchange+tick with the start aligned on the device (an NCCL barrier enqueued on
the stream right before the start event) and with a host-side barrier only; phases of the last tick."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import radio_sim_b200 as rsb
import bench

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
N, G = bench.N_NODES, world
x, y, src = bench.deployment()
lo, hi = rank * N // G, (rank + 1) * N // G
med = rsb.Medium("sinr", device=local, rx_base=lo, n_shards=G)
med.set_nodes(x[lo:hi], y[lo:hi])
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream); med.set_stream(stream.cuda_stream)
counts = [int(((src >= r * N // G) & (src < (r + 1) * N // G)).sum()) for r in range(G)]
tpad = max(counts)
mine = np.full(tpad, -1, dtype=np.int32); m = src[(src >= lo) & (src < hi)]; mine[:m.size] = m
rec_local = torch.zeros(tpad * 48, dtype=torch.uint8, device=dev)
med.build_tx_records(mine, rec_local.data_ptr())
blob = torch.frombuffer(bytearray(med.peer_export(G, rank, tpad)), dtype=torch.uint8).to(dev)
blobs = torch.empty(G * blob.numel(), dtype=torch.uint8, device=dev)
dist.all_gather_into_tensor(blobs, blob)
raw = blobs.cpu().numpy().tobytes()
med.peer_open([raw[i * 128:(i + 1) * 128] for i in range(G)])
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
token = torch.zeros(1, device=dev)

def step():
    med.peer_push_records(tpad, rec_local.data_ptr())
    med.tick_peers_resident()

def timeit(aligned, reps=30):
    ts = []
    for _ in range(reps):
        flush.fill_(1); torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if aligned:
            dist.all_reduce(token)           # stream-ordered: every GPU passes this point at the same device time
        e0.record(stream); step(); e1.record(stream); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return float(np.median(ts[5:]))

a, b = timeit(True), timeit(False)
tm = med.timing()
out = torch.tensor([a, b, tm["ms_prepare"] * 1e3, tm["ms_sweep"] * 1e3, tm["ms_resolve"] * 1e3, tm["ms_emit"] * 1e3, tm["ms_total"] * 1e3], device=dev)
allv = [torch.zeros_like(out) for _ in range(G)]
dist.all_gather(allv, out)
if rank == 0:
    print("per rank [us]: step(aligned start) step(host barrier) | prepare sweep resolve emit total")
    for r, v in enumerate(allv):
        print(r, " ".join(f"{float(q):8.1f}" for q in v))
med.close(); dist.barrier(); dist.destroy_process_group()
