"""Where does a sharded tick spend its time? (development aid; run under torchrun)"""
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
rec_all = torch.zeros(G * tpad * 48, dtype=torch.uint8, device=dev)
med.build_tx_records(mine, rec_local.data_ptr())

def timeit(fn, reps=30):
    ts = []
    for _ in range(reps):
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); e0.record(stream); fn(); e1.record(stream); torch.cuda.synchronize()
        ts.append((e0.elapsed_time(e1) * 1e3, (time.perf_counter() - t0) * 1e6))
    ts = ts[5:]
    return np.median([a for a, _ in ts]), np.median([b for _, b in ts])

ag = timeit(lambda: dist.all_gather_into_tensor(rec_all, rec_local))
tk = timeit(lambda: med.tick_resident(G * tpad, rec_all.data_ptr()))
both = timeit(lambda: (dist.all_gather_into_tensor(rec_all, rec_local), med.tick_resident(G * tpad, rec_all.data_ptr())))
tm = med.timing()
out = torch.tensor([ag[0], ag[1], tk[0], tk[1], both[0], both[1], tm["ms_prepare"] * 1e3, tm["ms_sweep"] * 1e3, tm["ms_resolve"] * 1e3, tm["ms_emit"] * 1e3], device=dev)
allv = [torch.zeros_like(out) for _ in range(G)]
dist.all_gather(allv, out)
if rank == 0:
    print("per rank: allgather(dev,wall) tick(dev,wall) both(dev,wall) | prepare sweep resolve emit  [us]")
    for r, v in enumerate(allv):
        print(r, " ".join(f"{float(a):8.1f}" for a in v))
med.close(); dist.barrier(); dist.destroy_process_group()
