"""2+ ranks: time hvcf_peer_push_grid alone (after one sharded frame of grid C, 2M lights), median of 20."""
import os, json
import numpy as np, torch, torch.distributed as dist
from hellovulkan_b200 import hvcf
from hellovulkan_b200.sharded import HvcfEngine, ShardedCuller, z_slabs
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
hvcf.load()
grid, screen, n = (64, 36, 96), (3840, 2160), 1 << 21
slabs = z_slabs(grid[2], world)
eng = HvcfEngine(hvcf, grid, n, slabs[rank], local)
sc = ShardedCuller(eng, grid, world=world, rank=rank, slabs=slabs, exchange="peer")
if rank == 0:
    host = np.zeros(n, hvcf.LIGHT_DT); hvcf.host_init_lights(n, out=host)
    eng.lights.copy_(torch.from_numpy(host.view(np.uint8).reshape(-1)))
ubo = hvcf.host_camera_ubo(grid, screen)
for f in range(2): eng.set_ubo(f, ubo)
sc.broadcast_lights(n)
tot, _ = sc.frame(0); sc.frame(1)
torch.cuda.synchronize(); dist.barrier()
res = {}
for what in ("push", "push+wait", "wait"):
    ts = []
    for i in range(20):
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if what != "wait": eng.peer_push_grid()
        if what != "push": eng.peer_wait_grid()
        e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    res[what] = float(np.median(ts))
print(json.dumps({"rank": rank, "totals": tot.tolist(), "blocks": os.environ.get("HVCF_PEER_BLOCKS"), "ms": res}), flush=True)
dist.barrier(); sc.close(); eng.close()
