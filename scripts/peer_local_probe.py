"""scratch: in-process sharded frame on one GPU (peer_attach_local), variants, to find what stalls at full size"""
import sys, time, json
import numpy as np, torch
sys.path.insert(0, "tests")
from hellovulkan_b200 import hvcf
from hellovulkan_b200.sharded import ShardedCuller
hvcf.load()
grid, screen = (64, 36, 96), (3840, 2160)
slabs = [(0, 40), (40, 61), (61, 70), (70, 96)]
def run(n, mode, nstreams=4):
    lights = np.zeros(n, hvcf.LIGHT_DT); hvcf.host_init_lights(n, out=lights)
    ubo = hvcf.host_camera_ubo(grid, screen)
    ctxs = [hvcf.Context(grid=grid, max_lights=n, z_begin=a, z_end=b) for a, b in slabs]
    streams = [torch.cuda.Stream() for _ in range(nstreams)]
    for c in ctxs:
        c.set_lights(lights); c.set_ubo(0, ubo)
    torch.cuda.synchronize()
    for r, c in enumerate(ctxs): c.peer_attach_local(r, ctxs)
    t0 = time.time()
    for i, c in enumerate(ctxs):
        sp = streams[i % nstreams].cuda_stream
        if mode == "sharded":
            c.cull_sharded(0, stream=sp)
        else:
            for b, e in ShardedCuller.chunk_bounds(n, 4):
                c.cull_begin_range(0, b, e, 0, stream=sp)
            c.peer_exchange_totals(stream=sp); c.cull_finish(0, c.peer_index_base_ptr(), stream=sp)
            c.peer_push_grid(stream=sp); c.peer_wait_grid(stream=sp)
        t_issue = time.time() - t0
    err = None
    try:
        for i, c in enumerate(ctxs): c.peer_status(stream=streams[i % nstreams].cuda_stream)
    except Exception as e:
        err = str(e)[:120]
    torch.cuda.synchronize()
    print(json.dumps({"n": n, "mode": mode, "issue_s": round(t_issue, 3), "total_s": round(time.time() - t0, 3), "err": err}), flush=True)
    for c in ctxs: c.peer_detach()
    for c in ctxs: c.close()
for n, mode in ((1 << 20, "ranges"), (1 << 20, "sharded"), (1 << 24, "sharded"), (1 << 24, "ranges")):
    run(n, mode)
