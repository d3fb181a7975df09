"""one z-slab of grid C with 16M lights on one GPU: ms of {transform, bin, count, finalize, write} for K x bin levels"""
import os, sys, json, subprocess
import numpy as np

def one(z0, z1, K, bin2):
    os.environ["HVCF_HIER_K"] = str(K); os.environ["HVCF_BIN2"] = str(bin2); os.environ["HVCF_KERNEL_TIMES"] = "1"
    from hellovulkan_b200 import hvcf
    hvcf.load()
    grid, screen, n = (64, 36, 96), (3840, 2160), 1 << 24
    lights = np.zeros(n, hvcf.LIGHT_DT); hvcf.host_init_lights(n, out=lights)
    with hvcf.Context(grid=grid, max_lights=n, z_begin=z0, z_end=z1) as ctx:
        ctx.set_lights(lights); ctx.set_ubo(0, hvcf.host_camera_ubo(grid, screen))
        acc = np.zeros(5)
        for i in range(4):
            ctx.cull(0); t = np.array(ctx.kernel_times(0))
            if i: acc += t
        print(json.dumps({"z": [z0, z1], "K": K, "bin2": bin2, "ms": [round(float(x), 3) for x in acc / 3], "sum": round(float(acc.sum() / 3), 3)}), flush=True)

if __name__ == "__main__":
    if len(sys.argv) > 1:
        one(*[int(a) for a in sys.argv[1:]])
    else:
        for z in ((69, 74), (74, 96), (0, 25), (53, 60)):
            for K in (1, 2, 4):
                for b in (0, 1):
                    subprocess.run([sys.executable, __file__, str(z[0]), str(z[1]), str(K), str(b)])
