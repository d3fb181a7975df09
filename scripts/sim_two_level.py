"""Workload probe for the two-level culling kernel: candidate rates of cluster-block union boxes and a
warp-instruction cost model (uses the oracle only to build the scene; not product code)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from cpu_checkers import Oracle

def run(grid, screen, L, shapes, sample=8192, cap=150):
    O = Oracle()
    ubo = O.camera_ubo(grid, screen)
    aabb = O.generate_aabbs(ubo).view(np.float32).reshape(-1, 8)
    lights = O.init_lights(L)
    lv = O.transform_lights(ubo, lights)  # (L,4) x y z r2 ?
    lv = np.asarray(lv).view(np.float32).reshape(-1, 4)
    X, Y, Z = grid
    lo = aabb[:, 0:3].reshape(Z, Y, X, 3); hi = aabb[:, 4:7].reshape(Z, Y, X, 3)
    S = min(sample, L)
    P = lv[:S, :3]; r2 = lv[:S, 3]
    def sq(lo_, hi_):  # lo_,hi_: (n,3) -> (n,S)
        d = np.maximum(np.maximum(lo_[:, None, :] - P[None], P[None] - hi_[:, None, :]), 0)
        return (d * d).sum(-1)
    # per-cluster hit rate
    hit = sq(lo.reshape(-1, 3), hi.reshape(-1, 3)) < r2[None]
    hr = hit.mean(1)
    print(f"grid {grid} L={L}: mean hit rate {hr.mean():.4f}; clusters reaching cap (est) {(hr*L>=cap).mean():.3f}")
    for (bx, by, bz) in shapes:
        nb = 0; costA = 0.0; costB = 0.0; brute = 0.0; exec_frac = 0.0
        for z0 in range(0, Z, bz):
            for y0 in range(0, Y, by):
                for x0 in range(0, X, bx):
                    l = lo[z0:z0+bz, y0:y0+by, x0:x0+bx].reshape(-1, 3); h = hi[z0:z0+bz, y0:y0+by, x0:x0+bx].reshape(-1, 3)
                    ul = l.min(0)[None]; uh = h.max(0)[None]
                    cand = (sq(ul, uh)[0] < r2).mean()
                    hrs = hr.reshape(Z, Y, X)[z0:z0+bz, y0:y0+by, x0:x0+bx].reshape(-1)
                    # lights scanned until every cluster of the block holds cap hits (early exit), as a fraction of L
                    need = np.where(hrs > 0, cap / np.maximum(hrs, 1e-12), np.inf)
                    frac = min(1.0, need.max() / L)
                    nb += 1
                    costA += frac * L / 32 * 12      # phase A: 12 warp-instr per 32 lights (packed)
                    costB += frac * L * cand * 24    # phase B: 24 warp-instr per candidate (32 clusters)
                    brute += frac * L * len(hrs) / 32 * 12
                    exec_frac += frac
        print(f"  block {bx}x{by}x{bz}: blocks {nb} warp-instr A {costA/1e6:.1f}M B {costB/1e6:.1f}M total {(costA+costB)/1e6:.1f}M  vs brute(early-exit per block) {brute/1e6:.1f}M  -> x{brute/(costA+costB):.2f}; mean scan frac {exec_frac/nb:.2f}")

if __name__ == "__main__":
    shapes = [(4, 4, 2), (8, 4, 1), (4, 4, 1), (4, 2, 4), (2, 2, 8), (4, 3, 2), (16, 1, 2), (8, 2, 2), (2, 2, 4), (4,2,2)]
    run((16, 9, 24), (2560, 1600), 65536, shapes)
    run((32, 18, 48), (3840, 2160), 1 << 20, [(4, 4, 2), (8, 4, 1), (4, 2, 4), (2, 2, 8), (4, 3, 2),(2,2,4)], sample=4096)
