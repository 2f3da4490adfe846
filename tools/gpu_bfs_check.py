"""GPU check of the warp-per-tile unit-weight field solver (pmp_bfs.cu) against the oracle on
random grids of awkward shapes, plus the certificate on a bigger grid.  Prints one line per case."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import python_motion_planning_b200 as pmp   # noqa: E402
from python_motion_planning_b200 import engine, _lib, workloads   # noqa: E402
from oracle import oracle   # noqa: E402


def env6(occ):
    env = pmp.Grid3D(*occ.shape, walls=False)
    env.motions = [m for m in env.motions if m.g == 1]
    env.update(occ)
    return env


def main():
    rng = np.random.default_rng(0)
    ok = True
    quick = "--quick" in sys.argv
    reps = 3
    if "--reps" in sys.argv:
        reps = int(sys.argv.pop(sys.argv.index("--reps") + 1))
    sizes = [int(a) for a in sys.argv[1:] if a.isdigit()] or [256, 512]
    keep = oracle.motion_costs(oracle.MOTIONS_3D) == 1.0
    motions, mg = oracle.MOTIONS_3D[keep], oracle.motion_costs(oracle.MOTIONS_3D)[keep]
    for shape, dens in () if quick else (((40, 24, 20), 0.1), ((48, 17, 9), 0.3), ((70, 33, 41), 0.35), ((33, 9, 9), 0.05),
                        ((130, 70, 50), 0.2), ((96, 96, 96), 0.25), ((16, 8, 8), 0.0), ((200, 40, 24), 0.4)):
        occ = rng.random(shape) < dens
        free = np.argwhere(~occ)
        src = tuple(int(v) for v in free[rng.integers(0, len(free))])
        ref = oracle.plan(occ, src, None, algo=oracle.DIJKSTRA, motions=motions, motion_g=mg,
                          want_field=True, want_expand=False).g_field
        f = pmp.cost_field(env6(occ), src, weight="u32")
        same = np.array_equal(f, ref)
        ok &= same
        print(f"field {shape} dens {dens}: {'OK' if same else 'MISMATCH'} reached {int(np.isfinite(ref).sum())}", flush=True)
        if not same:
            bad = np.argwhere(f != ref)
            print("  first bad cells:", bad[:5].tolist(), [(f[tuple(b)], ref[tuple(b)]) for b in bad[:5]])
    # blocked source -> bad query, nothing reached
    occ = np.zeros((40, 20, 20), dtype=bool)
    occ[3, 3, 3] = True
    f = pmp.cost_field(env6(occ), (3, 3, 3), weight="u32")
    print("blocked source:", "OK" if not np.isfinite(f).any() else "MISMATCH", flush=True)
    ok &= not np.isfinite(f).any()
    # bigger: the c5 block at 256^3, certificate + timing
    import torch
    for n in sizes:
        blk, _ = workloads.replicated_boxes3d(base=128, reps=n // 128)
        occ = workloads.replicated_slab(blk, (n, n, n), 0, n)
        env = env6(occ)
        kms = []
        for rep in range(reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ws, st = pmp.cost_field(env, (1, 1, 1), weight="u32", to_host=False)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            kms.append(float(ws.stats[8].item()) / 1e6)
        dg = env.device_grid()
        out = torch.zeros(3, dtype=torch.int64, device=dg.device)
        src = torch.tensor([1, 1, 1], dtype=torch.int32, device=dg.device)
        _lib.check(dg.L.pmp_verify_field(dg.handle, _lib.W_U32, ws.g.data_ptr(), src.data_ptr(), 0, n, out.data_ptr(), dg.stream()), "verify")
        bad, reached, pend = (int(v) for v in out.cpu())
        s = ws.stats.cpu().numpy()
        na = max(int(s[1]), 1)
        ph = {k: round(float(s[i]) / na / 1965.0, 2) for k, i in (("wait", 4), ("setup", 6), ("sweep", 11), ("extract", 3), ("publish", 0))}
        ks = sorted(kms[1:] or kms)
        print(f"c5 {n}^3: {dt*1e3:.2f} ms wall, kernel {s[8]/1e6:.2f} ms (median {ks[len(ks)//2]:.2f} min {ks[0]:.2f} of {len(ks)}), certificate violations {bad}, reached {reached}, pending {pend}, "
              f"activations {s[1]} (idle {s[13]}) lowered {s[2]} levels {s[12]} passes {s[14]} requeues {s[5]} err {s[15]} "
              f"status {ws.status.cpu().tolist()} us/activation/warp {ph if any(ph.values()) else '(build with -DBFS_CLOCKS=1)'}", flush=True)
        ok &= bad == 0 and pend == 0 and s[15] == 0
    print("ALL OK" if ok else "FAILED")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
