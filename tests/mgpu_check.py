"""Multi-GPU parity check (one process per GPU, NCCL): the block-decomposed ls1g_run against the single-rank run of the
same global system.  Launched by tests/test_multi_gpu.py through torch.distributed.run; prints MGPU_OK on success."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ls1g = importlib.import_module("ls1-mardyn_b200")
gen = importlib.import_module("ls1-mardyn_b200.generators")


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n_cells, steps, dt, rc = 20, 60, 0.004, 2.5
    ids, r, v = gen.fcc_block(n_cells, temperature=2.0)
    _, L = gen.fcc_box(n_cells)
    rng = np.random.default_rng(3)
    r = np.mod(r + rng.normal(scale=0.05, size=r.shape), L)
    n = len(ids)
    grid, coords = ls1g.decomp_grid(world, rank)
    bmin, bmax = ls1g.decomp_box(grid, coords, list(L))
    mine = np.all((r >= np.array(bmin)) & (r < np.array(bmax)), axis=1)
    results = {}
    for mode, skin in (("lists", None), ("per_step", "0")):
        if skin is None:
            os.environ.pop("LS1G_SKIN", None)
        else:
            os.environ["LS1G_SKIN"] = skin
        c = ls1g.Container(bmax, rc, box_min=bmin, global_length=list(L), rank_grid=grid, rank_coords=coords, device=local,
                           capacity=int(2.5 * mine.sum()) + 20000)
        c.set_components([ls1g.lj_component()])
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt = torch.tensor(list(ls1g.comm_unique_id()), dtype=torch.uint8, device="cuda")
        dist.broadcast(idt, 0)
        c.comm_init(world, rank, bytes(idt.cpu().tolist()))
        c.add_particles(ids[mine], np.zeros(mine.sum(), dtype=np.uint32), r[mine], v[mine])
        c.exchange()
        c.traverse_cells()
        rec = c.run(steps, dt)
        t = torch.tensor(rec[:, :5], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        out = c.download(("id", "r", "v", "F"))
        results[mode] = (t.cpu().numpy(), out)
        c.close()
    # single-rank reference of the global system on every rank's own GPU (cheap), per-step loop
    os.environ["LS1G_SKIN"] = "0"
    c = ls1g.Container(list(L), rc, device=local)
    c.set_components([ls1g.lj_component()])
    c.add_particles(ids, np.zeros(n, dtype=np.uint32), r, v)
    c.exchange()
    c.traverse_cells()
    ref_rec = c.run(steps, dt)[:, :5]
    ref = c.download(("id", "r", "v", "F"))
    c.close()
    order = np.argsort(ref["id"])
    ok = True
    for mode, (rec, out) in results.items():
        assert np.all(rec[:, 4] == n), (mode, "molecule count")
        err = np.abs(rec[:, :3] - ref_rec[:, :3]).max(axis=0) / np.abs(ref_rec[:, :3]).max(axis=0)
        pos = np.searchsorted(ref["id"][order], out["id"])
        rr, vv = ref["r"][order][pos], ref["v"][order][pos]
        dr = np.abs(out["r"] - rr)
        dr = np.minimum(dr, np.abs(dr - np.array(L)))            # a molecule exactly at the periodic boundary may sit on either side
        e_r, e_v = dr.max(), np.abs(out["v"] - vv).max()
        if rank == 0:
            print("mode %s: series rel err %s, max |dr| %.2e, max |dv| %.2e" % (mode, err, e_r, e_v), flush=True)
        ok = ok and err.max() < 1e-9 and e_r < 1e-8 and e_v < 1e-8
    t = torch.tensor([1.0 if ok else 0.0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("MGPU_OK" if t.item() == 1.0 else "MGPU_FAIL", flush=True)
    dist.destroy_process_group()
    return 0 if t.item() == 1.0 else 1


if __name__ == "__main__":
    sys.exit(main())
