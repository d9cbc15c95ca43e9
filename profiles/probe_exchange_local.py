"""Drives the multi-GPU exchange kernels (pack_kernel / unpack_kernel, csrc/comm.cu) on ONE GPU through the in-process
transport (ls1g_comm_init_local: two contexts = ranks of a 2x1x1 grid, records moved by device copies), so that ncu - which must
not wrap a multi-rank command - can time them.  Usage (GPU box): python profiles/probe_exchange_local.py [cells_per_rank=100]"""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ls1g = importlib.import_module("ls1-mardyn_b200")
gen = importlib.import_module("ls1-mardyn_b200.generators")
cells = int(sys.argv[1]) if len(sys.argv) > 1 else 100
world = 2
g0, _ = ls1g.decomp_grid(world, 0)
cg = [cells * g for g in g0]
_, L = gen.fcc_box(cg, 0.8442)
cs = []
n_tot = 0
for rank in range(world):
    grid, coords = ls1g.decomp_grid(world, rank)
    bmin, bmax = ls1g.decomp_box(grid, coords, list(L))
    ids, r, v = gen.fcc_block(cg, bmin, bmax, 0.8442, 1.44, seed=1)
    c = ls1g.Container(bmax, 2.5, box_min=bmin, global_length=list(L), rank_grid=grid, rank_coords=coords, capacity=int(1.4 * len(ids)) + 200000)
    c.set_components([ls1g.lj_component()])
    c.add_particles(ids, np.zeros(len(ids), dtype=np.uint32), r, v)
    cs.append(c); n_tot += len(ids)
ls1g.comm_init_local(cs)
for it in range(3):
    ls1g.exchange_group(cs)
print("molecules per rank", n_tot // world, "owned/halo", [c.count() for c in cs])
for c in cs:
    c.close()
