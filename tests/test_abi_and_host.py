"""CPU-only checks: the C-ABI library loads and exports every symbol include/ls1g.h declares, compute calls fail
loudly without a GPU (no CPU fallback), host-only helpers (block decomposition) follow the reference, and the
N>1 sharding of the synthetic workload is consistent across ranks (gloo, world_size 2)."""
import ctypes
import importlib
import os
import re
import socket

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ls1g = importlib.import_module("ls1-mardyn_b200")
gen = importlib.import_module("ls1-mardyn_b200.generators")


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "ls1g.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ls1g_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = ls1g.lib()
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(ls1g.EXPORTS) == names
    assert lib.ls1g_version() >= 100


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(ls1g.Ls1gError, match="no CPU fallback"):
        ls1g.Container([10.0, 10.0, 10.0], 2.5)
    # the product package never imports the oracle
    for f in os.listdir(os.path.join(ROOT, "ls1-mardyn_b200")):
        if f.endswith(".py"):
            assert "oracle" not in open(os.path.join(ROOT, "ls1-mardyn_b200", f)).read().replace("no oracle", "")


def test_decomp_grid_and_box_follow_reference():
    """MPI_Dims_create order + DomainDecomposition::getBoundingBoxMin/Max (DomainDecomposition.cpp:35-39,116-125)"""
    assert ls1g.decomp_grid(1, 0) == ([1, 1, 1], [0, 0, 0])
    assert ls1g.decomp_grid(2, 1) == ([2, 1, 1], [1, 0, 0])
    assert ls1g.decomp_grid(4, 3) == ([2, 2, 1], [1, 1, 0])
    assert ls1g.decomp_grid(8, 5) == ([2, 2, 2], [1, 0, 1])
    assert ls1g.decomp_grid(6, 0)[0] == [3, 2, 1]
    assert ls1g.decomp_grid(12, 0)[0] == [3, 2, 2]
    L = [10.0, 7.0, 3.0]
    cover = 0.0
    for rank in range(8):
        g, co = ls1g.decomp_grid(8, rank)
        lo, hi = ls1g.decomp_box(g, co, L)
        for d in range(3):
            assert lo[d] == co[d] * L[d] / g[d]
            assert hi[d] == (L[d] if co[d] == g[d] - 1 else (co[d] + 1) * L[d] / g[d])
        cover += np.prod(np.array(hi) - np.array(lo))
    assert abs(cover - np.prod(L)) < 1e-12
    with pytest.raises(ls1g.Ls1gError):
        ls1g.decomp_box([2, 1, 1], [2, 0, 0], L)


def test_fcc_generator_is_decomposition_invariant():
    ids, r, v = gen.fcc_block(6)
    assert len(ids) == 4 * 6 ** 3 and len(np.unique(ids)) == len(ids)
    a, L = gen.fcc_box(6)
    assert np.all(r >= 0) and np.all(r < L)
    assert np.allclose((v * v).sum(axis=1), 3 * 1.44)
    parts = []
    for rank in range(4):
        g, co = ls1g.decomp_grid(4, rank)
        lo, hi = ls1g.decomp_box(g, co, L)
        parts.append(gen.fcc_block(6, lo, hi))
    pid = np.concatenate([p[0] for p in parts])
    order = np.argsort(pid)
    assert np.array_equal(pid[order], np.sort(ids))
    o0 = np.argsort(ids)
    assert np.array_equal(np.concatenate([p[1] for p in parts])[order], r[o0])
    assert np.array_equal(np.concatenate([p[2] for p in parts])[order], v[o0])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _gloo_worker(rank, world, port, n_cells, q):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import bench
        shard = bench.make_shard(n_cells, world, rank)
        t = torch.tensor([float(len(shard["id"])), float(shard["id"].astype(np.float64).sum()),
                          float(shard["r"].sum()), float(shard["v"][:, 0].sum())], dtype=torch.float64)
        dist.all_reduce(t)
        vmean = bench.global_mean_velocity(shard["v"], dist)
        q.put((rank, t.tolist(), shard["grid"], shard["coords"], vmean.tolist()))
    finally:
        dist.destroy_process_group()


def test_bench_sharding_two_ranks_gloo():
    """world_size 2 over gloo: the shards of bench.py tile the global system exactly (weak-scaling path host logic)"""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    n_cells = [8, 4, 4]          # 2 ranks x 4^3 unit cells each
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, n_cells, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    ids, r, v = gen.fcc_block(n_cells)
    tot = res[0][1]
    assert tot[0] == len(ids) and tot[1] == float(ids.astype(np.float64).sum())
    assert abs(tot[2] - r.sum()) < 1e-6 and abs(tot[3] - v[:, 0].sum()) < 1e-9
    assert res[0][1] == res[1][1]
    assert res[0][2] == [2, 1, 1] and res[0][3] == [0, 0, 0] and res[1][3] == [1, 0, 0]
    assert np.allclose(res[0][4], v.mean(axis=0)) and res[0][4] == res[1][4]


def test_library_holds_sm100a_code_for_the_hot_kernels():
    """the product is native sm_100a code: every embedded cubin is built for sm_100a and the kernels of the hot path are in it"""
    import shutil
    import subprocess
    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not on PATH")
    so = ls1g.lib_path()
    elfs = subprocess.run(["cuobjdump", "-lelf", so], capture_output=True, text=True).stdout.split("\n")
    elfs = [l for l in elfs if "ELF file" in l]
    assert len(elfs) >= 10 and all("sm_100a" in l for l in elfs)
    sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
    for k in ("nl_force_kernel", "nl_build_kernel", "force_generic_kernel", "force_1clj_tile_kernel", "integrate_kernel", "bin_kernel",
              "scatter_kernel", "pack_kernel", "unpack_kernel", "halo_refresh_kernel", "selective_kernel", "flop_count_kernel", "rdf_kernel"):
        assert re.search(r"Function : \S*%s" % k, sass), k
    assert "DFMA" in sass and "MUFU.RCP64H" in sass
