#!/usr/bin/env python
"""Throughput of the hot path on the reference's SHIPPED examples (cfg A / C / D), next to bench.py's
headline workload (cfg B).  Supplementary measurement: bench.py keeps the driver's contract, this script answers "what
does a user of examples/Argon, examples/CO2, examples/EOX get".

    python bench_examples.py --workload co2 --replicas 8 --steps 20 --warmup 3

* input: the committed golden fixture of the example (tests/golden/ex_*.npz - positions, velocities, quaternions, angular
  momenta and component tables exactly as the reference read them from its .inp), replicated k^3 times like the reference's
  ReplicaGenerator; NVT with the example's own temperature, time step and cutoffs.
* GPU arm: Container.run (device resident) - the neighbour-list loop for single-site and rigid multi-site systems alike
  (LS1G_LIST_MULTISITE=0: the per-step cell scan, force_generic_kernel).  --gpus N (torch.distributed.run, NCCL): block decomposition
  of the replicated example, --replicas per dimension PER GPU (weak scaling).
* roofline: the reference's own FlopCounter total for this very configuration (fixture key "flops", dumped by the reference
  at t = 0) times k^3, over the pair-kernel time, against the measured FP64 peak.
* CPU arm (cpu_baseline, kind "reference"): oracle/_ref/MarDyn on a smaller replica of the same example, written as an
  ASCII phase space + XML by ls1-mardyn_b200/phasespace.py, all host threads.
One JSON line per run on stdout.
"""
import argparse
import importlib
import json
import os
import re
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import bench as B                                                     # noqa: E402  (host_cores, reference_binary, ClockSampler)

WORKLOADS = {
    # fixture, reduced temperature of the example's config.xml, name
    "argon": ("ex_Argon", 0.000633363365, "examples/Argon/200K_18mol_l (1 LJ site)"),
    "co2": ("ex_CO2", 0.00117172223, "examples/CO2/370K_15mol_l (3 LJ + 3 charges)"),
    "eox": ("ex_EOX", 0.0019000901, "examples/EOX/600K_15mol_l (3 LJ + 1 dipole)"),
    # SURVEY.md cfg C variant (CO2-Merker type, 3 LJ + 1 point quadrupole): the CO2 example's molecules with the three charges
    # replaced by the point quadrupole of the same charge distribution, Q = sum q_i z_i^2 (synthetic parameters, real geometry)
    "co2q": ("ex_CO2", 0.00117172223, "examples/CO2/370K_15mol_l geometry as a 3 LJ + 1 quadrupole model (CO2-Merker type)"),
}

XML = """<?xml version='1.0' encoding='UTF-8'?>
<mardyn version="20100525">
  <refunits type="SI"><length unit="nm">0.0529177</length><mass unit="u">1000</mass><energy unit="eV">27.2126</energy></refunits>
  <simulation type="MD">
    <integrator type="{integrator}"><timestep unit="reduced">{dt}</timestep></integrator>
    <run><currenttime>0</currenttime><production><steps>{steps}</steps></production></run>
    <ensemble type="NVT">
      <temperature unit="reduced">{temp}</temperature>
      <domain type="box"><lx>{L}</lx><ly>{L}</ly><lz>{L}</lz></domain>
      <components>
{components}
      </components>
      <phasespacepoint><file type="ASCII">sample.inp</file></phasespacepoint>
    </ensemble>
    <algorithm>
      <parallelisation type="DomainDecomposition"></parallelisation>
      <datastructure type="{container}"><cellsInCutoffRadius>1</cellsInCutoffRadius></datastructure>
      <cutoffs type="CenterOfMass"><defaultCutoff unit="reduced">{rc}</defaultCutoff><radiusLJ unit="reduced">{rclj}</radiusLJ></cutoffs>
      <electrostatic type="ReactionField"><epsilon>{epsrf}</epsilon></electrostatic>
    </algorithm>
    <output></output>
  </simulation>
</mardyn>
"""


def fixture(name, workload=None):
    d = dict(np.load(os.path.join(ROOT, "tests", "golden", name + ".npz")))
    if workload == "co2q":
        ch = d["charge_sites"]
        Q = float((ch[:, 4] * ch[:, 2] ** 2).sum())
        d["comp_counts"] = np.array([[3, 0, 0, 1, 2, 0]], dtype=d["comp_counts"].dtype)
        d["charge_sites"] = np.zeros((0, 5))
        d["quadrupole_sites"] = np.array([[0.0, 0.0, 0.0, 0.0, 0.0, 1.0, Q]])
        d.pop("flops", None)                 # no reference count for this variant: ls1g_flop_count supplies it
        d.pop("upot_virial", None)
    return d


def replicate(d, k):
    """ReplicaGenerator semantics (io/ReplicaGenerator.cpp): the box tiled k (or kx, ky, kz) times per dimension, every replica a
    shifted copy of the same molecules; ids renumbered"""
    kk = [int(k)] * 3 if np.isscalar(k) else [int(x) for x in k]
    box = d["box"]
    sh = np.array([(i, j, l) for i in range(kk[0]) for j in range(kk[1]) for l in range(kk[2])], dtype=np.float64) * box
    n, nrep = len(d["in_id"]), kk[0] * kk[1] * kk[2]
    rep = lambda a: np.tile(a, (nrep,) + (1,) * (a.ndim - 1))
    return dict(id=np.arange(1, n * nrep + 1, dtype=np.uint64), cid=rep(d["in_cid"]).astype(np.uint32),
                r=(d["in_r"][None, :, :] + sh[:, None, :]).reshape(-1, 3), v=rep(d["in_v"]), q=rep(d["in_q"]), D=rep(d["in_L"])), box * np.array(kk, dtype=np.float64)


def ascii_components(d, ls1g):
    """ABI component rows of the fixture -> the site tables of the ASCII header / component XML"""
    out = []
    for c in ls1g.components_from_tables(d["comp_counts"], d["comp_scalars"], d["lj_sites"], d["charge_sites"], d["dipole_sites"], d["quadrupole_sites"]):
        out.append(dict(lj=[list(r[:6]) + [0.0, 1.0 if r[6] != 0 else 0.0] for r in c["lj"]], charges=[list(r) for r in c["charge"]],
                        dipoles=[list(r) for r in c["dipole"]], quadrupoles=[list(r) for r in c["quadrupole"]], I=list(c["inertia"])))
    return out


def run_reference(d, temp, k, steps, threads, ls1g, gpu=False):
    """gpu=False: the reference executable (CPU arm).  gpu=True: the same executable with <datastructure type="GpuLinkedCells"> +
    <integrator type="GpuLeapfrog"> (MarDynGpu, ls1-mardyn_b200/host/) - the end-to-end arm: ASCII phase space -> host Molecule
    objects -> device -> steps driven by the reference's time loop -> host Molecule objects"""
    exe = B.MARDYN_GPU if gpu else B.reference_binary()
    if exe is None or not os.path.exists(exe):
        return None
    P = ls1g.phasespace
    m, box = replicate(d, k)
    comps = ascii_components(d, ls1g)
    ps = P.PhaseSpace(time=0.0, box=box, temperature=temp, components=comps, mixcoeff=list(d["mixcoeff"]), eps_rf=float(d["epsRF"][0]), **m)
    inner = "\n".join(l for l in P.components_xml(comps).splitlines() if not l.startswith("<?xml") and "components" not in l)
    with tempfile.TemporaryDirectory() as td:
        P.write_ascii(os.path.join(td, "sample.inp"), ps)
        with open(os.path.join(td, "cfg.xml"), "w") as f:
            f.write(XML.format(integrator="GpuLeapfrog" if gpu else "Leapfrog", container="GpuLinkedCells" if gpu else "LinkedCells",
                               dt=repr(float(d["dt"][0])), steps=steps, temp=repr(temp), L=repr(float(box[0])), components=inner,
                               rc=repr(float(d["rc"][0])), rclj=repr(float(d["rc"][1])), epsrf=repr(float(d["epsRF"][0]))))
        env = dict(os.environ, OMP_NUM_THREADS=str(threads), OMP_PROC_BIND="close", OMP_PLACES="cores")
        p = subprocess.run([exe, "cfg.xml", "--steps", str(steps), "--final-checkpoint=0"], cwd=td, env=env, capture_output=True, text=True, timeout=3000)
    out = p.stdout + p.stderr
    if p.returncode != 0 or "Exit code would have been" in out:
        raise RuntimeError("reference MarDyn failed (%d): %s" % (p.returncode, out[-3000:]))
    loop_s = float(re.search(r"Computation in main loop took:\s*([0-9.eE+-]+)", out).group(1))
    fm = re.search(r"Force calculation took:\s*([0-9.eE+-]+)", out)
    n = len(m["id"])
    res = dict(mmups=n * steps / loop_s / 1e6, n=n, loop_s=loop_s, force_s=float(fm.group(1)) if fm else None, binary=os.path.basename(exe))
    g = re.search(r"GpuLinkedCells: (\d+) device steps took ([0-9.eE+-]+) s; upload ([0-9.eE+-]+) s \((\d+) B\); (\d+) host mirror download\(s\) took ([0-9.eE+-]+) s \((\d+) B\)", out)
    if g:
        res.update(device_s=float(g.group(2)), upload_s=float(g.group(3)), h2d_bytes=int(g.group(4)), mirror_s=float(g.group(6)), d2h_bytes=int(g.group(7)) + 48 * steps,
                   selective=int(re.search(r"selective thermostat applied (\d+)", out).group(1)))
        res["mmups_e2e"] = n * steps / (loop_s + res["upload_s"]) / 1e6
    return res


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="co2")
    ap.add_argument("--gpus", type=int, default=1, help="ranks (one process per GPU, torch.distributed.run); weak scaling: --replicas per GPU")
    ap.add_argument("--replicas", type=int, default=8, help="replicas per dimension on the GPU (8 -> 1 048 576 molecules)")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--ref-replicas", type=int, default=3)
    ap.add_argument("--ref-steps", type=int, default=5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the MarDynGpu end-to-end arm")
    ap.add_argument("--e2e-replicas", type=int, default=4, help="replicas per dimension of the MarDynGpu arm (ASCII input: 4 -> 131 072 molecules)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    args = ap.parse_args(argv)
    ls1g = importlib.import_module("ls1-mardyn_b200")
    fx, temp, label = WORKLOADS[args.workload]
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d does not match WORLD_SIZE %d (launch with torch.distributed.run --nproc-per-node %d)" % (args.gpus, world, args.gpus))
    d = fixture(fx, args.workload)
    dt = float(d["dt"][0])
    threads = B.host_cores()
    metric = "MMUPS (million molecule-updates/s), %s replicated k^3, force+integrate step, NVT" % label

    if args.impl == "reference":
        if rank != 0:
            return 0
        res = run_reference(d, temp, args.ref_replicas, max(args.steps, 1), threads, ls1g)
        if res is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/MarDyn not built"})); return 0
        print(json.dumps({"impl": "reference", "metric": metric, "value": round(res["mmups"], 4), "unit": "MMUPS", "steps": args.steps,
                          "ms_per_step": round(1e3 * res["loop_s"] / max(args.steps, 1), 3), "higher_is_better": True, "dtype": "f64",
                          "config": {"workload": args.workload, "replicas": args.ref_replicas, "n_molecules": res["n"]},
                          "cpu_baseline": {"value": round(res["mmups"], 4), "unit": "MMUPS", "cores": threads, "kind": "reference",
                                           "sample": "%s, %d molecules x %d steps" % (res["binary"], res["n"], args.steps)}}), flush=True)
        return 0

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench_examples.py needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allsum(x):
        t = torch.tensor(np.atleast_1d(np.asarray(x, dtype=np.float64)), dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t)
        return t.cpu().numpy()

    # block decomposition (DomainDecomposition's grid) of the replicated example: --replicas per dimension PER GPU
    grid, coords = ls1g.decomp_grid(world, rank)
    m, box = replicate(d, [args.replicas * g for g in grid])
    bmin, bmax = ls1g.decomp_box(grid, coords, list(box))
    mine = np.all((m["r"] >= np.array(bmin)) & (m["r"] < np.array(bmax)), axis=1)
    n_local = int(mine.sum())
    n = int(allsum(n_local)[0])
    comps = ls1g.components_from_tables(d["comp_counts"], d["comp_scalars"], d["lj_sites"], d["charge_sites"], d["dipole_sites"], d["quadrupole_sites"])
    clocks = B.ClockSampler(local)
    with ls1g.Container(bmax, float(d["rc"][0]), box_min=bmin, lj_cutoff=float(d["rc"][1]), eps_rf=float(d["epsRF"][0]), global_length=list(box),
                        rank_grid=grid, rank_coords=coords, device=local, capacity=(int(2.2 * n_local) + 100000) if world > 1 else 0) as c:
        c.set_components(comps, d["mixcoeff"])
        if world > 1:
            idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
            if rank == 0:
                idt = torch.tensor(list(ls1g.comm_unique_id()), dtype=torch.uint8, device="cuda")
            dist.broadcast(idt, 0)
            c.comm_init(world, rank, bytes(idt.cpu().tolist()))
        c.add_particles(m["id"][mine], m["cid"][mine], m["r"][mine], m["v"][mine], q=m["q"][mine], L=m["D"][mine])
        del m
        c.exchange()
        fc = c.flop_count()               # the reference's FlopCounter on THIS replicated system (this rank's part), counted on the device
        u0, w0 = c.traverse_cells()
        if rank == 0:
            clocks.start()
            t_wait = time.time()
            while not clocks.rows and clocks.proc is not None and time.time() - t_wait < 10.0:
                time.sleep(0.05)
        if args.warmup > 0:
            c.run(args.warmup, dt, target_temperature=temp)
        c.timers(reset=True)
        rb0 = c.list_stats()["rebuilds"]
        stream = torch.cuda.ExternalStream(c.stream(), device=torch.device("cuda", local))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        tw0 = time.perf_counter()
        e0.record(stream)
        rec = c.run(args.steps, dt, target_temperature=temp)
        e1.record(stream)
        barrier()
        tw1 = time.perf_counter()
        mt = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(mt, op=dist.ReduceOp.MAX)
        ms = float(mt.item())
        if rank == 0:
            clocks.window(tw0, tw1)
        clk = clocks.stop() if rank == 0 else None
        tm = c.timers(reset=True)
        lists = c.list_stats()
        lists["rebuilds_in_window"] = lists["rebuilds"] - rb0
        n_pairs = int(allsum(c.last_pair_count())[0])
        peak = ls1g.fp64_peak_tflops(local) if rank == 0 else 0.0
    rec = allsum(rec.ravel()).reshape(rec.shape) if world > 1 else rec          # NCCL ranks report their own part (Domain::setLocal*)
    u0 = float(allsum(u0)[0])
    fc = dict(flops=float(allsum(fc["flops"])[0]) / world, kernel=allsum(fc["kernel"]) / world)      # per GPU: the kernel time is per GPU too
    flops = fc["flops"]
    model = ("the reference's FlopCounter (Newton-3 counts) on the replicated system at t = 0, counted on the device by ls1g_flop_count "
             "(bit-equal to the reference's own count on the golden fixtures)")
    fixture_flops = float(d["flops"][0]) * args.replicas ** 3 if "flops" in d else None      # per GPU
    if args.workload == "argon" and args.steps > 1:
        # single LJ site: the neighbour-list loop accumulates the per-molecule virial on the last step of a run only; the other
        # launches are priced like the reference prices that case (FlopCounter.cpp:33-39: 35 instead of 41 flop per pair)
        flops -= 6.0 * (n_pairs / world) * (args.steps - 1) / args.steps
        model += "; minus 6 flop per pair on the %d of %d launches without per-molecule virial" % (args.steps - 1, args.steps)
    force_ms = tm["force"]["ms"] / max(args.steps, 1)        # pair-kernel time per step
    achieved = flops / (force_ms * 1e-3) / 1e12 if force_ms > 0 else 0.0
    ref_u = float(d["upot_virial"][0]) / len(d["in_id"]) if "upot_virial" in d else u0 / n
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0
    line = {"metric": metric, "value": round(n * args.steps / (ms * 1e-3) / 1e6, 2), "unit": "MMUPS", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "reference example input (golden fixture), replicated",
            "config": {"workload": args.workload, "replicas_per_gpu": args.replicas, "rank_grid": grid, "n_molecules": n, "rc": float(d["rc"][0]), "dt": dt, "neighbour_lists": lists,
                       "target_T": temp, "upot_per_molecule_t0": u0 / n, "reference_upot_per_molecule_t0": ref_u,
                       "upot_rel_dev_t0": abs(u0 / n - ref_u) / abs(ref_u), "T_last": float((rec[-1, 2] + rec[-1, 3]) / (3 * rec[-1, 4] + rec[-1, 5]))},
            "gpu_launches": int(sum(t["launches"] for t in tm.values())),
            "roofline": {"bound": "fp64", "achieved": round(achieved, 3), "peak": round(peak, 3), "unit": "TFLOP/s",
                         "frac": round(achieved / peak, 4) if peak > 0 else None, "traffic": None, "flops_per_launch": flops, "kernel_ms": round(force_ms, 4),
                         "model": model, "reference_fixture_flops_x_replicas3": fixture_flops,
                         "kernel_calls": {k: float(v) for k, v in zip(("lj", "charge", "charge_dipole", "dipole", "charge_quadrupole", "dipole_quadrupole", "quadrupole"), fc["kernel"]) if v}},
            "kernel_ms_per_step": {k: round(t["ms"] / args.steps, 4) for k, t in tm.items()}, "clocks": clk}
    if not args.no_e2e and world == 1:
        md = run_reference(d, temp, args.e2e_replicas, args.steps, threads, ls1g, gpu=True)
        if md is not None:
            line["e2e"] = {"value": round(md["mmups_e2e"], 2), "unit": "MMUPS", "h2d_bytes_per_step": md["h2d_bytes"] // args.steps,
                           "d2h_bytes_per_step": md["d2h_bytes"] // args.steps, "steps": args.steps, "n_molecules": md["n"],
                           "what": "MarDynGpu: the reference executable + unmodified time loop with GpuLinkedCells/GpuLeapfrog on the same example replicated %d^3 "
                                   "(ASCII phase space read by the reference's own reader); time = upload + the reference's loop timer (last-step mirror inside)" % args.e2e_replicas,
                           "seconds": {"upload": round(md["upload_s"], 4), "loop": round(md["loop_s"], 4), "of_loop_device_calls": round(md["device_s"], 4),
                                       "of_loop_final_mirror": round(md["mirror_s"], 4)},
                           "selective_thermostat_steps": md["selective"]}
    if not args.no_cpu_baseline and world == 1:
        res = run_reference(d, temp, args.ref_replicas, args.ref_steps, threads, ls1g)
        if res is not None:
            line["cpu_baseline"] = {"value": round(res["mmups"], 4), "unit": "MMUPS", "cores": threads, "kind": "reference",
                                    "sample": "%s, same example replicated %d^3 = %d molecules x %d steps, OpenMP %d threads, loop %.2f s"
                                              % (res["binary"], args.ref_replicas, res["n"], args.ref_steps, threads, res["loop_s"])}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
