#!/usr/bin/env python
"""Throughput of the hot path on the reference's SHIPPED examples (cfg A / C / D), next to bench.py's
headline workload (cfg B).  Supplementary measurement: bench.py keeps the driver's contract, this script answers "what
does a user of examples/Argon, examples/CO2, examples/EOX get".

    python bench_examples.py --workload co2 --replicas 8 --steps 20 --warmup 3

* input: the committed golden fixture of the example (tests/golden/ex_*.npz - positions, velocities, quaternions, angular
  momenta and component tables exactly as the reference read them from its .inp), replicated k^3 times like the reference's
  ReplicaGenerator; NVT with the example's own temperature, time step and cutoffs.
* GPU arm: Container.run (device resident) - multi-site systems take the per-step path (sort + force_generic_kernel),
  single-site ones the neighbour-list loop when their lists fit.
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
    """ReplicaGenerator semantics (io/ReplicaGenerator.cpp): the box tiled k times per dimension, every replica a shifted
    copy of the same molecules; ids renumbered"""
    box = d["box"]
    sh = np.array([(i, j, l) for i in range(k) for j in range(k) for l in range(k)], dtype=np.float64) * box
    n = len(d["in_id"])
    rep = lambda a: np.tile(a, (k ** 3,) + (1,) * (a.ndim - 1))
    return dict(id=np.arange(1, n * k ** 3 + 1, dtype=np.uint64), cid=rep(d["in_cid"]).astype(np.uint32),
                r=(d["in_r"][None, :, :] + sh[:, None, :]).reshape(-1, 3), v=rep(d["in_v"]), q=rep(d["in_q"]), D=rep(d["in_L"])), box * k


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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="co2")
    ap.add_argument("--replicas", type=int, default=8, help="replicas per dimension on the GPU (8 -> 1 048 576 molecules)")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--ref-replicas", type=int, default=3)
    ap.add_argument("--ref-steps", type=int, default=5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the MarDynGpu end-to-end arm")
    ap.add_argument("--e2e-replicas", type=int, default=4, help="replicas per dimension of the MarDynGpu arm (ASCII input: 4 -> 131 072 molecules)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    args = ap.parse_args()
    ls1g = importlib.import_module("ls1-mardyn_b200")
    fx, temp, label = WORKLOADS[args.workload]
    d = fixture(fx, args.workload)
    dt = float(d["dt"][0])
    threads = B.host_cores()
    metric = "MMUPS (million molecule-updates/s), %s replicated k^3, force+integrate step, NVT" % label

    if args.impl == "reference":
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
    if not torch.cuda.is_available():
        raise SystemExit("bench_examples.py needs a CUDA device (there is no CPU fallback)")
    m, box = replicate(d, args.replicas)
    n = len(m["id"])
    comps = ls1g.components_from_tables(d["comp_counts"], d["comp_scalars"], d["lj_sites"], d["charge_sites"], d["dipole_sites"], d["quadrupole_sites"])
    clocks = B.ClockSampler(0)
    with ls1g.Container(box, float(d["rc"][0]), lj_cutoff=float(d["rc"][1]), eps_rf=float(d["epsRF"][0])) as c:
        c.set_components(comps, d["mixcoeff"])
        c.add_particles(m["id"], m["cid"], m["r"], m["v"], q=m["q"], L=m["D"])
        c.exchange()
        try:
            fc = c.flop_count()           # the reference's FlopCounter on THIS replicated system, counted on the device
        except AttributeError:            # LS1G_LIB pointing at an older build of the library (A/B timing only)
            fc = dict(flops=float(d["flops"][0]) * args.replicas ** 3, kernel=np.zeros(7))
        u0, w0 = c.traverse_cells()
        clocks.start()
        t_wait = time.time()
        while not clocks.rows and clocks.proc is not None and time.time() - t_wait < 10.0:
            time.sleep(0.05)
        if args.warmup > 0:
            c.run(args.warmup, dt, target_temperature=temp)
        c.timers(reset=True)
        stream = torch.cuda.ExternalStream(c.stream(), device=torch.device("cuda", 0))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        tw0 = time.perf_counter()
        e0.record(stream)
        rec = c.run(args.steps, dt, target_temperature=temp)
        e1.record(stream)
        torch.cuda.synchronize()
        tw1 = time.perf_counter()
        ms = e0.elapsed_time(e1)
        clocks.window(tw0, tw1)
        clk = clocks.stop()
        tm = c.timers(reset=True)
        n_pairs = c.last_pair_count()
        peak = ls1g.fp64_peak_tflops(0)
    flops = fc["flops"]
    model = ("the reference's FlopCounter (Newton-3 counts) on the replicated system at t = 0, counted on the device by ls1g_flop_count "
             "(bit-equal to the reference's own count on the golden fixtures)")
    fixture_flops = float(d["flops"][0]) * args.replicas ** 3 if "flops" in d else None
    if args.workload == "argon" and args.steps > 1:
        # single LJ site: the neighbour-list loop accumulates the per-molecule virial on the last step of a run only; the other
        # launches are priced like the reference prices that case (FlopCounter.cpp:33-39: 35 instead of 41 flop per pair)
        flops -= 6.0 * n_pairs * (args.steps - 1) / args.steps
        model += "; minus 6 flop per pair on the %d of %d launches without per-molecule virial" % (args.steps - 1, args.steps)
    force_ms = tm["force"]["ms"] / max(tm["force"]["launches"], 1)
    achieved = flops / (force_ms * 1e-3) / 1e12 if force_ms > 0 else 0.0
    ref_u = float(d["upot_virial"][0]) / len(d["in_id"]) if "upot_virial" in d else u0 / n
    line = {"metric": metric, "value": round(n * args.steps / (ms * 1e-3) / 1e6, 2), "unit": "MMUPS", "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True, "dtype": "f64",
            "data": "reference example input (golden fixture), replicated",
            "config": {"workload": args.workload, "replicas": args.replicas, "n_molecules": n, "rc": float(d["rc"][0]), "dt": dt,
                       "target_T": temp, "upot_per_molecule_t0": u0 / n, "reference_upot_per_molecule_t0": ref_u,
                       "upot_rel_dev_t0": abs(u0 / n - ref_u) / abs(ref_u), "T_last": float((rec[-1, 2] + rec[-1, 3]) / (3 * rec[-1, 4] + rec[-1, 5]))},
            "gpu_launches": int(sum(t["launches"] for t in tm.values())),
            "roofline": {"bound": "fp64", "achieved": round(achieved, 3), "peak": round(peak, 3), "unit": "TFLOP/s",
                         "frac": round(achieved / peak, 4) if peak > 0 else None, "traffic": None, "flops_per_launch": flops, "kernel_ms": round(force_ms, 4),
                         "model": model, "reference_fixture_flops_x_replicas3": fixture_flops,
                         "kernel_calls": {k: float(v) for k, v in zip(("lj", "charge", "charge_dipole", "dipole", "charge_quadrupole", "dipole_quadrupole", "quadrupole"), fc["kernel"]) if v}},
            "kernel_ms_per_step": {k: round(t["ms"] / args.steps, 4) for k, t in tm.items()}, "clocks": clk}
    if not args.no_e2e:
        md = run_reference(d, temp, args.e2e_replicas, args.steps, threads, ls1g, gpu=True)
        if md is not None:
            line["e2e"] = {"value": round(md["mmups_e2e"], 2), "unit": "MMUPS", "h2d_bytes_per_step": md["h2d_bytes"] // args.steps,
                           "d2h_bytes_per_step": md["d2h_bytes"] // args.steps, "steps": args.steps, "n_molecules": md["n"],
                           "what": "MarDynGpu: the reference executable + unmodified time loop with GpuLinkedCells/GpuLeapfrog on the same example replicated %d^3 "
                                   "(ASCII phase space read by the reference's own reader); time = upload + the reference's loop timer (last-step mirror inside)" % args.e2e_replicas,
                           "seconds": {"upload": round(md["upload_s"], 4), "loop": round(md["loop_s"], 4), "of_loop_device_calls": round(md["device_s"], 4),
                                       "of_loop_final_mirror": round(md["mirror_s"], 4)},
                           "selective_thermostat_steps": md["selective"]}
    if not args.no_cpu_baseline:
        res = run_reference(d, temp, args.ref_replicas, args.ref_steps, threads, ls1g)
        if res is not None:
            line["cpu_baseline"] = {"value": round(res["mmups"], 4), "unit": "MMUPS", "cores": threads, "kind": "reference",
                                    "sample": "%s, same example replicated %d^3 = %d molecules x %d steps, OpenMP %d threads, loop %.2f s"
                                              % (res["binary"], args.ref_replicas, res["n"], args.ref_steps, threads, res["loop_s"])}
    print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
