#!/usr/bin/env python
"""bench.py — MMUPS (million molecule-updates/s) of the ls1-MarDyn short-range force + integrate path.

  python bench.py --gpus N --steps K --warmup W            the B200 path (C ABI, libls1g.so)
  python bench.py --impl reference --steps K --warmup W    the reference's own CPU implementation (oracle/_ref/MarDyn)

Workload (BASELINE.json configs[1], SURVEY.md 8d cfg B): synthetic 1CLJ fcc lattice, eps=sigma=m=1, rho*=0.8442,
T*=1.44, rc=2.5 sigma, dt*=0.002, NVE, 136^3 unit cells = 10 061 824 molecules PER GPU (weak scaling; the rank grid
is DomainDecomposition's block grid, halo + leaving exchange over NCCL).  One step = one full time step of the path:
leapfrog pre -> wrap/halo/sort -> force pass (+calcFM, upot, virial) -> leapfrog post + kinetic sums.
MUPS is the reference's own definition, N_global * steps / time of the loop (src/MarDyn.cpp:264-266).

Prints ONE JSON line (rank 0).  `value`: device-resident loop, inputs in HBM.  `e2e` (N=1): the reference-facing drop-in end
to end - ls1-MarDyn's own executable and unmodified time loop with <datastructure type="GpuLinkedCells"> (MarDynGpu,
ls1-mardyn_b200/host/) on the same XML the reference arm runs: host Molecule objects -> device, K steps with the step
scalars read back every step, full state back into host Molecule objects; its `c_abi_full_state_roundtrip_every_step` entry
(and `e2e` itself at N>1, where the MPI-less reference executable cannot run) is the C-ABI worst case: every step uploads
id,cid,r,v,F from pinned host memory and downloads id,r,v,F."""
import argparse
import importlib
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

RHO, TEMP, RC, DT = 0.8442, 1.44, 2.5, 0.002
CELLS_PER_GPU = 136            # fcc unit cells per dimension per GPU: 4*136^3 = 10 061 824 molecules
METRIC = "MMUPS (million molecule-updates/s), 1CLJ fcc 10^7 molecules per B200, force+integrate step"


def workload_name(cells):
    """the same string in both arms (the driver compares the configs of the two lines)"""
    return ("cfg B: 1CLJ fcc lattice rho*=0.8442 T*=1.44 rc=2.5sigma dt*=0.002, %d^3 unit cells = %d molecules per GPU"
            % (cells, 4 * cells ** 3))


def _pkg():
    return importlib.import_module("ls1-mardyn_b200")


def _gen():
    return importlib.import_module("ls1-mardyn_b200.generators")


# ----------------------------------------------------------------------------------------------------------
# workload
# ----------------------------------------------------------------------------------------------------------
def make_shard(n_cells_global, world, rank):
    """this rank's block of the global fcc system (DomainDecomposition.cpp:116-125 boxes)"""
    ls1g, gen = _pkg(), _gen()
    grid, coords = ls1g.decomp_grid(world, rank)
    _, L = gen.fcc_box(n_cells_global, RHO)
    bmin, bmax = ls1g.decomp_box(grid, coords, list(L))
    ids, r, v = gen.fcc_block(n_cells_global, bmin, bmax, RHO, TEMP, seed=1)
    return dict(id=ids, r=r, v=v, grid=grid, coords=coords, box_min=bmin, box_max=bmax, L=list(L))


def global_cells(world, per_gpu):
    grid, _ = _pkg().decomp_grid(world, 0)
    return [per_gpu * g for g in grid]


def global_mean_velocity(v, dist=None):
    """IOHelpers::removeMomentum: global mean velocity (all-reduced when distributed)"""
    import torch
    t = torch.tensor([v[:, 0].sum(), v[:, 1].sum(), v[:, 2].sum(), float(len(v))], dtype=torch.float64)
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        if dist.get_backend() == "nccl":
            t = t.cuda()
        dist.all_reduce(t)
        t = t.cpu()
    return (t[:3] / t[3]).numpy()


def distance_checks(cell, cpd):
    """N_dist of the FlopCounter model: molecule-centre distance checks with Newton 3, every unordered pair of molecules
    in the same or adjacent cells once (halo-halo excluded; a pair with a halo image counts 1/2 from each side)."""
    nx, ny, nz = cpd
    cnt = np.bincount(cell.astype(np.int64), minlength=nx * ny * nz).astype(np.float64).reshape(nz, ny, nx)
    inner = cnt[1:-1, 1:-1, 1:-1]
    tot = (inner * (inner - 1.0)).sum() * 0.5
    for dz in (-1, 0, 1):
        for dy in (-1, 0, 1):
            for dx in (-1, 0, 1):
                if dx == 0 and dy == 0 and dz == 0:
                    continue
                nb = cnt[1 + dz:nz - 1 + dz, 1 + dy:ny - 1 + dy, 1 + dx:nx - 1 + dx]
                tot += 0.5 * (inner * nb).sum()
    return float(tot)


def lj_flops(n_dist, n_pairs, per_molecule_virial=True):
    """FlopCounter model (adapter/FlopCounter.cpp:19-75, FlopCounter.h:243-269) for one LJ site per molecule:
    8 per distance check + (8 + 15 + 12) per in-cutoff site pair + (4 + 2) macroscopic = 41 per pair.
    Steps that do not accumulate the per-molecule virial Vi are counted with the reference's own figure for that case
    (its reduced-memory mode, FlopCounter.cpp:33-39): (8 + 12 + 6) + (7 + 2) = 35 per pair."""
    return 8.0 * n_dist + (41.0 if per_molecule_virial else 35.0) * n_pairs


def hbm_kernels(tm, steps, n_local):
    """achieved HBM GB/s of the memory-bound kernels against the measured copy peak (MEASURED_PEAKS.json).
    integrate (fused post+pre, 1CLJ): read pos 32 + vel 24 + F 24 + cell 4, write pos 32 + vel 24 = 140 B per
    molecule-update; the list mode adds a 32-B read of the build position (not counted)."""
    peak, src = 6548.5, "fallback (B200_PROFILING.md)"
    try:
        peak, src = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json"
    except (OSError, KeyError, ValueError):
        pass
    if tm["integrate"]["launches"] < 2 * steps:
        # NVE steps of ls1g_run: the integrator runs in the epilogue of the force kernel, only the last step of a run
        # launches integrate_kernel - nothing to rate separately
        return {"integrate": {"bound": "hbm", "fused_into": "nl_force_kernel (lists.cuh, FUSE)", "separate_launches": int(tm["integrate"]["launches"]),
                              "bytes_per_molecule": 140, "peak": peak, "unit": "GB/s", "peak_source": src}}
    ms = tm["integrate"]["ms"] / max(steps, 1)
    gbs = 140.0 * n_local / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
    return {"integrate": {"bound": "hbm", "achieved": round(gbs, 1), "peak": peak, "unit": "GB/s", "frac": round(gbs / peak, 4),
                          "bytes_per_molecule": 140, "ms_per_step": round(ms, 4), "peak_source": src}}


# ----------------------------------------------------------------------------------------------------------
# clocks
# ----------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.rows, self.proc = device, [], None

    def start(self):
        if os.environ.get("LS1G_BENCH_NO_CLOCKS"):      # diagnosis only: a line without clocks is not a valid bench line
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")] + [time.perf_counter()])

    def window(self, t0, t1):
        """keep only the samples taken inside [t0, t1] (the timed region); the sampler itself is started before the
        warm-up because the first nvidia-smi query can stall the driver for a few hundred ms"""
        rows = [r for r in self.rows if t0 <= r[-1] <= t1]
        if rows:
            self.rows = rows

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) >= 7 and r[2 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ----------------------------------------------------------------------------------------------------------
# the reference's own CPU implementation (oracle/_ref/MarDyn, built from the unmodified sources by oracle/ref/Makefile)
# ----------------------------------------------------------------------------------------------------------
def mardyn_xml(steps, cells, gpu, devices=1):
    """the one configuration both arms run: `cells` = fcc unit cells per dimension (an int, or [nx, ny, nz] for the block grid of a
    multi-GPU run); gpu selects <datastructure type="GpuLinkedCells"> + <integrator type="GpuLeapfrog">, everything else is equal"""
    a = (4.0 / RHO) ** (1.0 / 3.0)
    cx, cy, cz = (cells, cells, cells) if isinstance(cells, int) else cells
    return MARDYN_XML.format(dt=DT, steps=steps, temp=TEMP, Lx=repr(cx * a), Ly=repr(cy * a), Lz=repr(cz * a), rho=RHO, rc=RC,
                             integrator="GpuLeapfrog" if gpu else "Leapfrog", container="GpuLinkedCells" if gpu else "LinkedCells",
                             container_options="<devices>%d</devices>" % devices if gpu and devices > 1 else "")


MARDYN_XML = """<?xml version='1.0' encoding='UTF-8'?>
<mardyn version="20100525">
  <refunits type="SI"><length unit="nm">0.1</length><mass unit="u">1</mass><energy unit="K">1</energy></refunits>
  <simulation type="MD">
    <integrator type="{integrator}"><timestep unit="reduced">{dt}</timestep></integrator>
    <run><currenttime>0</currenttime><production><steps>{steps}</steps></production></run>
    <ensemble type="NVT">
      <temperature unit="reduced">{temp}</temperature>
      <domain type="box"><lx>{Lx}</lx><ly>{Ly}</ly><lz>{Lz}</lz></domain>
      <components>
        <moleculetype id="1" name="1CLJ">
          <site type="LJ126" id="1" name="LJ"><coords><x>0.0</x><y>0.0</y><z>0.0</z></coords><mass>1.0</mass><sigma>1.0</sigma><epsilon>1.0</epsilon><shifted>false</shifted></site>
        </moleculetype>
      </components>
      <phasespacepoint>
        <generator name="MultiObjectGenerator">
          <objectgenerator>
            <filler type="GridFiller">
              <lattice system="cubic" centering="face">
                <vec id="a"><x>1</x><y>0</y><z>0</z></vec><vec id="b"><x>0</x><y>1</y><z>0</z></vec><vec id="c"><x>0</x><y>0</y><z>1</z></vec>
              </lattice>
              <basis><site><componentid>0</componentid><coordinate><x>0.25</x><y>0.25</y><z>0.25</z></coordinate></site></basis>
              <latticeOccupancy>1</latticeOccupancy>
              <density>{rho}</density>
            </filler>
            <object type="Cuboid"><lower><x>0</x><y>0</y><z>0</z></lower><upper><x>{Lx}</x><y>{Ly}</y><z>{Lz}</z></upper></object>
            <velocityAssigner type="EqualVelocityDistribution"></velocityAssigner>
          </objectgenerator>
        </generator>
      </phasespacepoint>
    </ensemble>
    <algorithm>
      <parallelisation type="DomainDecomposition"></parallelisation>
      <datastructure type="{container}"><cellsInCutoffRadius>1</cellsInCutoffRadius>{container_options}</datastructure>
      <cutoffs type="CenterOfMass"><defaultCutoff unit="reduced">{rc}</defaultCutoff><radiusLJ unit="reduced">{rc}</radiusLJ></cutoffs>
      <electrostatic type="ReactionField"><epsilon>1.0e+10</epsilon></electrostatic>
    </algorithm>
    <output></output>
  </simulation>
</mardyn>
"""


def host_cores():
    try:
        import psutil
        n = psutil.cpu_count(logical=False) or os.cpu_count()
    except Exception:
        n = os.cpu_count()
    try:
        n = min(n, len(os.sched_getaffinity(0)))
    except Exception:
        pass
    return max(int(n), 1)


def reference_binary():
    """oracle/_ref/MarDyn (AVX2 build); MarDyn_avx512 when present and the host CPU has avx512f"""
    ref = os.path.join(ROOT, "oracle", "_ref")
    cands = []
    try:
        flags = open("/proc/cpuinfo").read()
    except OSError:
        flags = ""
    if " avx512f" in flags:
        cands.append(os.path.join(ref, "MarDyn_avx512"))
    cands.append(os.path.join(ref, "MarDyn"))
    for c in cands:
        if os.path.exists(c) and os.access(c, os.X_OK):
            return c
    return None


_PICKED = {}


def run_reference_sample(sample_cells, steps, threads, exe=None):
    """runs the unmodified reference on a 1CLJ fcc sample of the bench workload (same rho, T, rc, dt; NVT thermostat
    as every shipped config) with all host threads; returns dict(mmups, n, loop_s, force_s, binary).  When both the AVX2
    and the AVX-512 build can run on this host, a short trial picks the faster one (the stronger baseline)."""
    if exe is None:
        exe = reference_binary()
        if exe is None:
            return None
        other = os.path.join(os.path.dirname(exe), "MarDyn")
        if os.path.basename(exe) != "MarDyn" and os.path.exists(other):
            if "exe" not in _PICKED:
                trial = {}
                for cand in (exe, other):
                    try:
                        trial[cand] = run_reference_sample(32, 3, threads, exe=cand)["mmups"]
                    except Exception:
                        trial[cand] = 0.0
                _PICKED["exe"] = max(trial, key=trial.get)
            exe = _PICKED["exe"]
    with tempfile.TemporaryDirectory() as td:
        cfg = os.path.join(td, "cfg.xml")
        with open(cfg, "w") as f:
            f.write(mardyn_xml(steps, sample_cells, gpu=False))
        env = dict(os.environ, OMP_NUM_THREADS=str(threads), OMP_PROC_BIND="close", OMP_PLACES="cores")
        t0 = time.time()
        p = subprocess.run([exe, cfg, "--steps", str(steps)], cwd=td, env=env, capture_output=True, text=True, timeout=3000)
        wall = time.time() - t0
    out = p.stdout + p.stderr
    if p.returncode != 0:
        raise RuntimeError("reference MarDyn failed (%d): %s" % (p.returncode, out[-2000:]))
    n = int(re.search(r"globally inserted molecules:\s*(\d+)", out).group(1))
    loop_s = float(re.search(r"Computation in main loop took:\s*([0-9.eE+-]+)", out).group(1))
    m = re.search(r"Force calculation took:\s*([0-9.eE+-]+)", out)
    vec = re.search(r"VectorizedCellProcessor: using (\S+) intrinsics", out)
    return dict(mmups=n * steps / loop_s / 1e6, n=n, loop_s=loop_s, force_s=float(m.group(1)) if m else None,
                binary=os.path.basename(exe), simd=vec.group(1) if vec else "?", wall_s=wall)


MARDYN_GPU = os.path.join(ROOT, "ls1-mardyn_b200", "host", "_build", "MarDynGpu")


def run_mardyn_gpu(cells, steps, threads, timeout=1500, devices=1):
    """The reference-facing drop-in, end to end: ls1-MarDyn's own executable and unmodified time loop with
    <datastructure type="GpuLinkedCells"> + <integrator type="GpuLeapfrog"> (ls1-mardyn_b200/host/), on the same XML the
    reference arm runs (its MultiObjectGenerator fcc lattice, NVT).  HOST buffers on both sides: the molecules start as the
    reference's host Molecule objects (uploaded once, in prepare_start) and end as host Molecule objects again (full-state
    mirror download at the last step, inside the loop timer).  Returns dict(mmups over upload + loop, ...) or None when the
    binary was not built (needs the reference tree at build time).
    devices > 1: <devices>N</devices> - the one process owns N device blocks (DomainDecomposition's grid), one per GPU, and steps them
    as a group (ls1g_step_group); cells is then [nx, ny, nz] of the whole box."""
    if not (os.path.exists(MARDYN_GPU) and os.access(MARDYN_GPU, os.X_OK)):
        return None
    with tempfile.TemporaryDirectory() as td:
        cfg = os.path.join(td, "cfg.xml")
        with open(cfg, "w") as f:
            f.write(mardyn_xml(steps, cells, gpu=True, devices=devices))
        env = dict(os.environ, OMP_NUM_THREADS=str(threads), OMP_PROC_BIND="close", OMP_PLACES="cores")
        t0 = time.time()
        p = subprocess.run([MARDYN_GPU, cfg, "--steps", str(steps), "--final-checkpoint=0"], cwd=td, env=env, capture_output=True, text=True, timeout=timeout)
        wall = time.time() - t0
    out = p.stdout + p.stderr
    if p.returncode != 0 or "Exit code would have been" in out:
        raise RuntimeError("MarDynGpu failed (%d): %s" % (p.returncode, out[-2000:]))
    n = int(re.search(r"globally inserted molecules:\s*(\d+)", out).group(1))
    loop_s = float(re.search(r"Computation in main loop took:\s*([0-9.eE+-]+)", out).group(1))
    g = re.search(r"GpuLinkedCells: (\d+) device steps took ([0-9.eE+-]+) s; upload ([0-9.eE+-]+) s \((\d+) B\); (\d+) host mirror download\(s\) took ([0-9.eE+-]+) s \((\d+) B\); context creation ([0-9.eE+-]+) s", out)
    dev_s, up_s, up_b, mirrors, mir_s, mir_b = float(g.group(2)), float(g.group(3)), int(g.group(4)), int(g.group(5)), float(g.group(6)), int(g.group(7))
    ctx_s = float(g.group(8))
    fs = re.search(r"first steps \[s\] ([0-9.eE+-]+) ([0-9.eE+-]+) ([0-9.eE+-]+) ([0-9.eE+-]+), slowest step ([0-9.eE+-]+)", out)
    first_steps = [float(fs.group(k)) for k in range(1, 5)] if fs else None
    slowest = float(fs.group(5)) if fs else None
    total = loop_s + up_s
    return dict(mmups=n * steps / total / 1e6, n=n, steps=steps, loop_s=loop_s, upload_s=up_s, device_s=dev_s, mirror_s=mir_s, mirrors=mirrors,
                h2d_bytes=up_b, d2h_bytes=mir_b + 48 * steps, wall_s=wall, context_s=ctx_s, loop_only_mmups=n * steps / loop_s / 1e6,
                first_steps_s=first_steps, slowest_step_s=slowest)


def run_oracle_port_sample(sample_cells, steps):
    """fallback when oracle/_ref/MarDyn is absent: the scalar C restatement (oracle/ls1_oracle.c), 1 thread"""
    from oracle import oracle as O
    gen = _gen()
    ids, r, v = gen.fcc_block(sample_cells, rho=RHO, temperature=TEMP)
    _, L = gen.fcc_box(sample_cells, RHO)
    n = len(ids)
    sysm = O.System.single_lj(rc=RC)
    g = O.grid([0, 0, 0], L, RC, 0)
    cid = np.zeros(n, dtype=np.uint32)
    q = np.tile([1.0, 0, 0, 0], (n, 1))
    z = np.zeros((n, 3))
    t0 = time.time()
    O.run(sysm, g, L, cid, r, v, q, z, z, z, steps, DT)
    dt = time.time() - t0
    return dict(mmups=n * steps / dt / 1e6, n=n, loop_s=dt)


def cpu_baseline(sample_cells=64, steps=10):
    """reported CPU baseline: the reference on a bounded sample (4*64^3 = 1 048 576 molecules, 10 steps)"""
    threads = host_cores()
    res = run_reference_sample(sample_cells, steps, threads)
    if res is not None:
        return {"value": round(res["mmups"], 4), "unit": "MMUPS", "cores": threads, "kind": "reference",
                "sample": "%s (%s), 1CLJ fcc %d molecules x %d steps, OpenMP %d threads, loop %.2f s (force %.2f s)"
                          % (res["binary"], res["simd"], res["n"], steps, threads, res["loop_s"], res["force_s"] or -1)}
    res = run_oracle_port_sample(16, 3)
    return {"value": round(res["mmups"], 4), "unit": "MMUPS", "cores": 1, "kind": "port",
            "sample": "oracle/ls1_oracle.c scalar port, 1CLJ fcc %d molecules x 3 steps" % res["n"]}


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = host_cores()
    steps = max(args.steps, 1)
    sample_cells = args.ref_cells
    res = run_reference_sample(sample_cells, steps, threads)
    kind = "reference"
    if res is None:
        sample_cells = 16
        res = run_oracle_port_sample(sample_cells, min(steps, 3))
        kind, threads, steps = "port", 1, min(steps, 3)
    val = res["mmups"]
    line = {"impl": "reference", "metric": METRIC, "value": round(val, 4), "unit": "MMUPS", "n_gpus": args.gpus, "steps": steps,
            "warmup": args.warmup, "ms_per_step": round(1e3 * res["loop_s"] / steps, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(sample_cells), "n_molecules": res["n"],
                       "ensemble": "NVT (velocity scaling every step): every shipped XML config is NVT and NVE cannot be selected from the reference's XML; "
                                   "the same XML drives the B200 arm's e2e figure (MarDynGpu), the B200 arm's device-resident `value` is NVE",
                       "input": "the reference's own MultiObjectGenerator fcc lattice (GridFiller)",
                       "warmup_note": "the reference's prepare_start force pass is its warm-up; its loop timer covers all steps"},
            "cpu_baseline": {"value": round(val, 4), "unit": "MMUPS", "cores": threads, "kind": kind,
                             "sample": "%s, %d molecules x %d steps" % (res.get("binary", "oracle port"), res["n"], steps)},
            "e2e": {"value": round(val, 4), "unit": "MMUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ----------------------------------------------------------------------------------------------------------
# the B200 arm
# ----------------------------------------------------------------------------------------------------------
def main_b200(args):
    import torch
    import torch.distributed as dist
    ls1g = _pkg()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node %d (one process per GPU)" % args.gpus)
        raise SystemExit("--gpus %d does not match WORLD_SIZE %d" % (args.gpus, world))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    cells_global = global_cells(world, args.cells)
    shard = make_shard(cells_global, world, rank)
    n_local = len(shard["id"])
    vmean = global_mean_velocity(shard["v"], dist if world > 1 else None)
    shard["v"] -= vmean
    nt = torch.tensor([float(n_local)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(nt)
    n_global = int(nt.item())

    cap = int(1.35 * n_local) + 200000
    c = ls1g.Container(shard["box_max"], RC, box_min=shard["box_min"], global_length=shard["L"], rank_grid=shard["grid"],
                       rank_coords=shard["coords"], device=local, capacity=cap)
    c.set_components([ls1g.lj_component()])
    if world > 1:
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt = torch.tensor(list(ls1g.comm_unique_id()), dtype=torch.uint8, device="cuda")
        dist.broadcast(idt, 0)
        c.comm_init(world, rank, bytes(idt.cpu().tolist()))

    # pinned host mirror (the e2e arm moves these every step)
    def pinned(shape, dtype):
        return torch.empty(shape, dtype=dtype, pin_memory=True).numpy()
    host = dict(id=pinned((cap,), torch.int64).view(np.uint64), cid=pinned((cap,), torch.int32).view(np.uint32),
                r=pinned((cap, 3), torch.float64), v=pinned((cap, 3), torch.float64), F=pinned((cap, 3), torch.float64))
    host["id"][:n_local] = shard["id"]
    host["cid"][:] = 0
    host["r"][:n_local] = shard["r"]
    host["v"][:n_local] = shard["v"]
    del shard["r"], shard["v"]

    c.add_particles(host["id"][:n_local], host["cid"][:n_local], host["r"][:n_local], host["v"][:n_local])
    c.exchange()
    u0, w0 = c.traverse_cells()
    stream = torch.cuda.ExternalStream(c.stream(), device=torch.device("cuda", local))

    # ---- device-resident loop ------------------------------------------------------------------------------
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
        t_wait = time.time()
        while not clocks.rows and time.time() - t_wait < 10.0:     # the first nvidia-smi query is slow on a cold box:
            time.sleep(0.05)                                       # let it finish before anything is timed
    if args.warmup > 0:
        c.run(args.warmup, DT)
    # ---- timed region: exactly K steps between barrier + synchronize, CUDA events on the container's stream, max over ranks.
    # The window is repeated (--windows, default 3; each is exactly K consecutive steps of the same trajectory) and `value` is
    # the MEDIAN window: a neighbour-list rebuild falls into a window or not (one every ~30 steps at this state point), which
    # moves a 20-step window by ~15 %; the spread and the rebuilds per window are reported next to it.
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    windows, tm, rec, clk_rows = [], None, None, []
    for w in range(max(args.windows, 1)):
        c.timers(reset=True)
        rb0 = c.list_stats()["rebuilds"]
        barrier()
        tw0 = time.perf_counter()
        e0.record(stream)
        rec_w = c.run(args.steps, DT)
        e1.record(stream)
        barrier()
        tw1 = time.perf_counter()
        ms_w = e0.elapsed_time(e1)
        mt = torch.tensor([ms_w], dtype=torch.float64, device="cuda")
        per_rank = [ms_w]
        if world > 1:
            gathered = [torch.zeros_like(mt) for _ in range(world)]
            dist.all_gather(gathered, mt)
            per_rank = [float(g.item()) for g in gathered]
        windows.append(dict(ms=max(per_rank), per_rank_ms=per_rank, rebuilds=c.list_stats()["rebuilds"] - rb0, tm=c.timers(reset=True), rec=rec_w,
                            wall=(tw0, tw1)))
    order = sorted(range(len(windows)), key=lambda k: windows[k]["ms"])
    med = windows[order[len(order) // 2]]
    ms_max, tm, rec = med["ms"], med["tm"], med["rec"]
    if rank == 0:
        clocks.window(windows[0]["wall"][0], windows[-1]["wall"][1])
    clk = clocks.stop() if rank == 0 else None
    n_pairs = c.last_pair_count()
    value = n_global * args.steps / (ms_max * 1e-3) / 1e6
    # ---- amortised over whole rebuild periods (reported, not the headline): one long window
    c.timers(reset=True)
    rb0 = c.list_stats()["rebuilds"]
    barrier()
    e0.record(stream)
    c.run(args.long_steps, DT, records=False)
    e1.record(stream)
    barrier()
    mt = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(mt, op=dist.ReduceOp.MAX)
    long_ms = float(mt.item())
    tm_long = c.timers(reset=True)
    amortised = {"steps": args.long_steps, "ms_per_step": round(long_ms / args.long_steps, 4), "list_rebuilds": c.list_stats()["rebuilds"] - rb0,
                 "value": round(n_global * args.long_steps / (long_ms * 1e-3) / 1e6, 2), "unit": "MMUPS",
                 "kernel_ms_per_step": {k: round(t["ms"] / args.long_steps, 4) for k, t in tm_long.items()},
                 "what": "one window of %d consecutive steps (several list-rebuild periods): the steady-state rate a long run sees" % args.long_steps}

    # energy conservation over the timed steps (NVE): reported, not timed
    et = torch.tensor(np.concatenate([rec[:, 0], 0.5 * rec[:, 2]]), dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(et)
    et = et.cpu().numpy()
    etot = et[:args.steps] + et[args.steps:]
    drift = float(abs(etot[-1] - etot[0]) / abs(etot[0])) if args.steps > 1 else 0.0

    # ---- roofline of the pair kernel ---------------------------------------------------------------------------
    cpd, _ = c.get_grid()
    c.exchange()                                      # halo copies back for the cell census (not timed)
    cells_all = c.download(("cell",), include_halo=True)["cell"]
    n_dist = distance_checks(cells_all, cpd)
    fc = c.flop_count()                               # the reference's FlopCounter on the same cells, counted on the device
    c.traverse_cells()
    n_pairs = c.last_pair_count()
    del cells_all
    # ls1g_run produces Vi on the last step of a run only (nothing can read it in between): K-1 launches at 35, one at 41 flop/pair
    vi_all = bool(int(os.environ.get("LS1G_VI_EVERY_STEP", "0") or 0))
    k_t = max(args.steps, 1)
    flops = lj_flops(n_dist, n_pairs, True) if vi_all else (lj_flops(n_dist, n_pairs, False) * (k_t - 1) + lj_flops(n_dist, n_pairs, True)) / k_t
    # pair-kernel time PER STEP (the overlapped multi-rank step launches the kernel twice: inner blocks, then boundary blocks)
    force_ms = tm["force"]["ms"] / max(args.steps, 1)
    traffic, traffic_src = None, None
    prof = os.path.join(ROOT, "profiles", "r02h_ncu_nl_force_10M_summary.csv")
    if os.path.exists(prof) and args.cells == CELLS_PER_GPU:      # dram bytes of ONE launch from the committed ncu --set full capture
        vals = {}
        for ln in open(prof):
            f = ln.strip().split(",")
            if len(f) >= 3 and f[0].startswith("dram__bytes_"):
                vals[f[0]] = float(f[1]) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}.get(f[2], 1.0)
        if len(vals) == 2:
            traffic, traffic_src = sum(vals.values()), "profiles/r02h_ncu_nl_force_10M_summary.csv (ncu --set full, one launch, same workload)"
    peak = ls1g.fp64_peak_tflops(local) if rank == 0 else 0.0
    achieved = flops / (force_ms * 1e-3) / 1e12 if force_ms > 0 else 0.0
    launches = sum(sum(t["launches"] for t in w["tm"].values()) for w in windows)      # all timed windows

    # ---- e2e, C ABI with host buffers: every step moves the FULL state host -> device -> host (worst case of the boundary) ----
    tparts = np.zeros(4)
    def e2e_step(n, acc=None):
        t = [time.perf_counter()]
        c.add_particles(host["id"][:n], host["cid"][:n], host["r"][:n], host["v"][:n]); t.append(time.perf_counter())
        c.set_forces(host["F"][:n]); t.append(time.perf_counter())
        c.run(1, DT, records=True); t.append(time.perf_counter())
        out = c.download(("id", "r", "v", "F"), out=host); t.append(time.perf_counter())
        if acc is not None:
            acc += np.diff(t)
        return len(out["id"])
    c.delete_outer_particles()
    n_cur = len(c.download(("id", "r", "v", "F"), out=host)["id"])
    for _ in range(max(1, min(args.warmup, 3))):
        n_cur = e2e_step(n_cur)
    k_e2e = max(1, min(args.steps, args.e2e_steps))
    h2d = d2h = 0
    barrier()
    t0 = time.perf_counter()
    e0.record(stream)
    for _ in range(k_e2e):
        h2d += n_cur * (8 + 4 + 24 + 24 + 24)
        n_cur = e2e_step(n_cur, tparts)
        d2h += n_cur * (8 + 24 + 24 + 24) + 48
    e1.record(stream)
    barrier()
    ms_e2e = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3)   # host-driven: wall clock and device clock agree
    mt = torch.tensor([ms_e2e], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(mt, op=dist.ReduceOp.MAX)
    e2e_val = n_global * k_e2e / (float(mt.item()) * 1e-3) / 1e6
    roundtrip = {"value": round(e2e_val, 2), "unit": "MMUPS", "h2d_bytes_per_step": h2d // k_e2e, "d2h_bytes_per_step": d2h // k_e2e, "steps": k_e2e,
                 "ms_per_step": {k: round(1e3 * v / k_e2e, 2) for k, v in zip(("upload", "upload_forces", "run_1_step", "download"), tparts)},
                 "what": "C ABI, full state both ways EVERY step: add_particles(id,cid,r,v)+set_forces(F) from pinned host, ls1g_run(1), download(id,r,v,F) to pinned host"}
    e2e = dict(roundtrip)
    c_list_stats = c.list_stats()
    c.close()

    # ---- e2e, the reference-facing drop-in: ls1-MarDyn's own executable + time loop with the GpuLinkedCells container ----------
    # N > 1: one MarDynGpu process owns all N GPUs (<devices>N</devices>, the blocks of DomainDecomposition's grid stepped as a group,
    # ls1g_step_group).  The other ranks have released their devices by then: the process group is left first, rank 0 runs it alone.
    md = None
    if world > 1:
        del host
        torch.cuda.empty_cache()
        barrier()
        dist.destroy_process_group()
    if rank == 0 and not args.no_mardyn_gpu:
        try:
            cells_e2e = args.e2e_cells or args.cells
            md = run_mardyn_gpu(cells_e2e if world == 1 else global_cells(world, cells_e2e), args.steps, host_cores(), devices=world)
        except Exception as ex:                                  # reported, never silently replaced
            e2e["mardyn_gpu_error"] = str(ex)[-400:]
    if md is not None:
        e2e = {"value": round(md["mmups"], 2), "unit": "MMUPS", "h2d_bytes_per_step": md["h2d_bytes"] // md["steps"],
               "d2h_bytes_per_step": md["d2h_bytes"] // md["steps"], "steps": md["steps"], "n_molecules": md["n"],
               "what": "MarDynGpu = the reference's executable and unmodified Simulation::simulate loop with <datastructure type=GpuLinkedCells> + "
                       "<integrator type=GpuLeapfrog> (ls1-mardyn_b200/host/): host Molecule objects -> device (once) -> %d steps, scalars to Domain every step "
                       "-> full state back into host Molecule objects at the last step; time = upload + the reference's own loop timer%s" % (md["steps"], "" if world == 1 else "; ONE process drives all %d GPUs (<devices>%d</devices>)" % (world, world)),
               "seconds": {"upload": round(md["upload_s"], 3), "loop": round(md["loop_s"], 3), "of_loop_device_calls": round(md["device_s"], 3),
                           "of_loop_final_mirror": round(md["mirror_s"], 3), "process_wall": round(md["wall_s"], 1),
                           "cuda_context_creation_not_counted": round(md["context_s"], 3),
                           "device_call_of_steps_1_to_4": md["first_steps_s"], "slowest_device_call": md["slowest_step_s"]},
               "loop_only_mmups": round(md["loop_only_mmups"], 2), "c_abi_full_state_roundtrip_every_step": roundtrip}

    if rank == 0:
        cb = cpu_baseline() if (world == 1 and not args.no_cpu_baseline) else None
        line = {"metric": METRIC, "value": round(value, 2), "unit": "MMUPS", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": round(ms_max / args.steps, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_name(args.cells), "n_molecules": n_global,
                           "ensemble": "NVE (`value`, ls1g_run); `e2e` runs the reference arm's XML, i.e. NVT velocity scaling every step",
                           "rank_grid": shard["grid"], "cells_per_dim": cpd,
                           "l2": "inputs larger than L2 (pos+vel+F = %.0f MB per GPU)" % (n_local * 80 / 1e6),
                           "nve_rel_energy_drift_over_timed_steps": drift, "upot_per_molecule_t0": u0 / max(n_local, 1),
                           "nve_note": "cfg B is the UNSHIFTED LJ: the energy jumps whenever a pair crosses rc, so this drift is the reference's own behaviour "
                                       "(reference: -1.905e-3 over 1000 steps at 1 048 576 molecules), not an integration error; with the shifted potential the same "
                                       "loop drifts 1.1e-6 over 1000 steps (reference: -1.37e-6): profiles/r01_nve_1000_steps.json, profiles/r02_ref_nve_1000_steps.json, "
                                       "tests/test_gpu_parity.py::test_nve_1000_steps_against_the_reference_record"},
                "e2e": e2e,
                "gpu_launches": int(launches),
                "roofline": {"bound": "fp64", "kernel": "nl_force_kernel (lists.cuh)", "achieved": round(achieved, 3), "peak": round(peak, 3), "unit": "TFLOP/s",
                             "frac": round(achieved / peak, 4) if peak > 0 else None, "traffic": traffic, "traffic_source": traffic_src,
                             "peak_source": "measured here: dependent-free DFMA kernel (ls1g_measure_fp64_peak); MEASURED_PEAKS.json has no FP64 figure",
                             "flops_per_launch": flops, "n_dist": n_dist, "n_pairs": n_pairs, "kernel_ms": round(force_ms, 4),
                             "flopcounter": {"flops": fc["flops"], "molecule_distances": fc["molecule_distances"], "lj_kernel_calls": float(fc["kernel"][0]),
                                             "lj_macro_calls": float(fc["macro"][0]),
                                             "note": "ls1g_flop_count = the reference's FlopCounter as a device cell processor (bit-equal to the reference on the golden fixtures); "
                                                     "it visits a pair across the periodic boundary from both images, so its total lies a few % above the Newton-3 count "
                                                     "(every pair once) that flops_per_launch and frac are quoted on"},
                             "model": "FlopCounter: 8*N_dist + 41*N_pairs with per-molecule virial (last step of the run), 8*N_dist + 35*N_pairs without (the other steps; the reference's own count for that case); Newton-3 counts, the owner-computes kernel executes ~2x these",
                             "note": "bound named by SURVEY 8d is the FP64 FMA pipe, not HBM/tensor.  ncu of this owner-computes kernel (it executes 2x the Newton-3 flops the fraction is quoted on): sm__pipe_fp64_cycles_active 55 %, shared-memory wavefronts 79 % of peak (the random 8 + 16 byte tile reads per listed partner, half of them bank-conflict replays), DRAM 3.3 GB per launch = 1.9 TB/s = 29 % of the copy peak (profiles/r02h_ncu_nl_force_10M_summary.csv)"},
                "timed_windows": {"n": len(windows), "value_is": "median window", "ms": [round(w["ms"], 3) for w in windows],
                                  "list_rebuilds": [int(w["rebuilds"]) for w in windows],
                                  "spread_rel": round((max(w["ms"] for w in windows) - min(w["ms"] for w in windows)) / ms_max, 4),
                                  "per_rank_ms_of_median_window": {"min": round(min(med["per_rank_ms"]), 3), "max": round(max(med["per_rank_ms"]), 3)}},
                "amortised": amortised,
                "neighbour_lists": c_list_stats,
                "kernel_ms_per_step": {k: round(t["ms"] / args.steps, 4) for k, t in tm.items()},
                "hbm_kernels": hbm_kernels(tm, args.steps, n_local),
                "clocks": clk}
        if cb is not None:
            line["cpu_baseline"] = cb
        print(json.dumps(line), flush=True)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cells", type=int, default=CELLS_PER_GPU, help="fcc unit cells per dimension per GPU")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--windows", type=int, default=3, help="repeats of the timed K-step window (value = the median window)")
    ap.add_argument("--long-steps", type=int, default=120, help="length of the extra window that spans whole list-rebuild periods (reported as `amortised`)")
    ap.add_argument("--ref-cells", type=int, default=CELLS_PER_GPU, help="fcc unit cells per dimension of the reference arm (default: the bench workload itself)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-mardyn-gpu", action="store_true", help="skip the MarDynGpu end-to-end arm (e2e then = the C-ABI round trip)")
    ap.add_argument("--e2e-cells", type=int, default=0, help="fcc unit cells per dimension of the MarDynGpu end-to-end run (0 = --cells)")
    ap.add_argument("--workload", default="1clj", choices=["1clj", "argon", "co2", "eox", "co2q"],
                    help="1clj = the headline workload (cfg B); the others are the reference's shipped examples replicated (bench_examples.py: "
                         "cfg A argon, cfg C co2 / co2q, cfg D eox), same JSON line, same launch line")
    ap.add_argument("--replicas", type=int, default=8, help="example workloads: replicas per dimension per GPU (8 -> 1 048 576 molecules)")
    args = ap.parse_args()
    if args.workload != "1clj":
        import bench_examples
        fwd = ["--workload", args.workload, "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup),
               "--replicas", str(args.replicas), "--impl", args.impl]
        if args.no_cpu_baseline:
            fwd.append("--no-cpu-baseline")
        if args.no_mardyn_gpu:
            fwd.append("--no-e2e")
        return bench_examples.main(fwd)
    if args.impl == "reference":
        return main_reference(args)
    return main_b200(args)


if __name__ == "__main__":
    sys.exit(main())
