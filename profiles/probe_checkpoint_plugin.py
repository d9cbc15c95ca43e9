"""MarDynGpu on the bench workload (cfg B) writing binary checkpoints during the run, two ways:
  records : <outputplugin name="GpuCheckpointWriter">  - packed records straight from the device (ls1g_download_records)
  mirror  : <outputplugin name="CheckpointWriter"><type>binary</type>  - the reference's writer over host Molecule objects, which the
            container refills on those steps (mirror steps)
Prints one JSON object with the reference's loop timer, the container's own summary and the checkpoint comparison.
Usage (GPU box): python profiles/probe_checkpoint_plugin.py [cells] [steps] [writefrequency] > gpurun_out/r03_checkpoint_plugin.json"""
import json, os, re, subprocess, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
cells = int(sys.argv[1]) if len(sys.argv) > 1 else 136
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 25
freq = int(sys.argv[3]) if len(sys.argv) > 3 else 10
PLUGINS = dict(records='<outputplugin name="GpuCheckpointWriter"><writefrequency>%d</writefrequency><outputprefix>cp</outputprefix></outputplugin>' % freq,
               mirror='<outputplugin name="CheckpointWriter"><type>binary</type><writefrequency>%d</writefrequency><outputprefix>cp</outputprefix></outputplugin>' % freq)
RECORD = np.dtype([("id", "<u8"), ("cid", "<u4"), ("r", "<f8", 3), ("v", "<f8", 3), ("q", "<f8", 4), ("D", "<f8", 3)])
out = dict(cells=cells, steps=steps, writefrequency=freq)
recs = {}
with tempfile.TemporaryDirectory() as td:
    for way, plugin in PLUGINS.items():
        d = os.path.join(td, way)
        os.mkdir(d)
        xml = bench.mardyn_xml(steps, cells, True)
        assert "<output></output>" in xml
        open(os.path.join(d, "cfg.xml"), "w").write(xml.replace("<output></output>", "<output>%s</output>" % plugin))
        env = dict(os.environ, OMP_NUM_THREADS=str(bench.host_cores()), OMP_PROC_BIND="close", OMP_PLACES="cores")
        t0 = time.perf_counter()
        p = subprocess.run([bench.MARDYN_GPU, "cfg.xml", "--steps", str(steps), "--final-checkpoint=0"], cwd=d, env=env, capture_output=True, text=True)
        log = p.stdout + p.stderr
        r = dict(returncode=p.returncode, wall_s=round(time.perf_counter() - t0, 2))
        m = re.search(r"Computation in main loop took:\s*([0-9.eE+-]+)", log)
        r["main_loop_s"] = float(m.group(1)) if m else None
        m = re.search(r"GpuLinkedCells: (\d+) device steps took ([0-9.eE+-]+) s; upload ([0-9.eE+-]+) s .*?; (\d+) host mirror download\(s\) took ([0-9.eE+-]+) s", log)
        if m:
            r.update(device_steps=int(m.group(1)), device_s=float(m.group(2)), upload_s=float(m.group(3)), mirrors=int(m.group(4)), mirror_s=float(m.group(5)))
        m = re.search(r"(\d+) record checkpoint\(s\) from the device took ([0-9.eE+-]+) s \((\d+) B\)", log)
        if m:
            r.update(record_checkpoints=int(m.group(1)), record_checkpoints_s=float(m.group(2)), record_bytes=int(m.group(3)))
        if p.returncode != 0:
            r["log_tail"] = log[-1500:]
        r["files"] = sorted(f for f in os.listdir(d) if f.startswith("cp-"))
        out[way] = r
        f = os.path.join(d, "cp-1.restart.dat")
        if os.path.exists(f):
            a = np.fromfile(f, dtype=RECORD)
            recs[way] = a[np.argsort(a["id"])]
            out[way]["checkpoint_1_records"] = len(a)
            out[way]["checkpoint_1_header"] = open(os.path.join(d, "cp-1.restart.header.xml")).read()
if len(recs) == 2:
    a, b = recs["records"], recs["mirror"]
    out["checkpoint_1_same_ids"] = bool(np.array_equal(a["id"], b["id"]))
    out["checkpoint_1_max_abs_diff"] = {k: float(np.abs(a[k] - b[k]).max()) for k in ("r", "v", "q", "D")}
    out["checkpoint_1_headers_equal"] = out["records"].pop("checkpoint_1_header") == out["mirror"].pop("checkpoint_1_header")
print(json.dumps(out))
