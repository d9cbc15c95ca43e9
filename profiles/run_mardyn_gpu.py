"""Runs MarDynGpu (the reference executable with the GpuLinkedCells container) on the bench workload and prints its own
timing lines.  Usage (GPU box): python profiles/run_mardyn_gpu.py [cells] [steps] [mode]"""
import os, re, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
cells = int(sys.argv[1]) if len(sys.argv) > 1 else 136
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 50
mode = sys.argv[3] if len(sys.argv) > 3 else "fused"
with tempfile.TemporaryDirectory() as td:
    xml = bench.mardyn_xml(steps, cells, True)
    xml = xml.replace("<cellsInCutoffRadius>1</cellsInCutoffRadius>", "<cellsInCutoffRadius>1</cellsInCutoffRadius><mode>%s</mode>" % mode)
    open(os.path.join(td, "cfg.xml"), "w").write(xml)
    env = dict(os.environ, OMP_NUM_THREADS=str(bench.host_cores()), OMP_PROC_BIND="close", OMP_PLACES="cores")
    p = subprocess.run([bench.MARDYN_GPU, "cfg.xml", "--steps", str(steps), "--final-checkpoint=0"], cwd=td, env=env, capture_output=True, text=True)
out = p.stdout + p.stderr
for l in out.splitlines():
    if re.search(r"GpuLinkedCells|took:|Simulation speed|ls1g:|ERROR|Simstep = (1|%d)\b" % steps, l) and " 0 sec" not in l:
        print(l)
