"""rebuild counts / energies of the multi-site list loop per fixture and time-step factor (diagnosis)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tests.helpers import load, initial_state
from tests.gpu_helpers import ls1g, container_of, by_id
for name in ("ex_CO2", "ex_EOX", "ex_CO2Merker", "per_MultiComponentMultiPotentials"):
    d = load(name); st = initial_state(d)
    for fac in (1.0, 2.0, 4.0):
        for mode in ("1", "0"):
            os.environ["LS1G_LIST_MULTISITE"] = mode
            try:
                with container_of(d) as c:
                    c.add_particles(st["id"], st["cid"], st["r"], st["v"], st["q"], st["L"]); c.exchange(); c.traverse_cells()
                    rec = c.run(60, fac * float(d["dt"][0]))
                    e = rec[:, 0] + 0.5 * rec[:, 2] + 0.5 * rec[:, 3]
                    print(name, "dt x", fac, "lists" if mode == "1" else "scan", "rebuilds", c.list_stats(), "E first/last %.6g %.6g" % (e[0], e[-1]), "finite", np.isfinite(rec).all(), flush=True)
            except Exception as ex:
                print(name, "dt x", fac, mode, "FAILED", str(ex)[:150], flush=True)
                break
