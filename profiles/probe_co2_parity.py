import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tests.helpers import load, vec_rel_err, initial_state
from tests.gpu_helpers import ls1g, container_of, by_id
for name in ("ex_CO2","ex_CO2_novec","ex_CO2Merker","ex_CO2Merker_novec","ex_EOX"):
    d=load(name); st=initial_state(d)
    with container_of(d) as c:
        c.add_particles(st["id"], st["cid"], st["r"], st["v"], st["q"], st["L"]); c.exchange()
        u,w=c.traverse_cells(); out=by_id(c.download(("id","F","M","Vi")))
        steps, dt = int(d["meta"][2]), float(d["dt"][0])
        series=c.run(steps, dt)[:, :4]
        post=by_id(c.download(("id","r","v","q","L","F","M")))
    own=d["f_owned"].astype(bool); order=np.argsort(d["f_id"][own])
    print(name, "F %.2e M %.2e Vi %.2e upot %.2e virial %.2e" % (vec_rel_err(out["F"], d["f_F"][own][order]), vec_rel_err(out["M"], d["f_M"][own][order]), vec_rel_err(out["Vi"], d["f_Vi"][own][order]), abs(u-d["upot_virial"][0])/abs(d["upot_virial"][0]), abs(w-d["upot_virial"][1])/abs(d["upot_virial"][1])),
          "| after %d steps:" % steps, " ".join("%s %.1e" % (k, np.abs(post[k]-d["post_"+k]).max()/max(np.abs(d["post_"+k]).max(),1e-300)) for k in ("r","v","q","L")), "F %.1e M %.1e" % (vec_rel_err(post["F"], d["post_F"]), vec_rel_err(post["M"], d["post_M"])),
          "series %.1e" % (np.abs(series-d["series"])/np.maximum(np.abs(d["series"]).max(axis=0),1e-300)).max())
