"""Times the binary-record boundary (ls1g_upload_records / ls1g_download_records, csrc/records.cu) at the bench size against the
array boundary it stands next to: 10 061 824 ICRVQD records (1.17 GB, the .dat file of a checkpoint of cfg B) in page-locked memory
-> device -> back, and the same molecules as separate arrays, including the host-side (de)composition of the records that the array
path needs (numpy structured-array field copies: what phasespace.read_binary / write_binary do).
Usage (GPU box): python profiles/probe_records.py [cells] > gpurun_out/r03_records.json"""
import importlib, json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ls1g = importlib.import_module("ls1-mardyn_b200")
gen = importlib.import_module("ls1-mardyn_b200.generators")
P = ls1g.phasespace
cells = int(sys.argv[1]) if len(sys.argv) > 1 else 136
ids, r, v = gen.fcc_block(cells, rho=0.8442, temperature=1.44, seed=1)
_, L = gen.fcc_box(cells, 0.8442)
n = len(ids)
def pinned(nbytes): return torch.empty(nbytes, dtype=torch.uint8, pin_memory=True).numpy()
raw = pinned(n * 116)
rec = raw.view(P.BINARY_RECORD)
rec["id"] = ids; rec["cid"] = 1; rec["r"] = r; rec["v"] = v; rec["q"] = [1.0, 0.0, 0.0, 0.0]; rec["D"] = 0.0
arr = dict(id=pinned(n * 8).view(np.uint64), cid=pinned(n * 4).view(np.uint32), r=pinned(n * 24).view(np.float64).reshape(n, 3),
           v=pinned(n * 24).view(np.float64).reshape(n, 3))
back = pinned(n * 116)
out = dict(n=n, bytes=n * 116, runs=[])
def timed(f):
    t0 = time.perf_counter(); x = f(); return x, 1e3 * (time.perf_counter() - t0)
with ls1g.Container(list(L), 2.5) as c:
    c.set_components([ls1g.lj_component()])
    c.reserve(n)
    for it in range(3):
        T = {}
        (kept, mx), T["upload_records_ms"] = timed(lambda: c.add_records(raw))
        assert kept == n and mx == int(ids.max())
        _, T["exchange_ms"] = timed(c.exchange)
        got, T["download_records_ms"] = timed(lambda: c.download_records(out=back))
        def decode():
            arr["id"][:] = rec["id"]; arr["cid"][:] = rec["cid"] - 1; arr["r"][:] = rec["r"]; arr["v"][:] = rec["v"]
        _, T["host_decode_ms"] = timed(decode)
        _, T["upload_arrays_ms"] = timed(lambda: c.add_particles(arr["id"], arr["cid"], arr["r"], arr["v"]))
        c.exchange()
        d, T["download_arrays_ms"] = timed(lambda: c.download(("id", "cid", "r", "v"), out=arr))
        def encode():
            b = back.view(P.BINARY_RECORD)
            b["id"] = d["id"]; b["cid"] = d["cid"] + 1; b["r"] = d["r"]; b["v"] = d["v"]; b["q"] = [1.0, 0.0, 0.0, 0.0]; b["D"] = 0.0
        _, T["host_encode_ms"] = timed(encode)
        T["upload_records_GBps"] = n * 116 / T["upload_records_ms"] / 1e6
        T["download_records_GBps"] = n * 116 / T["download_records_ms"] / 1e6
        out["runs"].append({k: round(x, 3) for k, x in T.items()})
    o = np.argsort(got.view(P.BINARY_RECORD)["id"].ravel())
    assert got[o].tobytes() == raw.reshape(n, 116)[np.argsort(ids)].tobytes()
    out["round_trip_byte_identical"] = True
print(json.dumps(out))
