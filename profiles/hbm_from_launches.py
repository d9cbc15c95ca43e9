"""Achieved HBM GB/s of the memory-bound kernels from an ncu launch list (gpu__time_duration per launch):
   python profiles/hbm_from_launches.py <launches.csv> <n_slots> <n_owned> [peak_GBs]
ALGORITHMIC bytes per launch (slots = owned molecules + halo copies at that launch):
  bin_kernel        read pos 32 + write cell 4 + slot key 8/record                                    44 B x slots
  scatter_kernel    read pos 32 + cell 4 + id 8, write slot records 17                                61 B x slots
  gather_kernel     read records 17 + payload (pos 32 + vel 24 + id 8), write payload + cell 4       149 B x slots
  integrate <0,1,0> (pre)   read pos 32 + vel 24 + F 24 + cell 4 (+ build pos 32 in list mode), write pos 32 + vel 24    140 B x owned
  integrate <1,0,0> (post)  read pos(cid) 32 + vel 24 + F 24 + cell 4, write vel 24                                      108 B x owned
  halo_refresh      read cell 4 per slot; per halo copy src 4 + code 1 + source pos 32, write pos 32                     4 B x slots + 69 B x halo
  pack_kernel       (multi-GPU, per dimension) read cell 4 + pos 32 per slot, 64-B record per exchanged molecule         36 B x slots (+ records)
  pack_pos / compact kernels of the host transfer: 24..32 B in + out per molecule
Times come from a serialised cold-cache profile, so they are upper bounds of the in-loop times; compare with the peak."""
import collections, csv, io, sys
path, slots, owned = sys.argv[1], float(sys.argv[2]), float(sys.argv[3])
peak = float(sys.argv[4]) if len(sys.argv) > 4 else 6455.9
halo = slots - owned
BYTES = {
    "bin_kernel": 44 * slots, "scatter_kernel": 61 * slots, "gather_kernel": 149 * slots,
    "void integrate_kernel<0, 1, 0>": 140 * owned, "void integrate_kernel<1, 0, 0>": 108 * owned, "void integrate_kernel<1, 1, 0>": 140 * owned,
    "halo_refresh_kernel": 4 * slots + 69 * halo, "pack_kernel": 36 * slots, "pack_pos_kernel": (24 + 4 + 32) * owned,
    "void compact_f64_kernel<3>": (24 + 24 + 4) * owned, "compact_pos_kernel": (32 + 24 + 4) * owned,
}
rows = [l for l in open(path) if l.startswith('"')]
t = collections.defaultdict(list)
for x in csv.DictReader(io.StringIO("".join(rows))):
    k = x["Kernel Name"].split("(")[0]
    t[k].append(float(x["Metric Value"].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}[x["Metric Unit"]])
print("kernel,launches,median_us,algorithmic_MB_per_launch,achieved_GBs,frac_of_measured_copy_peak_%.0f_GBs" % peak)
for k, b in BYTES.items():
    if k in t:
        v = sorted(t[k]); med = v[len(v) // 2]
        gbs = b / (med * 1e-6) / 1e9
        print('"%s",%d,%.1f,%.1f,%.0f,%.3f' % (k, len(v), med, b / 1e6, gbs, gbs / peak))
