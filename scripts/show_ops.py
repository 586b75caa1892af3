import json,sys
for f in sys.argv[1:]:
    for l in open(f):
        d=json.loads(l); print(f.split('/')[-1], round(d["ms_per_step"],2), round(d["rows_per_s"]/1e9,3), round(d["frac_of_measured_hbm"],3), [(k["kernel"], round(k["ms_per_step"],2), k["launches_per_step"]) for k in d["kernels"][:6]])
