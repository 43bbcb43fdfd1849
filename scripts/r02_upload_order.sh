cd $GRAFT_REPO_ROOT
N=${1:-64000000}
NB_B200_AOS_PERMUTE_MIN=4000000000 python scripts/r02_upload_order.py $N 2>&1 | tail -3
python scripts/r02_upload_order.py $N > gpurun_out/upload_plain.log 2>&1 && tail -3 gpurun_out/upload_plain.log && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'tree_finish|integrate_resort|permute_from_aos' --csv --log-file gpurun_out/upload_order_$N.csv python scripts/r02_upload_order.py $N > gpurun_out/upload_ncu.log 2>&1
echo rc=$?
NB_B200_AOS_PERMUTE_MIN=4000000000 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'tree_finish|integrate_resort|permute_from_aos' --csv --log-file gpurun_out/upload_order_off_$N.csv python scripts/r02_upload_order.py $N > gpurun_out/upload_ncu_off.log 2>&1
echo rc=$?
python - <<PY
import csv, collections
for tag in ("upload_order_$N", "upload_order_off_$N"):
    rows = [l for l in open(f"gpurun_out/{tag}.csv") if not l.startswith("==")]
    agg = collections.OrderedDict()
    for r in csv.DictReader(rows):
        k = (r["ID"], r["Kernel Name"].split("(")[0].replace("void nb::", ""))
        v = float(r["Metric Value"].replace(",", "")) * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1e3, "ms": 1e6}.get(r["Metric Unit"], 1)
        agg.setdefault(k, {})[r["Metric Name"]] = v
    print(tag)
    for (i, k), m in agg.items():
        print(f"  {k:45s} {m['gpu__time_duration.sum'] / 1e3:9.1f} us  DRAM {(m['dram__bytes_read.sum'] + m['dram__bytes_write.sum']) / 1e9:7.2f} GB")
PY
