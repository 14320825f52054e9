#!/bin/bash
# One `ncu --set full` capture of the N=1 update kernel of C2 / C3 mp-SGD / C3 mp-Adam (4th launch of
# the run), raw pages written to gpurun_out/, then profiles/traffic_<tag>.json stamped with the hash of
# the kernel sources (bench.py quotes `roofline.traffic` only while that hash matches).
#   gpurun -- 'bash tools/capture_traffic.sh r2b'      (then copy gpurun_out/<tag>_*_full_raw.csv and
#   gpurun_out/traffic_<tag>.json into profiles/)
tag=${1:-r2b}
mkdir -p gpurun_out
for w in resnet50 bert_base_sgd bert_base_adam; do
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:mta_kernel --launch-skip 3 --launch-count 1 \
    -f -o /tmp/${tag}_${w} python bench.py --workload $w --no-configs --no-cpu-baseline --no-parity --steps 3 --warmup 3 --e2e-steps 1 \
    > gpurun_out/${tag}_${w}_ncu.log 2>&1
  ncu -i /tmp/${tag}_${w}.ncu-rep --page raw --csv > gpurun_out/${tag}_${w}_full_raw.csv 2>/dev/null
done
python - "$tag" <<'PY'
import csv, json, sys
sys.path.insert(0, ".")
import bench
tag = sys.argv[1]
out = {}
units = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
tunits = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}
for w in ("resnet50", "bert_base_sgd", "bert_base_adam"):
    rows = list(csv.reader(open("gpurun_out/%s_%s_full_raw.csv" % (tag, w))))
    hdr, un, r = rows[0], rows[1], rows[2]
    def val(name, table):
        i = hdr.index(name)
        return float(r[i].replace(",", "")) * table[un[i]]
    out[w] = {"traffic": val("dram__bytes_read.sum", units) + val("dram__bytes_write.sum", units),
              "kernel": r[hdr.index("Kernel Name")],
              "duration_ms_under_ncu": val("gpu__time_duration.sum", tunits)}
out["kernel_source_sha16"] = bench.kernel_source_hash()
out["how"] = ("ncu --set full --clock-control none, one launch of the update kernel (4th launch of the run); "
              "dram__bytes_read.sum + dram__bytes_write.sum (tools/capture_traffic.sh)")
json.dump(out, open("gpurun_out/traffic_%s.json" % tag, "w"), indent=1)
print(json.dumps(out, indent=1))
PY
