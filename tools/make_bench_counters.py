"""profiles/bench_kernel_counters.json: per-launch counters of the bench workload's kernel, taken from an
`ncu --set full` capture (bench.py divides them by its own live CUDA-event time).
usage: python tools/make_bench_counters.py <workload>/<variant> <report.ncu-rep> [more pairs ...]"""
import csv, io, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
path = os.path.join(ROOT, "profiles", "bench_kernel_counters.json")
data = json.load(open(path)) if os.path.isfile(path) else {}
for key, rep in zip(sys.argv[1::2], sys.argv[2::2]):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[0]
    tot = {"dram": 0.0, "inst": 0.0, "thr": 0.0, "wf": 0.0, "ms": 0.0}
    names = []
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        f = lambda k: float(d[k].replace(",", ""))
        unit = dict(zip(hdr, rows[1]))
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        tot["dram"] += f("dram__bytes_read.sum") * scale[unit["dram__bytes_read.sum"]] + \
            f("dram__bytes_write.sum") * scale[unit["dram__bytes_write.sum"]]
        tot["inst"] += f("smsp__inst_executed.sum")
        tot["thr"] += f("smsp__thread_inst_executed_per_inst_executed.ratio") * f("smsp__inst_executed.sum")
        # the raw page has no .sum of the data-pipe wavefronts: pct of peak (1 / cycle / SM) x elapsed SM cycles
        tot["wf"] += f("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed") / 100.0 * f("sm__cycles_elapsed.sum")
        tunit = unit["gpu__time_duration.sum"]
        tot["ms"] += f("gpu__time_duration.sum") * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}[tunit]
        names.append(d["Kernel Name"])
    data[key] = {"kernels": names, "dram_bytes_per_launch": tot["dram"], "inst_executed_per_launch": tot["inst"],
                 "lanes_per_inst": tot["thr"] / tot["inst"], "l1_wavefronts_per_launch": tot["wf"],
                 "ncu_duration_ms": tot["ms"], "report": os.path.basename(rep)}
    print(key, data[key])
json.dump(data, open(path, "w"), indent=1)
