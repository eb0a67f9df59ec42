"""Summarise gpurun_out/ncu_northstar.csv (tools/ncu_northstar.sh): per kernel of one C2 step, the three
counters BASELINE.json's north_star names, each against its B200 peak.
usage: python tools/ncu_northstar.py gpurun_out/ncu_northstar.csv [hbm_peak_GBs]"""
import csv
import json
import os
import sys

path = sys.argv[1]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
try:
    hbm_peak = float(sys.argv[2]) if len(sys.argv) > 2 else json.load(open(os.path.join(root, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    hbm_peak = 6550.4
lines = [l for l in open(path) if l.startswith('"')]
rows = list(csv.DictReader(lines))
kern = {}
order = []
for r in rows:
    key = (r["ID"], r["Kernel Name"])
    if key not in kern:
        kern[key] = {}
        order.append(key)
    v = r["Metric Value"].replace(",", "")
    try:
        v = float(v)
    except ValueError:
        pass
    kern[key][r["Metric Name"]] = (v, r["Metric Unit"])


def val(m, name, unit_scale=None):
    v, u = m.get(name, (float("nan"), ""))
    if unit_scale and u in unit_scale:
        v *= unit_scale[u]
    return v


BYTES = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
TIME = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "nsecond": 1e-9, "usecond": 1e-6, "msecond": 1e-3, "second": 1.0}
print("HBM peak used: %.1f GB/s (MEASURED_PEAKS.json copy rate); FMA-pipe figures are ncu's own pct of peak" % hbm_peak)
print("%-28s %9s %9s %7s %9s %9s %9s %11s %9s" % ("kernel", "time ms", "HBM GB/s", "%peak", "FMA act%", "FMA ela%", "LSU wf%", "bank confl", "issue%"))
tot = 0.0
for key in order:
    m = kern[key]
    t = val(m, "gpu__time_duration.sum", TIME)
    b = val(m, "dram__bytes_read.sum", BYTES) + val(m, "dram__bytes_write.sum", BYTES)
    wf = val(m, "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum")
    bc = val(m, "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum")
    name = key[1].split("(")[0].replace("void ", "").replace("dau::", "").replace("(anonymous namespace)::", "").replace("unnamed>::", "")
    tot += t
    print("%-28s %9.3f %9.0f %6.1f%% %8.1f%% %8.1f%% %8.1f%% %10.3f%% %8.1f%%" % (
        name[:28], t * 1e3, b / t / 1e9, 100 * b / t / 1e9 / hbm_peak,
        val(m, "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
        val(m, "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed"),
        val(m, "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
        100 * bc / wf if wf else 0.0,
        val(m, "smsp__issue_active.avg.pct_of_peak_sustained_elapsed")))
print("sum of kernel times under ncu (cold caches, serialised): %.3f ms" % (tot * 1e3))
