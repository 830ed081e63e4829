"""Developer tool: condense one .ncu-rep (ncu --set full, one kernel launch) into the two text files kept under
profiles/: <out>_metrics.csv (selected raw metrics) and <out>_stalls.txt (stall reasons + hottest SASS lines).
Usage: python tools/ncu_export.py gpurun_out/x.ncu-rep profiles/r2_x "description" """
import csv
import io
import re
import subprocess
import sys

KEEP = re.compile(r"^(gpu__time_duration\.sum|dram__bytes_(read|write)\.sum|dram__bytes_(read|write)\.sum\.per_second|"
                  r"dram__throughput\.avg\.pct_of_peak_sustained_elapsed|lts__throughput\.avg\.pct_of_peak_sustained_elapsed|"
                  r"lts__t_bytes\.sum|launch__(grid_size|block_size|registers_per_thread|shared_mem_per_block_dynamic|waves_per_multiprocessor)|"
                  r"smsp__inst_executed\.sum|sm__inst_executed\.sum|smsp__issue_active\.avg\.pct_of_peak_sustained_active|"
                  r"sm__inst_executed_pipe_(alu|fma|lsu|xu|fp64|uniform)\.avg\.pct_of_peak_sustained_active|"
                  r"sm__pipe_tensor_cycles_active\.avg\.pct_of_peak_sustained_elapsed|"
                  r"sm__mem_tensor_cycles_active\.avg\.pct_of_peak_sustained_elapsed|sm__warps_active\.avg\.pct_of_peak_sustained_active|"
                  r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|sm__cycles_elapsed\.avg|sm__cycles_active\.avg|"
                  r"smsp__cycles_active\.avg|sm__throughput\.avg\.pct_of_peak_sustained_elapsed)$")


def main():
    rep, out, desc = sys.argv[1], sys.argv[2], sys.argv[3]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    with open(out + "_metrics.csv", "w") as f:
        w = csv.writer(f)
        w.writerow(["metric", "unit", "value (%s)" % desc])
        w.writerow(["Kernel Name", "", vals[hdr.index("Kernel Name")]])
        seen = set()
        for h, u, v in zip(hdr, units, vals):
            name = h.split("TriageCompute.")[-1]
            if KEEP.match(name) and name not in seen:
                seen.add(name)
                w.writerow([name, u, v])
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    tmp = out + "_src.tmp.csv"
    open(tmp, "w").write(src)
    txt = subprocess.run([sys.executable, __file__.replace("ncu_export.py", "ncu_src.py"), tmp, "40"], capture_output=True, text=True).stdout
    open(out + "_stalls.txt", "w").write("# %s\n%s" % (desc, txt))
    import os
    os.remove(tmp)
    print(open(out + "_metrics.csv").read())


if __name__ == "__main__":
    main()
