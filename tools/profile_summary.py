"""Turn an ncu report (gpurun_out/<tag>_full.ncu-rep) into the committed evidence under profiles/:
   <tag>_ncu_full_summary.csv  one row per captured launch with the metrics the roofline discussion uses
   <tag>_ncu_full_details.txt  ncu's own --page details text
   traffic.json                DRAM bytes per decode stage (only with --traffic; read by bench.py for roofline.traffic)
usage: python tools/profile_summary.py gpurun_out/r01d_full.ncu-rep profiles/r01d [--traffic N_IMAGES]"""
import csv
import io
import json
import subprocess
import sys

METRICS = ("dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,smsp__inst_executed.sum,"
           "smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio,"
           "sm__warps_active.avg.pct_of_peak_sustained_active,sm__throughput.avg.pct_of_peak_sustained_elapsed,"
           "l1tex__throughput.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed,"
           "launch__registers_per_thread,launch__occupancy_limit_registers,launch__occupancy_limit_shared_mem,"
           "launch__shared_mem_per_block_dynamic,launch__shared_mem_per_block_static")
STAGE_OF = {"k_unstuff_count": "unstuff", "k_unstuff_scan": "unstuff", "k_unstuff_write": "unstuff", "k_huff_sync": "huff_sync",
            "k_huff_sync_tail": "huff_tail", "k_huff_tail_round": "huff_tail", "k_subseq_scan": "huff_write", "k_huff_write": "huff_write", "k_dc_scan": "dc_scan",
            "k_idct": "idct", "k_color": "color"}


def main():
    rep, prefix = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--metrics", METRICS], capture_output=True, text=True).stdout
    open(prefix + "_ncu_full_summary.csv", "w").write(raw)
    det = subprocess.run(["ncu", "-i", rep, "--page", "details"], capture_output=True, text=True).stdout
    open(prefix + "_ncu_full_details.txt", "w").write(det)
    rows = list(csv.reader(io.StringIO(raw)))
    head, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(head)}
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tscale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}
    for r in data:
        name = r[col["Kernel Name"]].split("(")[0]
        rd = float(r[col["dram__bytes_read.sum"]]) * scale[units[col["dram__bytes_read.sum"]]]
        wr = float(r[col["dram__bytes_write.sum"]]) * scale[units[col["dram__bytes_write.sum"]]]
        us = float(r[col["gpu__time_duration.sum"]]) * tscale[units[col["gpu__time_duration.sum"]]]
        print("%-22s %9.1f us  dram %8.1f MB (%6.0f GB/s)  issue %5.1f%%" % (name, us, (rd + wr) / 1e6, (rd + wr) / us / 1e3,
                                                                            float(r[col["smsp__issue_active.avg.pct_of_peak_sustained_active"]])))
    if "--traffic" in sys.argv:
        n = int(sys.argv[sys.argv.index("--traffic") + 1])
        stages = {}
        for r in data:
            name = r[col["Kernel Name"]].split("(")[0]
            st = STAGE_OF.get(name)
            if not st:
                continue
            rd = float(r[col["dram__bytes_read.sum"]]) * scale[units[col["dram__bytes_read.sum"]]]
            wr = float(r[col["dram__bytes_write.sum"]]) * scale[units[col["dram__bytes_write.sum"]]]
            us = float(r[col["gpu__time_duration.sum"]]) * tscale[units[col["gpu__time_duration.sum"]]]
            e = stages.setdefault(st, {"dram_bytes": 0, "launches": 0, "kernels": [], "ncu_us": 0.0})
            e["dram_bytes"] += int(rd + wr)
            e["launches"] += 1
            e["kernels"].append(name)
            e["ncu_us"] += us
        out = {"source": "%s_ncu_full_summary.csv (ncu --set full --clock-control none, one warm decode step of tools/quick_bench.py %d 8: "
                         "%dx 1080p 4:2:0 q90, single stream)" % (prefix, n, n),
               "workload_images": n, "unit": "bytes per stage per step = dram__bytes_read.sum + dram__bytes_write.sum summed over the stage's launches",
               "stages": stages}
        json.dump(out, open("profiles/traffic.json", "w"), indent=1)


if __name__ == "__main__":
    main()
