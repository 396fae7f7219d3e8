"""Turns one profiling pass (tools/ncu_round.sh output in gpurun_out/) into the tracked summaries under profiles/:
launch-list tables (kernel, launches, average us, share of the listed time) and the --set full metrics per kernel.
Usage: python tools/ncu_summarise.py <suffix> "<note>"   (run where ncu is installed; reads gpurun_out/)"""
import csv
import io
import json
import re
import shutil
import subprocess
import sys
from collections import OrderedDict

OUT = "gpurun_out"
METRICS = ["sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
           "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed.avg.per_cycle_elapsed",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
           "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
           "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum",
           "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]


def short(name):
    name = re.sub(r"^void ", "", name)
    name = name.replace("<unnamed>::", "")
    return re.sub(r"\(.*$", "", name)


def launch_summary(path, header):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    h = rows[0]
    ki, vi = h.index("Kernel Name"), h.index("Metric Value")
    agg = OrderedDict()
    for r in rows[1:]:
        k = short(r[ki])
        n, t = agg.get(k, (0, 0.0))
        agg[k] = (n + 1, t + float(r[vi].replace(",", "")) / 1000.0)
    total = sum(t for _, t in agg.values())
    lines = [header, "%-52s %8s %10s %7s" % ("kernel", "launches", "avg us", "share")]
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        lines.append("%-52s %8d %10.1f %7.3f" % (k, n, t / n, t / total))
    return "\n".join(lines) + "\n"


def full_summary(rep, header):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, units = rows[0], rows[1]
    lines = [header]
    traffic = {}
    for r in rows[2:]:
        k = short(r[h.index("Kernel Name")])
        lines.append("kernel: " + k)
        for m in METRICS:
            if m in h:
                lines.append("  %-68s %s %s" % (m, r[h.index(m)], units[h.index(m)]))
        lines.append("")
        rd, wr = r[h.index("dram__bytes_read.sum")], r[h.index("dram__bytes_write.sum")]
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
        traffic[k] = int(float(rd.replace(",", "")) * scale[units[h.index("dram__bytes_read.sum")]] +
                         float(wr.replace(",", "")) * scale[units[h.index("dram__bytes_write.sum")]])
    return "\n".join(lines) + "\n", traffic, {short(r[h.index("Kernel Name")]): {m: r[h.index(m)] for m in METRICS if m in h} for r in rows[2:]}


def main():
    sfx, note = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
    for tag, csvname in (("seq", "launches_seq.csv"), ("train", "launches_trn.csv")):
        shutil.copy(f"{OUT}/{csvname}", f"profiles/r02_launches_{tag}_{sfx}.csv")
        cmd = "python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-extras" if tag == "seq" else \
            "python bench.py --workload train --steps 2 --warmup 3"
        open(f"profiles/r02_launches_{tag}_{sfx}_summary.txt", "w").write(launch_summary(
            f"{OUT}/{csvname}", f"ncu --metrics gpu__time_duration.sum --clock-control none -k <library kernels>, {cmd} ({note}): "
                                "model set-up + 5 steps"))
    text, traffic_all = "", {}
    for tag, rep in (("sequence step (120 frames, 100k Gaussians, 1024x1024)", "seq_full.ncu-rep"),
                     ("training step (4 subjects x 4 views, 100k Gaussians, 512x512): blend / sort / loss kernels of one step", "trn_full.ncu-rep")):
        t, traffic, facts = full_summary(f"{OUT}/{rep}", f"== ncu --set full --clock-control none, {tag} ({note}) ==")
        text += t + "\n"
        if rep.startswith("seq"):
            traffic_all, facts_seq = traffic, facts
    open(f"profiles/r02_ncu_full_summary_{sfx}.txt", "w").write(text)
    bf = facts_seq.get("blend_kernel<0, 0>", {})
    num = lambda k: float(bf[k].replace(",", "")) if k in bf else None
    blend_issue = {"issue_active_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                   "ipc": num("sm__inst_executed.avg.per_cycle_elapsed"),
                   "fma_pipe_active_pct": num("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
                   "alu_pipe_active_pct": num("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
                   "warp_instructions": num("smsp__inst_executed.sum")}
    names = {"amav_splat_blend": "blend_kernel<0, 0>", "amav_triplane_project": "project_kernel",
             "amav_splat_preprocess": "preprocess_kernel", "amav_sample_decode": "sample_decode_kernel"}
    json.dump({"source": f"profiles/r02_ncu_full_summary_{sfx}.txt (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch)",
               "workload": {"frames": 120, "gaussians": 100000, "size": 1024},
               "bytes_per_launch": {k: traffic_all[v] for k, v in names.items() if v in traffic_all},
               "blend_issue": blend_issue},
              open("profiles/ncu_traffic.json", "w"), indent=1)
    print(open(f"profiles/r02_launches_seq_{sfx}_summary.txt").read())
    print(open(f"profiles/r02_launches_train_{sfx}_summary.txt").read())


if __name__ == "__main__":
    main()
