"""Summarise an ncu report (--set full) and a launch list (gpu__time_duration.sum csv) into profiles/.

    python scripts/ncu_summary.py gpurun_out/prof.ncu-rep gpurun_out/launches.csv profiles/r1

writes <prefix>_ncu_summary.json / .md (per kernel: duration, DRAM bytes, tensor / MUFU / issue utilisation, registers)
and <prefix>_launches.md (per-kernel share of one layer forward).  Runs on the CPU box (ncu -i only)."""
import collections
import csv
import json
import subprocess
import sys

METRICS = {
    "duration_ms": "gpu__time_duration.sum",
    "dram_read_GB": "dram__bytes_read.sum",
    "dram_write_GB": "dram__bytes_write.sum",
    "dram_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "tensor_pct": "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "xu_pct": "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "issue_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "regs": "launch__registers_per_thread",
    "sm_mhz": "smsp__cycles_elapsed.avg.per_second",
}


def to_float(s):
    try:
        return float(s.replace(",", ""))
    except ValueError:
        return None


def main():
    rep, launches, prefix = sys.argv[1], sys.argv[2], sys.argv[3]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    head, units = rows[0], rows[1]
    kernels = collections.OrderedDict()
    for r in rows[2:]:
        name = r[head.index("Kernel Name")].split("(")[0].replace("void ", "").replace("haet::", "")
        d = {}
        for k, m in METRICS.items():
            if m in head:
                v = to_float(r[head.index(m)])
                u = units[head.index(m)]
                if v is not None:
                    if k == "duration_ms":
                        v *= {"us": 1e-3, "ms": 1.0, "ns": 1e-6, "s": 1e3}.get(u, 1.0)
                    if k.endswith("_GB"):
                        v *= {"Gbyte": 1.0, "Mbyte": 1e-3, "Kbyte": 1e-6, "byte": 1e-9}.get(u, 1.0)
                    if k == "sm_mhz":
                        v *= {"Ghz": 1e3, "Mhz": 1.0, "hz": 1e-6}.get(u, 1.0)
                d[k] = v
        d["traffic_bytes"] = int(((d.get("dram_read_GB") or 0) + (d.get("dram_write_GB") or 0)) * 1e9)
        kernels.setdefault(name, d)   # first captured launch of each kernel
    json.dump(kernels, open(prefix + "_ncu_summary.json", "w"), indent=1)
    with open(prefix + "_ncu_summary.md", "w") as f:
        f.write("ncu --set full --clock-control none, one launch per kernel, BASELINE config 4 layer "
                "(1M points, 256 hidden, 8 heads, 256 slices); cold-cache, serialised: compare shares.\n\n")
        f.write("| kernel | ms | DRAM read GB | DRAM write GB | DRAM % | tensor % | MUFU (xu) % | issue % | regs |\n|---|---|---|---|---|---|---|---|---|\n")
        for n, d in kernels.items():
            f.write("| %s | %.3f | %.3f | %.3f | %.1f | %.1f | %.1f | %.1f | %d |\n" % (
                n, d["duration_ms"], d["dram_read_GB"], d["dram_write_GB"], d["dram_pct"], d["tensor_pct"], d["xu_pct"],
                d["issue_pct"], d["regs"]))
    # launch list
    lines = [l for l in open(launches) if not l.startswith("==")]
    allr = list(csv.DictReader(lines))
    agg = collections.OrderedDict()
    for row in allr:
        k = row["Kernel Name"].split("(")[0].replace("void ", "").replace("haet::", "")
        v = to_float(row["Metric Value"])
        unit = row.get("Metric Unit", "ns")
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(unit, 1e-3)
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    ours = {k: a for k, a in agg.items() if not k.startswith("at::")}
    tot = sum(a[1] for a in ours.values())
    # kernels of ONE plain layer forward (haet_b200 PhysicsAttention at the bench shape) and how often each runs in it; the
    # LN / residual / MLP variants (gemm2_tc_kernel<1,1,0>, <0,0,1>, <2,1,0>, ln_fold) belong to the bench's fused-block leg
    layer = [("gemm_tc_pack_kernel", 2), ("gemm2_tc_kernel<1, 0, 0>", 1), ("slice_tc_kernel<1, 1>", 1), ("slice_merge_kernel", 2),
             ("slice_pos_kernel", 1), ("balltree_kernel", 1), ("erwin_gather_kernel", 1), ("erwin_block_kernel", 6),
             ("erwin_pool_kernel", 1), ("erwin_unpool_kernel", 1), ("erwin_scatter_kernel", 1), ("deslice_tc_kernel<1, 3, 3>", 1),
             ("gemm2_tc_kernel<0, 0, 0>", 1)]
    with open(prefix + "_launches.md", "w") as f:
        f.write("ncu --metrics gpu__time_duration.sum --clock-control none over `python bench.py --steps 2 --warmup 3 --no-configs` "
                "(the bench command: timed layer steps under graph replay, the e2e leg, the stage timings and the fused-block leg): "
                "every haet_b200 launch, summed per kernel.\n\n")
        f.write("| kernel | launches | total us | avg us | share of all |\n|---|---|---|---|---|\n")
        for k, a in ours.items():
            f.write("| %s | %d | %.1f | %.1f | %.1f%% |\n" % (k, a[0], a[1], a[1] / a[0], 100 * a[1] / tot))
        f.write("\nall haet_b200 kernels: %.1f us over %d launches\n" % (tot, sum(a[0] for a in ours.values())))
        have = [(k, n) for k, n in layer if k in ours]
        if have:
            lt = sum(n * ours[k][1] / ours[k][0] for k, n in have)
            f.write("\nOne layer forward (the timed step) = these kernels, at their average duration under ncu (cold cache, serialised):\n\n")
            f.write("| kernel | per layer | us | share of the step |\n|---|---|---|---|\n")
            for k, n in have:
                v = n * ours[k][1] / ours[k][0]
                f.write("| %s | %d | %.1f | %.1f%% |\n" % (k, n, v, 100 * v / lt))
            f.write("\nsum: %.1f us per layer forward\n" % lt)


if __name__ == "__main__":
    main()
