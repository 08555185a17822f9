"""Summarise .ncu-rep captures (ncu --set full --import-source on) into a text file for profiles/.
  python tools/ncu_summary.py out.txt name1=path1.ncu-rep [name2=path2.ncu-rep ...]
Per report: the headline counters (duration, DRAM bytes / throughput, L2, SM / tensor-pipe activity, occupancy, registers, shared
memory), the warp-stall breakdown from the source-level sampling, and the 12 SASS instructions that collected the most samples."""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_tensor.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__cluster_size", "smsp__inst_executed.sum"]


def ncu(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    out_path, items = sys.argv[1], sys.argv[2:]
    lines = ["# ncu --set full --clock-control none --import-source on, ONE launch each, captured inside a real run (see the script that"
             " made the .ncu-rep); serialised, cold-cache replays: shares and counters, not bench numbers"]
    for it in items:
        name, path = it.split("=", 1)
        raw = list(csv.reader(io.StringIO(ncu(["-i", path, "--page", "raw", "--csv"]))))
        if len(raw) < 3:
            lines.append(f"## {name}: unreadable report {path}")
            continue
        hdr, units, vals = raw[0], raw[1], raw[2]
        d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
        lines.append(f"## {name}")
        lines.append(f"{'Kernel Name':72s} {d.get('Kernel Name', ('?', ''))[0][:110]}")
        for k in KEYS:
            if k in d:
                lines.append(f"{k:72s} {d[k][0]} {d[k][1]}")
        stall = {k: float(v[0]) for k, v in d.items() if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio")}
        if stall:
            tot = sum(stall.values())
            top = sorted(stall.items(), key=lambda kv: -kv[1])[:8]
            lines.append("warp states per issue slot (smsp__average_warps_issue_stalled_*_per_issue_active), share of all warp-cycles:")
            for k, v in top:
                short = k[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]
                lines.append(f"    {short:28s} {v:8.3f}   {100 * v / tot:5.1f} %")
        src = list(csv.reader(io.StringIO(ncu(["-i", path, "--page", "source", "--csv", "--print-source", "sass"]))))
        hi = next((i for i, r in enumerate(src) if len(r) > 3 and r[1] == "Source"), None)
        if hi is not None:
            h = src[hi]
            ix = {x: i for i, x in enumerate(h)}
            rows = [r for r in src[hi + 1:] if len(r) == len(h)]
            if "# Samples" in ix:
                S = ix["# Samples"]
                tot = sum(int(r[S] or 0) for r in rows) or 1
                stalls = [x for x in h if x.startswith("stall_") and "Not Issued" not in x]
                agg = sorted(((sum(int(r[ix[x]] or 0) for r in rows), x) for x in stalls), reverse=True)[:6]
                lines.append(f"source-level warp sampling: {tot} samples; by reason: " + ", ".join(f"{x} {100 * n / tot:.1f} %" for n, x in agg))
                lines.append("SASS instructions with the most samples (samples, share, instruction, dominant reason):")
                for r in sorted(rows, key=lambda r: -int(r[S] or 0))[:12]:
                    n = int(r[S] or 0)
                    dom = max(((int(r[ix[x]] or 0), x) for x in stalls))
                    lines.append(f"    {n:6d} {100 * n / tot:5.1f} %  {r[ix['Source']].strip()[:64]:64s} {dom[1]}")
        lines.append("")
    open(out_path, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[:60]))


if __name__ == "__main__":
    main()
