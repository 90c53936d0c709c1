"""Summaries of ncu CSV output for profiles/ (round 2).

  python tools/ncu_summ.py traffic <launch-list.csv> <first_kernel_regex> <out.json>
      launch list taken with  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv
      -> DRAM bytes and kernel time per stage for ONE compress call (the launches from the last occurrence of <first_kernel_regex>
      that is followed by an emit_kernel, up to that emit_kernel), in the layout bench.py reads (profiles/r02_dram_traffic_512cube.json)
  python tools/ncu_summ.py full <raw.csv> <kernel_regex> <out.json>
      raw page of a --set full report (ncu -i x.ncu-rep --page raw --csv) -> the metrics DESIGN.md quotes, for the first launch that matches
"""
import csv
import json
import re
import sys

STAGE = [("range_kernel|first_equal", "range (host layer, before the job)"), ("sample_mean|opt_sample|opt_finish", "opt_intervals"),
         ("mean_kernel", "use_mean_mean"), ("fit_select|reg_flag|reg_gather", "fit_select"), ("scan_", "scan"),
         ("coeff_lattice", "coeff_chain"), ("prepass|wavefront", "wavefront_quant"), ("hist_kernel", "histogram"),
         ("huff_build", "coeff_trees_kernel_side_stream"), ("coef_encode", "coeff_encode"), ("enc_len|enc_write|indicator_pack", "encode"),
         ("emit_kernel", "emit")]


def rows_of(path):
    with open(path, newline="") as f:
        lines = [ln for ln in f if ln.startswith('"')]
    return list(csv.DictReader(lines))


def stage_of(name):
    for pat, st in STAGE:
        if re.search(pat, name):
            return st
    return "other"


def traffic(path, first_re, out):
    rows = rows_of(path)
    launches = {}
    for r in rows:
        i = int(r["ID"])
        d = launches.setdefault(i, {"name": r["Kernel Name"]})
        d[r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
    ids = sorted(launches)
    ends = [i for i in ids if re.search("emit_kernel", launches[i]["name"])]
    if not ends:
        raise SystemExit("no emit_kernel launch in " + path)
    end = ends[-1]
    starts = [i for i in ids if i < end and re.search(first_re, launches[i]["name"])]
    # the last start that belongs to this call: no other emit_kernel between it and `end`
    prev_end = ends[-2] if len(ends) > 1 else -1
    starts = [i for i in starts if i > prev_end]
    start = starts[0]
    by_stage, ms, n = {}, {}, 0
    main_tree_seen = False
    for i in ids:
        if i < start or i > end:
            continue
        L = launches[i]
        st = stage_of(L["name"])
        if st == "coeff_trees_kernel_side_stream":
            # two launches: the coefficient trees (grid 4 or 8) and the main tree (grid 1); the main one is the later, short one
            if L.get("gpu__time_duration.sum", 0) < 2e5:
                st = "huffman_tree"
        b = L.get("dram__bytes_read.sum", 0.0) + L.get("dram__bytes_write.sum", 0.0)
        by_stage[st] = by_stage.get(st, 0.0) + b
        ms[st] = ms.get(st, 0.0) + L.get("gpu__time_duration.sum", 0.0) * 1e-6
        n += 1
    res = {"source": "%s: ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none, tools/prof_one.py 512 2, "
                     "the launches of the second compress call (tools/ncu_summ.py traffic)" % path,
           "all_launches_of_one_step": sum(by_stage.values()), "launches": n, "by_stage": by_stage, "summed_kernel_ms_under_ncu": ms}
    with open(out, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res, indent=1))


KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__shared_mem_per_block_static", "launch__cluster_dim_x", "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio", "smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "lts__t_bytes.sum", "l1tex__t_bytes.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed"]


def full(path, kre, out):
    with open(path, newline="") as f:
        lines = [ln for ln in f if ln.startswith('"')]
    rd = csv.reader(lines)
    head = next(rd)
    units = next(rd)
    res = None
    for row in rd:
        d = dict(zip(head, row))
        if re.search(kre, d.get("Kernel Name", "")):
            res = {"kernel": d["Kernel Name"], "grid": d.get("Grid Size"), "block": d.get("Block Size"), "metrics": {}}
            u = dict(zip(head, units))
            for k in KEEP:
                if k in d and d[k] != "":
                    res["metrics"][k] = {"value": d[k], "unit": u.get(k, "")}
            break
    if res is None:
        raise SystemExit("no launch matches " + kre)
    with open(out, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    if len(sys.argv) == 5 and sys.argv[1] == "traffic":
        traffic(sys.argv[2], sys.argv[3], sys.argv[4])
    elif len(sys.argv) == 5 and sys.argv[1] == "full":
        full(sys.argv[2], sys.argv[3], sys.argv[4])
    else:
        raise SystemExit(__doc__)
