"""Turns the ncu captures of one steady-state tick into the tables bench.py and the judge read.
  python profiles/ncu_summarize.py <metrics.csv from --metrics ... --csv --log-file> <raw.csv from ncu -i rep --page raw --csv> <tag>
Writes profiles/roofline_traffic.json (dram bytes per launch, per kernel and per bench group), profiles/fp64_counts.json
(thread-level FP64 add / mul / fma instructions per launch) and prints a markdown table."""
import csv
import json
import sys
from collections import defaultdict

metrics_csv, raw_csv, tag = sys.argv[1:4]
GROUP = {"k_measure": "k_pre_plan", "k_pre": "k_pre_plan", "k_plan": "k_pre_plan", "k_replan_plan": "k_replan",
         "k_replan_deep": "k_replan", "k_replan_sample": "k_replan", "k_pfl_ltt": "k_replan", "k_replan_tail": "k_replan",
         "k_reach_verify": "k_reach_verify", "k_verify_stage": "k_reach_verify", "k_select": "k_select"}


def short(name):
    n = name.split("(")[0].replace("void ", "").replace("hrg::", "")
    base = n.split("<")[0]
    return base, n


# ---- pass 2: explicit metrics, possibly several ticks: average per launch INSTANCE of a tick (same kernel may launch twice)
rows = [r for r in csv.reader(l for l in open(metrics_csv) if l.startswith('"'))]
hdr = rows[0]
ik, im, iv, iid = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("ID")
per_id = defaultdict(dict)
for r in rows[1:]:
    per_id[int(r[iid])]["name"] = r[ik]
    per_id[int(r[iid])][r[im]] = float(r[iv].replace(",", ""))
ids = sorted(per_id)
names = [short(per_id[i]["name"])[1] for i in ids]
first = names[0]
n_ticks = sum(1 for n in names if n == first and True) if names.count(first) else 1
# occurrences of the first kernel mark tick boundaries (k_pre launches once per tick)
n_ticks = names.count("k_pre") or 1
agg = defaultdict(lambda: defaultdict(float))
order = []
for i in ids:
    base, full = short(per_id[i]["name"])
    key = full
    if key not in order:
        order.append(key)
    for m, v in per_id[i].items():
        if m != "name":
            agg[key][m] += v / n_ticks
traffic, fp64 = {"_source": f"ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none over {n_ticks} steady-state "
                            f"tick(s) of profiles/ncu_tick.py 65536 60 ({tag}); bytes per tick, summed over the launches of a name"}, \
    {"_source": f"ncu smsp__sass_thread_inst_executed_op_d{{add,mul,fma}}_pred_on.sum, same capture ({tag}); thread-level "
                "instructions per tick"}
print(f"| kernel (launches / tick) | us (ncu, serialised, cold) | DRAM read MB | DRAM write MB | FP64 thread inst (add / mul / fma) M |")
print("|---|---|---|---|---|")
for key in order:
    a = agg[key]
    base = short(key)[0]
    g = GROUP.get(base, "other")
    tr = a["dram__bytes_read.sum"] + a["dram__bytes_write.sum"]
    f = a["smsp__sass_thread_inst_executed_op_dadd_pred_on.sum"] + a["smsp__sass_thread_inst_executed_op_dmul_pred_on.sum"] + \
        a["smsp__sass_thread_inst_executed_op_dfma_pred_on.sum"]
    traffic.setdefault("per_kernel", {})[key] = tr
    traffic[g] = traffic.get(g, 0.0) + tr
    fp64.setdefault("per_kernel", {})[key] = f
    fp64[g] = fp64.get(g, 0.0) + f
    nl = sum(1 for n in names if n == key) / n_ticks
    print(f"| `{key}` ({nl:g}) | {a['gpu__time_duration.sum'] / 1e3:.1f} | {a['dram__bytes_read.sum'] / 1e6:.1f} | "
          f"{a['dram__bytes_write.sum'] / 1e6:.1f} | {a['smsp__sass_thread_inst_executed_op_dadd_pred_on.sum'] / 1e6:.1f} / "
          f"{a['smsp__sass_thread_inst_executed_op_dmul_pred_on.sum'] / 1e6:.1f} / {a['smsp__sass_thread_inst_executed_op_dfma_pred_on.sum'] / 1e6:.1f} |")
json.dump(traffic, open("profiles/roofline_traffic.json", "w"), indent=1)
json.dump(fp64, open("profiles/fp64_counts.json", "w"), indent=1)

# ---- pass 1: --set full of one tick
rows = list(csv.reader(open(raw_csv)))
hdr, data = rows[0], rows[2:]


def col(r, name, default=float("nan")):
    try:
        return float(r[hdr.index(name)].replace(",", ""))
    except (ValueError, IndexError):
        return default


print()
print("| kernel | us | regs | warps active % | issue active | lanes / inst | FP64 pipe % | top stalls (cycles per issue) |")
print("|---|---|---|---|---|---|---|---|")
st = [i for i, h in enumerate(hdr) if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio")
      and "not_issued" not in h]
for r in data:
    name = short(r[hdr.index("Kernel Name")])[1]
    tops = sorted(((float(r[i]), hdr[i].replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""))
                   for i in st), reverse=True)[:3]
    print(f"| `{name}` | {col(r, 'gpu__time_duration.sum'):.1f} | {col(r, 'launch__registers_per_thread'):.0f} | "
          f"{col(r, 'sm__warps_active.avg.pct_of_peak_sustained_active'):.1f} | {col(r, 'smsp__issue_active.avg.per_cycle_active'):.2f} | "
          f"{col(r, 'smsp__thread_inst_executed_per_inst_executed.ratio'):.1f} | "
          f"{col(r, 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active'):.1f} | "
          + ", ".join(f"{n} {v:.1f}" for v, n in tops) + " |")
