"""profiles/ncu_traffic.json + a launch-list summary from
   (a) the ncu CSV of `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c N --csv
       python bench.py --steps 1 --warmup 0 --no-cpu`, and
   (b) the JSON line of a normal `python bench.py` run of the same build (units per step).
usage: python tools/make_ncu_traffic.py <launches.csv> <bench.json> <tag>"""
import collections
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def klass(name):
    if "k_shade" in name or "k_bin_" in name:
        return "k_shade"
    if "k_trace" in name:
        return "k_trace"
    if "k_shadow" in name:
        return "k_shadow"
    return None


def main(csv_path, bench_path, tag):
    rows = list(csv.reader(open(csv_path)))
    hdr = None
    launches = collections.OrderedDict()   # id -> {name, metrics}
    for r in rows:
        if hdr is None:
            if "Kernel Name" in r:
                hdr = {h: i for i, h in enumerate(r)}
            continue
        if len(r) < len(hdr):
            continue
        i = int(r[hdr["ID"]])
        L = launches.setdefault(i, {"name": r[hdr["Kernel Name"]], "m": {}})
        v = float(r[hdr["Metric Value"]].replace(",", ""))
        unit = r[hdr["Metric Unit"]]
        scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "ns": 1.0, "us": 1e3, "ms": 1e6, "ns": 1.0}.get(unit, 1.0)
        L["m"][r[hdr["Metric Name"]]] = v * scale
    # one bench step = the launches up to the n-th k_film_add (n = waves per step)
    bench = json.loads(open(bench_path).read().strip().splitlines()[-1])
    samples_step = bench["config"]["camera_samples_per_step"]
    waves = int(round(bench.get("waves_per_step", -(-samples_step // (1 << 20)))))
    step, films, build = [], 0, []
    for i, L in launches.items():
        if any(t in L["name"] for t in ("k_tri_setup", "k_morton", "cub::", "DeviceRadixSort", "k_hierarchy", "k_refit", "k_emit", "k_tri_shade", "k_pad_boxes", "k_collapse")):
            build.append(L)
            continue
        step.append(L)
        if "k_film_add" in L["name"]:
            films += 1
            if films == waves:
                break
    agg = collections.defaultdict(lambda: {"launches": 0, "ns": 0.0, "dram": 0.0})
    byname = collections.defaultdict(lambda: {"launches": 0, "ns": 0.0, "dram": 0.0})
    for L in step:
        ns = L["m"].get("gpu__time_duration.sum", 0.0)
        dram = L["m"].get("dram__bytes_read.sum", 0.0) + L["m"].get("dram__bytes_write.sum", 0.0)
        short = L["name"].split("(")[0].replace("void ", "").replace("yg::", "")
        for d, k in ((agg, klass(L["name"])), (byname, short)):
            if k is None:
                continue
            d[k]["launches"] += 1; d[k]["ns"] += ns; d[k]["dram"] += dram
    kern = bench["roofline"]["kernels"]
    side_steps = min(bench["steps"], 3)
    out = {"tag": tag, "source": os.path.basename(csv_path), "waves_per_step": waves, "complete_step": films == waves}
    for k, a in agg.items():
        units_step = kern[k]["units"] / side_steps
        out[k] = {"dram_bytes_per_step": a["dram"], "units_per_step": units_step, "dram_bytes_per_unit": a["dram"] / max(1.0, units_step),
                  "alg_bytes_per_unit": kern[k]["alg_bytes_per_unit"], "gpu_ms_per_step_under_ncu": a["ns"] / 1e6, "launches_per_step": a["launches"]}
    with open(os.path.join(ROOT, "profiles", "ncu_traffic.json"), "w") as fh:
        json.dump(out, fh, indent=1)
    tot = sum(v["ns"] for v in byname.values())
    lines = ["# %s — ncu launch list of one bench step (%s, %d wave%s per step)" % (tag, bench["config"]["workload"], waves, "" if waves == 1 else "s"), "",
             "Command: `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c N --csv python bench.py --steps 2 --warmup 1 --no-cpu`",
             "(per-launch times under ncu are serialised and cold-cache: compare SHARES with bench.py's `roofline.kernels[*].share_of_step`, not absolutes).", "",
             "| kernel | launches | GPU time ms | share | DRAM MB |", "|---|---|---|---|---|"]
    for k, v in sorted(byname.items(), key=lambda kv: -kv[1]["ns"]):
        lines.append("| `%s` | %d | %.3f | %.1f %% | %.1f |" % (k, v["launches"], v["ns"] / 1e6, 100 * v["ns"] / tot, v["dram"] / 1e6))
    lines += ["", "Total %.2f ms of GPU time in the step under ncu; bench.py measured %.2f ms per step without the profiler." % (tot / 1e6, bench["ms_per_step"]), "",
              "| kernel class | DRAM bytes / unit (ncu) | algorithmic bytes / unit (bench.py) | ratio |", "|---|---|---|---|"]
    for k in agg:
        o = out[k]
        lines.append("| %s (per %s) | %.1f | %.1f | %.2f |" % (k, kern[k]["unit"], o["dram_bytes_per_unit"], o["alg_bytes_per_unit"], o["dram_bytes_per_unit"] / max(1e-9, o["alg_bytes_per_unit"])))
    # the upload that precedes the step: BVH build kernels (the radix sort is the one library primitive on the path)
    btot = sum(L["m"].get("gpu__time_duration.sum", 0.0) for L in build)
    bsort = sum(L["m"].get("gpu__time_duration.sum", 0.0) for L in build if "cub::" in L["name"] or "DeviceRadixSort" in L["name"])
    if build:
        bn = collections.defaultdict(lambda: [0, 0.0])
        for L in build:
            short = L["name"].split("(")[0].replace("void ", "").replace("yg::", "")[:70]
            bn[short][0] += 1; bn[short][1] += L["m"].get("gpu__time_duration.sum", 0.0)
        lines += ["", "## BVH build kernels of the uploads in the capture (%d launches, %.2f ms under ncu; cub::DeviceRadixSort share %.1f %%)" % (len(build), btot / 1e6, 100 * bsort / max(1.0, btot)), "",
                  "| kernel | launches | GPU time ms | share |", "|---|---|---|---|"]
        for k, v in sorted(bn.items(), key=lambda kv: -kv[1][1]):
            lines.append("| `%s` | %d | %.3f | %.1f %% |" % (k, v[0], v[1] / 1e6, 100 * v[1] / btot))
    with open(os.path.join(ROOT, "profiles", "%s_launches.md" % tag), "w") as fh:
        fh.write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main(*sys.argv[1:4])
