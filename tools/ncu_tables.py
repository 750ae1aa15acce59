"""Summarise the ncu CSVs of tools/gpu_prof8.sh: launch list (per kernel totals / shares) and per-launch frame metrics.
usage: python tools/ncu_tables.py gpurun_out/v8   ->  prints markdown tables, writes <prefix>_traffic.json"""
import csv, re, sys, json, collections

pre = sys.argv[1]


def rows(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    return list(csv.DictReader(lines))


def ms(v, u):
    return v / 1e6 if u in ("ns", "nsecond") else v / 1e3 if u in ("us", "usecond") else v if u in ("ms", "msecond") else v * 1e3


def gb(v, u):
    return v * {"byte": 1e-9, "Kbyte": 1e-6, "Mbyte": 1e-3, "Gbyte": 1.0}[u]


def short(k):
    return re.sub(r"\(.*", "", k).replace("void ", "").replace("b200rt::", "")


agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows(pre + "_launches.csv"):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    k = short(r["Kernel Name"])
    agg[k][0] += 1
    agg[k][1] += ms(float(r["Metric Value"].replace(",", "")), r["Metric Unit"])
tot = sum(v[1] for k, v in agg.items() if "probe" not in k and "at::" not in k)
print("| kernel | launches | total ms | share of render kernels |\n|---|---|---|---|")
for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
    if "at::" in k:
        continue
    print(f"| `{k}` | {v[0]} | {v[1]:.3f} | {'-' if 'probe' in k else f'{100 * v[1] / tot:.2f} %'} |")
print(f"\nrender kernels total {tot:.3f} ms\n")

L = collections.OrderedDict()
for r in rows(pre + "_frame.csv"):
    d = L.setdefault(int(r["ID"]), {"k": short(r["Kernel Name"])})
    d[r["Metric Name"]] = (float(r["Metric Value"].replace(",", "")), r["Metric Unit"])
print("| # | kernel | ms | FMA pipe % | issue % | DRAM GB | warp instr M |\n|---|---|---|---|---|---|---|")
first_frame_isect = []
seen_raygen = 0
for i, d in L.items():
    t = ms(*d["gpu__time_duration.sum"])
    g = gb(*d["dram__bytes_read.sum"]) + gb(*d["dram__bytes_write.sum"])
    if d["k"].startswith("k_raygen"):
        seen_raygen += 1
    if seen_raygen == 2 and ("k_isect3" in d["k"] or "k_shadow_shade" in d["k"]):          # second frame of the process = the TIME_ONLY frame: production kernels, launches serialised
        first_frame_isect.append((t, g))
    if seen_raygen == 2:
        print(f"| {i} | `{d['k']}` | {t:.3f} | {d['sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed'][0]:.1f} | "
              f"{d['smsp__issue_active.avg.pct_of_peak_sustained_active'][0]:.1f} | {g:.3f} | {d['smsp__inst_executed.sum'][0] / 1e6:.1f} |")
if first_frame_isect:
    out = {"traffic_bytes_per_frame": sum(g for _, g in first_frame_isect) * 1e9, "launches": len(first_frame_isect),
           "source": "ncu dram__bytes_read.sum + dram__bytes_write.sum over the k_isect3 + k_shadow_shade launches (closest-hit and shadow queries) of one production C4 frame (" + pre.split('/')[-1] + "_frame.csv)",
           "kernel_ms_under_ncu": sum(t for t, _ in first_frame_isect)}
    json.dump(out, open(pre + "_traffic.json", "w"), indent=1)
    print("\n", json.dumps(out))
