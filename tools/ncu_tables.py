"""Summarise the ncu CSVs of tools/gpu_profile.sh: launch list (per-kernel totals / shares) and the per-launch metrics of one
production frame, with per-class DRAM traffic.
usage: python tools/ncu_tables.py gpurun_out/r02   ->  prints markdown tables, writes <prefix>_traffic.json"""
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
    m = re.search(r"\b(k_\w+(?:<[^>]*>)?)", k)
    if m:
        return m.group(1).replace("(bool)", "")
    k = k.replace("void ", "")
    m = re.match(r"([\w:]+)", k)
    return m.group(1) if m else k


def klass(k):
    if k.startswith("k_isect3") or k.startswith("k_closest_grid") or k.startswith("k_intersect"):
        return "closest_hit"
    if k.startswith("k_shadow_shade") or k.startswith("k_setup_shade"):      # (bench.py's live class "shadow_shade" times k_setup_shade as a whole)
        return "shadow_shade"
    return "other"


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
# frames = runs of launches starting at k_scene_cmp; a production frame has no diagnostic (<..., 1>) k_isect3 instantiation
frames, cur = [], []
for i, d in L.items():
    if d["k"].startswith("k_set_cam") and cur:
        frames.append(cur); cur = []
    cur.append((i, d))
if cur:
    frames.append(cur)
prod = [f for f in frames if any(d["k"].startswith("k_resolve") for _, d in f) and not any(re.search(r"k_isect3<\d, \d, 1>", d["k"]) for _, d in f)]
frame = prod[0] if prod else frames[0]
print(f"production frame: launches {frame[0][0]}..{frame[-1][0]} of the process ({len(frames)} frames / tiles captured)\n")
print("| # | kernel | ms | FMA pipe % | issue % | threads / instr | DRAM GB | warp instr M |\n|---|---|---|---|---|---|---|---|")
per = collections.defaultdict(lambda: [0.0, 0.0])
for i, d in frame:
    t = ms(*d["gpu__time_duration.sum"])
    g = gb(*d["dram__bytes_read.sum"]) + gb(*d["dram__bytes_write.sum"])
    per[klass(d["k"])][0] += t; per[klass(d["k"])][1] += g
    if t < 0.004 or d["k"].startswith("at::"):
        continue
    print(f"| {i} | `{d['k']}` | {t:.3f} | {d['sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed'][0]:.1f} | "
          f"{d['smsp__issue_active.avg.pct_of_peak_sustained_active'][0]:.1f} | {d.get('smsp__thread_inst_executed_per_inst_executed.ratio', (0,))[0]:.1f} | {g:.3f} | {d['smsp__inst_executed.sum'][0] / 1e6:.1f} |")
print("\n| class | ms (under ncu, cold) | DRAM GB |\n|---|---|---|")
for k, v in per.items():
    print(f"| {k} | {v[0]:.3f} | {v[1]:.3f} |")
print(f"| frame | {sum(v[0] for v in per.values()):.3f} | {sum(v[1] for v in per.values()):.3f} |")
out = {"traffic_bytes_per_frame": sum(v[1] for v in per.values()) * 1e9, "dram_bytes_per_class": {k: v[1] * 1e9 for k, v in per.items()},
       "launches": len(frame), "what": "ncu dram__bytes_read.sum + dram__bytes_write.sum of every launch of one production tile/frame (" + pre.split('/')[-1] + "_frame.csv)"}
json.dump(out, open(pre + "_traffic.json", "w"))
