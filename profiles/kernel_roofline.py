"""Per-kernel HBM roofline table from an ncu launch list (gpu__time_duration.sum, --clock-control none): average
duration of the level-0 (largest-grid) launches of every kernel against the bytes that kernel must move
(DESIGN.md section 5 "pass" accounting, 513^3 nodes / 512^3 cells) and MEASURED_PEAKS.json hbm_gbs.
ncu times are cold-cache and serialised: use the shares, and bench.py's in-situ events for absolute numbers.
usage: python profiles/kernel_roofline.py <launches.csv> <const|var>"""
import collections, csv, json, os, re, sys

p, mode = sys.argv[1], sys.argv[2]
peak = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
N = 513 ** 3
C = 512 ** 3
sig = 0 if mode == "const" else 8
# bytes per launch on level 0: (reads + writes) of whole arrays, each counted once
BYTES = {
    "k_gs2": (8 + 8 + sig + 8) * N,                    # x, rhs, (sigma), write x   [zero-input launches read one array less, add-to ones one more]
    "k_stencil": (8 + 8 + sig + 8) * N,                # x, b, (sigma), write y
    "k_restrict": 8 * N + 8 * (257 ** 3),
    "k_resid_restrict": (8 + 8 + sig) * N + 8 * (257 ** 3),   # cor, res, (sigma) in; coarse res out (the fine residual is stored on the shell only)
    "k_interp_add_w2": 16 * N + 8 * C + 8 * (257 ** 3),
    "k_interp_add_tri": 16 * N + 8 * (257 ** 3),
    "k_interp_add_w": 16 * N + 8 * C + 8 * (257 ** 3),
    "k_divu": 24 * (516 ** 3) + 8 * N,
    "k_fill_boundary": None, "k_minmax_cell": 8 * C,
}
rows = list(csv.DictReader(l for l in open(p) if not l.startswith("==")))
agg = collections.defaultdict(list)
for r in rows:
    name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "")
    name = re.sub(r"<.*", "", name)
    v = float(r["Metric Value"].replace(",", ""))
    v = v / 1e6 if r["Metric Unit"] in ("ns", "nsecond") else v / 1e3 if r["Metric Unit"] in ("us", "usecond") else v
    g = tuple(int(x) for x in re.findall(r"\d+", r["Grid Size"]))
    agg[name].append((g[0] * g[1] * g[2], v))
print(f"| kernel (level-0 launches) | launches | avg ms | bytes / launch | GB/s | % of measured HBM peak ({peak:.0f} GB/s) |")
print("|---|---:|---:|---:|---:|---:|")
for name, lst in sorted(agg.items(), key=lambda kv: -sum(v for _, v in kv[1])):
    if name not in BYTES or BYTES[name] is None:
        continue
    big = max(g for g, _ in lst)
    sel = [v for g, v in lst if g == big]
    ms = sum(sel) / len(sel)
    gbs = BYTES[name] / (ms * 1e-3) / 1e9
    print(f"| {name} | {len(sel)} | {ms:.3f} | {BYTES[name] / 1e9:.2f} GB | {gbs:.0f} | {100 * gbs / peak:.0f} % |")
