"""Per-kernel totals of an ncu launch list (`ncu --metrics gpu__time_duration.sum --clock-control none --csv`).
usage: python tools/launch_summary.py <launches.csv>  -> JSON on stdout (kernel family -> launches, time_ms, share)"""
import collections, csv, json, re, sys
rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 14 and r[0].isdigit()]
agg = collections.OrderedDict()
def family(name):
    if "DeviceRadixSort" in name or "RadixSort" in name or "radix" in name.lower(): return "cub radix sort"
    m = re.search(r"(bswk::)?(\w+_kernel)(<[^>]*>)?", name)
    if m and ("bswk" in name or m.group(2).startswith(("bsw_", "kswv_", "prep_", "proto_", "class_", "ref_", "global2_", "restore_", "dpx_"))):
        fam = m.group(2)
        if fam == "bsw_packed_kernel": fam += "<true>" if re.search(r"bsw_packed_kernel<\(bool\)1|bsw_packed_kernel<true", name) else "<false>"
        return fam
    return "other (torch / cub helpers)"
tot = 0.0
for r in rows:
    f = family(r[4]); ns = float(r[14].replace(",", ""))
    a = agg.setdefault(f, [0, 0.0]); a[0] += 1; a[1] += ns; tot += ns
out = {k: {"launches": v[0], "time_ms": round(v[1] * 1e-6, 3), "share": round(v[1] / tot, 4)} for k, v in sorted(agg.items(), key=lambda x: -x[1][1])}
out["_total_ms"] = round(tot * 1e-6, 3); out["_launches"] = len(rows)
print(json.dumps(out, indent=1))
