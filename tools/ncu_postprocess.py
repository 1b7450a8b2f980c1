"""Turn the ncu outputs of tools/profile_r02x.sh into the two small files the bench line cites.

    python tools/ncu_postprocess.py shares  gpurun_out/r02x_launches.csv        > profiles/r02x_launch_shares.csv
    python tools/ncu_postprocess.py traffic gpurun_out/r02x_ncu_bench_step_raw.csv k_ptrace_fused > profiles/r02x_traffic.json
"""
import csv, io, json, re, sys


def rows_of(path):
    text = "".join(l for l in open(path, errors="replace") if not l.startswith("=="))
    return list(csv.reader(io.StringIO(text)))


def shares(path):
    rows = rows_of(path)
    hdr = rows[0]
    ix = {h: i for i, h in enumerate(hdr)}
    per = {}
    for r in rows[1:]:
        if len(r) < len(hdr) or r[ix["Metric Name"]] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("(bool)", "").replace("(int)", "")
        v = float(r[ix["Metric Value"]].replace(",", ""))
        unit = r[ix["Metric Unit"]]
        us = v * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1.0)
        n, t = per.get(name, (0, 0.0))
        per[name] = (n + 1, t + us)
    total = sum(t for _, t in per.values())
    print("kernel, launches, total_us, share (ncu --metrics gpu__time_duration.sum --clock-control none, bench.py --steps 1 --warmup 1: "
          "warm-up step + timed step + e2e steps + profile step; cold-cache, serialised)")
    for name, (n, t) in sorted(per.items(), key=lambda kv: -kv[1][1]):
        print("%s, %d, %.1f, %.3f" % (name, n, t, t / total))


def traffic(path, kernel):
    rows = rows_of(path)
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}

    def val(r, k):
        v = float(r[ix[k]].replace(",", ""))
        u = units[ix[k]]
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1.0)
    sel = [r for r in rows[2:] if len(r) >= len(hdr) and kernel in r[ix["Kernel Name"]]]
    n = len(sel)
    print(json.dumps({"kernel": kernel, "launches": n,
                      "dram_bytes_read": sum(val(r, "dram__bytes_read.sum") for r in sel) / n,
                      "dram_bytes_write": sum(val(r, "dram__bytes_write.sum") for r in sel) / n,
                      "duration_ms": sum(val(r, "gpu__time_duration.sum") for r in sel) / n,
                      "capture": "%s (ncu --set full, bench.py --steps 1 --warmup 1, the %d %s launches of the captured step; per-launch average, "
                                 "carried over to the bench run's average launch by duration)" % (path.replace("gpurun_out/", "profiles/").replace("_raw.csv", "_summary.txt"), n, kernel),
                      "note": "cold L2 under ncu replay; the scene (9.2 MB) is L1/L2-resident, the DRAM traffic is the ray queues in and the hit records / "
                              "occlusion bytes out -- far below the algorithmic bytes, which count every node and primitive fetch"}, indent=1))


if __name__ == "__main__":
    if sys.argv[1] == "shares":
        shares(sys.argv[2])
    else:
        traffic(sys.argv[2], sys.argv[3])
