"""Turn the ncu reports of a record run (gpurun_out/final/*.ncu-rep, read here with `ncu -i`) into the summaries
committed under profiles/: raw metric lists, source regions / top lines of k_obs, SASS listings, the launch list, and
profiles/k_obs_traffic.json (dram bytes per launch, which bench.py reports as roofline.traffic).
usage: python tools/make_profiles.py [gpurun_out/final] [r02]"""
import csv, json, os, re, shutil, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "final")
TAG = sys.argv[2] if len(sys.argv) > 2 else "r02"
OUT = os.path.join(ROOT, "profiles")
LIB = os.path.join(ROOT, "gpudrive_lab_b200", "libgpudrive_b200.so")

KEEP = ["dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__t_sector_hit_rate.pct",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__grid_size", "launch__block_size", "launch__waves_per_multiprocessor",
        "lts__t_sector_hit_rate.pct", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active"]


def ncu(rep, *args):
    return subprocess.run(["ncu", "-i", rep, *args], capture_output=True, text=True).stdout


def raw(rep, out):
    rows = list(csv.reader(ncu(rep, "--page", "raw", "--csv").splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[-1]
    m = {h: (u, v) for h, u, v in zip(hdr, units, vals)}
    lines = [f"kernel {m['Kernel Name'][1]}"]
    for k in KEEP:
        if k in m:
            lines.append(f"{k} {m[k][0]} {m[k][1]}")
    for h in sorted(m):
        mm = re.match(r"smsp__average_warps_issue_stalled_(\w+)_per_issue_active\.ratio", h)
        if mm:
            lines.append(f"stall {mm.group(1)} {m[h][1]}")
    open(out, "w").write("\n".join(lines) + "\n")
    return m


def to_bytes(u, v):
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    return float(v) * scale


def main():
    os.makedirs(OUT, exist_ok=True)
    traffic = {}
    tpath = os.path.join(OUT, "k_obs_traffic.json")
    if os.path.exists(tpath):
        traffic = json.load(open(tpath))
    for name, key in ((f"{TAG}_k_obs", "config3_u64"), (f"{TAG}_k_obs_u512", "config3_u512"), (f"{TAG}_k_obs_u8192", "config3_u8192"),
                      (f"{TAG}_k_agents", None)):
        rep = os.path.join(SRC, name + ".ncu-rep")
        if not os.path.exists(rep):
            continue
        m = raw(rep, os.path.join(OUT, name + "_raw.txt"))
        if key:
            b = to_bytes(*m["dram__bytes_read.sum"]) + to_bytes(*m["dram__bytes_write.sum"])
            traffic[key] = {"bytes": b, "read": to_bytes(*m["dram__bytes_read.sum"]), "write": to_bytes(*m["dram__bytes_write.sum"]),
                            "lts_hit_rate_pct": float(m["lts__t_sector_hit_rate.pct"][1]),
                            "source": f"profiles/{name}_raw.txt (ncu --set full: dram__bytes_read.sum + dram__bytes_write.sum)"}
        src_csv = os.path.join(SRC, name + "_source.csv")
        open(src_csv, "w").write(ncu(rep, "--page", "source", "--csv", "--print-source", "sass,cuda"))
        lines = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_lines.py"), src_csv, "60"], capture_output=True, text=True).stdout
        open(os.path.join(OUT, name + "_lines.txt"), "w").write(lines)
        if "k_obs" in name:
            reg = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_regions.py"), src_csv], capture_output=True, text=True).stdout
            open(os.path.join(OUT, name + "_regions.txt"), "w").write(reg)
    json.dump(traffic, open(tpath, "w"), indent=1)
    lc = os.path.join(SRC, f"{TAG}_launches.csv")
    if os.path.exists(lc):
        shutil.copy(lc, os.path.join(OUT, f"{TAG}_launches.csv"))
    # SASS listings of the shipped library (cuobjdump here; the same file travelled to the GPU box)
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    blocks = re.split(r"(?=\n\s*Function : )", sass)
    for kern, fname in (("k_obsILb0", f"{TAG}_k_obs.sass"), ("k_agents", f"{TAG}_k_agents.sass"), ("k_lidar_fix", f"{TAG}_k_lidar_fix.sass")):
        sel = [b for b in blocks if kern in b.split("\n", 2)[1] if len(b.split("\n", 2)) > 1]
        if sel:
            open(os.path.join(OUT, fname), "w").write(sel[0].strip() + "\n")
    for f in sorted(os.listdir(SRC)):
        if f.startswith("bench_") and f.endswith(".json") and os.path.getsize(os.path.join(SRC, f)) > 0:
            shutil.copy(os.path.join(SRC, f), os.path.join(OUT, f))
    print("profiles written:", sorted(f for f in os.listdir(OUT) if f.startswith(TAG) or f.startswith("bench_" + TAG) or f == "k_obs_traffic.json"))


if __name__ == "__main__":
    main()
