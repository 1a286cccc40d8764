"""profiles/r2_* from the reports and logs of scripts/gpu_evidence.sh (gpurun_out/): ncu summaries of the hot kernels, the DRAM / L2
traffic per launch that bench.py quotes (r2_traffic.json, r2_l2_bytes.json), the launch list and the bench lines.
usage: python scripts/make_profiles.py [commit]"""
import csv, io, json, os, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
commit = sys.argv[1] if len(sys.argv) > 1 else subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()

REPORTS = {"k_primary": "config 3, fifth frame of the table, chunk 5", "k_shadow_grid": "config 3, fifth frame of the table, chunk 4",
           "k_shade": "config 3, fifth frame of the table, chunk 4", "k_shadow_grid_c3b": "config 3b (distance 576), fifth frame, chunk 4",
           "k_shadow_grid_c4": "config 4 (64 bunnies, one chunk per frame), fifth frame"}


def raw_metrics(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h, units, r = rows[0], rows[1], rows[2]

    def val(name):
        if name not in h:
            return None
        v, u = float(r[h.index(name)].replace(",", "")), units[h.index(name)]
        scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1e-3, "ms": 1.0, "ns": 1e-6, "s": 1e3}
        return v * scale.get(u, 1.0)
    return val


traffic, l2 = {}, {}
for k, what in REPORTS.items():
    rep = os.path.join(G, "r2_%s.ncu-rep" % k)
    if not os.path.exists(rep):
        print("missing", rep); continue
    txt = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "ncu_summary.py"), rep], capture_output=True, text=True).stdout
    val = raw_metrics(rep)
    ms, rd, wr = val("gpu__time_duration.sum"), val("dram__bytes_read.sum"), val("dram__bytes_write.sum")
    sect, pct = val("lts__t_sectors_srcunit_tex.sum"), val("lts__t_sectors.avg.pct_of_peak_sustained_elapsed")
    lts = sect * 32.0 if sect is not None else None
    with open(os.path.join(P, "r2_ncu_%s.txt" % k), "w") as f:
        f.write("# r2: ncu --set full --clock-control none of %s, %s; scripts/gpu_ncu_r2.sh at %s\n" % (k.split("_c")[0], what, commit))
        f.write(txt)
        if lts is not None:
            f.write("## L2: lts__t_sectors_srcunit_tex.sum %.0f sectors x 32 B = %.1f GB/s over the launch; lts__t_sectors %.2f %% of peak sustained (ncu)\n"
                    % (sect, lts / (ms * 1e-3) / 1e9, pct or 0.0))
    traffic[k] = {"launch_ms": ms, "dram_read_bytes": rd, "dram_write_bytes": wr, "dram_gbs": (rd + wr) / (ms * 1e-3) / 1e9, "lts_t_bytes": lts,
                  "l2_gbs": lts / (ms * 1e-3) / 1e9 if lts else None, "l1tex_t_bytes": None, "report": "profiles/r2_ncu_%s.txt" % k,
                  "lts_t_sectors_pct_of_peak_sustained": pct}
    l2[k] = {"launch_ms": ms, "lts_t_sectors_srcunit_tex": sect, "lts_t_bytes": lts, "l2_gbs": traffic[k]["l2_gbs"],
             "lts_t_sectors_pct_of_peak_sustained": pct, "source": "profiles/r2_ncu_%s.txt" % k}
if "k_shadow_grid" in traffic:
    t = traffic["k_shadow_grid"]
    top = {"kernel": "k_shadow_grid (the largest launch of the chain: chunk 4 of 8 of the 4K/16spp frame)", "dram_bytes_read": t["dram_read_bytes"],
           "dram_bytes_write": t["dram_write_bytes"], "bytes": t["dram_read_bytes"] + t["dram_write_bytes"], "launch_ms": t["launch_ms"],
           "source": "profiles/r2_ncu_k_shadow_grid.txt (ncu --set full, one launch)", "all_kernels": traffic}
    json.dump(top, open(os.path.join(P, "r2_traffic.json"), "w"), indent=1)
    json.dump(l2, open(os.path.join(P, "r2_l2_bytes.json"), "w"), indent=1)

lst = os.path.join(G, "r2_launches.csv")
if os.path.exists(lst):
    txt = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "launch_summary.py"), lst], capture_output=True, text=True).stdout
    with open(os.path.join(P, "r2_launches.txt"), "w") as f:
        f.write("# r2: ncu launch list of 'python bench.py --steps 2 --warmup 3 --no-cpu-baseline' at commit %s (config 3, 1 GPU; includes the untimed "
                "counters / profile / e2e legs and rtb_measure_peaks)\n# ncu --metrics gpu__time_duration.sum --clock-control none -c 900; per-launch "
                "times are cold-cache/serialised: compare shares\n" % commit)
        f.write(txt)
for f in sorted(os.listdir(G)):
    if f.startswith("r2_bench_c") and f.endswith(".json") and os.path.getsize(os.path.join(G, f)) > 0:
        shutil.copy(os.path.join(G, f), os.path.join(P, f))
        print("copied", f)
