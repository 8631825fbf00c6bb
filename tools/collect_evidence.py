"""Turns the raw outputs of tools/evidence.sh (gpurun_out/<tag>_*) into the tracked summaries under profiles/:
bench lines, launch shares, the `ncu --set full` summary, per-line attribution of K-trace and the DRAM traffic file
that bench.py reads for `roofline.traffic`.   usage: python tools/collect_evidence.py <tag>"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT, PROF = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")


def sh(cmd, out=None):
    r = subprocess.run(cmd, shell=True, capture_output=True, text=True, cwd=ROOT)
    if out:
        open(out, "w").write(r.stdout + r.stderr)
    return r.stdout


def main():
    tag = sys.argv[1]
    rows = []
    for w in ("c1", "c2", "c3", "c4", "c5", "ref_c2"):
        src = os.path.join(OUT, f"{tag}_bench_{w}.json")
        if not os.path.exists(src):
            continue
        line = open(src).read().strip().splitlines()[-1]
        open(os.path.join(PROF, f"{tag}_bench_{w}.json"), "w").write(line + "\n")
        d = json.loads(line)
        r = d.get("roofline") or {}
        rows.append((w, d["value"], (d.get("e2e") or {}).get("value"), d["ms_per_step"], d.get("spp_per_s"), r.get("trace_mrays_per_s"),
                     r.get("nodes_per_ray"), r.get("tris_per_ray"), r.get("frac"), r.get("trace_ms_per_step"), r.get("shade_ms_per_step"),
                     r.get("other_ms_per_step"), (d.get("cpu_baseline") or {}).get("value")))
    for row in rows:
        print(" | ".join("-" if v is None else (f"{v:.3f}" if isinstance(v, float) else str(v)) for v in row))
    lc = os.path.join(OUT, f"{tag}_launches_c2.csv")
    if os.path.exists(lc):
        sh(f"cp {lc} {PROF}/{tag}_launches_c2.csv")
        sh(f"python tools/ncu_launch_shares.py {PROF}/{tag}_launches_c2.csv", os.path.join(PROF, f"{tag}_launch_shares_c2.txt"))
    rep = os.path.join(OUT, f"{tag}_full.ncu-rep")
    if os.path.exists(rep):
        raw = f"/tmp/{tag}_raw.csv"
        open(raw, "w").write(sh(f"ncu -i {rep} --page raw --csv"))
        sh(f"python tools/ncu_summary.py {raw}", os.path.join(PROF, f"{tag}_ncu_full_summary.txt"))
        text = ""
        for k, name in ((2, "primary wavefront"), (3, "bounce wavefront")):
            cs = f"/tmp/{tag}_trace_cs{k}.csv"
            open(cs, "w").write(sh(f"ncu -i {rep} --page source --csv --print-source cuda,sass --kernel-name regex:k_trace_persistent "
                                   f"--launch-skip {k} --launch-count 1"))
            text += f"== K-trace, {name} (second captured frame)\n" + sh(f"python tools/ncu_lines.py {cs} 0 22")
            text += "\n".join(sh(f"python tools/ncu_util_buckets.py {cs}").splitlines()[:8]) + "\n\n"
        open(os.path.join(PROF, f"{tag}_ncu_lines_trace.txt"), "w").write(text)
        rr = list(csv.reader(open(raw)))
        hdr, units = rr[0], rr[1]
        ni, ri, wi = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}
        kt = [r for r in rr[2:] if "k_trace_persistent" in r[ni]][-2:]      # the last captured frame: primary + bounce wavefront
        last = [float(r[ri]) * scale[units[ri]] + float(r[wi]) * scale[units[wi]] for r in kt]
        ii, li = hdr.index("smsp__inst_executed.sum"), hdr.index("smsp__thread_inst_executed_per_inst_executed.ratio")
        inst = [float(r[ii]) for r in kt]
        lanes = sum(float(r[li]) * n for r, n in zip(kt, inst)) / sum(inst)
        json.dump({"workload": "c2", "flags": 0,
                   "source": f"profiles/{tag}_ncu_full_summary.txt (ncu --set full --clock-control none, bench.py --steps 1 --warmup 3)",
                   "k_trace_persistent": {"dram_bytes_per_launch": sum(last) / len(last), "launches": last,
                                          "warp_inst_per_frame": sum(inst), "warp_inst_per_launch": inst, "lanes_per_inst": lanes}},
                  open(os.path.join(PROF, "ncu_traffic.json"), "w"), indent=1)
        print("traffic per K-trace launch (bytes):", last, "warp instructions per frame:", sum(inst), "lanes per instruction:", lanes)


if __name__ == "__main__":
    main()
