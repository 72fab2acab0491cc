"""Rewrite profiles/traffic.json from an `ncu --set full` report of one pathtrace_megakernel launch (run where ncu is installed):
    python benchmarks/update_traffic.py gpurun_out/X.ncu-rep "source note"
bench.py reads the DRAM bytes and the warp-instruction count of the dominant kernel from that file (roofline.traffic / roofline.issue)."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main(rep, note):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    d = dict(zip(hdr, rows[2]))
    assert "pathtrace_megakernel" in d["Kernel Name"], d["Kernel Name"]

    def val(key):
        v = float(d[key].replace(",", ""))
        u = units[hdr.index(key)].lower()
        return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)


    # the bench line of the same command, run without ncu just before the capture (gpu_bench_and_profile.sh: <tag>_plain2.log): the
    # steps per ray of the tree the captured launch walked
    steps = tests = None
    plain = rep.replace("_megakernel.ncu-rep", "_plain2.log")
    if os.path.exists(plain):
        for line in open(plain):
            if line.startswith("{"):
                r = json.loads(line)["roofline"]
                steps, tests = r["node_steps_per_ray_product_order"], r["tri_tests_per_ray_product_order"]
    path = os.path.join(ROOT, "profiles", "traffic.json")
    sys.path.insert(0, ROOT)
    import bench
    t = json.load(open(path)) if os.path.exists(path) else {}
    t["pathtrace_megakernel"] = {
        "dram_bytes_per_launch": int(val("dram__bytes_read.sum") + val("dram__bytes_write.sum")),
        "warp_instructions_per_launch": int(val("smsp__inst_executed.sum")),
        "threads_per_instruction": round(val("smsp__thread_inst_executed_per_inst_executed.ratio"), 2),
        "l2_bytes_per_launch": int(val("lts__t_sectors.sum")) * 32,
        "l2_hit_pct": round(val("lts__t_sector_hit_rate.pct"), 2),
        "dram_read_bytes": int(val("dram__bytes_read.sum")), "dram_write_bytes": int(val("dram__bytes_write.sum")),
        "issue_active_pct": round(val("smsp__issue_active.avg.pct_of_peak_sustained_active"), 2),
        "duration_under_ncu": d["gpu__time_duration.sum"] + " " + units[hdr.index("gpu__time_duration.sum")],
        "csrc_hash": bench.csrc_hash(),
        "node_steps_per_ray": steps, "tri_tests_per_ray": tests,
        "source": note,
    }
    json.dump(t, open(path, "w"), indent=1)
    print(json.dumps(t["pathtrace_megakernel"]))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else os.path.basename(sys.argv[1]))
