"""Physical memory traffic of the trace kernel over whole frames (the roofline's `traffic`).

Step 1 (GPU box, under ncu):
    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,gpu__time_duration.sum \
        --clock-control none -k regex:k_pt_trace --csv --log-file gpurun_out/TAG_trace_metrics_C2.csv \
        python tools/trace_traffic.py run C2 4 gpurun_out/TAG_trace_rays_C2.json
    (renders FRAMES frames of the workload at the given spp and writes the ray count)
Step 2 (anywhere):
    python tools/trace_traffic.py parse gpurun_out/TAG_trace_metrics_C2.csv \
        gpurun_out/TAG_trace_rays_C2.json profiles/TAG_trace_traffic_C2.json
    sums the metrics over every k_pt_trace* launch and divides by the rays: DRAM and L2 bytes per
    ray, their rates against the measured HBM peak.  bench.py reads the newest
    profiles/*_trace_traffic_<workload>.json and names it in roofline.traffic_source.
Durations under ncu are serialised and cold-clock: the rates are lower bounds, the BYTES are exact."""
import collections
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
FRAMES = 2


def run(workload, spp, out):
    from bench import make_workload
    from gpu_pathtracer_b200.api import PathTracer
    sc = make_workload(workload, spp)
    pt = PathTracer(sc)
    pt.build()
    rays = 0
    for _ in range(FRAMES):
        pt.render_pathtrace()
        rays += pt.stats()["rays_traced"]
    json.dump({"workload": workload, "spp": spp, "frames": FRAMES, "rays": rays,
               "triangles": sc.num_triangles}, open(out, "w"))
    print("rays", rays)


def parse(csv_path, rays_path, out):
    info = json.load(open(rays_path))
    rows = [r for r in csv.reader(open(csv_path)) if len(r) > 5]
    hdr = rows[0]
    ki, mi, vi, ui = (hdr.index(k) for k in ("Kernel Name", "Metric Name", "Metric Value", "Metric Unit"))
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9,
             "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "usecond": 1e-6, "nsecond": 1e-9, "msecond": 1e-3, "second": 1.0}
    tot = collections.Counter()
    kernels = collections.Counter()
    for r in rows[1:]:
        if "k_pt_trace" not in r[ki]:
            continue
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        tot[r[mi]] += v * scale.get(r[ui], 1.0)
        if r[mi] == "gpu__time_duration.sum":
            kernels[r[ki].split("(")[0][:60]] += 1
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6550.7))
    n = sum(kernels.values())
    dram = tot["dram__bytes_read.sum"] + tot["dram__bytes_write.sum"]
    t = tot["gpu__time_duration.sum"]
    res = {
        **info, "source_csv": os.path.basename(csv_path), "trace_launches": n, "kernels": dict(kernels),
        "dram_read_bytes": tot["dram__bytes_read.sum"], "dram_write_bytes": tot["dram__bytes_write.sum"],
        "l2_bytes": tot["lts__t_bytes.sum"], "duration_s_under_ncu": t,
        "dram_bytes_per_ray": dram / info["rays"], "l2_bytes_per_ray": tot["lts__t_bytes.sum"] / info["rays"],
        "dram_gbs_under_ncu": dram / t / 1e9, "l2_gbs_under_ncu": tot["lts__t_bytes.sum"] / t / 1e9,
        "hbm_peak_gbs": peak, "physical_frac_under_ncu": dram / t / 1e9 / peak,
    }
    json.dump(res, open(out, "w"), indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    if sys.argv[1] == "run":
        run(sys.argv[2], int(sys.argv[3]), sys.argv[4])
    else:
        parse(*sys.argv[2:5])
