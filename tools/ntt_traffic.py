"""DRAM traffic of the NTT kernels per C2 case, for the `roofline.traffic` field of bench.py's NTT rows.

  run     (on the GPU box, under ncu):
          ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv \
              --log-file gpurun_out/ntt_traffic.csv python tools/ntt_traffic.py run gpurun_out/ntt_traffic_manifest.json
          one INTT and one NTT of a 256 MiB batch per case; the manifest records how many kernels each phase launched
          (gpu_kernel_launch_count), so the parser can attribute the CSV's launches in order.
  parse   (anywhere): python tools/ntt_traffic.py parse gpurun_out/ntt_traffic.csv gpurun_out/ntt_traffic_manifest.json profiles/r02_ntt_traffic.json
"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
BATCH = 256 << 20


def run(manifest_path):
    from bench import NTT_CASES
    from mxx_b200 import GpuDCRTPolyMatrix, GpuDCRTPolyParams, ffi, standard_moduli

    L = ffi.lib()
    rows = []
    for n, depth, bits in NTT_CASES:
        moduli = standard_moduli(n, depth, bits)
        params = GpuDCRTPolyParams(n, moduli, 14)
        W = 4 if max(moduli) < (1 << 30) else 8
        P = max(1, BATCH // (W * depth * n))
        c0 = L.gpu_kernel_launch_count()
        m = GpuDCRTPolyMatrix.sample_distribution(params, 1, P, 0, 0.0, [n, depth, bits, 1])
        params.synchronize()
        c1 = L.gpu_kernel_launch_count()
        m.intt_all_in_place()
        params.synchronize()
        c2 = L.gpu_kernel_launch_count()
        m.ntt_in_place()
        params.synchronize()
        c3 = L.gpu_kernel_launch_count()
        rows.append({"key": [n, depth, bits], "word_bytes": W, "polys": P, "coeffs": P * depth * n, "setup": c1 - c0, "intt": c2 - c1, "ntt": c3 - c2})
        del m, params
    with open(manifest_path, "w") as f:
        json.dump(rows, f)
    print("manifest written:", len(rows), "cases")


def parse(csv_path, manifest_path, out_path):
    with open(manifest_path) as f:
        manifest = json.load(f)
    launches = {}
    order = []
    with open(csv_path) as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    rd = csv.DictReader(lines)
    for r in rd:
        i = int(r["ID"])
        if i not in launches:
            launches[i] = {"name": r["Kernel Name"]}
            order.append(i)
        val = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(unit, 1)
        launches[i][r["Metric Name"]] = val * scale
    pos = 0
    rows = []
    for c in manifest:
        pos += c["setup"]
        phase = {}
        for ph in ("intt", "ntt"):
            ks = [launches[order[pos + j]] for j in range(c[ph])]
            pos += c[ph]
            phase[ph] = {"kernels": [k["name"].split("(")[0][:90] for k in ks],
                         "dram_bytes": int(sum(k["dram__bytes_read.sum"] + k["dram__bytes_write.sum"] for k in ks)),
                         "us": round(sum(k["gpu__time_duration.sum"] for k in ks), 2)}
        algo = 2 * c["word_bytes"] * c["coeffs"]
        rows.append({"key": c["key"], "word_bytes": c["word_bytes"], "coeffs": c["coeffs"], "algorithmic_bytes_per_direction": algo,
                     "intt": phase["intt"], "ntt": phase["ntt"],
                     "dram_bytes_per_roundtrip": phase["intt"]["dram_bytes"] + phase["ntt"]["dram_bytes"],
                     "traffic_over_algorithmic": round((phase["intt"]["dram_bytes"] + phase["ntt"]["dram_bytes"]) / (2 * algo), 3)})
    assert pos <= len(order), "the CSV holds fewer launches than the manifest"
    with open(out_path, "w") as f:
        json.dump({"how": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none over tools/ntt_traffic.py "
                          "(one INTT + one NTT of a 256 MiB batch per C2 case; launch times under ncu are cold-cache and serialised)", "rows": rows}, f, indent=1)
    for r in rows:
        print(r["key"], "traffic/algorithmic", r["traffic_over_algorithmic"], "|", r["ntt"]["kernels"], r["ntt"]["us"], "us |", r["intt"]["kernels"], r["intt"]["us"], "us")


if __name__ == "__main__":
    if sys.argv[1] == "run":
        run(sys.argv[2])
    else:
        parse(*sys.argv[2:5])
