"""DRAM traffic per launch of the hot kernels, from `ncu --set full` reports, as the json bench.py reads for
roofline.traffic (profiles/r2/ncu_traffic.json).  The json records a hash of the kernel sources the reports were
captured from; bench.py reports traffic only while the sources still hash to it.

    python tools/ncu_traffic.py <tag> <forward micro-batch> <train batch>      (reports: gpurun_out/<tag>_prof_*.ncu-rep)
"""
import csv
import glob
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

KIND = {  # (section, substring of the kernel name) -> bench.py kernel kind
    "forward": [("dwpw_kernel", "gemm_dwpw_fused"), ("gemm_tc_kernel<256, 1, 0>", "gemm_conv1x1"),
                ("updc1_kernel", "gemm_update_conv1x1"), ("resid_update", "resid_update"),
                ("encoder_fwd", "encoder"), ("decoder_fwd", "decoder")],
    "train": [("wgrad_tc_kernel", "bwd_wgrad"), ("bwd_dw_march", "bwd_dwconv"), ("dwconv_march", "fwd_dwconv"),
              ("dwpw_kernel", "fwd_gemm_dwpw_fused"), ("updc1_kernel", "fwd_gemm_update_conv1x1")],
}


def launches(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    idx = {h: i for i, h in enumerate(rows[0])}
    units = rows[1]
    out = []
    for r in rows[2:]:
        tot = 0.0
        for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[idx[m]]]
            tot += float(r[idx[m]].replace(",", "")) * scale
        out.append((r[idx["Kernel Name"]], tot, float(r[idx["gpu__time_duration.sum"]].replace(",", ""))))
    return out


def main():
    tag, mb, tb = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    doc = {"kernel_source_hash": bench.kernel_source_hash(), "captured_with": "ncu --set full --clock-control none",
           "tag": tag}
    for section, batch, pat in (("forward", mb, "fwd"), ("train", tb, "train")):
        kernels = {}
        for rep in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "%s_prof_%s_*.ncu-rep" % (tag, pat)))):
            for name, nbytes, _ in launches(rep):
                for sub, kind in KIND[section]:
                    if sub in name:
                        k = kernels.setdefault(kind, {"kernel": name.split("(")[0], "samples": []})
                        k["samples"].append(nbytes)
        for k in kernels.values():
            k["traffic_bytes"] = sum(k["samples"]) / len(k["samples"])
        doc[section] = {"batch": batch, "kernels": kernels}
    json.dump(doc, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main()
